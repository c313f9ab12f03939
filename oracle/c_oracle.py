"""ctypes access to oracle/paged_attn_ref.c (liboracle_ref.so) — TEST INFRASTRUCTURE / CPU baseline only."""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import torch

HERE = Path(__file__).resolve().parent
LIB = HERE / "_build" / "liboracle_ref.so"
_DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2, torch.uint8: 3, torch.float8_e4m3fn: 3}
_lib = None


def build() -> Path:
    """gcc the C restatement (same recipe as vllm_metax_b200.build.build_oracle, kept here so that the
    oracle can be rebuilt without the product package)."""
    import subprocess

    src = HERE / "paged_attn_ref.c"
    LIB.parent.mkdir(exist_ok=True)
    if not LIB.exists() or LIB.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["gcc", "-O3", "-mavx2", "-mfma", "-fopenmp", "-fPIC", "-shared", "-std=c11", str(src), "-o",
                        str(LIB), "-lm"], check=True)
    return LIB


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(os.fspath(LIB))
        _lib.ref_paged_attention.restype = C.c_int
        _lib.ref_f32_to_e4m3.restype = C.c_uint8
        _lib.ref_f32_to_e4m3.argtypes = [C.c_float]
    return _lib


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def paged_attention(query, key_cache, value_cache, block_tables, seq_lens, scale, alibi_slopes=None,
                    kv_cache_dtype="auto", k_scale=1.0, v_scale=1.0, num_threads=None) -> torch.Tensor:
    assert query.device.type == "cpu" and key_cache.is_contiguous() or True
    S, H, D = query.shape
    Hkv, B = value_cache.shape[1], value_cache.shape[3]
    out = torch.empty((S, H, D), dtype=query.dtype)
    nt = num_threads or os.cpu_count() or 1
    rc = lib().ref_paged_attention(
        _p(out), _p(query), _p(key_cache), _p(value_cache), _p(block_tables), _p(seq_lens), _p(alibi_slopes),
        C.c_float(scale), C.c_float(float(k_scale)), C.c_float(float(v_scale)), C.c_int(S), C.c_int(H), C.c_int(Hkv),
        C.c_int(D), C.c_int(B), C.c_int(block_tables.shape[1]), C.c_int64(query.stride(0)),
        C.c_int64(key_cache.stride(0)), C.c_int64(key_cache.stride(1)), C.c_int(_DT[query.dtype]),
        C.c_int(_DT[key_cache.dtype]), C.c_int(nt))
    if rc != 0:
        raise MemoryError("ref_paged_attention: allocation failed")
    return out


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale) -> None:
    T = slot_mapping.shape[0]
    H, D = key.shape[1], key.shape[2]
    lib().ref_reshape_and_cache(
        _p(key), _p(value), _p(key_cache), _p(value_cache), _p(slot_mapping), C.c_float(float(k_scale)),
        C.c_float(float(v_scale)), C.c_int(T), C.c_int(H), C.c_int(D), C.c_int(key_cache.shape[3]),
        C.c_int(key_cache.shape[4]), C.c_int64(key.stride(0)), C.c_int64(value.stride(0)), C.c_int(_DT[key.dtype]),
        C.c_int(_DT[key_cache.dtype]))


def reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale) -> None:
    T = slot_mapping.shape[0]
    H, D = key.shape[1], key.shape[2]
    lib().ref_reshape_and_cache_flash(
        _p(key), _p(value), _p(key_cache), _p(value_cache), _p(slot_mapping), C.c_float(float(k_scale)),
        C.c_float(float(v_scale)), C.c_int(T), C.c_int(H), C.c_int(D), C.c_int(key_cache.shape[1]),
        C.c_int64(key.stride(0)), C.c_int64(value.stride(0)), C.c_int64(key_cache.stride(0)),
        C.c_int64(key_cache.stride(1)), C.c_int64(key_cache.stride(2)), C.c_int(_DT[key.dtype]),
        C.c_int(_DT[key_cache.dtype]))


def convert_fp8(dst, src, scale=1.0) -> None:
    lib().ref_convert_fp8(_p(dst), _p(src), C.c_int64(src.numel()), C.c_float(scale), C.c_int(_DT[dst.dtype]),
                          C.c_int(_DT[src.dtype]))
