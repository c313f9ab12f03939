"""Generate the golden vectors that pin the oracle to the reference (run in the BUILD container only).

    python tests/golden/make_golden.py        # needs /root/reference; writes tests/golden/*.pt

What is executed is the reference's OWN Python oracle code, unmodified: the source text of
`ref_masked_attention` and `ref_single_query_cached_kv_attention` is cut out of
/root/reference/tests/kernels/attention/test_attention.py (lines 50-118) with `ast` and exec'd with only
`torch` / `Optional` in scope.  (The module itself cannot be imported here: it pulls in vllm._C, xformers
and a CUDA platform.)  The cache-writer goldens execute the scatter loops that the reference keeps inline
in its tests (tests/kernels/attention/test_cache.py:190-200 for the v1/v2 layout, :325-337 for the flash
layouts); they are transcribed below next to their line numbers because they are not separate functions.

The fixtures hold inputs and outputs, so the GPU box (which has no /root/reference) can replay them.
FP8 cases use caches dequantised by oracle.torch_oracle.fp8_e4m3_dequant before the reference oracle runs,
as the reference test dequantises with convert_fp8 before calling it (test_attention.py:303-319).
"""
from __future__ import annotations

import ast
import os
import random
import sys
from typing import Optional

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF_TEST = "/root/reference/tests/kernels/attention/test_attention.py"

from oracle import torch_oracle as orc  # noqa: E402


def load_reference_oracle():
    src = open(REF_TEST).read()
    tree = ast.parse(src)
    wanted = {"ref_masked_attention", "ref_single_query_cached_kv_attention"}
    ns = {"torch": torch, "Optional": Optional}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in wanted:
            code = compile(ast.Module(body=[node], type_ignores=[]), REF_TEST, "exec")
            exec(code, ns)
    assert wanted <= set(ns), "reference oracle functions not found"
    return ns["ref_single_query_cached_kv_attention"]


ATTN_CASES = [
    # name, S, H, Hkv, D, B, max_len, NB, dtype, kv, alibi
    ("attn_bf16_gqa4", 3, 8, 2, 128, 16, 150, 40, torch.bfloat16, "auto", False),
    ("attn_bf16_mha_alibi", 3, 4, 4, 64, 16, 97, 24, torch.bfloat16, "auto", True),
    ("attn_bf16_gqa8_b32", 2, 8, 1, 128, 32, 130, 12, torch.bfloat16, "auto", False),
    ("attn_fp8_gqa4", 3, 8, 2, 128, 16, 120, 32, torch.bfloat16, "fp8", False),
    ("attn_f32_d80", 2, 4, 2, 80, 16, 70, 16, torch.float32, "auto", True),
    ("attn_f16_d32_b8", 2, 4, 2, 32, 8, 50, 16, torch.float16, "auto", False),
]


def make_attention(ref_fn):
    for idx, (name, S, H, Hkv, D, B, L, NB, dtype, kv, alibi) in enumerate(ATTN_CASES):
        d = orc.make_decode_inputs(S, H, Hkv, D, B, L, NB, dtype=dtype, kv_cache_dtype=kv, seed=100 + idx,
                                   use_alibi=alibi)
        kc, vc = d["key_cache"], d["value_cache"]
        if kv != "auto":
            d["k_scale"], d["v_scale"] = 0.5, 2.0  # exact powers of two: dequant stays exact in bf16
            kc = orc.fp8_e4m3_dequant(kc, d["k_scale"], dtype)
            vc = orc.fp8_e4m3_dequant(vc, d["v_scale"], dtype)
        out = torch.empty_like(d["query"])
        ref_fn(out, d["query"], H // Hkv, kc, vc, d["block_tables"], d["seq_lens"], d["scale"], d["alibi_slopes"])
        d["ref_out"] = out
        torch.save(d, os.path.join(HERE, name + ".pt"))
        print("wrote", name, tuple(out.shape), float(out.float().abs().max()))


CACHE_CASES = [
    # name, T, H, D, B, NB, dtype
    ("cache_bf16_d64_b16", 11, 2, 64, 16, 8, torch.bfloat16),
    ("cache_f32_d80_b8", 7, 2, 80, 8, 8, torch.float32),
]


def make_cache():
    for name, T, H, D, B, NB, dtype in CACHE_CASES:
        rnd = random.Random(7)
        g = torch.Generator().manual_seed(11)
        slots = rnd.sample(range(B * NB), T)
        slot_mapping = torch.tensor(slots, dtype=torch.long)
        key = torch.randn(T, H, D, generator=g).to(dtype)
        value = torch.randn(T, H, D, generator=g).to(dtype)
        x = 16 // torch.tensor([], dtype=dtype).element_size()
        key_cache = torch.randn(NB, H, D // x, B, x, generator=g).to(dtype)
        value_cache = torch.randn(NB, H, D, B, generator=g).to(dtype)
        nhd_k = torch.randn(NB, B, H, D, generator=g).to(dtype)
        nhd_v = torch.randn(NB, B, H, D, generator=g).to(dtype)
        hnd_k = torch.randn(NB, H, B, D, generator=g).to(dtype)
        hnd_v = torch.randn(NB, H, B, D, generator=g).to(dtype)
        fix = dict(key=key, value=value, slot_mapping=slot_mapping, key_cache=key_cache.clone(),
                   value_cache=value_cache.clone(), nhd_k=nhd_k.clone(), nhd_v=nhd_v.clone(), hnd_k=hnd_k.clone(),
                   hnd_v=hnd_v.clone())
        # reference: tests/kernels/attention/test_cache.py:190-200
        reshaped_key = key.reshape(T, *key_cache[0, :, :, 0, :].shape)
        block_indices = torch.div(slot_mapping, B, rounding_mode="floor").tolist()
        block_offsets = (slot_mapping % B).tolist()
        for i in range(T):
            key_cache[block_indices[i], :, :, block_offsets[i], :] = reshaped_key[i]
            value_cache[block_indices[i], :, :, block_offsets[i]] = value[i]
        # reference: tests/kernels/attention/test_cache.py:325-337 (NHD, then HND)
        for i in range(T):
            nhd_k[block_indices[i], block_offsets[i], :, :] = key[i]
            nhd_v[block_indices[i], block_offsets[i], :, :] = value[i]
            hnd_k[block_indices[i], :, block_offsets[i], :] = key[i]
            hnd_v[block_indices[i], :, block_offsets[i], :] = value[i]
        fix.update(ref_key_cache=key_cache, ref_value_cache=value_cache, ref_nhd_k=nhd_k, ref_nhd_v=nhd_v,
                   ref_hnd_k=hnd_k, ref_hnd_v=hnd_v)
        torch.save(fix, os.path.join(HERE, name + ".pt"))
        print("wrote", name)


if __name__ == "__main__":
    if not os.path.exists(REF_TEST):
        sys.exit("the reference is not mounted at /root/reference: goldens can only be regenerated in the build container")
    make_attention(load_reference_oracle())
    make_cache()
