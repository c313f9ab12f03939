"""ctypes binding of the C ABI in include/b200_paged_attn.h.

Used by the bench harness and the parity tests to call the library without the torch
dispatcher (raw device pointers, sizes, strides, a stream handle).  Struct layouts below must
match the header field for field; tests/test_abi.py checks sizeof() and the exported symbols.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

LIB_PATH = Path(os.environ.get("B200_PAGED_ATTN_LIB") or Path(__file__).resolve().parent / "libb200_paged_attn.so")  # override: development A/B runs

B200_OK = 0
ERR_INVALID_ARGUMENT, ERR_UNSUPPORTED, ERR_CUDA, ERR_WORKSPACE = 1, 2, 3, 4
DTYPE_F32, DTYPE_F16, DTYPE_BF16, DTYPE_U8 = 0, 1, 2, 3
KV_AUTO, KV_FP8_E4M3 = 0, 1
VARIANT_AUTO, VARIANT_GENERIC, VARIANT_STREAM_MMA, VARIANT_UMMA_FP8 = 0, 1, 2, 3
PARTITION_SIZE = 512

# every symbol include/b200_paged_attn.h declares
EXPORTED_SYMBOLS = (
    "b200_paged_attention_v1",
    "b200_paged_attention_v2",
    "b200_paged_attention_flash",
    "b200_paged_attention_workspace_bytes",
    "b200_paged_attention_workspace_reset",
    "b200_paged_attention_v1_static",
    "b200_paged_attention_static_scratch_bytes",
    "b200_reshape_and_cache",
    "b200_reshape_and_cache_flash",
    "b200_convert_fp8",
    "b200_rotary_embedding",
    "b200_rotary_embedding_and_cache_flash",
    "b200_merge_attn_states",
    "b200_copy_blocks",
    "b200_copy_blocks_mla",
    "b200_concat_and_cache_mla",
    "b200_mla_decode",
    "b200_mla_decode_workspace_bytes",
    "b200_swap_blocks",
    "b200_gather_cache",
    "b200_dcp_exchange_fold",
    "b200_dcp_gather_query",
    "b200_peer_alloc",
    "b200_peer_open",
    "b200_peer_close",
    "b200_peer_free",
    "b200_kv_dtype_from_string",
    "b200_last_error",
    "b200_abi_version",
    "b200_set_attention_variant",
    "b200_last_attention_variant",
    "b200_set_stream_dynamic",
    "b200_launch_count",
)


class PagedAttnParams(C.Structure):
    _fields_ = [
        ("out", C.c_void_p),
        ("query", C.c_void_p),
        ("key_cache", C.c_void_p),
        ("value_cache", C.c_void_p),
        ("block_tables", C.c_void_p),
        ("seq_lens", C.c_void_p),
        ("alibi_slopes", C.c_void_p),
        ("k_scale", C.c_void_p),
        ("v_scale", C.c_void_p),
        ("exp_sums", C.c_void_p),
        ("max_logits", C.c_void_p),
        ("tmp_out", C.c_void_p),
        ("workspace", C.c_void_p),
        ("workspace_bytes", C.c_uint64),
        ("q_stride", C.c_int64),
        ("kv_block_stride", C.c_int64),
        ("kv_head_stride", C.c_int64),
        ("num_seqs", C.c_int32),
        ("num_heads", C.c_int32),
        ("num_kv_heads", C.c_int32),
        ("head_size", C.c_int32),
        ("block_size", C.c_int32),
        ("max_seq_len", C.c_int32),
        ("max_num_blocks_per_seq", C.c_int32),
        ("dtype", C.c_int32),
        ("kv_dtype", C.c_int32),
        ("scale", C.c_float),
        ("tp_rank", C.c_int32),
        ("blocksparse_local_blocks", C.c_int32),
        ("blocksparse_vert_stride", C.c_int32),
        ("blocksparse_block_size", C.c_int32),
        ("blocksparse_head_sliding_step", C.c_int32),
        ("out_lse", C.c_void_p),
        ("kv_page_stride", C.c_int64),
        ("kv_num_blocks", C.c_int64),
    ]


class CacheWriteParams(C.Structure):
    _fields_ = [
        ("key", C.c_void_p),
        ("value", C.c_void_p),
        ("key_cache", C.c_void_p),
        ("value_cache", C.c_void_p),
        ("slot_mapping", C.c_void_p),
        ("k_scale", C.c_void_p),
        ("v_scale", C.c_void_p),
        ("key_stride", C.c_int64),
        ("value_stride", C.c_int64),
        ("block_stride", C.c_int64),
        ("page_stride", C.c_int64),
        ("head_stride", C.c_int64),
        ("num_tokens", C.c_int32),
        ("num_heads", C.c_int32),
        ("head_size", C.c_int32),
        ("block_size", C.c_int32),
        ("x", C.c_int32),
        ("dtype", C.c_int32),
        ("kv_dtype", C.c_int32),
    ]


class MergeAttnStatesParams(C.Structure):
    _fields_ = [
        ("output", C.c_void_p),
        ("output_lse", C.c_void_p),
        ("prefix_output", C.c_void_p),
        ("prefix_lse", C.c_void_p),
        ("suffix_output", C.c_void_p),
        ("suffix_lse", C.c_void_p),
        ("output_stride", C.c_int64),
        ("prefix_stride", C.c_int64),
        ("suffix_stride", C.c_int64),
        ("num_tokens", C.c_int32),
        ("num_heads", C.c_int32),
        ("head_size", C.c_int32),
        ("dtype", C.c_int32),
    ]


class GatherCacheParams(C.Structure):
    _fields_ = [
        ("src_cache", C.c_void_p),
        ("dst", C.c_void_p),
        ("block_table", C.c_void_p),
        ("cu_seq_lens", C.c_void_p),
        ("seq_starts", C.c_void_p),
        ("scale", C.c_void_p),
        ("block_table_stride", C.c_int64),
        ("cache_block_stride", C.c_int64),
        ("cache_entry_stride", C.c_int64),
        ("dst_entry_stride", C.c_int64),
        ("batch_size", C.c_int32),
        ("block_size", C.c_int32),
        ("entry_size", C.c_int32),
        ("src_dtype", C.c_int32),
        ("dst_dtype", C.c_int32),
        ("kv_dtype", C.c_int32),
        ("token_granular_starts", C.c_int32),
    ]


class RotaryParams(C.Structure):
    _fields_ = [
        ("positions", C.c_void_p),
        ("query", C.c_void_p),
        ("key", C.c_void_p),
        ("cos_sin_cache", C.c_void_p),
        ("query_stride", C.c_int64),
        ("key_stride", C.c_int64),
        ("head_stride", C.c_int64),
        ("num_tokens", C.c_int32),
        ("num_heads", C.c_int32),
        ("num_kv_heads", C.c_int32),
        ("head_size", C.c_int32),
        ("rot_dim", C.c_int32),
        ("is_neox", C.c_int32),
        ("dtype", C.c_int32),
        ("num_cache_tokens", C.c_int32),
        ("value", C.c_void_p),
        ("key_cache", C.c_void_p),
        ("value_cache", C.c_void_p),
        ("slot_mapping", C.c_void_p),
        ("k_scale", C.c_void_p),
        ("v_scale", C.c_void_p),
        ("value_stride", C.c_int64),
        ("block_stride", C.c_int64),
        ("page_stride", C.c_int64),
        ("cache_head_stride", C.c_int64),
        ("block_size", C.c_int32),
        ("kv_dtype", C.c_int32),
    ]


class ConcatMlaParams(C.Structure):
    _fields_ = [
        ("kv_c", C.c_void_p),
        ("k_pe", C.c_void_p),
        ("kv_cache", C.c_void_p),
        ("slot_mapping", C.c_void_p),
        ("scale", C.c_void_p),
        ("kv_c_stride", C.c_int64),
        ("k_pe_stride", C.c_int64),
        ("block_stride", C.c_int64),
        ("entry_stride", C.c_int64),
        ("num_tokens", C.c_int32),
        ("kv_lora_rank", C.c_int32),
        ("pe_dim", C.c_int32),
        ("block_size", C.c_int32),
        ("dtype", C.c_int32),
        ("kv_dtype", C.c_int32),
    ]


class MlaDecodeParams(C.Structure):
    _fields_ = [
        ("out", C.c_void_p),
        ("lse", C.c_void_p),
        ("q_nope", C.c_void_p),
        ("q_pe", C.c_void_p),
        ("kv_cache", C.c_void_p),
        ("seq_lens", C.c_void_p),
        ("page_table", C.c_void_p),
        ("workspace", C.c_void_p),
        ("workspace_bytes", C.c_uint64),
        ("out_seq_stride", C.c_int64),
        ("out_head_stride", C.c_int64),
        ("q_nope_seq_stride", C.c_int64),
        ("q_nope_head_stride", C.c_int64),
        ("q_pe_seq_stride", C.c_int64),
        ("q_pe_head_stride", C.c_int64),
        ("block_stride", C.c_int64),
        ("entry_stride", C.c_int64),
        ("page_table_stride", C.c_int64),
        ("lse_stride", C.c_int64),
        ("num_blocks", C.c_int64),
        ("scale", C.c_float),
        ("num_seqs", C.c_int32),
        ("num_heads", C.c_int32),
        ("page_size", C.c_int32),
        ("max_pages", C.c_int32),
        ("dtype", C.c_int32),
        ("num_kv_splits", C.c_int32),
        ("debug_logits", C.c_void_p),
        ("kv_dtype", C.c_int32),
        ("reserved", C.c_int32),
    ]


class DcpFoldParams(C.Structure):
    _fields_ = [
        ("peer_out", C.POINTER(C.c_void_p)),
        ("peer_lse", C.POINTER(C.c_void_p)),
        ("peer_flags", C.POINTER(C.c_void_p)),
        ("out", C.c_void_p),
        ("out_lse", C.c_void_p),
        ("rank", C.c_int32),
        ("world", C.c_int32),
        ("num_seqs", C.c_int32),
        ("heads_local", C.c_int32),
        ("head_size", C.c_int32),
        ("dtype", C.c_int32),
        ("epoch", C.c_uint32),
        ("step_counter", C.c_void_p),
    ]


class DcpGatherParams(C.Structure):
    _fields_ = [
        ("peer_q", C.POINTER(C.c_void_p)),
        ("peer_flags", C.POINTER(C.c_void_p)),
        ("query", C.c_void_p),
        ("cta_counter", C.c_void_p),
        ("q_stride", C.c_int64),
        ("rank", C.c_int32),
        ("world", C.c_int32),
        ("num_seqs", C.c_int32),
        ("heads_local", C.c_int32),
        ("head_size", C.c_int32),
        ("dtype", C.c_int32),
        ("epoch", C.c_uint32),
        ("step_counter", C.c_void_p),
    ]


DCP_MAX_WORLD = 16
COPY_D2D, COPY_D2H, COPY_H2D = 0, 1, 2

_lib = None


class B200Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(msg)
        self.code = code


def load() -> C.CDLL:
    """dlopen the C-ABI library; fail loudly (no fallback) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            "%s is missing: the CUDA kernels are not built. Run `python -m vllm_metax_b200.build` "
            "(there is no CPU fallback for this path)." % LIB_PATH
        )
    lib = C.CDLL(os.fspath(LIB_PATH), mode=C.RTLD_GLOBAL)
    P = C.POINTER
    lib.b200_paged_attention_v1.argtypes = [P(PagedAttnParams), C.c_void_p]
    lib.b200_paged_attention_v1.restype = C.c_int
    lib.b200_paged_attention_v2.argtypes = [P(PagedAttnParams), C.c_void_p]
    lib.b200_paged_attention_v2.restype = C.c_int
    lib.b200_paged_attention_flash.argtypes = [P(PagedAttnParams), C.c_void_p]
    lib.b200_paged_attention_flash.restype = C.c_int
    lib.b200_paged_attention_workspace_bytes.argtypes = [P(PagedAttnParams)]
    lib.b200_paged_attention_workspace_bytes.restype = C.c_uint64
    lib.b200_paged_attention_workspace_reset.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
    lib.b200_paged_attention_workspace_reset.restype = C.c_int
    lib.b200_paged_attention_v1_static.argtypes = [C.POINTER(PagedAttnParams), C.c_void_p]
    lib.b200_paged_attention_v1_static.restype = C.c_int
    lib.b200_paged_attention_static_scratch_bytes.argtypes = []
    lib.b200_paged_attention_static_scratch_bytes.restype = C.c_uint64
    lib.b200_reshape_and_cache.argtypes = [P(CacheWriteParams), C.c_void_p]
    lib.b200_reshape_and_cache.restype = C.c_int
    lib.b200_reshape_and_cache_flash.argtypes = [P(CacheWriteParams), C.c_void_p]
    lib.b200_reshape_and_cache_flash.restype = C.c_int
    lib.b200_convert_fp8.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_int32, C.c_int32, C.c_int32,
                                     C.c_void_p]
    lib.b200_convert_fp8.restype = C.c_int
    lib.b200_rotary_embedding.argtypes = [P(RotaryParams), C.c_void_p]
    lib.b200_rotary_embedding.restype = C.c_int
    lib.b200_rotary_embedding_and_cache_flash.argtypes = [P(RotaryParams), C.c_void_p]
    lib.b200_rotary_embedding_and_cache_flash.restype = C.c_int
    lib.b200_merge_attn_states.argtypes = [P(MergeAttnStatesParams), C.c_void_p]
    lib.b200_merge_attn_states.restype = C.c_int
    lib.b200_copy_blocks.argtypes = [P(C.c_void_p), P(C.c_void_p), C.c_int32, C.c_void_p, C.c_int32, C.c_int64,
                                     C.c_void_p]
    lib.b200_copy_blocks.restype = C.c_int
    lib.b200_copy_blocks_mla.argtypes = [P(C.c_void_p), C.c_int32, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p]
    lib.b200_copy_blocks_mla.restype = C.c_int
    lib.b200_concat_and_cache_mla.argtypes = [P(ConcatMlaParams), C.c_void_p]
    lib.b200_concat_and_cache_mla.restype = C.c_int
    lib.b200_mla_decode.argtypes = [P(MlaDecodeParams), C.c_void_p]
    lib.b200_mla_decode.restype = C.c_int
    lib.b200_mla_decode_workspace_bytes.argtypes = [C.c_int32, C.c_int32, C.c_int64, C.c_int32]
    lib.b200_mla_decode_workspace_bytes.restype = C.c_uint64
    lib.b200_swap_blocks.argtypes = [C.c_void_p, C.c_void_p, P(C.c_int64), C.c_int64, C.c_int64, C.c_int32, C.c_void_p]
    lib.b200_swap_blocks.restype = C.c_int
    lib.b200_gather_cache.argtypes = [P(GatherCacheParams), C.c_void_p]
    lib.b200_gather_cache.restype = C.c_int
    lib.b200_dcp_exchange_fold.argtypes = [P(DcpFoldParams), C.c_void_p]
    lib.b200_dcp_exchange_fold.restype = C.c_int
    lib.b200_dcp_gather_query.argtypes = [P(DcpGatherParams), C.c_void_p]
    lib.b200_dcp_gather_query.restype = C.c_int
    lib.b200_peer_alloc.argtypes = [C.c_uint64, P(C.c_void_p), C.c_void_p]
    lib.b200_peer_alloc.restype = C.c_int
    lib.b200_peer_open.argtypes = [C.c_void_p, P(C.c_void_p)]
    lib.b200_peer_open.restype = C.c_int
    lib.b200_peer_close.argtypes = [C.c_void_p]
    lib.b200_peer_close.restype = C.c_int
    lib.b200_peer_free.argtypes = [C.c_void_p]
    lib.b200_peer_free.restype = C.c_int
    lib.b200_kv_dtype_from_string.argtypes = [C.c_char_p]
    lib.b200_kv_dtype_from_string.restype = C.c_int
    lib.b200_last_error.argtypes = []
    lib.b200_last_error.restype = C.c_char_p
    lib.b200_abi_version.argtypes = []
    lib.b200_abi_version.restype = C.c_int
    lib.b200_set_attention_variant.argtypes = [C.c_int]
    lib.b200_set_attention_variant.restype = C.c_int
    lib.b200_last_attention_variant.argtypes = []
    lib.b200_last_attention_variant.restype = C.c_int
    lib.b200_launch_count.argtypes = []
    lib.b200_launch_count.restype = C.c_uint64
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != B200_OK:
        raise B200Error(rc, load().b200_last_error().decode())


def launch_count() -> int:
    return int(load().b200_launch_count())


def set_attention_variant(v: int) -> None:
    check(load().b200_set_attention_variant(v))


def last_attention_variant() -> int:
    return int(load().b200_last_attention_variant())


def set_stream_dynamic(min_chunk_tiles: int) -> None:
    """Experiment switch of the flat schedule's dynamic tail region (0 = static split, the default)."""
    lib = load()
    lib.b200_set_stream_dynamic.argtypes = [C.c_int]
    lib.b200_set_stream_dynamic.restype = C.c_int
    check(lib.b200_set_stream_dynamic(int(min_chunk_tiles)))


# ---- torch-tensor conveniences (pointers only; torch is plumbing here) ----------------------
def _torch_dtype_code(t) -> int:
    import torch

    return {torch.float32: DTYPE_F32, torch.float16: DTYPE_F16, torch.bfloat16: DTYPE_BF16,
            torch.uint8: DTYPE_U8, torch.float8_e4m3fn: DTYPE_U8}[t]


def kv_dtype_code(s: str) -> int:
    v = load().b200_kv_dtype_from_string(s.encode())
    if v < 0:
        raise B200Error(ERR_UNSUPPORTED, "Unsupported data type of kv cache: %s" % s)
    return v


def attn_params(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
                max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale, exp_sums=None, max_logits=None,
                tmp_out=None, workspace=None, out_lse=None) -> PagedAttnParams:
    p = PagedAttnParams()
    p.out = out.data_ptr()
    p.query = query.data_ptr()
    p.key_cache = key_cache.data_ptr()
    p.value_cache = value_cache.data_ptr()
    p.block_tables = block_tables.data_ptr()
    p.seq_lens = seq_lens.data_ptr()
    p.alibi_slopes = alibi_slopes.data_ptr() if alibi_slopes is not None else None
    p.k_scale = k_scale.data_ptr()
    p.v_scale = v_scale.data_ptr()
    if exp_sums is not None:
        p.exp_sums, p.max_logits, p.tmp_out = exp_sums.data_ptr(), max_logits.data_ptr(), tmp_out.data_ptr()
    if workspace is not None:
        p.workspace, p.workspace_bytes = workspace.data_ptr(), workspace.numel() * workspace.element_size()
    p.q_stride = query.stride(0)
    p.kv_block_stride = key_cache.stride(0)
    p.kv_head_stride = key_cache.stride(1)
    p.num_seqs, p.num_heads, p.head_size = query.shape
    p.num_kv_heads = num_kv_heads
    p.block_size = block_size
    p.max_seq_len = max_seq_len
    p.max_num_blocks_per_seq = block_tables.shape[1]
    p.dtype = _torch_dtype_code(query.dtype)
    p.kv_dtype = kv_dtype_code(kv_cache_dtype)
    p.scale = scale
    p.blocksparse_block_size = 64
    if out_lse is not None:
        p.out_lse = out_lse.data_ptr()
    return p


def workspace_bytes(p: PagedAttnParams) -> int:
    return int(load().b200_paged_attention_workspace_bytes(C.byref(p)))


def paged_attention_v1(p: PagedAttnParams, stream: int) -> None:
    check(load().b200_paged_attention_v1(C.byref(p), C.c_void_p(stream)))


def paged_attention_v1_static(p: PagedAttnParams, stream: int) -> None:
    """v1 with the library's static device scratch (no allocation; calls on one device must not overlap)."""
    check(load().b200_paged_attention_v1_static(C.byref(p), C.c_void_p(stream)))


def static_scratch_bytes() -> int:
    return int(load().b200_paged_attention_static_scratch_bytes())


def paged_attention_v2(p: PagedAttnParams, stream: int) -> None:
    check(load().b200_paged_attention_v2(C.byref(p), C.c_void_p(stream)))


def cache_params(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale,
                 flash: bool) -> CacheWriteParams:
    p = CacheWriteParams()
    p.key, p.value = key.data_ptr(), value.data_ptr()
    p.key_cache, p.value_cache = key_cache.data_ptr(), value_cache.data_ptr()
    p.slot_mapping = slot_mapping.data_ptr()
    p.k_scale, p.v_scale = k_scale.data_ptr(), v_scale.data_ptr()
    p.key_stride, p.value_stride = key.stride(0), value.stride(0)
    p.num_tokens = slot_mapping.shape[0]
    p.num_heads, p.head_size = key.shape[1], key.shape[2]
    p.dtype = _torch_dtype_code(key.dtype)
    p.kv_dtype = kv_dtype_code(kv_cache_dtype)
    if flash:
        p.block_size = key_cache.shape[1]
        p.block_stride, p.page_stride, p.head_stride = key_cache.stride(0), key_cache.stride(1), key_cache.stride(2)
    else:
        p.block_size = key_cache.shape[3]
        p.x = key_cache.shape[4]
    return p


def reshape_and_cache(p: CacheWriteParams, stream: int) -> None:
    check(load().b200_reshape_and_cache(C.byref(p), C.c_void_p(stream)))


def reshape_and_cache_flash(p: CacheWriteParams, stream: int) -> None:
    check(load().b200_reshape_and_cache_flash(C.byref(p), C.c_void_p(stream)))


def merge_attn_states(output, prefix_output, prefix_lse, suffix_output, suffix_lse, output_lse, stream: int) -> None:
    p = MergeAttnStatesParams()
    p.output = output.data_ptr()
    p.output_lse = output_lse.data_ptr() if output_lse is not None else None
    p.prefix_output, p.prefix_lse = prefix_output.data_ptr(), prefix_lse.data_ptr()
    p.suffix_output, p.suffix_lse = suffix_output.data_ptr(), suffix_lse.data_ptr()
    p.output_stride, p.prefix_stride, p.suffix_stride = output.stride(0), prefix_output.stride(0), suffix_output.stride(0)
    p.num_tokens, p.num_heads, p.head_size = output.shape
    p.dtype = _torch_dtype_code(output.dtype)
    check(load().b200_merge_attn_states(C.byref(p), C.c_void_p(stream)))


def flash_attn_params(out, query, key_cache, value_cache, scale, block_tables, seq_lens, max_seq_len, alibi_slopes=None,
                      workspace=None, out_lse=None) -> PagedAttnParams:
    """Parameters of b200_paged_attention_flash: caches [num_blocks, block_size, num_kv_heads, head_size]."""
    p = PagedAttnParams()
    p.out, p.query = out.data_ptr(), query.data_ptr()
    p.key_cache, p.value_cache = key_cache.data_ptr(), value_cache.data_ptr()
    p.block_tables, p.seq_lens = block_tables.data_ptr(), seq_lens.data_ptr()
    p.alibi_slopes = alibi_slopes.data_ptr() if alibi_slopes is not None else None
    if workspace is not None:
        p.workspace, p.workspace_bytes = workspace.data_ptr(), workspace.numel() * workspace.element_size()
    if out_lse is not None:
        p.out_lse = out_lse.data_ptr()
    p.q_stride = query.stride(0)
    p.kv_block_stride, p.kv_page_stride, p.kv_head_stride = key_cache.stride(0), key_cache.stride(1), key_cache.stride(2)
    p.num_seqs, p.num_heads, p.head_size = query.shape
    p.num_kv_heads, p.block_size = key_cache.shape[2], key_cache.shape[1]
    p.kv_num_blocks = key_cache.shape[0]
    p.max_seq_len = max_seq_len
    p.max_num_blocks_per_seq = block_tables.shape[1]
    p.dtype = _torch_dtype_code(query.dtype)
    p.kv_dtype = KV_AUTO
    p.scale = scale
    p.blocksparse_block_size = 64
    return p


def paged_attention_flash(p: PagedAttnParams, stream: int) -> None:
    check(load().b200_paged_attention_flash(C.byref(p), C.c_void_p(stream)))


def peer_alloc(nbytes: int) -> tuple[int, bytes]:
    """(device pointer, 64-byte IPC handle) of a zero-filled cudaMalloc on the current device."""
    ptr = C.c_void_p()
    handle = C.create_string_buffer(64)
    check(load().b200_peer_alloc(C.c_uint64(nbytes), C.byref(ptr), handle))
    return ptr.value, handle.raw


def peer_open(handle: bytes) -> int:
    ptr = C.c_void_p()
    buf = C.create_string_buffer(handle, 64)
    check(load().b200_peer_open(buf, C.byref(ptr)))
    return ptr.value


def peer_close(ptr: int) -> None:
    check(load().b200_peer_close(C.c_void_p(ptr)))


def peer_free(ptr: int) -> None:
    check(load().b200_peer_free(C.c_void_p(ptr)))


def dcp_fold_params(peer_out, peer_lse, peer_flags, rank, num_seqs, heads_local, head_size, dtype,
                    step_counter: int | None = None) -> DcpFoldParams:
    """Everything of a fold launch that does not change from step to step (peer_*: `world` device addresses).
    step_counter: device address of the step word the paired gather advances (then `epoch` is ignored)."""
    world = len(peer_out)
    arr = C.c_void_p * world
    p = DcpFoldParams()
    p._keep = (arr(*peer_out), arr(*peer_lse), arr(*peer_flags))  # the struct only holds their addresses
    p.peer_out, p.peer_lse, p.peer_flags = p._keep
    p.rank, p.world, p.num_seqs, p.heads_local, p.head_size = rank, world, num_seqs, heads_local, head_size
    p.dtype = _torch_dtype_code(dtype)
    p.step_counter = step_counter
    return p


def dcp_fold_launch(p: DcpFoldParams, out, out_lse, epoch: int, stream: int) -> None:
    p.out = out.data_ptr()
    p.out_lse = out_lse.data_ptr() if out_lse is not None else None
    p.epoch = epoch
    check(load().b200_dcp_exchange_fold(C.byref(p), C.c_void_p(stream)))


def dcp_exchange_fold(peer_out, peer_lse, peer_flags, out, out_lse, rank, num_seqs, heads_local, head_size, epoch,
                      stream: int) -> None:
    """peer_out / peer_lse / peer_flags: sequences of `world` device addresses (ints)."""
    p = dcp_fold_params(peer_out, peer_lse, peer_flags, rank, num_seqs, heads_local, head_size, out.dtype)
    dcp_fold_launch(p, out, out_lse, epoch, stream)


def dcp_gather_params(peer_q, peer_flags, cta_counter: int, rank, num_seqs, heads_local, head_size,
                      dtype, step_counter: int | None = None) -> DcpGatherParams:
    world = len(peer_q)
    arr = C.c_void_p * world
    p = DcpGatherParams()
    p._keep = (arr(*peer_q), arr(*peer_flags))
    p.peer_q, p.peer_flags = p._keep
    p.cta_counter = cta_counter
    p.rank, p.world, p.num_seqs, p.heads_local, p.head_size = rank, world, num_seqs, heads_local, head_size
    p.dtype = _torch_dtype_code(dtype)
    p.step_counter = step_counter
    return p


def dcp_gather_launch(p: DcpGatherParams, query, epoch: int, stream: int) -> None:
    p.query, p.q_stride, p.epoch = query.data_ptr(), query.stride(0), epoch
    check(load().b200_dcp_gather_query(C.byref(p), C.c_void_p(stream)))


def dcp_gather_query(peer_q, peer_flags, query, cta_counter: int, rank, epoch, stream: int) -> None:
    """query [S, H_local, D] with contiguous heads; peer_q / peer_flags: `world` device addresses."""
    S, Hl, D = query.shape
    p = dcp_gather_params(peer_q, peer_flags, cta_counter, rank, S, Hl, D, query.dtype)
    dcp_gather_launch(p, query, epoch, stream)
