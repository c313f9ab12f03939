"""Python op wrappers of the hot path, with the argument lists callers of the reference use.

The reference's kernel tests and attention backends reach the ops through these names
(`ops.paged_attention_v1`, `ops.reshape_and_cache_flash`, ... — reference:
tests/kernels/attention/test_attention.py:200-267, tests/kernels/attention/test_cache.py:177-182,301-306,
vllm_metax/v1/attention/backends/flash_attn.py:678-687 via vllm_metax/attention/utils/fa_utils.py:12).
Each wrapper only forwards to the torch-dispatcher op registered by `vllm_metax_b200._C`
(schemas: reference csrc/torch_bindings.cpp:45-69,394-426); defaults follow the wrappers the reference
imports (`tp_rank=0`, block-sparse `(0, 0, 64, 0)`, `convert_fp8(..., scale=1.0, kv_dtype="fp8")`).
"""
from __future__ import annotations

import torch

from . import import_kernels

import_kernels()


def paged_attention_v1(
    out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: torch.Tensor | None,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
    tp_rank: int = 0,
    blocksparse_local_blocks: int = 0,
    blocksparse_vert_stride: int = 0,
    blocksparse_block_size: int = 64,
    blocksparse_head_sliding_step: int = 0,
) -> None:
    torch.ops._C.paged_attention_v1(
        out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len,
        alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank, blocksparse_local_blocks, blocksparse_vert_stride,
        blocksparse_block_size, blocksparse_head_sliding_step)


def paged_attention_v2(
    out: torch.Tensor,
    exp_sum: torch.Tensor,
    max_logits: torch.Tensor,
    tmp_out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: torch.Tensor | None,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
    tp_rank: int = 0,
    blocksparse_local_blocks: int = 0,
    blocksparse_vert_stride: int = 0,
    blocksparse_block_size: int = 64,
    blocksparse_head_sliding_step: int = 0,
) -> None:
    torch.ops._C.paged_attention_v2(
        out, exp_sum, max_logits, tmp_out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
        seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank,
        blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size, blocksparse_head_sliding_step)


def reshape_and_cache(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    torch.ops._C_cache_ops.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype,
                                             k_scale, v_scale)


def reshape_and_cache_flash(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    torch.ops._C_cache_ops.reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype,
                                                   k_scale, v_scale)


def convert_fp8(output: torch.Tensor, input: torch.Tensor, scale: float = 1.0, kv_dtype: str = "fp8") -> None:
    torch.ops._C_cache_ops.convert_fp8(output, input, scale, kv_dtype)


def merge_attn_states(
    output: torch.Tensor,
    prefix_output: torch.Tensor,
    prefix_lse: torch.Tensor,
    suffix_output: torch.Tensor,
    suffix_lse: torch.Tensor,
    output_lse: torch.Tensor | None = None,
) -> None:
    """Argument order of the wrapper the reference dispatches to (reference:
    vllm_metax/attention/ops/merge_attn_states.py:9-16,39-43; op schema csrc/torch_bindings.cpp:74-82)."""
    torch.ops._C.merge_attn_states(output, output_lse, prefix_output, prefix_lse, suffix_output, suffix_lse)


def copy_blocks(key_caches: list[torch.Tensor], value_caches: list[torch.Tensor], block_mapping: torch.Tensor) -> None:
    """reference: tests/kernels/attention/test_cache.py:104 -> csrc/cache_kernels.cu:113-163"""
    torch.ops._C_cache_ops.copy_blocks(key_caches, value_caches, block_mapping)


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: torch.Tensor) -> None:
    """reference: tests/kernels/attention/test_cache.py:422-425 -> csrc/cache_kernels.cu:18-60"""
    torch.ops._C_cache_ops.swap_blocks(src, dst, block_mapping)


def gather_and_maybe_dequant_cache(
    src_cache: torch.Tensor,
    dst: torch.Tensor,
    block_table: torch.Tensor,
    cu_seq_lens: torch.Tensor,
    batch_size: int,
    kv_cache_dtype: str,
    scale: torch.Tensor,
    seq_starts: torch.Tensor | None = None,
) -> None:
    """reference: tests/kernels/attention/test_cache.py:787-789 -> csrc/cache_kernels.cu:702-773"""
    torch.ops._C_cache_ops.gather_and_maybe_dequant_cache(src_cache, dst, block_table, cu_seq_lens, batch_size,
                                                          kv_cache_dtype, scale, seq_starts)


def cp_gather_cache(
    src_cache: torch.Tensor,
    dst: torch.Tensor,
    block_table: torch.Tensor,
    cu_seq_lens: torch.Tensor,
    batch_size: int,
    seq_starts: torch.Tensor | None = None,
) -> None:
    """reference: tests/kernels/attention/test_cache.py:861 -> csrc/cache_kernels.cu:1075-1134"""
    torch.ops._C_cache_ops.cp_gather_cache(src_cache, dst, block_table, cu_seq_lens, batch_size, seq_starts)


def paged_attention_v1_lse(
    out: torch.Tensor,
    lse: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: torch.Tensor | None,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    """paged_attention_v1 that also writes lse[num_heads, num_seqs] (natural-log sum-exp of the scaled logits, -inf for
    an empty sequence) — the operand merge_attn_states needs to merge this decode result with another partial result.
    Not a reference op (namespace `_C_b200`)."""
    torch.ops._C_b200.paged_attention_v1_lse(out, lse, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
                                             seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale,
                                             v_scale)


def paged_attention_flash(
    out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    max_seq_len: int,
    alibi_slopes: torch.Tensor | None = None,
    lse: torch.Tensor | None = None,
) -> None:
    """Decode attention over the flash cache layout [num_blocks, block_size, num_kv_heads, head_size] (NHD or HND
    strides) written by reshape_and_cache_flash; one query token per sequence.  Not a reference op (namespace
    `_C_b200`): the reference gets this from flash_attn_with_kvcache of the external flash-attention library
    (reference: vllm_metax/v1/attention/backends/flash_attn.py:742-757); see attention/flash_decode.py for the wrapper
    with that signature."""
    torch.ops._C_b200.paged_attention_flash(out, lse, query, key_cache, value_cache, scale, block_tables, seq_lens,
                                            max_seq_len, alibi_slopes)


def rotary_embedding(
    positions: torch.Tensor,
    query: torch.Tensor,
    key: torch.Tensor | None,
    head_size: int,
    cos_sin_cache: torch.Tensor,
    is_neox: bool,
) -> None:
    """In-place rotary position embedding of query (and key).  reference: csrc/torch_bindings.cpp:216-219,
    csrc/pos_encoding_kernels.cu:126-209; wrapper as upstream `ops.rotary_embedding`."""
    torch.ops._C.rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox)


def rotary_embedding_and_cache_flash(
    positions: torch.Tensor,
    query: torch.Tensor,
    key: torch.Tensor,
    value: torch.Tensor,
    head_size: int,
    cos_sin_cache: torch.Tensor,
    is_neox: bool,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    """rotary_embedding(positions, query, key, ...) followed by reshape_and_cache_flash(key, value, ...) in ONE launch
    (same results, byte for byte, as the two reference ops back to back).  Not a reference op (namespace `_C_b200`)."""
    torch.ops._C_b200.rotary_embedding_and_cache_flash(positions, query, key, value, head_size, cos_sin_cache, is_neox,
                                                       key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale,
                                                       v_scale)


def concat_and_cache_mla(
    kv_c: torch.Tensor,
    k_pe: torch.Tensor,
    kv_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    scale: torch.Tensor,
) -> None:
    """reference: tests/kernels/attention/test_cache.py:563-564 -> csrc/cache_kernels.cu:490-539"""
    torch.ops._C_cache_ops.concat_and_cache_mla(kv_c, k_pe, kv_cache, slot_mapping, kv_cache_dtype, scale)


def copy_blocks_mla(kv_caches: list[torch.Tensor], block_mapping: torch.Tensor) -> None:
    """reference: tests/kernels/attention/test_cache.py:642 -> csrc/cache_kernels.cu:165-201"""
    torch.ops._C_cache_ops.copy_blocks_mla(kv_caches, block_mapping)


def reserve_workspace(query_like: torch.Tensor, max_seq_len: int, reset: bool = False) -> None:
    """Size (and optionally re-zero) the current stream's split-KV workspace for decode batches shaped like
    `query_like` [num_seqs, num_heads, head_size] with contexts up to `max_seq_len`.  Not a reference op: the
    reference's paged_attention_v1 never splits and so has no workspace (csrc/attention/paged_attention_v1.cu:43-125).
    Call it once at warm-up on every stream that will run (or capture) paged_attention_v1 / paged_attention_flash, so
    that those ops never allocate; without it their first call on a stream allocates ~27 MB lazily, and raises if that
    first call happens under CUDA-graph capture."""
    torch.ops._C_b200.reserve_workspace(query_like, max_seq_len, reset)


def cutlass_mla_decode(
    out: torch.Tensor,
    q_nope: torch.Tensor,
    q_pe: torch.Tensor,
    kv_c_and_k_pe_cache: torch.Tensor,
    seq_lens: torch.Tensor,
    page_table: torch.Tensor,
    scale: float,
) -> torch.Tensor:
    """reference: csrc/torch_bindings.cpp:287-290 -> csrc/attention/mla/cutlass_mla_entry.cu:28-38.  q_nope
    [B, H, 512], q_pe [B, H, 64], cache [num_blocks, page, 576], out [B, H, 512]; H <= 128 (the reference needs 128)."""
    torch.ops._C.cutlass_mla_decode(out, q_nope, q_pe, kv_c_and_k_pe_cache, seq_lens, page_table, scale)
    return out


def sm100_cutlass_mla_decode(
    out: torch.Tensor,
    lse: torch.Tensor,
    q_nope: torch.Tensor,
    q_pe: torch.Tensor,
    kv_c_and_k_pe_cache: torch.Tensor,
    seq_lens: torch.Tensor,
    page_table: torch.Tensor,
    workspace: torch.Tensor,
    scale: float,
    num_kv_splits: int,
) -> torch.Tensor:
    """the call the reference's test makes (tests/kernels/attention/test_cutlass_mla_decode.py:121-123; upstream
    wrapper _custom_ops.py:3017-3041): also writes lse [B, H] (natural log), workspace from
    sm100_cutlass_mla_get_workspace_size, num_kv_splits -1 = choose / 1 = no token split."""
    torch.ops._C.sm100_cutlass_mla_decode(out, lse, q_nope, q_pe, kv_c_and_k_pe_cache, seq_lens, page_table, workspace,
                                          scale, num_kv_splits)
    return out


def sm100_cutlass_mla_get_workspace_size(max_seq_len: int, num_batches: int, sm_count: int, num_kv_splits: int) -> int:
    """reference test :112-113; upstream wrapper _custom_ops.py:3044-3049"""
    return torch.ops._C.sm100_cutlass_mla_get_workspace_size(max_seq_len, num_batches, sm_count, num_kv_splits)
