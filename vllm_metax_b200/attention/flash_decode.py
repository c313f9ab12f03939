"""`flash_attn_with_kvcache` for the decode step, on the B200 flash-layout kernel.

The reference's serving path does not call paged_attention_v1/v2: it writes the flash cache layout with
reshape_and_cache_flash and decodes with `flash_attn_with_kvcache` imported from the external (MetaX build of the)
flash-attention library (reference: vllm_metax/attention/utils/fa_utils.py:12;
vllm_metax/v1/attention/backends/flash_attn.py:742-757).  This module keeps that call's names and argument meaning for
the case the backend uses it for — one query token per sequence over a paged cache — so the backend's decode branch
can bind it unchanged.  Anything the decode branch never passes (new k/v to append, rotary, several query tokens,
sliding window, soft cap) raises NotImplementedError instead of falling back.
"""
from __future__ import annotations

import torch

from .. import _custom_ops as ops


def flash_attn_with_kvcache(
    q: torch.Tensor,
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    k: torch.Tensor | None = None,
    v: torch.Tensor | None = None,
    rotary_cos: torch.Tensor | None = None,
    rotary_sin: torch.Tensor | None = None,
    cache_seqlens: torch.Tensor | None = None,
    cache_batch_idx: torch.Tensor | None = None,
    cache_leftpad: torch.Tensor | None = None,
    block_table: torch.Tensor | None = None,
    softmax_scale: float | None = None,
    causal: bool = False,
    window_size: tuple[int, int] | list[int] | None = (-1, -1),
    softcap: float = 0.0,
    rotary_interleaved: bool = True,
    alibi_slopes: torch.Tensor | None = None,
    num_splits: int = 0,
    return_softmax_lse: bool = False,
    out: torch.Tensor | None = None,
    max_seqlen_k: int | None = None,
    **unused,
):
    """q [batch, 1, num_heads, head_size]; k_cache / v_cache [num_blocks, block_size, num_kv_heads, head_size] (NHD or
    HND strides); cache_seqlens [batch] int32; block_table [batch, max_blocks] int32.  Returns out [batch, 1, num_heads,
    head_size] (and softmax_lse [batch, num_heads, 1] when return_softmax_lse).  `causal` is irrelevant for one query
    token at the end of the sequence.  `max_seqlen_k` (not a flash-attn argument) saves the device read of
    cache_seqlens.max(); without it the block table's capacity is used, which costs nothing: the kernel schedules from
    cache_seqlens on the device."""
    if k is not None or v is not None or rotary_cos is not None or rotary_sin is not None:
        raise NotImplementedError("flash_attn_with_kvcache: appending k/v or rotary is not part of the decode path")
    if cache_batch_idx is not None or cache_leftpad is not None:
        raise NotImplementedError("flash_attn_with_kvcache: cache_batch_idx / cache_leftpad are not supported")
    if block_table is None or cache_seqlens is None:
        raise NotImplementedError("flash_attn_with_kvcache: only the paged cache (block_table + cache_seqlens) is supported")
    if q.dim() != 4 or q.shape[1] != 1:
        raise NotImplementedError("flash_attn_with_kvcache: the decode kernel takes one query token per sequence")
    if window_size is not None and tuple(window_size) != (-1, -1):
        raise NotImplementedError("flash_attn_with_kvcache: sliding window is not supported by the decode kernel")
    if softcap:
        raise NotImplementedError("flash_attn_with_kvcache: logits soft cap is not supported by the decode kernel")
    batch, _, num_heads, head_size = q.shape
    scale = float(head_size ** -0.5) if softmax_scale is None else float(softmax_scale)
    q3 = q.view(batch, num_heads, head_size) if q.is_contiguous() else q[:, 0]
    result = torch.empty((batch, 1, num_heads, head_size), dtype=q.dtype, device=q.device) if out is None else out
    out3 = result.view(batch, num_heads, head_size)
    lse = torch.empty((num_heads, batch), dtype=torch.float32, device=q.device) if return_softmax_lse else None
    if cache_seqlens.dtype != torch.int32:
        cache_seqlens = cache_seqlens.to(torch.int32)
    if max_seqlen_k is None:
        max_seqlen_k = block_table.shape[1] * k_cache.shape[1]
    if alibi_slopes is not None and alibi_slopes.dim() == 2:
        raise NotImplementedError("flash_attn_with_kvcache: per-batch alibi slopes are not supported")
    ops.paged_attention_flash(out3, q3, k_cache, v_cache, scale, block_table, cache_seqlens, int(max_seqlen_k),
                              alibi_slopes, lse)
    if return_softmax_lse:
        return result, lse.t().unsqueeze(-1)
    return result
