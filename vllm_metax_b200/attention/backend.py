"""Decode branch of the reference's FlashAttention backend, on the B200 ops.

Mirrors the pieces of `vllm_metax/v1/attention/backends/flash_attn.py` that sit directly on the hot path, with the same
names and argument meaning, so that the reference's `AttentionImpl` can delegate its decode branch here:

  * `MacaFlashAttentionBackend.get_kv_cache_shape / get_kv_cache_stride_order` (:101-132): `(2, num_blocks, block_size,
    num_kv_heads, head_size)`, NHD or HND storage;
  * `FlashAttentionMetadata` (:160-198) reduced to the fields the decode branch reads: `num_actual_tokens`,
    `num_decode_tokens`, `slot_mapping`, `decode_block_table`, `decode_seq_lens`, `max_seq_len`;
  * `FlashAttentionImpl.forward` (:600-765) for a batch of decode tokens: `reshape_and_cache_flash` of the step's K/V
    (slot_mapping decides the token count, :671-687), then `flash_attn_with_kvcache` over the paged cache (:742-762);
  * the split + `merge_attn_states` combination that cascade attention / decode context parallel apply to partial
    results (:766-872): `forward_decode_split` attends to a prefix and a suffix range of every sequence and merges them.

This is host-side glue only; the engine's class hierarchy (`AttentionBackend`, `AttentionMetadataBuilder`, scheduler
metadata, prefill, encoder attention) is upstream vLLM's and out of scope.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

from .. import _custom_ops as ops
from .flash_decode import flash_attn_with_kvcache


@dataclass
class DecodeMetadata:
    """Fields of the reference's FlashAttentionMetadata that the decode branch reads (flash_attn.py:160-198)."""
    num_actual_tokens: int
    num_decode_tokens: int
    slot_mapping: torch.Tensor        # [num_actual_tokens] int64
    decode_block_table: torch.Tensor  # [num_decodes, max_blocks] int32
    decode_seq_lens: torch.Tensor     # [num_decodes] int32
    max_seq_len: int


class B200FlashDecodeBackend:
    @staticmethod
    def get_supported_head_sizes() -> list[int]:
        return [128]

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int) -> tuple[int, ...]:
        if block_size % 16 != 0:  # flash_attn.py:108-109
            raise ValueError("Block size must be a multiple of 16.")
        return (2, num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def get_kv_cache_stride_order(layout: str = "NHD") -> tuple[int, ...]:
        if layout == "NHD":
            return (0, 1, 2, 3, 4)
        if layout == "HND":
            return (0, 1, 3, 2, 4)
        raise ValueError("Unknown cache layout format %s." % layout)


class B200FlashDecodeImpl:
    """`FlashAttentionImpl` for decode tokens (flash_attn.py:532-765)."""

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: int,
                 alibi_slopes: list[float] | None = None, sliding_window: int | None = None,
                 kv_cache_dtype: str = "auto", logits_soft_cap: float | None = None):
        if sliding_window is not None or logits_soft_cap:
            raise NotImplementedError("sliding window / logits soft cap are not served by the decode kernel")
        if kv_cache_dtype != "auto":
            # the reference's flash path has no fp8 cache either (attention/utils/fa_utils.py:26-30)
            raise NotImplementedError("the flash-layout decode kernel takes a bf16 cache")
        self.num_heads, self.head_size, self.scale, self.num_kv_heads = num_heads, head_size, float(scale), num_kv_heads
        self.kv_cache_dtype = kv_cache_dtype
        self._alibi_list = alibi_slopes
        self._alibi: torch.Tensor | None = None
        self._one: torch.Tensor | None = None

    def _consts(self, device):
        if self._one is None or self._one.device != device:
            self._one = torch.ones(1, dtype=torch.float32, device=device)
            self._alibi = (torch.tensor(self._alibi_list, dtype=torch.float32, device=device)
                           if self._alibi_list is not None else None)
        return self._one, self._alibi

    def forward(self, query: torch.Tensor, key: torch.Tensor | None, value: torch.Tensor | None, kv_cache: torch.Tensor,
                attn_metadata: DecodeMetadata, output: torch.Tensor) -> torch.Tensor:
        """query [num_tokens, num_heads, head_size]; key/value [num_tokens, num_kv_heads, head_size] (may be padded past
        slot_mapping, :671-677); kv_cache [2, num_blocks, block_size, num_kv_heads, head_size]; output like query."""
        one, alibi = self._consts(query.device)
        key_cache, value_cache = kv_cache.unbind(0)
        if key is not None and value is not None:
            ops.reshape_and_cache_flash(key, value, key_cache, value_cache, attn_metadata.slot_mapping,
                                        self.kv_cache_dtype, one, one)
        n = attn_metadata.num_decode_tokens
        if n > 0:
            flash_attn_with_kvcache(q=query[:n].unsqueeze(1), k_cache=key_cache, v_cache=value_cache,
                                    block_table=attn_metadata.decode_block_table,
                                    cache_seqlens=attn_metadata.decode_seq_lens, softmax_scale=self.scale, causal=True,
                                    alibi_slopes=alibi, out=output[:n].unsqueeze(1),
                                    max_seqlen_k=attn_metadata.max_seq_len)
        return output

    def forward_decode_split(self, query: torch.Tensor, kv_cache: torch.Tensor, block_table: torch.Tensor,
                             seq_lens: torch.Tensor, split_lens: torch.Tensor, max_seq_len: int,
                             output: torch.Tensor, output_lse: torch.Tensor | None = None) -> torch.Tensor:
        """Attend to tokens [0, split) and [split, len) of every sequence separately and combine the two partial
        results with merge_attn_states — the combination rule of cascade attention and decode context parallel
        (flash_attn.py:766-872).  `split_lens` must be multiples of the block size (the suffix starts at a block)."""
        _, alibi = self._consts(query.device)
        if alibi is not None:
            raise NotImplementedError("ALiBi positions of a suffix range are not expressible through seq_lens alone")
        key_cache, value_cache = kv_cache.unbind(0)
        block_size = key_cache.shape[1]
        S, H, _ = query.shape
        o1, o2 = torch.empty_like(query), torch.empty_like(query)
        l1 = torch.empty(H, S, dtype=torch.float32, device=query.device)
        l2 = torch.empty_like(l1)
        ops.paged_attention_flash(o1, query, key_cache, value_cache, self.scale, block_table, split_lens, max_seq_len,
                                  None, l1)
        # the suffix of sequence i starts at block split_lens[i] / block_size of its table
        cols = torch.arange(block_table.shape[1], device=query.device, dtype=torch.int64).unsqueeze(0)
        shifted = (cols + (split_lens.to(torch.int64) // block_size).unsqueeze(1)).clamp_(max=block_table.shape[1] - 1)
        suffix_table = torch.gather(block_table, 1, shifted)
        ops.paged_attention_flash(o2, query, key_cache, value_cache, self.scale, suffix_table,
                                  (seq_lens - split_lens).to(torch.int32), max_seq_len, None, l2)
        ops.merge_attn_states(output, o1, l1, o2, l2, output_lse)
        return output
