"""Host-side helpers around the decode ops, shaped like the callers the reference relies on.

`PagedAttention` mirrors the helper the reference's op users call (upstream
`vllm/v1/attention/ops/paged_attn.py`: `split_kv_cache`, `write_to_paged_cache`; the reference's
backends reach the cache writer the same way, vllm_metax/v1/attention/backends/flash_attn.py:661-687) and
adds `forward_decode`, which picks paged_attention_v1 or _v2 the way the callers of these ops did
(v1 when the context is short or the grid is already large, else the split-K v2 with caller-allocated
scratch).  `shard_kv_heads` is the tensor-parallel partition rule of the path
(SURVEY.md §8e): ranks own contiguous KV heads and their query heads; there is no collective inside.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

from .. import _custom_ops as ops

_PARTITION_SIZE = 512  # layout constant of the v2 scratch tensors (reference csrc/attention/paged_attention_v2.cu:45)


class PagedAttention:
    @staticmethod
    def get_supported_head_sizes() -> list[int]:
        # reference csrc/attention/paged_attention_v1.cu:90-124
        return [32, 64, 80, 96, 112, 120, 128, 192, 256]

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int) -> tuple[int, ...]:
        return (2, num_blocks, block_size * num_kv_heads * head_size)

    @staticmethod
    def split_kv_cache(kv_cache: torch.Tensor, num_kv_heads: int, head_size: int) -> tuple[torch.Tensor, torch.Tensor]:
        """[2, num_blocks, B*Hkv*D] -> key_cache [NB,Hkv,D/x,B,x], value_cache [NB,Hkv,D,B], x = 16 / elt size."""
        x = 16 // kv_cache.element_size()
        num_blocks = kv_cache.shape[1]
        key_cache = kv_cache[0].view(num_blocks, num_kv_heads, head_size // x, -1, x)
        value_cache = kv_cache[1].view(num_blocks, num_kv_heads, head_size, -1)
        return key_cache, value_cache

    @staticmethod
    def write_to_paged_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale) -> None:
        ops.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping.flatten(), kv_cache_dtype, k_scale,
                              v_scale)

    @staticmethod
    def forward_decode(query, key_cache, value_cache, block_tables, seq_lens, max_seq_len: int, kv_cache_dtype: str,
                       num_kv_heads: int, scale: float, alibi_slopes, k_scale, v_scale, output=None,
                       force_version: str | None = None) -> torch.Tensor:
        """One decode attention call over the paged cache; returns [num_seqs, num_heads, head_size]."""
        out = torch.empty_like(query) if output is None else output
        block_size = value_cache.shape[3]
        num_seqs, num_heads, head_size = query.shape
        max_num_partitions = (max_seq_len + _PARTITION_SIZE - 1) // _PARTITION_SIZE
        use_v1 = max_seq_len <= 8192 and (max_num_partitions == 1 or num_seqs * num_heads > 512)
        if force_version is not None:
            use_v1 = force_version == "v1"
        if use_v1:
            ops.paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                                   block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale)
        else:
            tmp_out = torch.empty((num_seqs, num_heads, max_num_partitions, head_size), dtype=out.dtype, device=out.device)
            exp_sums = torch.empty((num_seqs, num_heads, max_num_partitions), dtype=torch.float32, device=out.device)
            max_logits = torch.empty_like(exp_sums)
            ops.paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache, num_kv_heads,
                                   scale, block_tables, seq_lens, block_size, max_seq_len, alibi_slopes,
                                   kv_cache_dtype, k_scale, v_scale)
        return out


def shard_kv_heads(num_heads: int, num_kv_heads: int, tp_size: int, tp_rank: int) -> tuple[slice, slice]:
    """(query-head slice, kv-head slice) owned by `tp_rank`: contiguous KV heads and the query heads of their
    GQA groups, as QKVParallelLinear shards them; KV heads are replicated when tp_size > num_kv_heads."""
    if num_heads % num_kv_heads or tp_size <= 0 or not 0 <= tp_rank < tp_size:
        raise ValueError("bad head / rank configuration")
    group = num_heads // num_kv_heads
    if tp_size <= num_kv_heads:
        if num_kv_heads % tp_size:
            raise ValueError("num_kv_heads must be divisible by tp_size")
        per = num_kv_heads // tp_size
        kv = slice(tp_rank * per, (tp_rank + 1) * per)
        return slice(kv.start * group, kv.stop * group), kv
    if tp_size % num_kv_heads or num_heads % tp_size:
        raise ValueError("tp_size must be a multiple of num_kv_heads and divide num_heads")
    rep = tp_size // num_kv_heads  # ranks sharing one kv head
    qper = num_heads // tp_size
    kvh = tp_rank // rep
    return slice(tp_rank * qper, (tp_rank + 1) * qper), slice(kvh, kvh + 1)


@dataclass
class _Field:
    name: str
    shape: tuple
    dtype: torch.dtype
    offset: int
    nbytes: int


class PinnedDecodeInputs:
    """Per-step host inputs of a decode layer-call packed into ONE pinned buffer, so that the step costs a
    single H2D copy: query, new key/value rows, slot_mapping, block_tables, seq_lens.  `host(name)` gives the
    pinned views to fill, `upload()` enqueues the copy on the current stream, `dev(name)` gives device views."""

    def __init__(self, fields: dict[str, tuple[tuple, torch.dtype]], device):
        off = 0
        self._fields: dict[str, _Field] = {}
        for name, (shape, dtype) in fields.items():
            n = 1
            for s in shape:
                n *= s
            nbytes = n * torch.tensor([], dtype=dtype).element_size()
            self._fields[name] = _Field(name, tuple(shape), dtype, off, nbytes)
            off = (off + nbytes + 255) // 256 * 256
        self.nbytes = off
        self._host = torch.empty(off, dtype=torch.uint8).pin_memory()
        self._dev = torch.empty(off, dtype=torch.uint8, device=device)

    def _view(self, buf, f: _Field):
        return buf[f.offset:f.offset + f.nbytes].view(f.dtype).view(f.shape)

    def host(self, name: str) -> torch.Tensor:
        return self._view(self._host, self._fields[name])

    def dev(self, name: str) -> torch.Tensor:
        return self._view(self._dev, self._fields[name])

    def upload(self) -> None:
        self._dev.copy_(self._host, non_blocking=True)
