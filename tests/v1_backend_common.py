"""Shared pieces of the v1-backend tests, shaped like the reference's tests/v1/attention/utils.py (BatchSpec,
create_common_attn_metadata) and tests/v1/attention/test_attention_backends.py (BATCH_SPECS, the prepopulated paged cache,
MockAttentionLayer) — without the model download `create_vllm_config` needs: the builder reads only head counts, head
size and the block size, which a stand-in config provides."""
from __future__ import annotations

from dataclasses import dataclass
from types import SimpleNamespace

import torch


@dataclass
class BatchSpec:  # tests/v1/attention/utils.py:36-52
    seq_lens: list[int]
    query_lens: list[int]

    @property
    def batch_size(self):
        return len(self.seq_lens)

    def compute_num_tokens(self):
        return sum(self.query_lens)


# tests/v1/attention/test_attention_backends.py:46-70, reordered decode-first where the spec mixes the two (the model
# runner reorders the batch for backends with reorder_batch_threshold; the reference's own mixed specs are already so)
BATCH_SPECS = {
    "small_decode": BatchSpec(seq_lens=[32, 40], query_lens=[1, 1]),
    "small_prefill": BatchSpec(seq_lens=[32, 40], query_lens=[8, 8]),
    "mixed_small": BatchSpec(seq_lens=[32, 40, 48, 56], query_lens=[1, 1, 5, 5]),
    "medium_decode": BatchSpec(seq_lens=[128, 256, 512, 1024, 128, 256, 512, 1024], query_lens=[1] * 8),
    "medium_prefill": BatchSpec(seq_lens=[256, 512, 1024, 2048], query_lens=[16, 16, 16, 16]),
    "mixed_medium": BatchSpec(seq_lens=[512, 1024, 2048, 512, 1024, 2048], query_lens=[1, 1, 1, 7, 7, 7]),
    "large_decode": BatchSpec(seq_lens=[2048] * 32, query_lens=[1] * 32),
    "large_prefill": BatchSpec(seq_lens=[4096] * 8, query_lens=[32] * 8),
    "single_decode": BatchSpec(seq_lens=[1024], query_lens=[1]),
    "single_prefill": BatchSpec(seq_lens=[1024], query_lens=[64]),
}


def stand_in_config(num_heads=32, num_kv_heads=8, head_size=128, block_size=16, tp=1):
    """The attributes of VllmConfig the builder reads (flash_attn.py:232-245), for Llama-3-8B's attention shape."""
    model_config = SimpleNamespace(
        get_num_attention_heads=lambda pc: num_heads // pc.tensor_parallel_size,
        get_num_kv_heads=lambda pc: max(1, num_kv_heads // pc.tensor_parallel_size),
        get_head_size=lambda: head_size, dtype=torch.bfloat16)
    return SimpleNamespace(model_config=model_config, parallel_config=SimpleNamespace(tensor_parallel_size=tp),
                           cache_config=SimpleNamespace(block_size=block_size, cache_dtype="auto"))


def create_common_attn_metadata(spec: BatchSpec, block_size: int, device, max_block_idx: int = 1000):
    """tests/v1/attention/utils.py:55-108 (arange_block_indices=False: random table, overwritten by the cache builder)."""
    from vllm.v1.attention.backend import CommonAttentionMetadata

    qsl = torch.zeros(spec.batch_size + 1, dtype=torch.int32)
    qsl[1:] = torch.tensor(spec.query_lens, dtype=torch.int32).cumsum(0)
    seq_lens_cpu = torch.tensor(spec.seq_lens, dtype=torch.int32)
    num_tokens = spec.compute_num_tokens()
    max_blocks = (max(spec.seq_lens) + block_size - 1) // block_size
    block_table = torch.randint(0, max_block_idx, (spec.batch_size, max_blocks), dtype=torch.int32, device=device)
    slot_mapping = torch.randint(0, max_block_idx, (num_tokens,), dtype=torch.int64, device=device)
    ctx = [s - q for s, q in zip(spec.seq_lens, spec.query_lens)]
    return CommonAttentionMetadata(
        query_start_loc=qsl.to(device), query_start_loc_cpu=qsl, seq_lens=seq_lens_cpu.to(device),
        _seq_lens_cpu=seq_lens_cpu, _num_computed_tokens_cpu=torch.tensor(ctx, dtype=torch.int32),
        num_reqs=spec.batch_size, num_actual_tokens=num_tokens, max_query_len=max(spec.query_lens),
        max_seq_len=max(spec.seq_lens), block_table_tensor=block_table, slot_mapping=slot_mapping, causal=True)


class MockAttentionLayer:  # tests/v1/attention/test_attention_backends.py:176-186
    def __init__(self, device):
        self._q_scale = torch.tensor(1.0, device=device)
        self._k_scale = torch.tensor(1.0, device=device)
        self._v_scale = torch.tensor(1.0, device=device)
        self._k_scale_float = 1.0
        self._v_scale_float = 1.0
