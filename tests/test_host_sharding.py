"""Host logic of the multi-GPU path: KV-head sharding with no collective inside the path.

CPU-only.  The world_size-2 test runs two gloo ranks; each rank takes its `shard_kv_heads` slice of the query
heads / KV heads / cache, evaluates the decode attention of ITS shard with the oracle (the checker stands in for the
kernel here: no GPU in this tier of the suite), and the gathered per-rank outputs concatenated along heads must
equal the unsharded oracle bit for bit — the property that makes the path shard with zero exchange.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import torch_oracle as orc
from vllm_metax_b200.attention.paged_attn import PagedAttention, shard_kv_heads


@pytest.mark.parametrize("H,Hkv,tp", [(32, 8, 1), (32, 8, 2), (32, 8, 4), (32, 8, 8), (64, 8, 8), (64, 8, 16), (8, 1, 4)])
def test_shard_kv_heads_partitions_heads(H, Hkv, tp):
    G = H // Hkv
    q_seen = []
    for r in range(tp):
        qs, ks = shard_kv_heads(H, Hkv, tp, r)
        q_heads = list(range(H))[qs]
        kv_heads = list(range(Hkv))[ks]
        assert q_heads and kv_heads
        # every query head of the shard belongs to a KV head of the shard, at the right local index
        for i, h in enumerate(q_heads):
            assert h // G in kv_heads
            g_local = len(q_heads) // len(kv_heads)
            assert kv_heads[i // g_local] == h // G
        q_seen += q_heads
    assert q_seen == list(range(H))  # disjoint, ordered, complete


def test_shard_kv_heads_rejects_bad_configs():
    with pytest.raises(ValueError):
        shard_kv_heads(32, 8, 3, 0)
    with pytest.raises(ValueError):
        shard_kv_heads(32, 8, 2, 2)
    with pytest.raises(ValueError):
        shard_kv_heads(30, 8, 2, 0)


def test_split_kv_cache_views():
    NB, B, Hkv, D = 5, 16, 2, 64
    for dtype, x in ((torch.bfloat16, 8), (torch.uint8, 16), (torch.float32, 4)):
        kv = torch.zeros(PagedAttention.get_kv_cache_shape(NB, B, Hkv, D), dtype=dtype)
        k, v = PagedAttention.split_kv_cache(kv, Hkv, D)
        assert k.shape == (NB, Hkv, D // x, B, x) and v.shape == (NB, Hkv, D, B)
        assert k.data_ptr() == kv[0].data_ptr() and v.data_ptr() == kv[1].data_ptr()


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank: int, world: int, port: int, kvd: str, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        S, H, Hkv, D, B, L, NB = 3, 16, 4, 64, 16, 100, 64
        d = orc.make_decode_inputs(S, H, Hkv, D, B, L, NB, kv_cache_dtype=kvd, seed=7, unique_blocks=True)
        ks = torch.tensor(d.get("k_scale", 1.0))
        vs = torch.tensor(d.get("v_scale", 1.0))
        qs, kvs = shard_kv_heads(H, Hkv, world, rank)
        out_local = orc.paged_attention(d["query"][:, qs].contiguous(), d["key_cache"][:, kvs].contiguous(),
                                        d["value_cache"][:, kvs].contiguous(), d["block_tables"], d["seq_lens"],
                                        d["scale"], None, kvd, ks, vs)
        # the harness (not the path) gathers the shards to check them
        parts = [torch.empty_like(out_local) for _ in range(world)]
        dist.all_gather(parts, out_local.contiguous())
        if rank == 0:
            full = orc.paged_attention(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"],
                                       d["scale"], None, kvd, ks, vs)
            ret["equal"] = bool(torch.equal(torch.cat(parts, dim=1), full))
            ret["shape"] = tuple(full.shape)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kvd", ["auto", "fp8"])
def test_world2_gloo_sharded_equals_unsharded(kvd):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_rank_main, args=(r, 2, port, kvd, ret)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0
        assert ret.get("equal") is True, dict(ret)
        assert ret["shape"] == (3, 16, 64)
