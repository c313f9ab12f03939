"""Decode context parallel, host logic on CPU (gloo, world_size 2 and 3): the query all-gather, the packed all-to-all of
(output, lse) head slices and the fold with the merge rule reproduce attention over the whole, unsharded sequences.
The attention / merge callables are the CPU oracle here (the checker stands in for the kernels in this tier of the
suite); tests/test_dcp_gpu.py runs the same function on the CUDA kernels over NCCL."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import torch_oracle as orc
from dcp_common import make_problem, shard_for_rank
from vllm_metax_b200.attention.dcp import dcp_local_seq_lens, dcp_token_owner, forward_decode_dcp


def test_local_seq_lens_match_token_ownership():
    for world in (1, 2, 3, 8):
        lens = torch.tensor([0, 1, 2, 7, 8, 9, 4096, 4097], dtype=torch.int32)
        counts = torch.zeros(world, len(lens), dtype=torch.int32)
        for i, L in enumerate(lens.tolist()):
            for t in range(L):
                r, local = dcp_token_owner(t, world)
                assert local == counts[r, i]  # tokens of a rank are dense, in order
                counts[r, i] += 1
        for r in range(world):
            assert torch.equal(dcp_local_seq_lens(lens, world, r), counts[r])


@pytest.mark.parametrize("world", [2, 3, 8])
def test_fold_of_token_shards_is_attention_over_the_whole_sequence(world):
    """The N-way combination rule (what b200_dcp_exchange_fold computes) on the oracle: per-rank attention over the
    token shards, folded at once, equals attention over the unsharded sequences and equals the chain of pairwise
    merge_attn_states calls; a rank without tokens of a sequence (lse -inf) contributes nothing."""
    d = make_problem(S=4, H=8, Hkv=2, D=64, L=90, dtype=torch.float32)
    outs, lses = [], []
    for r in range(world):
        lk, lv, table, lens = shard_for_rank(d, world, r)
        outs.append(orc.paged_attention_flash(d["query"], lk, lv, table, lens, d["scale"]))
        lses.append(orc.paged_attention_flash_lse(d["query"], lk, lv, table, lens, d["scale"]))
    assert torch.isneginf(lses[-1][:, 0]).all()  # sequence 0 has one token: only rank 0 holds it
    out, lse = orc.fold_attn_partials(outs, lses)
    full = orc.paged_attention_flash(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"], d["scale"])
    full_lse = orc.paged_attention_flash_lse(d["query"], d["key_cache"], d["value_cache"], d["block_tables"],
                                             d["seq_lens"], d["scale"])
    assert (out - full).abs().max().item() <= 1e-5 and (lse - full_lse).abs().max().item() <= 1e-5
    co, cl = outs[0], lses[0]
    for r in range(1, world):
        co, cl = orc.merge_attn_states(co, cl, outs[r], lses[r])
    assert (out - co).abs().max().item() <= 1e-5 and (lse - cl).abs().max().item() <= 1e-5


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _oracle_attend(query, key_cache, value_cache, scale, block_table, seq_lens, max_seq_len):
    out = orc.paged_attention_flash(query, key_cache, value_cache, block_table, seq_lens, scale)
    lse = orc.paged_attention_flash_lse(query, key_cache, value_cache, block_table, seq_lens, scale)
    return out, lse


def _oracle_merge(po, pl, so, sl):
    return orc.merge_attn_states(po, pl, so, sl)


def _rank_main(rank: int, world: int, port: int, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        H = 4 * world
        d = make_problem(H=H, Hkv=2, D=64, L=150, dtype=torch.float32)
        Hl = H // world
        lk, lv, table, local_lens = shard_for_rank(d, world, rank)
        assert torch.equal(local_lens, dcp_local_seq_lens(d["seq_lens"], world, rank))
        q_local = d["query"][:, rank * Hl:(rank + 1) * Hl].contiguous()
        out, lse = forward_decode_dcp(q_local, lk, lv, table, local_lens, int(local_lens.max()), d["scale"],
                                      attend=_oracle_attend, merge=_oracle_merge)
        full = orc.paged_attention_flash(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"],
                                         d["scale"])
        full_lse = orc.paged_attention_flash_lse(d["query"], d["key_cache"], d["value_cache"], d["block_tables"],
                                                 d["seq_lens"], d["scale"])
        err = (out - full[:, rank * Hl:(rank + 1) * Hl]).abs().max().item()
        lerr = (lse - full_lse[rank * Hl:(rank + 1) * Hl]).abs().max().item()
        ret[rank] = (err, lerr)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_dcp_decode_equals_unsharded_attention(world):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_rank_main, args=(r, world, port, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(180)
            assert p.exitcode == 0
        assert len(ret) == world
        for r in range(world):
            err, lerr = ret[r]
            assert err <= 1e-5 and lerr <= 1e-5, (r, err, lerr)


def test_reference_local_length_rule_over_reads_when_world_divides_length():
    """The reference computes the per-rank length as L // N + (rank <= (L - 1) % N)
    (vllm_metax/v1/attention/backends/flash_attn.py:425-429).  With token t stored on rank t % N, that is the ownership
    count except when N divides L, where it is ONE TOO MANY ON EVERY RANK: the extra position is the slot the NEXT token
    will be written to — stale cache today.  This test demonstrates the over-read on a concrete shard (the deliberate
    divergence of dcp_local_seq_lens from the reference formula) and that attention over the over-read shard differs
    from attention over the sequence."""
    for world in (2, 4, 8):
        for L in (world, 4 * world, 64 * world):
            lens = torch.tensor([L], dtype=torch.int32)
            for rank in range(world):
                owned = sum(1 for t in range(L) if dcp_token_owner(t, world)[0] == rank)
                ours = int(dcp_local_seq_lens(lens, world, rank))
                ref_rule = L // world + (1 if rank <= (L - 1) % world else 0)
                assert ours == owned == L // world
                assert ref_rule == owned + 1, "the reference rule reads one slot past the rank's tokens when N | L"
        # one off a multiple: the two rules agree
        lens = torch.tensor([4 * world + 1], dtype=torch.int32)
        for rank in range(world):
            L = int(lens[0])
            assert int(dcp_local_seq_lens(lens, world, rank)) == L // world + (1 if rank <= (L - 1) % world else 0)
    # effect on the result: a shard padded by one stale slot changes the attention output
    world, L, D = 2, 8, 16
    g = torch.Generator().manual_seed(0)
    q = torch.randn(1, D, generator=g)
    kv = torch.randn(L, D, generator=g)
    stale = torch.full((1, D), 3.0)  # what an earlier request left in the next slot of rank 0
    shard0 = kv[0::2]
    exact = torch.softmax(q @ shard0.T / D ** 0.5, -1) @ shard0
    over = torch.cat([shard0, stale])
    wrong = torch.softmax(q @ over.T / D ** 0.5, -1) @ over
    assert (exact - wrong).abs().max() > 1e-2
