"""Decode context parallel on the CUDA kernels.

`test_dcp_kernels_one_gpu`: the per-rank attention launches (paged_attention_flash with LSE over each rank's token
shard, all heads) and the merge_attn_states fold, with the exchange done in-process — runs on a 1-GPU box and covers what
DCP asks of the kernels (empty local shards -> out 0 / lse -inf, fold in fp32).
`test_dcp_nccl`: `forward_decode_dcp` itself, one process per GPU over NCCL (skipped below 2 GPUs)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import torch_oracle as orc
from dcp_common import make_problem, shard_for_rank

pytestmark = pytest.mark.gpu

TOL = 2e-3  # bf16 attention tolerance of the suite (tests/test_attention_gpu.py)


def _full(d):
    out = orc.paged_attention_flash(d["query"].float(), d["key_cache"].float(), d["value_cache"].float(),
                                    d["block_tables"], d["seq_lens"], d["scale"])
    lse = orc.paged_attention_flash_lse(d["query"].float(), d["key_cache"].float(), d["value_cache"].float(),
                                        d["block_tables"], d["seq_lens"], d["scale"])
    return out, lse


@pytest.mark.parametrize("world", [2, 4, 8])
def test_dcp_kernels_one_gpu(world):
    from vllm_metax_b200 import _custom_ops as ops
    from vllm_metax_b200.attention.dcp import _cuda_attend, _cuda_merge

    dev = "cuda:0"
    d = make_problem(S=5, H=8, Hkv=2, D=128, L=333)
    d["seq_lens"][2] = world - 1 if world > 2 else 1  # the last rank(s) hold nothing of this sequence
    q = d["query"].to(dev)
    out = lse = None
    for r in range(world):
        lk, lv, table, lens = shard_for_rank(d, world, r)
        o, l = _cuda_attend(q, lk.to(dev), lv.to(dev), d["scale"], table.to(dev), lens.to(dev), int(lens.max()))
        empty = (lens == 0).nonzero().flatten().tolist()
        for s in empty:
            assert torch.isneginf(l[:, s]).all() and (o[s] == 0).all()
        o = o.float()
        out, lse = (o, l) if out is None else _cuda_merge(out, lse, o, l)
    full, full_lse = _full(d)
    assert (out.cpu() - full).abs().max().item() <= TOL
    assert (lse.cpu() - full_lse).abs().max().item() <= TOL
    del ops


@pytest.mark.parametrize("world", [2, 4, 8])
def test_dcp_peer_kernels_with_emulated_peers(world):
    """b200_dcp_gather_query / b200_dcp_exchange_fold at world > 1 on ONE GPU: the `world` peers' buffers are distinct
    local allocations and the launches of the ranks run one after the other, with the arrival flags preset (5) above the
    epoch (3) so that no launch waits for a rank that has not run yet.  Checks the addressing of the pushes and pulls,
    the N-way fold against attention over the unsharded sequences, and where each rank's signal lands."""
    from vllm_metax_b200 import _native as native
    from vllm_metax_b200.attention.dcp import _cuda_attend

    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream(dev).cuda_stream
    H = 4 * world
    Hl = H // world
    d = make_problem(S=6, H=H, Hkv=2, D=128, L=500)
    S, D = d["query"].shape[0], 128
    epoch, preset = 3, 5

    # gather: every rank pushes its head slice into every rank's gathered query
    qbuf = [torch.zeros(S, H, D, dtype=torch.bfloat16, device=dev) for _ in range(world)]
    gflags = torch.full((world, 16), preset, dtype=torch.int32, device=dev)
    counter = torch.zeros(1, dtype=torch.int32, device=dev)
    for r in range(world):
        q_local = d["query"][:, r * Hl:(r + 1) * Hl].contiguous().to(dev)
        native.dcp_gather_query([b.data_ptr() for b in qbuf], [gflags[t].data_ptr() for t in range(world)], q_local,
                                counter.data_ptr(), r, epoch, stream)
    torch.cuda.synchronize()
    for b in qbuf:
        assert torch.equal(b.cpu(), d["query"])
    assert counter.item() == 0
    assert (gflags[:, :world] == epoch).all() and (gflags[:, world:] == preset).all()

    # partial attention of every rank over its token shard, all heads
    partial = []
    for r in range(world):
        lk, lv, table, lens = shard_for_rank(d, world, r)
        partial.append(_cuda_attend(qbuf[r], lk.to(dev), lv.to(dev), d["scale"], table.to(dev), lens.to(dev), int(lens.max())))
    full, full_lse = _full(d)
    # the fold kernel against the oracle's N-way fold of the SAME partials: fp32 arithmetic, one bf16 rounding
    ref_out, ref_lse = orc.fold_attn_partials([o.cpu() for o, _ in partial], [l.cpu() for _, l in partial])

    # exchange + fold of every rank
    fflags = torch.full((world, 16), preset, dtype=torch.int32, device=dev)
    for r in range(world):
        out = torch.empty(S, Hl, D, dtype=torch.bfloat16, device=dev)
        lse = torch.empty(Hl, S, dtype=torch.float32, device=dev)
        native.dcp_exchange_fold([o.data_ptr() for o, _ in partial], [l.data_ptr() for _, l in partial],
                                 [fflags[t].data_ptr() for t in range(world)], out, lse, r, S, Hl, D, epoch, stream)
        torch.cuda.synchronize()
        assert (out.float().cpu() - full[:, r * Hl:(r + 1) * Hl]).abs().max().item() <= 2 * TOL
        assert (lse.cpu() - full_lse[r * Hl:(r + 1) * Hl]).abs().max().item() <= TOL
        assert (out.float().cpu() - ref_out[:, r * Hl:(r + 1) * Hl].float()).abs().max().item() <= 2 ** -8  # <= 1 bf16 ulp below 1.0
        assert (lse.cpu() - ref_lse[r * Hl:(r + 1) * Hl]).abs().max().item() <= 1e-5
        assert (fflags[:, r] == epoch).all()


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank: int, world: int, port: int, fused: bool, ret):
    from vllm_metax_b200.attention.dcp import DcpExchange, dcp_local_seq_lens, forward_decode_dcp

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        H = 4 * world
        d = make_problem(S=6, H=H, Hkv=2, D=128, L=1000)
        Hl = H // world
        lk, lv, table, local_lens = shard_for_rank(d, world, rank)
        assert torch.equal(local_lens, dcp_local_seq_lens(d["seq_lens"], world, rank))
        lk, lv, table, lens_dev = lk.to(dev), lv.to(dev), table.to(dev), local_lens.to(dev)
        max_local = int(local_lens.max())
        xch = DcpExchange(8, H, 128, torch.bfloat16) if fused else None
        errs = []
        # several steps with different queries: the fused path reuses its one set of buffers and its flags grow
        for step in range(5 if fused else 1):
            q = d["query"] if step == 0 else torch.randn(d["query"].shape, generator=torch.Generator().manual_seed(step)).bfloat16()
            d_step = dict(d, query=q)
            q_local = q[:, rank * Hl:(rank + 1) * Hl].contiguous().to(dev)
            if fused:
                out, lse = xch.forward_decode(q_local, lk, lv, table, lens_dev, max_local, d["scale"])
            else:
                out, lse = forward_decode_dcp(q_local, lk, lv, table, lens_dev, max_local, d["scale"])
            torch.cuda.synchronize()
            full, full_lse = _full(d_step)
            err = (out.float().cpu() - full[:, rank * Hl:(rank + 1) * Hl]).abs().max().item()
            lerr = (lse.cpu() - full_lse[rank * Hl:(rank + 1) * Hl]).abs().max().item()
            errs.append((err, lerr))
        if fused:
            xch.close()
        ret[rank] = (max(e for e, _ in errs), max(l for _, l in errs))
    finally:
        dist.destroy_process_group()


def _soak_main(rank: int, world: int, port: int, steps: int, ret):
    """The fused peer-memory step under skew: every rank delays a random amount (up to ~200 us of torch.cuda._sleep)
    before each step, alternately before the gather and before the fold; queries cycle through 4 seeds and every
    step's output is compared bit for bit with the first time that query was run."""
    import random as _random

    from vllm_metax_b200.attention.dcp import DcpExchange

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        H = 4 * world
        d = make_problem(S=6, H=H, Hkv=2, D=128, L=1000)
        Hl = H // world
        lk, lv, table, local_lens = shard_for_rank(d, world, rank)
        lk, lv, table, lens_dev = lk.to(dev), lv.to(dev), table.to(dev), local_lens.to(dev)
        max_local = int(local_lens.max())
        xch = DcpExchange(8, H, 128, torch.bfloat16)
        rnd = _random.Random(1000 + rank)
        queries = [torch.randn(d["query"].shape, generator=torch.Generator().manual_seed(k)).bfloat16()[:, rank * Hl:(rank + 1) * Hl]
                   .contiguous().to(dev) for k in range(4)]
        first = {}
        mismatches = 0
        for step in range(steps):
            k = step % 4
            torch.cuda._sleep(rnd.randint(0, 400000))
            out, lse = xch.forward_decode(queries[k], lk, lv, table, lens_dev, max_local, d["scale"])
            if step % 50 == 49 or step < 8:
                torch.cuda.synchronize()
            if k not in first:
                first[k] = (out.clone(), lse.clone())
            elif not (torch.equal(out, first[k][0]) and torch.equal(lse, first[k][1])):
                mismatches += 1
        torch.cuda.synchronize()
        xch.close()
        ret[rank] = mismatches
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_dcp_soak_skewed_ranks(world):
    """B200_DCP_SOAK_STEPS steps (default 200; the round-2 evidence run used 2000)."""
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    steps = int(os.environ.get("B200_DCP_SOAK_STEPS", "200"))
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_soak_main, args=(r, world, port, steps, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(600)
            assert p.exitcode == 0
        assert all(ret[r] == 0 for r in range(world)), dict(ret)


@pytest.mark.parametrize("fused", [False, True], ids=["nccl_all_to_all", "peer_memory_fold"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_dcp_nccl(world, fused):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_rank_main, args=(r, world, port, fused, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(300)
            assert p.exitcode == 0
        for r in range(world):
            err, lerr = ret[r]
            # bf16 output: one rounding of the merged value on top of the kernel tolerance
            assert err <= 2 * TOL and lerr <= TOL, (r, err, lerr)


def test_dcp_exchange_world1():
    """The fused exchange on one GPU (group of one): flags, buffer reuse and the fold with a single partial."""
    from vllm_metax_b200.attention.dcp import DcpExchange

    port = _free_port()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, rank=0, world_size=1, device_id=dev)
    try:
        d = make_problem(S=5, H=8, Hkv=2, D=128, L=333)
        xch = DcpExchange(8, 8, 128, torch.bfloat16)
        kc, vc = d["key_cache"].to(dev), d["value_cache"].to(dev)
        table, lens = d["block_tables"].to(dev), d["seq_lens"].to(dev)
        full, full_lse = _full(d)
        for _ in range(3):
            out, lse = xch.forward_decode(d["query"].to(dev), kc, vc, table, lens, int(d["seq_lens"].max()), d["scale"])
            assert (out.float().cpu() - full).abs().max().item() <= TOL
            assert (lse.cpu() - full_lse).abs().max().item() <= TOL
        xch.close()
    finally:
        dist.destroy_process_group()


def test_dcp_host_epoch_launches_refuse_cuda_graph_capture():
    """A step number passed as a host-side kernel argument would be frozen by a graph: a replay would wait for a step
    that has long passed.  Both C entry points return B200_ERR_UNSUPPORTED under capture in that mode; with
    the step number in device memory (step_counter, what DcpExchange uses) they are capturable — next tests."""
    from vllm_metax_b200 import _native as native

    dev = torch.device("cuda", 0)
    S, Hl, D, world = 2, 4, 128, 1
    q = torch.zeros(S, Hl, D, dtype=torch.bfloat16, device=dev)
    qbuf = torch.zeros(S, Hl, D, dtype=torch.bfloat16, device=dev)
    flags = torch.zeros(1, 16, dtype=torch.int32, device=dev)
    counter = torch.zeros(1, dtype=torch.int32, device=dev)
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        torch.cuda.synchronize()
        g.capture_begin()
        try:
            with pytest.raises(RuntimeError, match="CUDA graph"):
                native.dcp_gather_query([qbuf.data_ptr()], [flags[0].data_ptr()], q, counter.data_ptr(), 0, 1, side.cuda_stream)
            out = torch.empty(S, Hl, D, dtype=torch.bfloat16, device=dev)
            lse = torch.empty(Hl, S, dtype=torch.float32, device=dev)
            with pytest.raises(RuntimeError, match="CUDA graph"):
                native.dcp_exchange_fold([qbuf.data_ptr()], [lse.data_ptr()], [flags[0].data_ptr()], out, lse, 0, S, Hl, D, 1,
                                         side.cuda_stream)
        finally:
            g.capture_end()
    torch.cuda.synchronize()


@pytest.mark.parametrize("world", [2, 4])
def test_dcp_device_step_counter_with_emulated_peers(world):
    """The device-side step number on ONE GPU: `world` emulated ranks (distinct local buffers, one stream each, all
    resident at once: the launches are tiny) run gather -> fold for several steps with NO host-side epoch.  Every
    rank's step word must advance by one per step, the flags must follow it, and the folded result must be the N-way
    merge of the partials the ranks hold."""
    from vllm_metax_b200 import _native as native

    dev = torch.device("cuda", 0)
    S, Hl, D = 3, 2, 128
    H = Hl * world
    g = torch.Generator(device=dev).manual_seed(5)
    q_all_buf = [torch.zeros(S, H, D, dtype=torch.bfloat16, device=dev) for _ in range(world)]
    part_out = [torch.randn(S, H, D, device=dev, generator=g).bfloat16() for _ in range(world)]
    part_lse = [torch.randn(H, S, device=dev, generator=g) for _ in range(world)]
    gflags = torch.zeros(world, 16, dtype=torch.int32, device=dev)
    fflags = torch.zeros(world, 16, dtype=torch.int32, device=dev)
    counters = torch.zeros(world, dtype=torch.int32, device=dev)
    steps = torch.zeros(world, dtype=torch.int32, device=dev)
    streams = [torch.cuda.Stream() for _ in range(world)]
    torch.cuda.synchronize()
    gp, fp, outs, lses = [], [], [], []
    for r in range(world):
        gp.append(native.dcp_gather_params([b.data_ptr() for b in q_all_buf], [gflags[t].data_ptr() for t in range(world)],
                                           counters[r:].data_ptr(), r, S, Hl, D, torch.bfloat16, steps[r:].data_ptr()))
        fp.append(native.dcp_fold_params([b.data_ptr() for b in part_out], [b.data_ptr() for b in part_lse],
                                         [fflags[t].data_ptr() for t in range(world)], r, S, Hl, D, torch.bfloat16,
                                         steps[r:].data_ptr()))
        outs.append(torch.empty(S, Hl, D, dtype=torch.bfloat16, device=dev))
        lses.append(torch.empty(Hl, S, dtype=torch.float32, device=dev))
    # reference fold (fp32, merge_attn_states rule over all ranks at once)
    lse_all = torch.stack(part_lse)                      # [world, H, S]
    m = lse_all.max(dim=0).values
    w = torch.exp(lse_all - m)                           # [world, H, S]
    o_all = torch.stack([p.float() for p in part_out])   # [world, S, H, D]
    want = (o_all * w.permute(0, 2, 1).unsqueeze(-1)).sum(0) / w.sum(0).t().unsqueeze(-1)
    want_lse = m + torch.log(w.sum(0))
    for step in range(1, 5):
        qs = [torch.randn(S, Hl, D, device=dev, generator=g).bfloat16() for _ in range(world)]
        torch.cuda.synchronize()
        for r in range(world):   # every emulated rank enqueues its whole step; the kernels wait for each other on the GPU
            with torch.cuda.stream(streams[r]):
                native.dcp_gather_launch(gp[r], qs[r], 0, streams[r].cuda_stream)
                native.dcp_fold_launch(fp[r], outs[r], lses[r], 0, streams[r].cuda_stream)
        torch.cuda.synchronize()
        assert (steps == step).all(), steps
        assert (gflags[:, :world] == step).all() and (fflags[:, :world] == step).all()
        assert (counters == 0).all()
        for r in range(world):
            for t in range(world):
                assert torch.equal(q_all_buf[t][:, r * Hl:(r + 1) * Hl], qs[r])
            torch.testing.assert_close(outs[r].float(), want[:, r * Hl:(r + 1) * Hl], atol=2e-2, rtol=2e-2)
            torch.testing.assert_close(lses[r], want_lse[r * Hl:(r + 1) * Hl], atol=1e-5, rtol=1e-5)


def test_dcp_exchange_graph_capture_world1():
    """DcpExchange.forward_decode captured once and replayed on new queries (group of one, one GPU): the step word
    advances on the device with every replay."""
    from vllm_metax_b200 import _custom_ops as ops
    from vllm_metax_b200.attention.dcp import DcpExchange

    port = _free_port()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, rank=0, world_size=1, device_id=dev)
    try:
        d = make_problem(S=5, H=8, Hkv=2, D=128, L=333)
        xch = DcpExchange(8, 8, 128, torch.bfloat16)
        kc, vc = d["key_cache"].to(dev), d["value_cache"].to(dev)
        table, lens = d["block_tables"].to(dev), d["seq_lens"].to(dev)
        q = d["query"].to(dev).clone()
        L = int(d["seq_lens"].max())
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            ops.reserve_workspace(torch.empty(5, 8, 128, dtype=torch.bfloat16, device=dev), L)
            xch.forward_decode(q, kc, vc, table, lens, L, d["scale"])   # warm-up: step 1
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            out, lse = xch.forward_decode(q, kc, vc, table, lens, L, d["scale"])
        assert xch.step() == 1   # capture runs nothing
        for k in range(3):
            qk = torch.randn(d["query"].shape, generator=torch.Generator().manual_seed(40 + k)).bfloat16()
            q.copy_(qk)
            graph.replay()
            torch.cuda.synchronize()
            full, full_lse = _full(dict(d, query=qk))
            assert (out.float().cpu() - full).abs().max().item() <= TOL
            assert (lse.cpu() - full_lse).abs().max().item() <= TOL
            assert xch.step() == 2 + k
        xch.close()
    finally:
        dist.destroy_process_group()


def _graph_main(rank: int, world: int, port: int, ret):
    """Every rank captures one fused step and replays it on new queries, interleaved with eager steps."""
    from vllm_metax_b200 import _custom_ops as ops
    from vllm_metax_b200.attention.dcp import DcpExchange

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        H = 4 * world
        d = make_problem(S=6, H=H, Hkv=2, D=128, L=1000)
        Hl = H // world
        lk, lv, table, local_lens = shard_for_rank(d, world, rank)
        lk, lv, table, lens_dev = lk.to(dev), lv.to(dev), table.to(dev), local_lens.to(dev)
        max_local = int(local_lens.max())
        xch = DcpExchange(8, H, 128, torch.bfloat16)
        q_local = torch.zeros(6, Hl, 128, dtype=torch.bfloat16, device=dev)
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            ops.reserve_workspace(torch.empty(6, H, 128, dtype=torch.bfloat16, device=dev), max_local)
            xch.forward_decode(q_local, lk, lv, table, lens_dev, max_local, d["scale"])
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            out, lse = xch.forward_decode(q_local, lk, lv, table, lens_dev, max_local, d["scale"])
        errs = []
        for step in range(6):
            q = torch.randn(d["query"].shape, generator=torch.Generator().manual_seed(70 + step)).bfloat16()
            q_local.copy_(q[:, rank * Hl:(rank + 1) * Hl])
            if step % 3 == 2:   # an eager step between replays: both advance the same device-side step word
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    o, l = xch.forward_decode(q_local, lk, lv, table, lens_dev, max_local, d["scale"])
            else:
                torch.cuda._sleep(100000 * (rank + 1))   # skew the ranks
                graph.replay()
                o, l = out, lse
            torch.cuda.synchronize()
            full, full_lse = _full(dict(d, query=q))
            errs.append(((o.float().cpu() - full[:, rank * Hl:(rank + 1) * Hl]).abs().max().item(),
                         (l.cpu() - full_lse[rank * Hl:(rank + 1) * Hl]).abs().max().item()))
        steps_done = xch.step()
        xch.close()
        ret[rank] = (max(e for e, _ in errs), max(l for _, l in errs), steps_done)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_dcp_exchange_graph_replay(world):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_graph_main, args=(r, world, port, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(300)
            assert p.exitcode == 0
        for r in range(world):
            err, lerr, steps_done = ret[r]
            assert err <= 2 * TOL and lerr <= TOL, (r, err, lerr)
            assert steps_done == 7, steps_done   # warm-up + 6
