"""Developer benchmark of decode context parallel (vllm_metax_b200/attention/dcp.py): a long-context decode step whose
TOKENS are sharded over the ranks (token t on rank t % N), strong scaling of one fixed problem.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/dcp_bench.py [--seqs 8 --ctx 131072 --heads 32 --kv-heads 8]

Every rank holds ctx / N tokens of every sequence for all KV heads (flash NHD layout, bf16) and H / N of the query heads.
Timed with CUDA events on the current stream, max over ranks, median of --iters steps:
  step    = forward_decode_dcp (query all-gather, attention with LSE, pack, all-to-all, merge fold)
  fused   = DcpExchange.forward_decode (query all-gather, attention with LSE into peer-mapped memory, one
            exchange + fold launch over NVLink peer loads)
  gather_launch / fold_launch = back-to-back launches of b200_dcp_gather_query / b200_dcp_exchange_fold alone; at these
            sizes the interval is the HOST's issue time per launch (python + ctypes), an upper bound on the kernel
  attend  = the attention launch alone on the gathered query (what the step would cost with a free exchange)
GB/s = the KV bytes of the WHOLE problem (all ranks) / time, comparable with the 1-GPU number of the same problem."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_metax_b200.attention import dcp  # noqa: E402


def timed(fn, iters, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for s, e in evs:
        s.record()
        fn()
        e.record()
    torch.cuda.synchronize()
    ts = sorted(s.elapsed_time(e) for s, e in evs)
    t = torch.tensor([ts[len(ts) // 2]], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seqs", type=int, default=8)
    ap.add_argument("--ctx", type=int, default=131072)
    ap.add_argument("--heads", type=int, default=32)
    ap.add_argument("--kv-heads", type=int, default=8)
    ap.add_argument("--iters", type=int, default=50)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    saved = os.dup(1)
    os.dup2(2, 1)  # NCCL prints its banner on stdout
    dist.init_process_group("nccl", device_id=dev)
    S, H, Hkv, D, B = a.seqs, a.heads, a.kv_heads, 128, 16
    assert H % world == 0
    Hl = H // world
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    lens = dcp.dcp_local_seq_lens(torch.full((S,), a.ctx, dtype=torch.int32), world, rank).to(dev)
    max_local = int(lens.max())
    nblk = (max_local + B - 1) // B
    NB = S * nblk + 8
    kv = torch.empty(2, NB, B, Hkv, D, dtype=torch.bfloat16, device=dev)
    kv.normal_(generator=g)
    table = torch.randperm(NB, device=dev, generator=g)[: S * nblk].view(S, nblk).to(torch.int32)
    q = torch.randn(S, Hl, D, device=dev, generator=g).bfloat16()
    scale = D ** -0.5
    q_all = torch.randn(S, H, D, device=dev, generator=g).bfloat16()

    step = timed(lambda: dcp.forward_decode_dcp(q, kv[0], kv[1], table, lens, max_local, scale), a.iters)
    xch = dcp.DcpExchange(S, H, D, torch.bfloat16)
    fused = timed(lambda: xch.forward_decode(q, kv[0], kv[1], table, lens, max_local, scale), a.iters)
    def fold_only():
        xch._fold(S, Hl, D, dev)  # a fold of the current step over whatever the partial buffers hold (flags already there)

    # the same step captured once and replayed (the step number lives in device memory)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        xch.forward_decode(q, kv[0], kv[1], table, lens, max_local, scale)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    dist.barrier()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):  # the stream whose attention workspace the warm-up call sized
        g_out, g_lse = xch.forward_decode(q, kv[0], kv[1], table, lens, max_local, scale)
    ref_out, ref_lse = xch.forward_decode(q, kv[0], kv[1], table, lens, max_local, scale)
    graph.replay()
    torch.cuda.synchronize()
    graph_ok = bool(torch.equal(g_out, ref_out) and torch.equal(g_lse, ref_lse))
    dist.barrier()
    fused_graph = timed(graph.replay, a.iters)

    gather = timed(lambda: xch._gather(q), a.iters)
    fold = timed(fold_only, a.iters)
    attend = timed(lambda: dcp._cuda_attend(q_all, kv[0], kv[1], scale, table, lens, max_local), a.iters)
    kv_bytes = 2 * S * a.ctx * Hkv * D * 2
    if rank == 0:
        os.dup2(saved, 1)
        print(json.dumps({
            "bench": "dcp_decode", "n_gpus": world, "seqs": S, "ctx": a.ctx, "heads": H, "kv_heads": Hkv,
            "step_us": round(step * 1e3, 1), "fused_step_us": round(fused * 1e3, 1),
            "fused_graph_step_us": round(fused_graph * 1e3, 1), "graph_replay_bit_identical": graph_ok,
            "fused_step_gbs": round(kv_bytes / fused / 1e6), "attend_us": round(attend * 1e3, 1),
            "gather_launch_us": round(gather * 1e3, 1), "fold_launch_us": round(fold * 1e3, 1),
            "exchange_us": round((step - attend) * 1e3, 1),
            "step_gbs": round(kv_bytes / step / 1e6), "attend_gbs": round(kv_bytes / attend / 1e6),
            "kv_bytes_total": kv_bytes, "timing": "cuda events, median of %d, max over ranks" % a.iters}), flush=True)
    xch.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
