#!/usr/bin/env python
"""Benchmark of the paged-attention decode hot path (BASELINE.json metric) — contract in the task prompt.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|...] [--no-extras]

One "step" = one decode layer-call of the hot path over one batch:
    reshape_and_cache(new K/V of the batch -> paged cache)  +  paged_attention_v1(query, cache)
Headline workload (config.workload): BASELINE.json configs[1] — Llama-3-8B decode attention, GQA 32q/8kv, head 128,
block 16, batch 64, context 4096, bf16 KV, synthetic random cache (seed 0), unique random block tables.
N > 1 (torchrun, one rank per GPU): every rank runs the same per-GPU workload on its own sequences — the
path has no collective — so scaling is weak and value is the sum over ranks.

value      device-resident inputs, CUDA-event timed, max over ranks      (GB/s of algorithmic bytes)
e2e        the same steps through the public op wrappers with HOST inputs: pinned H2D of the step's
           q/k/v/slot_mapping/block_tables/seq_lens and D2H of the attention output inside the timed region
roofline   the attention launch alone (CUDA events on the launching stream) against MEASURED_PEAKS.json
parity_sample   after the timed regions: sampled sequences of the timed output compared with the CPU oracle on the same
           cache bytes (max-abs, tolerance 2e-3) — the oracle is the checker here, never the thing measured
extra.configs   (N = 1) the other named configs of BASELINE.json, attention launch alone, each with its own roofline
           fraction and parity sample: c1 (cold L2), c3 (fp8), c4 (Llama-3-70B full layer), c5 bf16 / fp8 through
           paged_attention_v2 (P = 256 partitions)
tp         (N > 1) BASELINE.json configs[3]: the Llama-3-70B layer (batch 128, ctx 8192) sharded by KV head over the N
           ranks (shard_kv_heads), all ranks timed together, max over ranks; strong-scaling efficiency against the
           full layer timed on one GPU in the same run
cpu_baseline / --impl reference: the CPU restatement (oracle/paged_attn_ref.c, OpenMP, all host threads) on the same
           workload.  The oracle is used here only as the timed CPU baseline.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

WORKLOADS = {
    # name: (S, H, Hkv, D, B, L, kv_cache_dtype, op, description)
    "c1": (8, 32, 32, 128, 16, 1024, "auto", "v1",
           "paged_attention_v1 kernel-test shape: 8 seqs, 32 heads (MHA), head 128, block 16, ctx 1024, bf16"),
    "c2": (64, 32, 8, 128, 16, 4096, "auto", "v1",
           "Llama-3-8B decode attention, GQA 32q/8kv, head 128, block 16, batch 64, ctx 4096, bf16 KV"),
    "c3": (64, 32, 8, 128, 16, 4096, "fp8", "v1",
           "Llama-3-8B decode attention, GQA 32q/8kv, head 128, block 16, batch 64, ctx 4096, FP8-e4m3 KV"),
    # the serving path of the reference: flash cache layout [NB, B, Hkv, D] (NHD), reshape_and_cache_flash +
    # flash-layout decode (flash_attn_with_kvcache in the reference).  Same shapes and bytes as c2.
    "c2_flash": (64, 32, 8, 128, 16, 4096, "auto", "flash",
                 "Llama-3-8B decode attention, GQA 32q/8kv, head 128, block 16, batch 64, ctx 4096, bf16 KV, flash NHD cache layout"),
    "c4": (128, 64, 8, 128, 16, 8192, "auto", "v1",
           "Llama-3-70B decode attention, GQA 64q/8kv, head 128, block 16, batch 128, ctx 8192, bf16 KV (whole layer; KV-head shards under --gpus N)"),
    "c4_fp8": (128, 64, 8, 128, 16, 8192, "fp8", "v1",
               "Llama-3-70B decode attention, GQA 64q/8kv, batch 128, ctx 8192, FP8-e4m3 KV"),
    "c5": (8, 32, 8, 128, 16, 131072, "auto", "v2",
           "long-context paged_attention_v2 split-K, batch 8, ctx 131072, partition 512 (P=256), GQA 32q/8kv, bf16 KV"),
    "c5_fp8": (8, 32, 8, 128, 16, 131072, "fp8", "v2",
               "long-context paged_attention_v2 split-K, batch 8, ctx 131072, partition 512 (P=256), GQA 32q/8kv, FP8-e4m3 KV"),
}
N_LAYERS = {"c4": 80, "c4_fp8": 80}  # Llama-3-70B; everything else Llama-3-8B (32); tokens/s = S / (n_layers * t_step)
METRIC = "paged-attn decode HBM GB/s"
UNIT = "GB/s"
PARITY_TOL = 2e-3  # BASELINE.json north_star: max-abs, bf16 output, fp32 accumulate
L2_BYTES = 126 * 1024 * 1024


def algorithmic_bytes(S, H, Hkv, D, B, L, kv_eb, with_cache_write=True, v2=False):
    """every KV byte once + q/out + block-table entries + seq_lens (+ the v2 scratch written by the
    main kernel and read by the reduce) (+ the decode cache write)."""
    attn = S * L * Hkv * D * 2 * kv_eb + S * H * D * 4 + S * math.ceil(L / B) * 4 + 4 * S
    if v2:
        P = math.ceil(L / 512)
        attn += S * H * P * (D * 2 * 2 + 16)
    write = S * Hkv * D * 2 * (2 + kv_eb) + S * 8 if with_cache_write else 0
    return attn, write


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md 6.65 TB/s)"


def measured_traffic(name):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the attention kernel from the committed
    `ncu --set full` captures: profiles/traffic.json (written by tools/ncu_summary.py --traffic from the .ncu-rep;
    it names the capture and the commit).  Read at run time, never a literal here; null when no capture exists."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        e = t.get(name)
        return (e["dram_bytes"], e["source"]) if e else (None, None)
    except Exception:
        return None, None


# ---- clocks under load -----------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost"}

    def __init__(self, index: int):
        self.samples, self.mask, self.stop_flag, self.h, self.max_mhz = [], 0, threading.Event(), None, None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.h = None
        self.thread = None

    def _run(self):
        while not self.stop_flag.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                self.mask |= int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.h is not None:
            self.stop_flag.clear()
            self.thread = threading.Thread(target=self._run, daemon=True)
            self.thread.start()

    def stop(self):
        if self.thread is not None:
            self.stop_flag.set()
            self.thread.join()
            self.thread = None

    def summary(self, note):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0, "note": "NVML unavailable"}
        s = sorted(self.samples)
        reasons = [n for bit, n in self.REASONS.items() if self.mask & bit]
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(s), "note": note}


def physical_gpu_index(local: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local])
        except Exception:
            return local
    return local


# ---- synthetic workload ------------------------------------------------------------------------------
def make_workload(name, dev, seed=0, kv_heads=None):
    """Random cache / query / block tables of a named config on the device.  kv_heads = (slice of query heads, slice of
    kv heads) builds the shard a tensor-parallel rank owns: its own cache [NB, Hkv/TP, ...] and query heads, the SAME
    block tables and sequence lengths on every rank (generated before anything head-dependent from the same seed)."""
    S, H, Hkv, D, B, L, kvd, op, _ = WORKLOADS[name]
    nblk = (L + B - 1) // B
    NB = int(S * nblk * 1.25)
    g = torch.Generator(device=dev).manual_seed(seed)
    scale = D ** -0.5
    fp8 = kvd != "auto"
    block_tables = torch.randperm(NB, device=dev, generator=g)[: S * nblk].view(S, nblk).to(torch.int32)
    if kv_heads is not None:
        qs, ks = kv_heads
        H, Hkv = qs.stop - qs.start, ks.stop - ks.start
        g = torch.Generator(device=dev).manual_seed(seed * 1000 + 17 + ks.start)

    def uni(shape, dtype):
        # generated in slabs of the leading dimension: the fp32 staging of a 5 GB cache would double its footprint
        out = torch.empty(shape, dtype=dtype, device=dev)
        step = max(1, (1 << 28) // max(1, out[0].numel()))
        for i in range(0, shape[0], step):
            n = min(step, shape[0] - i)
            r = (torch.rand((n,) + tuple(shape[1:]), device=dev, generator=g) * 2 - 1) * scale
            out[i:i + n] = r.to(torch.float8_e4m3fn).view(torch.uint8) if dtype == torch.uint8 else r.to(dtype)
        return out

    x = 16 if fp8 else 8
    flash = op == "flash"
    cdt = torch.uint8 if fp8 else torch.bfloat16
    if flash:
        kc, vc = uni((NB, B, Hkv, D), cdt), uni((NB, B, Hkv, D), cdt)
    else:
        kc, vc = uni((NB, Hkv, D // x, B, x), cdt), uni((NB, Hkv, D, B), cdt)
    w = dict(
        name=name, op=op, flash=flash, S=S, H=H, Hkv=Hkv, D=D, B=B, L=L, kvd=kvd, scale=scale, kv_eb=1 if fp8 else 2,
        key_cache=kc, value_cache=vc,
        query=uni((S, H, D), torch.bfloat16), key=uni((S, Hkv, D), torch.bfloat16), value=uni((S, Hkv, D), torch.bfloat16),
        block_tables=block_tables,
        seq_lens=torch.full((S,), L, dtype=torch.int32, device=dev),
        k_scale=torch.ones(1, dtype=torch.float32, device=dev), v_scale=torch.ones(1, dtype=torch.float32, device=dev),
    )
    # the decode step writes the K/V of position L-1 of every sequence
    w["slot_mapping"] = block_tables[:, (L - 1) // B].to(torch.int64) * B + (L - 1) % B
    w["out"] = torch.empty_like(w["query"])
    if op == "v2":
        P = (L + 511) // 512
        w["exp_sums"] = torch.empty(S, H, P, dtype=torch.float32, device=dev)
        w["max_logits"] = torch.empty_like(w["exp_sums"])
        w["tmp_out"] = torch.empty(S, H, P, D, dtype=torch.bfloat16, device=dev)
    return w


def make_ops(ops, w):
    """(cache_write, attention) closures of a workload: the public op wrappers with the reference's argument lists."""
    Hkv, B, L = w["Hkv"], w["B"], w["L"]

    def cache_write(key, value, slots):
        if w["flash"]:
            ops.reshape_and_cache_flash(key, value, w["key_cache"], w["value_cache"], slots, w["kvd"], w["k_scale"], w["v_scale"])
        else:
            ops.reshape_and_cache(key, value, w["key_cache"], w["value_cache"], slots, w["kvd"], w["k_scale"], w["v_scale"])

    def attention(q, bt, lens, out):
        if w["op"] == "flash":
            ops.paged_attention_flash(out, q, w["key_cache"], w["value_cache"], w["scale"], bt, lens, L)
        elif w["op"] == "v2":
            ops.paged_attention_v2(out, w["exp_sums"], w["max_logits"], w["tmp_out"], q, w["key_cache"], w["value_cache"],
                                   Hkv, w["scale"], bt, lens, B, L, None, w["kvd"], w["k_scale"], w["v_scale"])
        else:
            ops.paged_attention_v1(out, q, w["key_cache"], w["value_cache"], Hkv, w["scale"], bt, lens, B, L, None,
                                   w["kvd"], w["k_scale"], w["v_scale"])

    return cache_write, attention


def parity_sample(w, n_seqs):
    """Compare sampled sequences of the timed output w["out"] with the CPU oracle on the same cache bytes: the blocks
    of each sampled sequence are gathered on the device, copied to the host as a compact cache, and
    oracle.torch_oracle.paged_attention (paged_attention_flash for the flash layout) is evaluated on them.
    The oracle is test infrastructure; here it checks the product's output, it is never timed or shipped."""
    from oracle import torch_oracle as orc

    S = w["S"]
    picks = sorted({int(round(i * (S - 1) / max(1, n_seqs - 1))) for i in range(n_seqs)}) if n_seqs > 1 else [S - 1]
    worst = 0.0
    nblk = w["block_tables"].shape[1]
    for s in picks:
        ids = w["block_tables"][s].long()
        kc, vc = w["key_cache"][ids].cpu(), w["value_cache"][ids].cpu()
        bt = torch.arange(nblk, dtype=torch.int32).view(1, nblk)
        lens = w["seq_lens"][s:s + 1].cpu()
        q = w["query"][s:s + 1].cpu()
        if w["flash"]:
            ref = orc.paged_attention_flash(q, kc, vc, bt, lens, w["scale"])
        else:
            ref = orc.paged_attention(q, kc, vc, bt, lens, w["scale"], None, w["kvd"], 1.0, 1.0)
        got = w["out"][s:s + 1].cpu()
        worst = max(worst, (got.float() - ref.float()).abs().max().item())
    return {"n_seqs": len(picks), "seqs": picks, "heads": w["H"], "max_abs": worst, "tol": PARITY_TOL,
            "ok": bool(worst <= PARITY_TOL), "oracle": "oracle/torch_oracle.py on the sampled sequences' cache blocks (CPU)"}


def time_attention(attention, w, steps, warmup, flush=None):
    """Mean CUDA-event time (ms) of `steps` attention launches after `warmup`; with `flush` (a buffer larger than L2)
    the buffer is rewritten before every timed launch, so each launch starts with a cold L2."""
    for _ in range(max(warmup, 3)):
        attention(w["query"], w["block_tables"], w["seq_lens"], w["out"])
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * steps)]
    for i in range(steps):
        if flush is not None:
            flush.add_(1)
        ev[2 * i].record()
        attention(w["query"], w["block_tables"], w["seq_lens"], w["out"])
        ev[2 * i + 1].record()
    torch.cuda.synchronize()
    return sum(ev[2 * i].elapsed_time(ev[2 * i + 1]) for i in range(steps)) / steps


def extra_config(ops, name, dev, steps, warmup, pk, n_parity, kv_heads=None, seed=0):
    """One named config, attention launch alone: time, GB/s of algorithmic bytes, fraction of the measured HBM peak and
    a parity sample.  Returns (entry, ms)."""
    w = make_workload(name, dev, seed=seed, kv_heads=kv_heads)
    _, attention = make_ops(ops, w)
    attn_bytes, _ = algorithmic_bytes(w["S"], w["H"], w["Hkv"], w["D"], w["B"], w["L"], w["kv_eb"], False, w["op"] == "v2")
    flush = None
    if attn_bytes < 4 * L2_BYTES:  # working set not ≫ L2: flush explicitly, report cold
        flush = torch.zeros(2 * L2_BYTES // 4, dtype=torch.int32, device=dev)
    ms = time_attention(attention, w, steps, warmup, flush)
    entry = {"name": name, "op": "paged_attention_" + w["op"], "attention_us": ms * 1e3,
             "gbs": attn_bytes / (ms * 1e-3) / 1e9, "frac": attn_bytes / (ms * 1e-3) / 1e9 / pk["hbm_gbs"],
             "algorithmic_bytes": attn_bytes, "kv_heads": w["Hkv"], "heads": w["H"],
             "l2": "flushed before every timed launch (cold)" if flush is not None else "working set larger than L2"}
    if n_parity > 0:
        ps = parity_sample(w, n_parity)
        entry["max_abs"], entry["parity_seqs"], entry["parity_ok"] = ps["max_abs"], ps["n_seqs"], ps["ok"]
    del w
    torch.cuda.empty_cache()
    return entry, ms


MLA_SHAPES = [  # (name, batch, ctx, heads): DeepSeek-style MLA decode, latent 512 + rope 64, page 64, bf16
    ("mla_h128", 128, 8192, 128), ("mla_h16_tp8", 128, 8192, 16),
]


def mla_config(ops, name, b, L, H, dev, steps, warmup, pk, n_parity):
    """MLA decode attention: the reference op cutlass_mla_decode / sm100_cutlass_mla_decode on the
    test's shapes (tests/kernels/attention/test_cutlass_mla_decode.py:41-52).  bound: the op sits on the tensor/HBM
    ridge at 128 heads (242 flop/B), so both fractions are reported: HBM (bytes = sum L*1152 + q + out) and tensor
    (flops = 2*L*H*(576+512) per sequence) against MEASURED_PEAKS.json."""
    from oracle import torch_oracle as orc
    page = 64
    g = torch.Generator(device=dev).manual_seed(b + L + H)
    pages = L // page
    cache = torch.randn(b * pages, page, 576, generator=g, device=dev, dtype=torch.bfloat16)
    q = torch.randn(b, H, 576, generator=g, device=dev, dtype=torch.bfloat16)
    table = torch.randperm(b * pages, generator=g, device=dev).to(torch.int32).view(b, pages)
    lens = torch.full((b,), L, dtype=torch.int32, device=dev)
    qn, qp = q[..., :512].contiguous(), q[..., 512:].contiguous()
    out = torch.empty(b, H, 512, dtype=q.dtype, device=dev)
    lse = torch.empty(b, H, dtype=torch.float32, device=dev)
    ws = torch.empty(max(16, ops.sm100_cutlass_mla_get_workspace_size(L, b, 148, -1)), dtype=torch.uint8, device=dev)
    scale = 1.0 / math.sqrt(576)

    def run():
        ops.sm100_cutlass_mla_decode(out, lse, qn, qp, cache, lens, table, ws, scale, -1)

    for _ in range(max(warmup, 3)):
        run()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * steps)]
    for i in range(steps):
        ev[2 * i].record()
        run()
        ev[2 * i + 1].record()
    torch.cuda.synchronize()
    ms = sum(ev[2 * i].elapsed_time(ev[2 * i + 1]) for i in range(steps)) / steps
    byt = b * L * 1152 + b * H * (576 + 512) * 2
    flops = 2.0 * b * L * H * (576 + 512)
    e = {"name": name, "op": "sm100_cutlass_mla_decode", "batch": b, "ctx": L, "heads": H, "page": page,
         "attention_us": ms * 1e3, "gbs": byt / ms / 1e6, "frac_hbm": byt / ms / 1e6 / pk["hbm_gbs"],
         "tflops": flops / ms / 1e9, "frac_tensor": flops / ms / 1e9 / pk.get("bf16_tflops", 1674.0),
         "algorithmic_bytes": byt, "l2": "working set larger than L2"}
    e["traffic"], e["traffic_source"] = measured_traffic(name)
    if n_parity > 0:
        picks = sorted(set(int(round(i * (b - 1) / max(n_parity - 1, 1))) for i in range(n_parity)))
        worst = 0.0
        for s_ in picks:
            ref, _ = orc.mla_decode(q[s_:s_ + 1].cpu(), cache[table[s_].long()].cpu(), scale,
                                    torch.arange(pages, dtype=torch.int32).view(1, -1), lens[s_:s_ + 1].cpu())
            worst = max(worst, (out[s_:s_ + 1].float().cpu() - ref).abs().max().item())
        e["max_abs"], e["parity_seqs"], e["parity_ok"] = worst, len(picks), bool(worst <= PARITY_TOL)
    del cache, q, table, out, ws
    torch.cuda.empty_cache()
    return e


def run_ours(args):
    from vllm_metax_b200 import _custom_ops as ops  # loads the native libraries; raises if they are missing
    from vllm_metax_b200 import _native
    from vllm_metax_b200.attention import shard_kv_heads

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    saved_stdout = None
    if world > 1:
        import torch.distributed as dist

        # NCCL prints its version banner to stdout when the first communicator comes up; stdout carries the ONE JSON
        # line, so file descriptor 1 points at stderr until the line is printed (restore_stdout below)
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    pk, pk_src = peaks()
    w = make_workload(args.workload, dev, seed=rank)
    S, H, Hkv, D, B, L = w["S"], w["H"], w["Hkv"], w["D"], w["B"], w["L"]
    cache_write, attention = make_ops(ops, w)

    def step():
        cache_write(w["key"], w["value"], w["slot_mapping"])
        attention(w["query"], w["block_tables"], w["seq_lens"], w["out"])

    # ---- device-resident timing ("value") -------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(physical_gpu_index(local))
    launches0 = _native.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * args.steps + 2)]
    sampler.start()
    barrier()
    ev[0].record()
    for i in range(args.steps):
        cache_write(w["key"], w["value"], w["slot_mapping"])
        ev[2 + 2 * i].record()
        attention(w["query"], w["block_tables"], w["seq_lens"], w["out"])
        ev[3 + 2 * i].record()
    ev[1].record()
    barrier()
    sampler.stop()
    launches = _native.launch_count() - launches0
    total_ms = ev[0].elapsed_time(ev[1])
    attn_ms = sum(ev[2 + 2 * i].elapsed_time(ev[3 + 2 * i]) for i in range(args.steps)) / args.steps
    clock_note = "sampled by NVML every 2 ms inside the timed region"
    if len(sampler.samples) < 8:  # timed region too short for a meaningful median: keep the same load running
        sampler.start()
        t_end = time.time() + 0.6
        while time.time() < t_end:
            for _ in range(50):
                step()
            torch.cuda.synchronize()
        sampler.stop()
        clock_note = "timed region shorter than 8 samples; sampling continued for 0.6 s of the same steps right after it"
    clocks = sampler.summary(clock_note)
    variant = {1: "generic_fma", 2: "stream_tma_mma", 3: "umma_fp8"}.get(_native.last_attention_variant(), "?")

    # ---- parity of the timed output (rank 0's own sequences) ------------------------------------------------
    parity = None
    if rank == 0 and not args.no_parity:
        step()
        torch.cuda.synchronize()
        parity = parity_sample(w, 4)

    # ---- end to end through the public op wrappers, host inputs ------------------------------------------
    from vllm_metax_b200.attention import PinnedDecodeInputs

    names = ("query", "key", "value", "slot_mapping", "block_tables", "seq_lens")
    staged = PinnedDecodeInputs({k: (tuple(w[k].shape), w[k].dtype) for k in names}, dev)  # one pinned buffer
    for k in names:
        staged.host(k).copy_(w[k].cpu())
    host_out = torch.empty(w["out"].shape, dtype=w["out"].dtype).pin_memory()
    h2d = staged.nbytes
    d2h = host_out.numel() * host_out.element_size()

    def e2e_step():
        staged.upload()  # host -> device: this step's q, k, v, slot_mapping, block_tables, seq_lens
        cache_write(staged.dev("key"), staged.dev("value"), staged.dev("slot_mapping"))
        attention(staged.dev("query"), staged.dev("block_tables"), staged.dev("seq_lens"), w["out"])
        host_out.copy_(w["out"], non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the caller reads the result

    for _ in range(max(args.warmup, 3)):
        e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    del staged

    # ---- reduce over ranks (max time) -----------------------------------------------------------------------
    t = torch.tensor([total_ms, e2e_ms, attn_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, attn_ms = t.tolist()
    attn_bytes, write_bytes = algorithmic_bytes(S, H, Hkv, D, B, L, w["kv_eb"], True, w["op"] == "v2")
    step_bytes = attn_bytes + write_bytes
    ms_per_step = total_ms / args.steps
    value = world * step_bytes / (ms_per_step * 1e-3) / 1e9
    e2e_value = world * step_bytes / (e2e_ms / args.steps * 1e-3) / 1e9
    achieved = attn_bytes / (attn_ms * 1e-3) / 1e9
    n_layers = N_LAYERS.get(args.workload, 32)
    kv_eb_headline, flash_headline, op_headline = w["kv_eb"], w["flash"], w["op"]
    del w, cache_write, attention
    torch.cuda.empty_cache()

    extra = {}
    if rank == 0:
        extra["tokens_per_s_attention_only"] = world * S / (n_layers * ms_per_step * 1e-3)
        extra["attention_ms"] = attn_ms
        extra["frac_of_nominal_8TBps"] = achieved / 8000.0
        extra["variant"] = variant

    # ---- the other named configs (N = 1) ----------------------------------------------------------------------
    if world == 1 and not args.no_extras:
        cfgs = []
        for name, npar in (("c1", 4), ("c3", 4), ("c4", 2), ("c4_fp8", 2), ("c5", 1), ("c5_fp8", 1)):
            if name == args.workload:
                continue
            try:
                e, _ = extra_config(ops, name, dev, min(args.steps, 20), args.warmup, pk, 0 if args.no_parity else npar)
                e["frac_of_nominal_8TBps"] = e["gbs"] / 8000.0
            except Exception as ex:  # an extra must never take the headline line down with it
                e = {"name": name, "error": "%s: %s" % (type(ex).__name__, ex)}
                torch.cuda.empty_cache()
            cfgs.append(e)
        extra["configs"] = cfgs
        mla = []
        for name, b, L, H in MLA_SHAPES:
            try:
                mla.append(mla_config(ops, name, b, L, H, dev, min(args.steps, 20), args.warmup, pk, 0 if args.no_parity else 2))
            except Exception as ex:
                mla.append({"name": name, "error": "%s: %s" % (type(ex).__name__, ex)})
                torch.cuda.empty_cache()
        extra["mla"] = mla

    # ---- tensor parallel: BASELINE.json configs[3], KV-head shards over the ranks (N > 1) --------------------------
    tp = None
    if world > 1 and not args.no_extras:
        tp = {"config": WORKLOADS["c4"][8], "tp": world, "partition": "shard_kv_heads(64, 8, tp, rank): contiguous KV heads and "
              "their query heads per rank, replicated block tables / seq_lens, no collective in the path"}
        for name in ("c4", "c4_fp8"):
            Sx, Hx, Hkvx = WORKLOADS[name][0], WORKLOADS[name][1], WORKLOADS[name][2]
            shard = shard_kv_heads(Hx, Hkvx, world, rank)
            barrier()
            e, ms = extra_config(ops, name, dev, min(args.steps, 20), args.warmup, pk, 0 if (args.no_parity or rank) else 1,
                                 kv_heads=shard, seed=0)
            barrier()
            ef, ms_full = extra_config(ops, name, dev, min(args.steps, 20), args.warmup, pk, 0, seed=0)  # whole layer, one GPU
            tt = torch.tensor([ms, ms_full], dtype=torch.float64, device=dev)
            gb = torch.tensor([e["gbs"]], dtype=torch.float64, device=dev)
            allgb = [torch.zeros_like(gb) for _ in range(world)]
            dist.all_gather(allgb, gb)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms, ms_full = tt.tolist()
            agg = world * e["algorithmic_bytes"] / (ms * 1e-3) / 1e9
            tp[name] = {
                "shard_attention_us": ms * 1e3, "per_rank_gbs": [float(x.item()) for x in allgb],
                "aggregate_gbs": agg, "frac_per_gpu": agg / world / pk["hbm_gbs"],
                "tokens_per_s_attention_only_80_layers": Sx / (80 * ms * 1e-3),
                "full_layer_one_gpu_us": ms_full * 1e3,
                "strong_scaling_efficiency": ms_full / (world * ms),
                "kv_heads_per_rank": e["kv_heads"], "heads_per_rank": e["heads"],
                "max_abs": e.get("max_abs"), "parity_ok": e.get("parity_ok"),
            }

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(args.workload, budget_s=12.0)

    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    traffic, traffic_src = measured_traffic(args.workload)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if kv_eb_headline == 2 else "bf16 (q/out) + fp8-e4m3 (KV)", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload][8], "name": args.workload,
                   "step": {"flash": "reshape_and_cache_flash + paged_attention_flash", "v2": "reshape_and_cache + paged_attention_v2",
                            "v1": "reshape_and_cache + paged_attention_v1"}[op_headline],
                   "l2": "inputs larger than L2 (KV working set %.2f GB vs 126 MB L2)" % (attn_bytes / 1e9),
                   "parallelism": "dp%d (independent sequences per GPU, no collective)%s" % (
                       world, "; the KV-head-sharded Llama-3-70B layer is in `tp`" if tp else ""),
                   "block_tables": "unique random permutation, 1.25x over-provisioned cache", "seed": 0},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                     "frac": achieved / pk["hbm_gbs"], "traffic": traffic, "traffic_source": traffic_src,
                     "kernel": "paged_attention_stream_kernel" + (" (flash layout, tiled TMA)" if flash_headline else " (+ split reduce)"),
                     "algorithmic_bytes": attn_bytes, "peak_source": pk_src},
        "parity_sample": parity,
        "cpu_baseline": cpu,
        "extra": extra,
    }
    if tp is not None:
        line["tp"] = tp
    if saved_stdout is not None:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)


# ---- CPU arm --------------------------------------------------------------------------------------------
def cpu_sample(workload, n_seqs, seed=0):
    """The workload for the CPU arm: its first n_seqs sequences (n_seqs = S: the whole batch).  The C restatement
    reads the v1/v2 cache layout; the flash workload is the same arithmetic over the same bytes, so it is timed on that."""
    S, H, Hkv, D, B, L, kvd, _, _ = WORKLOADS[workload]
    nblk = (L + B - 1) // B
    NB = n_seqs * nblk
    g = torch.Generator().manual_seed(seed)
    scale = D ** -0.5
    x = 16 if kvd != "auto" else 8

    def uni(shape, dtype):
        out = torch.empty(shape, dtype=dtype)
        step = max(1, (1 << 26) // max(1, out[0].numel()))
        for i in range(0, shape[0], step):
            n = min(step, shape[0] - i)
            r = (torch.rand((n,) + tuple(shape[1:]), generator=g) * 2 - 1) * scale
            out[i:i + n] = r.to(torch.float8_e4m3fn).view(torch.uint8) if dtype == torch.uint8 else r.to(dtype)
        return out

    cdt = torch.bfloat16 if kvd == "auto" else torch.uint8
    kc, vc = uni((NB, Hkv, D // x, B, x), cdt), uni((NB, Hkv, D, B), cdt)
    q = uni((n_seqs, H, D), torch.bfloat16)
    bt = torch.randperm(NB, generator=g).view(n_seqs, nblk).to(torch.int32)
    lens = torch.full((n_seqs,), L, dtype=torch.int32)
    return dict(q=q, kc=kc, vc=vc, bt=bt, lens=lens, scale=scale, kvd=kvd, n=n_seqs,
                bytes=algorithmic_bytes(n_seqs, H, Hkv, D, B, L, 1 if kvd != "auto" else 2, False)[0])


def cpu_step(c, threads):
    from oracle import c_oracle  # the timed CPU baseline: the one use of oracle/ outside tests and the parity samples

    return c_oracle.paged_attention(c["q"], c["kc"], c["vc"], c["bt"], c["lens"], c["scale"], None, c["kvd"], 1.0, 1.0,
                                    num_threads=threads)


def cpu_seqs(workload):
    """Sequences per CPU step: the whole batch when one step is ~0.1 s of CPU work (c1/c2/c3), else a stated subset."""
    S, _, Hkv, D, _, L, kvd, _, _ = WORKLOADS[workload]
    per_seq = L * Hkv * D * 2 * (1 if kvd != "auto" else 2)
    return max(1, min(S, int(1.2e9 // per_seq)))


def cpu_baseline(workload, budget_s):
    cores = os.cpu_count() or 1
    n = cpu_seqs(workload)
    c = cpu_sample(workload, n)
    cpu_step(c, cores)  # warm-up (builds / loads the C library, touches the pages)
    t0 = time.time()
    reps = 0
    while True:
        cpu_step(c, cores)
        reps += 1
        if time.time() - t0 > budget_s or reps >= 20:
            break
    dt = (time.time() - t0) / reps
    return {"value": c["bytes"] / dt / 1e9, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d of %d sequences of the workload (ctx %d), attention only, %d repetitions, OpenMP over (seq, kv head)"
                      % (n, WORKLOADS[workload][0], WORKLOADS[workload][5], reps)}


def run_reference(args):
    """CPU arm: the reference has no CPU implementation of this path and its CUDA/MACA sources do not build
    here, so the arm times the C restatement of its algorithm on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n = cpu_seqs(args.workload)
    c = cpu_sample(args.workload, n)
    for _ in range(max(1, min(args.warmup, 2))):
        cpu_step(c, cores)
    steps = min(args.steps, 40)  # ~70 ms per step at c2 on 16 threads; bounded so that any --steps ends within minutes
    t0 = time.time()
    for _ in range(steps):
        cpu_step(c, cores)
    dt = (time.time() - t0) / steps
    value = c["bytes"] / dt / 1e9
    sample = "%d of %d sequences of the workload per step, attention only, OpenMP over (seq, kv head)%s" % (
        n, WORKLOADS[args.workload][0], "" if steps == args.steps else "; %d of the %d requested steps timed" % (steps, args.steps))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload][8], "name": args.workload, "step": "paged attention (CPU restatement)",
                   "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip extra.configs (N = 1) / the tp block (N > 1)")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle comparison of sampled outputs")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
