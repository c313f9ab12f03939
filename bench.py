#!/usr/bin/env python
"""Benchmark of the paged-attention decode hot path (BASELINE.json metric) — contract in the task prompt.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3]

One "step" = one decode layer-call of the hot path over one batch:
    reshape_and_cache(new K/V of the batch -> paged cache)  +  paged_attention_v1(query, cache)
Workload (config.workload): BASELINE.json configs[1] — Llama-3-8B decode attention, GQA 32q/8kv, head 128,
block 16, batch 64, context 4096, bf16 KV, synthetic random cache (seed 0), unique random block tables.
N > 1 (torchrun, one rank per GPU): every rank runs the same per-GPU workload on its own sequences — the
path has no collective — so scaling is weak and value is the sum over ranks.

value      device-resident inputs, CUDA-event timed, max over ranks      (GB/s of algorithmic bytes)
e2e        the same steps through the public op wrappers with HOST inputs: pinned H2D of the step's
           q/k/v/slot_mapping/block_tables/seq_lens and D2H of the attention output inside the timed region
roofline   the attention launch alone (CUDA events on the launching stream) against MEASURED_PEAKS.json
cpu_baseline / --impl reference: the CPU restatement (oracle/paged_attn_ref.c, OpenMP, all host threads) on a
           bounded sample of the same workload.  The oracle is used here only as the timed CPU baseline.
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
    # name: (S, H, Hkv, D, B, L, kv_cache_dtype, description)
    "c2": (64, 32, 8, 128, 16, 4096, "auto",
           "Llama-3-8B decode attention, GQA 32q/8kv, head 128, block 16, batch 64, ctx 4096, bf16 KV"),
    "c3": (64, 32, 8, 128, 16, 4096, "fp8",
           "Llama-3-8B decode attention, GQA 32q/8kv, head 128, block 16, batch 64, ctx 4096, FP8-e4m3 KV"),
    # the serving path of the reference: flash cache layout [NB, B, Hkv, D] (NHD), reshape_and_cache_flash +
    # flash-layout decode (flash_attn_with_kvcache in the reference).  Same shapes and bytes as c2.
    "c2_flash": (64, 32, 8, 128, 16, 4096, "auto",
                 "Llama-3-8B decode attention, GQA 32q/8kv, head 128, block 16, batch 64, ctx 4096, bf16 KV, flash NHD cache layout"),
}
N_LAYERS = 32  # Llama-3-8B; tokens/s = S / (n_layers * t_step)  
METRIC = "paged-attn decode HBM GB/s"
UNIT = "GB/s"


def algorithmic_bytes(S, H, Hkv, D, B, L, kv_eb, with_cache_write=True):
    """every KV byte once + q/out + block-table entries + seq_lens (+ the decode cache write)."""
    attn = S * L * Hkv * D * 2 * kv_eb + S * H * D * 4 + S * math.ceil(L / B) * 4 + 4 * S
    write = S * Hkv * D * 2 * (2 + kv_eb) + S * 8 if with_cache_write else 0
    return attn, write


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md 6.65 TB/s)"


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
def make_workload(name, dev, seed=0):
    S, H, Hkv, D, B, L, kvd, _ = WORKLOADS[name]
    nblk = (L + B - 1) // B
    NB = int(S * nblk * 1.25)
    g = torch.Generator(device=dev).manual_seed(seed)
    scale = D ** -0.5
    fp8 = kvd != "auto"

    def uni(shape):
        return (torch.rand(shape, device=dev, generator=g) * 2 - 1) * scale

    x = 16 if fp8 else 8
    flash = name.endswith("_flash")
    if flash:
        kc = uni((NB, B, Hkv, D)).to(torch.bfloat16)
        vc = uni((NB, B, Hkv, D)).to(torch.bfloat16)
    elif fp8:
        kc = uni((NB, Hkv, D // x, B, x)).to(torch.float8_e4m3fn).view(torch.uint8)
        vc = uni((NB, Hkv, D, B)).to(torch.float8_e4m3fn).view(torch.uint8)
    else:
        kc = uni((NB, Hkv, D // x, B, x)).to(torch.bfloat16)
        vc = uni((NB, Hkv, D, B)).to(torch.bfloat16)
    w = dict(
        name=name, flash=flash, S=S, H=H, Hkv=Hkv, D=D, B=B, L=L, kvd=kvd, scale=scale, kv_eb=1 if fp8 else 2,
        key_cache=kc, value_cache=vc,
        query=uni((S, H, D)).to(torch.bfloat16), key=uni((S, Hkv, D)).to(torch.bfloat16),
        value=uni((S, Hkv, D)).to(torch.bfloat16),
        block_tables=torch.randperm(NB, device=dev, generator=g)[: S * nblk].view(S, nblk).to(torch.int32),
        seq_lens=torch.full((S,), L, dtype=torch.int32, device=dev),
        k_scale=torch.ones(1, dtype=torch.float32, device=dev), v_scale=torch.ones(1, dtype=torch.float32, device=dev),
    )
    # the decode step writes the K/V of position L-1 of every sequence
    last = w["block_tables"][:, (L - 1) // B].to(torch.int64) * B + (L - 1) % B
    w["slot_mapping"] = last
    w["out"] = torch.empty_like(w["query"])
    return w


def run_ours(args):
    from vllm_metax_b200 import _custom_ops as ops  # loads the native libraries; raises if they are missing
    from vllm_metax_b200 import _native

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

    w = make_workload(args.workload, dev, seed=rank)
    S, H, Hkv, D, B, L = w["S"], w["H"], w["Hkv"], w["D"], w["B"], w["L"]

    def cache_write(key, value, slots):
        if w["flash"]:
            ops.reshape_and_cache_flash(key, value, w["key_cache"], w["value_cache"], slots, w["kvd"], w["k_scale"],
                                        w["v_scale"])
        else:
            ops.reshape_and_cache(key, value, w["key_cache"], w["value_cache"], slots, w["kvd"], w["k_scale"], w["v_scale"])

    def attention(q, bt, lens, out):
        if w["flash"]:
            ops.paged_attention_flash(out, q, w["key_cache"], w["value_cache"], w["scale"], bt, lens, L)
        else:
            ops.paged_attention_v1(out, q, w["key_cache"], w["value_cache"], Hkv, w["scale"], bt, lens, B, L, None,
                                   w["kvd"], w["k_scale"], w["v_scale"])

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

    # ---- reduce over ranks (max time) -----------------------------------------------------------------------
    t = torch.tensor([total_ms, e2e_ms, attn_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, attn_ms = t.tolist()
    attn_bytes, write_bytes = algorithmic_bytes(S, H, Hkv, D, B, L, w["kv_eb"])
    step_bytes = attn_bytes + write_bytes
    ms_per_step = total_ms / args.steps
    value = world * step_bytes / (ms_per_step * 1e-3) / 1e9
    e2e_value = world * step_bytes / (e2e_ms / args.steps * 1e-3) / 1e9
    pk, pk_src = peaks()
    achieved = attn_bytes / (attn_ms * 1e-3) / 1e9

    extra = {}
    if rank == 0:
        extra["tokens_per_s_attention_only"] = world * S / (N_LAYERS * ms_per_step * 1e-3)
        extra["attention_ms"] = attn_ms
        extra["frac_of_nominal_8TBps"] = achieved / 8000.0
        extra["variant"] = {1: "generic_fma", 2: "stream_tma_mma"}.get(_native.last_attention_variant(), "?")

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(args.workload, budget_s=12.0)

    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if w["kvd"] == "auto" else "bf16 (q/out) + fp8-e4m3 (KV)", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload][7], "name": args.workload,
                   "step": "reshape_and_cache_flash + paged_attention_flash" if w["flash"] else "reshape_and_cache + paged_attention_v1",
                   "l2": "inputs larger than L2 (KV working set %.2f GB vs 126 MB L2)" % (attn_bytes / 1e9),
                   "parallelism": "dp%d (independent sequences per GPU, no collective)" % world,
                   "block_tables": "unique random permutation, 1.25x over-provisioned cache", "seed": 0},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                     "frac": achieved / pk["hbm_gbs"], "traffic": TRAFFIC_BYTES.get(args.workload),
                     "kernel": "paged_attention_stream_kernel" + (" (flash layout, tiled TMA)" if w["flash"] else " (+ split reduce)"), "algorithmic_bytes": attn_bytes,
                     "peak_source": pk_src},
        "cpu_baseline": cpu,
        "extra": extra,
    }
    if saved_stdout is not None:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the attention kernel, from the ncu --set full
# captures committed under profiles/ (r1_c2_bf16_ncu_summary.txt, r1_c3_fp8_ncu_summary.txt, r1_c2_flash_ncu_summary.txt); null where no capture exists.
TRAFFIC_BYTES = {"c2": 1075123000 + 4585728, "c3": 538070016 + 4122624, "c2_flash": 1075082000 + 7046144}


# ---- CPU arm --------------------------------------------------------------------------------------------
def cpu_sample(workload, n_seqs, seed=0):
    """A bounded sample of the workload for the CPU arm: the first n_seqs sequences (same shapes).  The C restatement
    reads the v1/v2 cache layout; the flash workload is the same arithmetic over the same bytes, so it is timed on that."""
    S, H, Hkv, D, B, L, kvd, _ = WORKLOADS[workload]
    nblk = (L + B - 1) // B
    NB = n_seqs * nblk
    g = torch.Generator().manual_seed(seed)
    scale = D ** -0.5
    x = 16 if kvd != "auto" else 8

    def uni(shape):
        return (torch.rand(shape, generator=g) * 2 - 1) * scale

    if kvd == "auto":
        kc, vc = uni((NB, Hkv, D // x, B, x)).to(torch.bfloat16), uni((NB, Hkv, D, B)).to(torch.bfloat16)
    else:
        kc = uni((NB, Hkv, D // x, B, x)).to(torch.float8_e4m3fn).view(torch.uint8)
        vc = uni((NB, Hkv, D, B)).to(torch.float8_e4m3fn).view(torch.uint8)
    q = uni((n_seqs, H, D)).to(torch.bfloat16)
    bt = torch.randperm(NB, generator=g).view(n_seqs, nblk).to(torch.int32)
    lens = torch.full((n_seqs,), L, dtype=torch.int32)
    return dict(q=q, kc=kc, vc=vc, bt=bt, lens=lens, scale=scale, kvd=kvd, n=n_seqs,
                bytes=algorithmic_bytes(n_seqs, H, Hkv, D, B, L, 1 if kvd != "auto" else 2, False)[0])


def cpu_step(c, threads):
    from oracle import c_oracle  # the timed CPU baseline: the one use of oracle/ outside tests

    return c_oracle.paged_attention(c["q"], c["kc"], c["vc"], c["bt"], c["lens"], c["scale"], None, c["kvd"], 1.0, 1.0,
                                    num_threads=threads)


def cpu_baseline(workload, budget_s):
    cores = os.cpu_count() or 1
    n = 16
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
    n = 16
    c = cpu_sample(args.workload, n)
    for _ in range(max(1, min(args.warmup, 2))):
        cpu_step(c, cores)
    t0 = time.time()
    for _ in range(args.steps):
        cpu_step(c, cores)
    dt = (time.time() - t0) / args.steps
    value = c["bytes"] / dt / 1e9
    sample = "%d of %d sequences of the workload per step, attention only, OpenMP over (seq, kv head)" % (
        n, WORKLOADS[args.workload][0])
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload][7], "name": args.workload, "step": "paged attention (CPU restatement)",
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
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
