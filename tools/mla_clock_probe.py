"""Developer probe: SM clock, power and throttle reasons while the MLA decode kernel runs back to back (NVML sampled
from a side thread), next to the same for the paged-attention headline kernel.  python tools/mla_clock_probe.py [--fp8]"""
import argparse
import os
import sys
import threading
import time

import pynvml
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from oracle import torch_oracle as orc  # noqa: E402  (input generator only)
from quick_bench import make  # noqa: E402
from vllm_metax_b200 import _custom_ops as ops  # noqa: E402


def sample_while(fn, seconds=1.5):
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(int(os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0]))
    rows, stop = [], threading.Event()

    def loop():
        while not stop.is_set():
            rows.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1e3,
                         pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)))
            time.sleep(0.005)

    t = threading.Thread(target=loop)
    t.start()
    t_end = time.time() + seconds
    n = 0
    while time.time() < t_end:
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
        n += 20
    stop.set()
    t.join()
    rows = rows[len(rows) // 3:]  # the first third ramps up
    clk = sorted(r[0] for r in rows)
    pw = sorted(r[1] for r in rows)
    reasons = 0
    for r in rows:
        reasons |= r[2]
    names = [k[len("nvmlClocksThrottleReason"):] for k in dir(pynvml)
             if k.startswith("nvmlClocksThrottleReason") and isinstance(getattr(pynvml, k), int) and getattr(pynvml, k) & reasons
             and k != "nvmlClocksThrottleReasonAll"]
    return {"launches": n, "sm_mhz_median": clk[len(clk) // 2], "sm_mhz_min": clk[0], "power_w_median": round(pw[len(pw) // 2]),
            "power_limit_w": pynvml.nvmlDeviceGetEnforcedPowerLimit(h) / 1e3, "reasons": names, "samples": len(rows)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--fp8", action="store_true")
    a = ap.parse_args()
    dev = "cuda"
    for H in (16, 64, 128):
        d = orc.make_mla_inputs(128, H, 64, 8192, seed=1, varlen=False, shuffle_pages=True)
        q, kv = d["query"].to(dev), d["kv_cache"].to(dev)
        if a.fp8:
            q, kv = q.to(torch.float8_e4m3fn), kv.to(torch.float8_e4m3fn)
        qn, qp = q[..., :512].contiguous(), q[..., 512:].contiguous()
        bt, sl = d["block_tables"].to(dev), d["seq_lens"].to(dev)
        out = torch.empty(128, H, 512, dtype=torch.bfloat16, device=dev)
        lse = torch.empty(128, H, dtype=torch.float32, device=dev)
        ws = torch.empty(ops.sm100_cutlass_mla_get_workspace_size(8192, 128, 148, -1), dtype=torch.uint8, device=dev)
        fn = lambda: ops.sm100_cutlass_mla_decode(out, lse, qn, qp, kv, sl, bt, ws, d["scale"], -1)
        print("mla %s H=%d:" % ("fp8" if a.fp8 else "bf16", H), sample_while(fn), flush=True)
        del q, kv, qn, qp, out, ws
    q, kc, vc, bt, lens, scale = make(64, 32, 8, 128, 16, 4096, False)
    out = torch.empty_like(q)
    one = torch.ones(1, dtype=torch.float32, device=dev)
    fn = lambda: ops.paged_attention_v1(out, q, kc, vc, 8, scale, bt, lens, 16, 4096, None, "auto", one, one)
    print("paged_attention_v1 C2:", sample_while(fn), flush=True)


if __name__ == "__main__":
    main()
