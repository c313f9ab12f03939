"""Developer probe: where does a streaming-kernel launch spend its time?  Needs a -DB200_TIMELINE build:

    B200_EXTRA_NVCC_FLAGS=-DB200_TIMELINE python -m vllm_metax_b200.build   (after touching attention_stream.cu)
    python tools/timeline_probe.py --S 64 --H 32 --Hkv 8 --L 4096 [--fp8] [--v2]

Prints, relative to the earliest CTA entry of the LAST launch: per-CTA entry, schedule-ready, first-tile, last-tile and
exit times (min / median / max over CTAs), next to the CUDA-event time of the launch."""
import argparse
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from quick_bench import make  # noqa: E402
from vllm_metax_b200 import _custom_ops as ops  # noqa: E402
from vllm_metax_b200 import _native  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    for k, v in (("S", 64), ("H", 32), ("Hkv", 8), ("D", 128), ("B", 16), ("L", 4096)):
        ap.add_argument("--" + k, type=int, default=v)
    ap.add_argument("--fp8", action="store_true")
    ap.add_argument("--v2", action="store_true")
    a = ap.parse_args()
    q, kc, vc, bt, lens, scale = make(a.S, a.H, a.Hkv, a.D, a.B, a.L, a.fp8)
    out = torch.empty_like(q)
    one = torch.ones(1, dtype=torch.float32, device="cuda")
    kvd = "fp8" if a.fp8 else "auto"
    P = (a.L + 511) // 512
    if a.v2:
        tmp = torch.empty(a.S, a.H, P, a.D, dtype=q.dtype, device="cuda")
        es = torch.empty(a.S, a.H, P, dtype=torch.float32, device="cuda")
        ml = torch.empty_like(es)
        fn = lambda: ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, a.Hkv, scale, bt, lens, a.B, a.L, None, kvd, one, one)
    else:
        fn = lambda: ops.paged_attention_v1(out, q, kc, vc, a.Hkv, scale, bt, lens, a.B, a.L, None, kvd, one, one)
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush = torch.zeros(64 << 20, dtype=torch.int32, device="cuda")
    for rep in range(3):
        flush.add_(1)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        lib = _native.load()
        buf = (C.c_ulonglong * (256 * 8))()
        lib.b200_debug_timeline.argtypes = [C.c_void_p]
        assert lib.b200_debug_timeline(buf) == 0
        t = np.frombuffer(buf, dtype=np.uint64).reshape(256, 8).astype(np.float64)
        n = min(148, int((t[:, 0] > 0).sum()))
        t = t[:n]
        t0 = t[:, 0].min()
        names = ["entry", "sched ready", "first tile", "last tile consumed", "merged / exit"]
        print("rep %d: event time %.1f us, %d CTAs" % (rep, e0.elapsed_time(e1) * 1e3, n))
        for i, nm in enumerate(names):
            c = (t[:, i] - t0) / 1e3
            print("   %-20s min %7.2f  med %7.2f  max %7.2f us" % (nm, c.min(), np.median(c), c.max()))
        stream = (t[:, 3] - t[:, 2]) / 1e3
        print("   streaming phase per CTA: min %.2f med %.2f max %.2f us" % (stream.min(), np.median(stream), stream.max()))


if __name__ == "__main__":
    main()
