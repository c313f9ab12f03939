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
    ap.add_argument("--per-sm", action="store_true", help="per-CTA streaming time against %%smid, and how stable it is across launches")
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
    per_rep = []
    for rep in range(4 if a.per_sm else 3):
        flush.add_(1)
        if a.per_sm:  # a new block placement per launch, so that what repeats is the SM and not the data
            bt.copy_(torch.randperm(kc.shape[0], device="cuda")[: bt.numel()].view_as(bt).to(torch.int32))
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
        per_rep.append((t[:, 5].astype(np.int64), stream, (t[:, 3] - t0) / 1e3))
    if a.per_sm:
        sm0 = per_rep[0][0]
        print("blockIdx -> smid identical across launches:", all((r[0] == sm0).all() for r in per_rep))
        by_sm = []
        for smid, stream, _ in per_rep:
            d = np.full(int(max(r[0].max() for r in per_rep)) + 1, np.nan)
            d[smid] = stream
            by_sm.append(d)
        by_sm = np.array(by_sm)
        ok = ~np.isnan(by_sm).any(axis=0)
        rel = by_sm[:, ok] / np.median(by_sm[:, ok], axis=1, keepdims=True)
        for i in range(len(rel)):
            for j in range(i + 1, len(rel)):
                print("   correlation of per-SM streaming time, launch %d vs %d: %.3f" % (i, j, np.corrcoef(rel[i], rel[j])[0, 1]))
        mean = rel.mean(axis=0)
        ids = np.nonzero(ok)[0]
        print("   per-SM streaming time / median (mean of the launches), by smid:")
        for k in range(0, len(ids), 16):
            print("   smid %3d..: %s" % (ids[k], " ".join("%.3f" % v for v in mean[k:k + 16])))
        print("   spread of the per-SM means: min %.3f max %.3f; launch-to-launch std of one SM: %.3f" % (mean.min(), mean.max(), rel.std(axis=0).mean()))


if __name__ == "__main__":
    main()
