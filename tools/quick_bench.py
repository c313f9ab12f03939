"""Developer micro-benchmark: time paged_attention_v1/v2 on synthetic caches (CUDA events)."""
import argparse
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_metax_b200 import _custom_ops as ops  # noqa: E402
from vllm_metax_b200 import _native  # noqa: E402


def make(S, H, Hkv, D, B, L, fp8, dev="cuda", slack=1.25):
    nblk_seq = (L + B - 1) // B
    NB = int(S * nblk_seq * slack) + 1
    g = torch.Generator(device=dev).manual_seed(0)
    scale = D ** -0.5
    if fp8:
        x = 16
        kc = torch.randint(0, 120, (NB, Hkv, D // x, B, x), dtype=torch.uint8, device=dev, generator=g)
        vc = torch.randint(0, 120, (NB, Hkv, D, B), dtype=torch.uint8, device=dev, generator=g)
    else:
        x = 8
        kc = ((torch.rand((NB, Hkv, D // x, B, x), device=dev, generator=g) * 2 - 1) * scale).to(torch.bfloat16)
        vc = ((torch.rand((NB, Hkv, D, B), device=dev, generator=g) * 2 - 1) * scale).to(torch.bfloat16)
    q = ((torch.rand((S, H, D), device=dev, generator=g) * 2 - 1) * scale).to(torch.bfloat16)
    perm = torch.randperm(NB, device=dev, generator=g)[: S * nblk_seq].view(S, nblk_seq).to(torch.int32)
    lens = torch.full((S,), L, dtype=torch.int32, device=dev)
    return q, kc, vc, perm, lens, scale


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--S", type=int, default=64)
    ap.add_argument("--H", type=int, default=32)
    ap.add_argument("--Hkv", type=int, default=8)
    ap.add_argument("--D", type=int, default=128)
    ap.add_argument("--B", type=int, default=16)
    ap.add_argument("--L", type=int, default=4096)
    ap.add_argument("--fp8", action="store_true")
    ap.add_argument("--v2", action="store_true")
    ap.add_argument("--flash", action="store_true", help="flash cache layout [NB, B, Hkv, D] (paged_attention_flash)")
    ap.add_argument("--hnd", action="store_true", help="with --flash: HND storage [NB, Hkv, B, D]")
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--check", action="store_true", help="compare against the generic FMA kernel")
    a = ap.parse_args()
    q, kc, vc, bt, lens, scale = make(a.S, a.H, a.Hkv, a.D, a.B, a.L, a.fp8)
    os.environ["B200_DEBUG_NUM_BLOCKS"] = str(kc.shape[0])
    out = torch.empty_like(q)
    one = torch.ones(1, dtype=torch.float32, device="cuda")
    kvd = "fp8" if a.fp8 else "auto"
    _native.set_attention_variant(a.variant)
    P = (a.L + 511) // 512
    if a.flash:
        # same values, re-laid out: [NB, Hkv, D/x, B, x] -> [NB, B, Hkv, D]
        NB = kc.shape[0]
        kf = kc.permute(0, 3, 1, 2, 4).reshape(NB, a.B, a.Hkv, a.D)
        vf = vc.permute(0, 3, 1, 2)
        if a.hnd:
            kf = kf.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
            vf = vf.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
        else:
            kf, vf = kf.contiguous(), vf.contiguous()
        fn = lambda: ops.paged_attention_flash(out, q, kf, vf, scale, bt, lens, a.L)
        if a.check:
            ref = torch.empty_like(q)
            ops.paged_attention_v1(ref, q, kc, vc, a.Hkv, scale, bt, lens, a.B, a.L, None, kvd, one, one)
            fn()
            torch.cuda.synchronize()
            print("check flash vs v1 layout: max abs diff %.3e (ref max %.3e)" % (
                (ref.float() - out.float()).abs().max().item(), ref.float().abs().max().item()))
            a.check = False
    elif a.v2:
        tmp = torch.empty(a.S, a.H, P, a.D, dtype=q.dtype, device="cuda")
        es = torch.empty(a.S, a.H, P, dtype=torch.float32, device="cuda")
        ml = torch.empty_like(es)
        fn = lambda: ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, a.Hkv, scale, bt, lens, a.B, a.L, None, kvd, one, one)
    else:
        fn = lambda: ops.paged_attention_v1(out, q, kc, vc, a.Hkv, scale, bt, lens, a.B, a.L, None, kvd, one, one)
    fn()
    torch.cuda.synchronize()
    if a.check:
        res = out.clone()
        _native.set_attention_variant(1)
        fn()
        torch.cuda.synchronize()
        print("check vs generic: max abs diff %.3e (ref max %.3e)" % ((res.float() - out.float()).abs().max().item(), out.float().abs().max().item()))
        _native.set_attention_variant(a.variant)
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.iters)]
    for s, e in evs:
        s.record()
        fn()
        e.record()
    torch.cuda.synchronize()
    ts = sorted(s.elapsed_time(e) for s, e in evs)
    med = ts[len(ts) // 2]
    eb = 1 if a.fp8 else 2
    kvbytes = a.S * a.L * a.Hkv * a.D * 2 * eb
    total = kvbytes + a.S * a.H * a.D * 4 + a.S * math.ceil(a.L / a.B) * 4 + 4 * a.S
    print(("flash %s " % ("hnd" if a.hnd else "nhd") if a.flash else "") + "cfg S=%d H=%d Hkv=%d D=%d B=%d L=%d fp8=%d v2=%d variant=%d: median %.1f us  min %.1f us  -> %.0f GB/s (median) %.0f GB/s (best)  [%.3f GB]" % (
        a.S, a.H, a.Hkv, a.D, a.B, a.L, a.fp8, a.v2, _native.last_attention_variant(), med * 1e3, ts[0] * 1e3,
        total / med / 1e6, total / ts[0] / 1e6, total / 1e9))


if __name__ == "__main__":
    main()
