"""Developer check: tcgen05 fp8 kernel (variant 3) against the generic FMA kernel (variant 1) on a list of shapes."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_metax_b200 import _custom_ops as ops  # noqa: E402
from vllm_metax_b200 import _native  # noqa: E402


def make(S, H, Hkv, L, seed, ragged, dev="cuda", D=128, B=16):
    g = torch.Generator(device=dev).manual_seed(seed)
    nblk = (L + B - 1) // B
    NB = S * nblk + 7
    scale = D ** -0.5
    kc = ((torch.rand((NB, Hkv, D // 16, B, 16), device=dev, generator=g) * 2 - 1) * 3.0).to(torch.float8_e4m3fn).view(torch.uint8)
    vc = ((torch.rand((NB, Hkv, D, B), device=dev, generator=g) * 2 - 1) * 3.0).to(torch.float8_e4m3fn).view(torch.uint8)
    q = ((torch.rand((S, H, D), device=dev, generator=g) * 2 - 1) * 1.5).to(torch.bfloat16)
    bt = torch.randperm(NB, device=dev, generator=g)[: S * nblk].view(S, nblk).to(torch.int32)
    if ragged:
        lens = torch.randint(1, L + 1, (S,), device=dev, generator=g, dtype=torch.int32)
        lens[-1] = L
    else:
        lens = torch.full((S,), L, dtype=torch.int32, device=dev)
    return q, kc, vc, bt, lens, scale


def run(variant, v2, q, kc, vc, bt, lens, scale, Hkv, L, ks, vs, alibi):
    _native.set_attention_variant(variant)
    out = torch.full_like(q, float("nan"))
    S, H, D = q.shape
    if v2:
        P = (L + 511) // 512
        tmp = torch.empty(S, H, P, D, dtype=q.dtype, device=q.device)
        es = torch.empty(S, H, P, dtype=torch.float32, device=q.device)
        ml = torch.empty_like(es)
        ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, lens, 16, L, alibi, "fp8", ks, vs)
    else:
        ops.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, lens, 16, L, alibi, "fp8", ks, vs)
    torch.cuda.synchronize()
    _native.set_attention_variant(0)
    return out


def main():
    cases = [
        # S, H, Hkv, L, ragged, v2, alibi
        (1, 8, 1, 32, False, False, False),
        (1, 8, 1, 128, False, False, False),
        (1, 4, 1, 100, False, False, False),
        (2, 8, 2, 7, False, False, False),
        (3, 32, 8, 700, True, False, False),
        (3, 32, 8, 700, True, True, False),
        (5, 16, 2, 1500, True, False, True),
        (64, 32, 8, 4096, False, False, False),
        (64, 32, 8, 4096, True, True, False),
        (8, 8, 1, 20000, True, False, False),
        (2, 2, 2, 3000, True, True, True),
    ]
    only = int(sys.argv[1]) if len(sys.argv) > 1 else None
    ks = torch.tensor([0.02], dtype=torch.float32, device="cuda")
    vs = torch.tensor([0.05], dtype=torch.float32, device="cuda")
    bad = 0
    for i, (S, H, Hkv, L, ragged, v2, use_alibi) in enumerate(cases):
        if only is not None and i != only:
            continue
        q, kc, vc, bt, lens, scale = make(S, H, Hkv, L, i, ragged)
        alibi = (torch.rand(H, device="cuda") * 0.1).float() if use_alibi else None
        ks = torch.tensor([2.0 if i % 2 else 0.02], dtype=torch.float32, device="cuda")
        ref = run(1, v2, q, kc, vc, bt, lens, scale, Hkv, L, ks, vs, alibi)
        got = run(3, v2, q, kc, vc, bt, lens, scale, Hkv, L, ks, vs, alibi)
        got2 = run(3, v2, q, kc, vc, bt, lens, scale, Hkv, L, ks, vs, alibi)
        err = (got.float() - ref.float()).abs().max().item()
        det = torch.equal(got, got2)
        nan = int(torch.isnan(got.float()).sum())
        print("case %2d S=%d H=%d Hkv=%d L=%d ragged=%d v2=%d alibi=%d: max|umma-generic| %.3e (ref max %.3e) nan=%d deterministic=%s"
              % (i, S, H, Hkv, L, ragged, v2, use_alibi, err, ref.float().abs().max().item(), nan, det), flush=True)
        if not (err <= 2e-3) or nan or not det:
            bad += 1
    print("FAILED %d" % bad if bad else "ALL OK")


if __name__ == "__main__":
    main()
