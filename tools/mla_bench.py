"""MLA decode timing on a GPU box: this repository's kernel (--impl ours) or the stock sm_100a CUTLASS kernel that
upstream vLLM 0.22 ships (--impl upstream; the reference's csrc/attention/mla/cutlass_mla_kernels.cu is the same
kernel).  The two run in separate processes on identical seeded inputs; each writes a JSON line per shape and a small
output sample so the results can be compared afterwards (--compare a.pt b.pt).

    python tools/mla_bench.py --impl ours --out gpurun_out/mla_ours
    python tools/mla_bench.py --impl upstream --out gpurun_out/mla_upstream
    python tools/mla_bench.py --compare gpurun_out/mla_ours.pt gpurun_out/mla_upstream.pt
"""
import argparse
import json
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHAPES = [  # batch, ctx, heads
    (128, 4096, 128), (128, 8192, 128), (16, 32768, 128), (1, 131072, 128),
    (128, 4096, 16), (128, 8192, 16), (128, 8192, 64), (1, 131072, 16),
]
PAGE = 64


VARLEN = False


def make(b, L, H, dev):
    g = torch.Generator(device=dev).manual_seed(b * 7 + L + H)
    pages = L // PAGE
    nb = b * pages
    cache = torch.randn(nb, PAGE, 576, generator=g, device=dev, dtype=torch.bfloat16)
    q = torch.randn(b, H, 576, generator=g, device=dev, dtype=torch.bfloat16)
    table = torch.randperm(nb, generator=g, device=dev).to(torch.int32).view(b, pages)
    lens = torch.full((b,), L, dtype=torch.int32, device=dev)
    if VARLEN:  # ragged batch like the reference test's varlen case (normal around the mean, sigma = mean / 2), capped at L
        gc = torch.Generator().manual_seed(b + L)
        lens = torch.empty(b).normal_(0.6 * L, 0.3 * L, generator=gc).clamp_(128, L).to(torch.int32).to(dev)
    return q, cache, table, lens


def time_it(fn, reps, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", choices=["ours", "upstream"])
    ap.add_argument("--out", default="gpurun_out/mla")
    ap.add_argument("--compare", nargs=2)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--splits", type=int, default=-1)
    ap.add_argument("--only", type=int, default=-1)
    ap.add_argument("--fp8", action="store_true", help="float8_e4m3fn query and cache (descale 1), bf16 output")
    ap.add_argument("--varlen", action="store_true", help="ragged sequence lengths (mean 0.6 ctx, sigma 0.3 ctx, capped at ctx)")
    ap.add_argument("--witness", help="(--impl upstream) run the stock kernel on small CPU-seeded inputs and save inputs + outputs")
    args = ap.parse_args()
    global VARLEN
    VARLEN = args.varlen
    if args.compare:
        a, b = torch.load(args.compare[0]), torch.load(args.compare[1])
        for k in sorted(set(a) & set(b)):
            x, y = a[k].double(), b[k].double()
            cd = 1 - 2 * (x * y).sum().item() / max((x * x + y * y).sum().item(), 1e-12)
            print("%s: max-abs diff %.3e cos-diff %.2e" % (k, (x - y).abs().max().item(), cd))
        return
    dev = "cuda:0"
    if args.impl == "ours":
        sys.path.insert(0, ROOT)
        import vllm_metax_b200
        vllm_metax_b200.import_kernels()
    else:
        import vllm  # noqa: F401
        try:
            import vllm._C_stable_libtorch  # noqa: F401
        except Exception as e:
            print(json.dumps({"impl": "upstream", "unavailable": "%s: %s" % (type(e).__name__, e)}))
            return
    op = torch.ops._C.sm100_cutlass_mla_decode
    ws_size = torch.ops._C.sm100_cutlass_mla_get_workspace_size
    if args.witness:
        # small fixture for tests/golden: CPU-seeded inputs (no project code), stock kernel outputs
        g = torch.Generator().manual_seed(2026)
        b, H, page, pages = 3, 128, 64, 6
        lens = torch.tensor([384, 201, 77], dtype=torch.int32)
        nb = b * pages + 2
        cache = torch.randn(nb, page, 576, generator=g).to(torch.bfloat16)
        q = torch.randn(b, H, 576, generator=g).to(torch.bfloat16)
        table = torch.randperm(nb, generator=g)[:b * pages].to(torch.int32).view(b, pages)
        out = torch.empty(b, H, 512, dtype=torch.bfloat16, device=dev)
        lse = torch.empty(b, H, dtype=torch.float32, device=dev)
        ws = torch.empty(max(16, ws_size(pages * page, b, 148, 1)), dtype=torch.uint8, device=dev)
        scale = 1.0 / math.sqrt(576)
        op(out, lse, q[..., :512].contiguous().to(dev), q[..., 512:].contiguous().to(dev), cache.to(dev), lens.to(dev),
           table.to(dev), ws, scale, 1)
        torch.cuda.synchronize()
        torch.save(dict(query=q, kv_cache=cache, block_tables=table, seq_lens=lens, scale=scale, upstream_out=out.cpu(),
                        upstream_lse=lse.cpu(), producer="vLLM 0.22.0 torch.ops._C.sm100_cutlass_mla_decode on B200"),
                   args.witness)
        print("witness written:", args.witness)
        return
    samples = {}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for idx, (b, L, H) in enumerate(SHAPES):
        if args.only >= 0 and idx != args.only:
            continue
        q, cache, table, lens = make(b, L, H, dev)
        Hk = H if args.impl == "ours" else 128   # the stock kernel needs exactly 128 heads: the caller pads
        qn = torch.zeros(b, Hk, 512, dtype=q.dtype, device=dev)
        qp = torch.zeros(b, Hk, 64, dtype=q.dtype, device=dev)
        qn[:, :H], qp[:, :H] = q[..., :512], q[..., 512:]
        if args.fp8:
            qn, qp, cache = qn.to(torch.float8_e4m3fn), qp.to(torch.float8_e4m3fn), cache.to(torch.float8_e4m3fn)
        out = torch.empty(b, Hk, 512, dtype=torch.bfloat16, device=dev)
        lse = torch.empty(b, Hk, dtype=torch.float32, device=dev)
        ws = torch.empty(max(16, ws_size(L, b, 148, args.splits)), dtype=torch.uint8, device=dev)
        scale = 1.0 / math.sqrt(576)

        def fn():
            op(out, lse, qn, qp, cache, lens, table, ws, scale, args.splits)

        try:
            med, mn = time_it(fn, args.reps, flush if b * L * 1152 < (300 << 20) else None)
        except Exception as e:
            print(json.dumps({"impl": args.impl, "shape": [b, L, H], "error": "%s: %s" % (type(e).__name__, str(e)[:300])}), flush=True)
            continue
        tot = int(lens.sum().item())
        eb = 1 if args.fp8 else 2
        kv_bytes = tot * 576 * eb
        byt = kv_bytes + b * H * (576 * eb + 512 * 2)
        flops = 2.0 * tot * H * (576 + 512)
        print(json.dumps({"impl": args.impl, "shape": {"batch": b, "ctx": L, "heads": H, "page": PAGE}, "us": med * 1e3,
                          "us_min": mn * 1e3, "gbs": byt / med / 1e6, "tflops": flops / med / 1e9,
                          "heads_computed": Hk, "varlen": VARLEN, "tokens": tot, "fp8": bool(args.fp8)}), flush=True)
        samples["out_b%d_L%d_H%d" % (b, L, H)] = out[:2, :H].float().cpu()
        samples["lse_b%d_L%d_H%d" % (b, L, H)] = lse[:2, :H].cpu()
        del q, cache, table, lens, qn, qp, out, lse, ws
        torch.cuda.empty_cache()
    torch.save(samples, args.out + ".pt")


if __name__ == "__main__":
    main()
