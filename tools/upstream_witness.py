#!/usr/bin/env python
"""Upstream witness (SURVEY.md F5 / §8(d) "secondary GPU comparator" / App. C.1): run on the GPU box, in its OWN
process, against the kernels UPSTREAM vLLM ships compiled for sm_100 in `vllm/_C.abi3.so` — the same SIMT algorithm as
the reference's csrc/attention/attention_kernels.cuh and csrc/cache_kernels.cu with the real `__nv_cvt` FP8 path the
MetaX fork stubs out (csrc/quantization/fp8/metax/quant_utils.cuh:13-23).

    python tools/upstream_witness.py [--out gpurun_out/witness] [--no-timing]

Nothing of this repository's product or oracle is imported: the only project file used is
tests/golden/witness_inputs.py (deterministic inputs).  Outputs:

  <out>/fp8_upstream_convert.pt     convert_fp8 quantisation of every bf16 / f16 pattern and of fp32 rounding-boundary
                                    values at several scales; dequantisation of every e4m3 code to bf16 / f16 / f32
  <out>/fp8_upstream_cache.pt       reshape_and_cache / reshape_and_cache_flash / concat_and_cache_mla with fp8 caches
                                    and non-unit scales: the cache bytes
  <out>/fp8_upstream_attn.pt        paged_attention_v1 / _v2 outputs (bf16 and fp8 caches, non-unit scales)
  <out>/summary.json                import status, versions, and the timing of the stock kernels on C1..C5

Copy the .pt files to tests/golden/ to pin the oracle (CPU test) and the kernels (-m gpu test) to them.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

import torch  # noqa: E402

import witness_inputs as wi  # noqa: E402


def import_upstream(summary):
    try:
        import vllm  # noqa: F401
        import vllm._C  # noqa: F401

        summary["vllm_version"] = getattr(vllm, "__version__", "?")
        summary["import_vllm_C"] = "ok"
        for name in ("paged_attention_v1", "paged_attention_v2"):
            summary["has_" + name] = hasattr(torch.ops._C, name)
        for name in ("reshape_and_cache", "reshape_and_cache_flash", "convert_fp8", "concat_and_cache_mla",
                     "gather_and_maybe_dequant_cache"):
            summary["has_" + name] = hasattr(torch.ops._C_cache_ops, name)
        return True
    except Exception as e:  # recorded verbatim for BASELINE.md
        summary["import_vllm_C"] = "failed: %s: %s" % (type(e).__name__, e)
        summary["import_traceback"] = traceback.format_exc()[-2000:]
        return False


def convert_witness(dev):
    C = torch.ops._C_cache_ops
    rec = {"quant_scales": list(wi.QUANT_SCALES), "dequant_scales": list(wi.DEQUANT_SCALES)}
    srcs = {"bf16": wi.all_bf16_patterns(), "f16": wi.all_f16_patterns(), "f32": wi.tricky_f32(), "f32_special": wi.special_f32()}
    for name, src in srcs.items():
        for sc in wi.QUANT_SCALES:
            dst = torch.zeros(src.shape, dtype=torch.uint8, device=dev)
            C.convert_fp8(dst, src.to(dev), float(sc), "fp8")
            torch.cuda.synchronize()
            rec["quant_%s_%r" % (name, sc)] = dst.cpu()
    for sc in wi.QUANT_SCALES:
        src = wi.tricky_f32_scaled(sc)
        dst = torch.zeros(src.shape, dtype=torch.uint8, device=dev)
        C.convert_fp8(dst, src.to(dev), float(sc), "fp8")
        torch.cuda.synchronize()
        rec["quant_f32_scaled_%r" % sc] = dst.cpu()
    codes = wi.all_bytes()
    for dname, dt in (("bf16", torch.bfloat16), ("f16", torch.float16), ("f32", torch.float32)):
        for sc in wi.DEQUANT_SCALES:
            dst = torch.zeros(codes.shape, dtype=dt, device=dev)
            C.convert_fp8(dst, codes.to(dev), float(sc), "fp8")
            torch.cuda.synchronize()
            rec["dequant_%s_%r" % (dname, sc)] = dst.cpu()
    return rec


def cache_witness(dev):
    C = torch.ops._C_cache_ops
    rec = {}
    for name in wi.CACHE_CASES_V1:
        c = wi.cache_case(name)
        x = 16
        kc = torch.zeros(c["NB"], c["Hkv"], c["D"] // x, c["B"], x, dtype=torch.uint8, device=dev)
        vc = torch.zeros(c["NB"], c["Hkv"], c["D"], c["B"], dtype=torch.uint8, device=dev)
        ks = torch.tensor([c["k_scale"]], dtype=torch.float32, device=dev)
        vs = torch.tensor([c["v_scale"]], dtype=torch.float32, device=dev)
        C.reshape_and_cache(c["key"].to(dev), c["value"].to(dev), kc, vc, c["slot_mapping"].to(dev), "fp8", ks, vs)
        torch.cuda.synchronize()
        rec[name] = {"key_cache": kc.cpu(), "value_cache": vc.cpu()}
    for name in wi.CACHE_CASES_FLASH:
        c = wi.cache_case(name)
        kc = torch.zeros(c["NB"], c["B"], c["Hkv"], c["D"], dtype=torch.uint8, device=dev)
        vc = torch.zeros_like(kc)
        ks = torch.tensor([c["k_scale"]], dtype=torch.float32, device=dev)
        vs = torch.tensor([c["v_scale"]], dtype=torch.float32, device=dev)
        C.reshape_and_cache_flash(c["key"].to(dev), c["value"].to(dev), kc, vc, c["slot_mapping"].to(dev), "fp8", ks, vs)
        torch.cuda.synchronize()
        rec[name] = {"key_cache": kc.cpu(), "value_cache": vc.cpu()}
    try:
        m = wi.mla_case()
        cache = torch.zeros(m["NB"], m["B"], m["LORA"] + m["ROPE"], dtype=torch.uint8, device=dev)
        sc = torch.tensor([m["scale"]], dtype=torch.float32, device=dev)
        C.concat_and_cache_mla(m["kv_c"].to(dev), m["k_pe"].to(dev), cache, m["slot_mapping"].to(dev), "fp8", sc)
        torch.cuda.synchronize()
        rec["mla"] = {"kv_cache": cache.cpu()}
    except Exception as e:
        rec["mla_error"] = "%s: %s" % (type(e).__name__, e)
    return rec


def run_attention(c, dev, version):
    A = torch.ops._C
    q = c["query"].to(dev)
    kc, vc = c["key_cache"].to(dev), c["value_cache"].to(dev)
    bt, lens = c["block_tables"].to(dev), c["seq_lens"].to(dev)
    ks = torch.tensor([c["k_scale"]], dtype=torch.float32, device=dev)
    vs = torch.tensor([c["v_scale"]], dtype=torch.float32, device=dev)
    out = torch.zeros_like(q)
    L = int(c["max_len"])
    if version == 1:
        A.paged_attention_v1(out, q, kc, vc, c["Hkv"], c["scale"], bt, lens, c["B"], L, None, c["kv"], ks, vs, 0, 0, 0, 64, 0)
    else:
        P = (L + 511) // 512
        es = torch.empty(c["S"], c["H"], P, dtype=torch.float32, device=dev)
        ml = torch.empty_like(es)
        tmp = torch.empty(c["S"], c["H"], P, c["D"], dtype=q.dtype, device=dev)
        A.paged_attention_v2(out, es, ml, tmp, q, kc, vc, c["Hkv"], c["scale"], bt, lens, c["B"], L, None, c["kv"], ks, vs,
                             0, 0, 0, 64, 0)
    torch.cuda.synchronize()
    return out.cpu()


def attn_witness(dev):
    rec = {}
    for name in wi.ATTN_CASES:
        c = wi.attn_case(name)
        rec[name] = {"v1": run_attention(c, dev, 1), "v2": run_attention(c, dev, 2)}
    return rec


# ---- timing of the stock kernels on the named configs ------------------------------------------------------
CONFIGS = {
    # name: (S, H, Hkv, D, B, L, kv, version)
    "c1": (8, 32, 32, 128, 16, 1024, "auto", 1),
    "c2": (64, 32, 8, 128, 16, 4096, "auto", 1),
    "c2_v2": (64, 32, 8, 128, 16, 4096, "auto", 2),
    "c3": (64, 32, 8, 128, 16, 4096, "fp8", 1),
    "c3_v2": (64, 32, 8, 128, 16, 4096, "fp8", 2),
    "c4_tp8": (128, 8, 1, 128, 16, 8192, "auto", 1),
    "c4_tp8_fp8": (128, 8, 1, 128, 16, 8192, "fp8", 1),
    "c5_bf16": (8, 32, 8, 128, 16, 131072, "auto", 2),
    "c5_fp8": (8, 32, 8, 128, 16, 131072, "fp8", 2),
}


def time_config(name, dev, reps=10):
    S, H, Hkv, D, B, L, kv, version = CONFIGS[name]
    A = torch.ops._C
    g = torch.Generator(device=dev).manual_seed(0)
    fp8 = kv != "auto"
    x = 16 if fp8 else 8
    nblk = (L + B - 1) // B
    NB = int(S * nblk * 1.25)
    sc = D ** -0.5

    def uni(shape):
        return (torch.rand(shape, device=dev, generator=g) * 2 - 1) * sc

    if fp8:
        kc = uni((NB, Hkv, D // x, B, x)).to(torch.float8_e4m3fn).view(torch.uint8)
        vc = uni((NB, Hkv, D, B)).to(torch.float8_e4m3fn).view(torch.uint8)
    else:
        kc = uni((NB, Hkv, D // x, B, x)).to(torch.bfloat16)
        vc = uni((NB, Hkv, D, B)).to(torch.bfloat16)
    q = uni((S, H, D)).to(torch.bfloat16)
    bt = torch.randperm(NB, device=dev, generator=g)[: S * nblk].view(S, nblk).to(torch.int32)
    lens = torch.full((S,), L, dtype=torch.int32, device=dev)
    one = torch.ones(1, dtype=torch.float32, device=dev)
    out = torch.empty_like(q)
    P = (L + 511) // 512
    if version == 2:
        es = torch.empty(S, H, P, dtype=torch.float32, device=dev)
        ml = torch.empty_like(es)
        tmp = torch.empty(S, H, P, D, dtype=q.dtype, device=dev)

    def call():
        if version == 1:
            A.paged_attention_v1(out, q, kc, vc, Hkv, sc, bt, lens, B, L, None, kv, one, one, 0, 0, 0, 64, 0)
        else:
            A.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, sc, bt, lens, B, L, None, kv, one, one, 0, 0, 0, 64, 0)

    for _ in range(3):
        call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    eb = 1 if fp8 else 2
    nbytes = S * L * Hkv * D * 2 * eb + S * H * D * 4 + S * nblk * 4 + 4 * S
    return {"ms": ms, "algorithmic_bytes": nbytes, "gbs": nbytes / (ms * 1e-3) / 1e9, "reps": reps,
            "op": "paged_attention_v%d" % version, "shape": dict(S=S, H=H, Hkv=Hkv, D=D, B=B, L=L, kv=kv)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "witness"))
    ap.add_argument("--no-timing", action="store_true")
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    summary = {"torch": torch.__version__, "cuda_available": torch.cuda.is_available(), "when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime())}
    if torch.cuda.is_available():
        summary["gpu"] = torch.cuda.get_device_name(0)
    ok = torch.cuda.is_available() and import_upstream(summary)
    if ok:
        dev = torch.device("cuda", 0)
        for name, fn in (("convert", convert_witness), ("cache", cache_witness), ("attn", attn_witness)):
            try:
                rec = fn(dev)
                torch.save(rec, os.path.join(args.out, "fp8_upstream_%s.pt" % name))
                summary["witness_" + name] = "ok (%d entries)" % len(rec)
            except Exception as e:
                summary["witness_" + name] = "failed: %s: %s" % (type(e).__name__, e)
                summary["witness_%s_traceback" % name] = traceback.format_exc()[-1500:]
        if not args.no_timing:
            summary["stock_kernel_timing"] = {}
            for name in CONFIGS:
                try:
                    summary["stock_kernel_timing"][name] = time_config(name, dev)
                except Exception as e:
                    summary["stock_kernel_timing"][name] = {"error": "%s: %s" % (type(e).__name__, e)}
                torch.cuda.empty_cache()
    with open(os.path.join(args.out, "summary.json"), "w") as f:
        json.dump(summary, f, indent=1)
    print(json.dumps({k: v for k, v in summary.items() if "traceback" not in k}, indent=1))


if __name__ == "__main__":
    main()
