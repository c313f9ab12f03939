"""Developer micro-benchmark of the ops either side of the decode path: merge_attn_states, copy_blocks,
gather_and_maybe_dequant_cache (auto and fp8), cp_gather_cache.  CUDA events, median of --iters launches; GB/s counts
the algorithmic bytes (every byte read once + every byte written once)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_metax_b200 import _custom_ops as ops  # noqa: E402


def timeit(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for s, e in evs:
        s.record()
        fn()
        e.record()
    torch.cuda.synchronize()
    ts = sorted(s.elapsed_time(e) for s, e in evs)
    return ts[len(ts) // 2], ts[0]


def report(name, nbytes, med, best):
    print("%-58s %9.1f us  %7.0f GB/s (median)  %7.0f GB/s (best)  [%.1f MB]" % (
        name, med * 1e3, nbytes / med / 1e6, nbytes / best / 1e6, nbytes / 1e6))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=30)
    a = ap.parse_args()
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)

    # merge_attn_states: prefill-sized (8192 tokens) and decode-sized (64 tokens), Llama-3-8B heads
    for N in (8192, 64):
        H, D = 32, 128
        p = torch.randn(N, H, D, device=dev, generator=g).bfloat16()
        s = torch.randn(N, H, D, device=dev, generator=g).bfloat16()
        pl, sl = torch.randn(H, N, device=dev, generator=g), torch.randn(H, N, device=dev, generator=g)
        o, ol = torch.empty_like(p), torch.empty_like(pl)
        med, best = timeit(lambda: ops.merge_attn_states(o, p, pl, s, sl, ol), a.iters)
        report("merge_attn_states bf16 N=%d H=32 D=128" % N, 3 * p.numel() * 2 + 3 * pl.numel() * 4, med, best)

    # copy_blocks: 32 layers, 256 pairs, Llama-3-8B blocks (16 x 8 x 128 bf16 = 32 KiB per K or V block)
    layers, NB, pairs = 32, 1024, 256
    kcs = [torch.zeros(NB, 8, 16, 16, 8, dtype=torch.bfloat16, device=dev) for _ in range(layers)]
    vcs = [torch.zeros(NB, 8, 128, 16, dtype=torch.bfloat16, device=dev) for _ in range(layers)]
    perm = torch.randperm(NB, device=dev, generator=g)
    mapping = torch.stack([perm[:pairs], perm[pairs:2 * pairs]], dim=1).contiguous()
    med, best = timeit(lambda: ops.copy_blocks(kcs, vcs, mapping), a.iters)
    report("copy_blocks 32 layers x 256 pairs x 2 x 32 KiB", layers * pairs * 2 * 32768 * 2, med, best)
    del kcs, vcs

    # gathers: MLA-shaped entries (576), 64 sequences x 4096 tokens
    B, E, S, L = 16, 576, 64, 4096
    nblk = L // B
    NB = S * nblk + 64
    table = torch.randperm(NB, device=dev, generator=g)[: S * nblk].view(S, nblk).to(torch.int32)
    cu = (torch.arange(S + 1, device=dev) * L).to(torch.int32)
    scale = torch.full((1,), 0.1, device=dev)
    src16 = torch.randn(NB, B, E, device=dev, generator=g).bfloat16()
    dst16 = torch.empty(S * L, E, dtype=torch.bfloat16, device=dev)
    med, best = timeit(lambda: ops.gather_and_maybe_dequant_cache(src16, dst16, table, cu, S, "auto", scale, None), a.iters)
    report("gather_and_maybe_dequant_cache auto bf16 64x4096x576", 2 * dst16.numel() * 2, med, best)
    med, best = timeit(lambda: ops.cp_gather_cache(src16, dst16, table, cu, S), a.iters)
    report("cp_gather_cache bf16 64x4096x576", 2 * dst16.numel() * 2, med, best)
    src8 = torch.randint(0, 120, (NB, B, E), dtype=torch.uint8, device=dev, generator=g)
    med, best = timeit(lambda: ops.gather_and_maybe_dequant_cache(src8, dst16, table, cu, S, "fp8", scale, None), a.iters)
    report("gather_and_maybe_dequant_cache fp8 -> bf16 64x4096x576", dst16.numel() * 3, med, best)
    rotary_bench(a.iters)


def rotary_bench(iters):
    """rotary_embedding + reshape_and_cache_flash as two launches vs the fused op (Llama-3-8B heads, bf16)."""
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(1)
    H, Hkv, D, B = 32, 8, 128, 16
    for T in (64, 8192):
        NB = T // B + 64
        cache = torch.randn(8192, D, device=dev, generator=g).bfloat16()
        pos = torch.randint(0, 8192, (T,), device=dev, generator=g)
        qkv = torch.randn(T, (H + 2 * Hkv) * D, device=dev, generator=g).bfloat16()
        q, k, v = qkv.split([H * D, Hkv * D, Hkv * D], dim=-1)
        k3, v3 = k.view(T, Hkv, D), v.view(T, Hkv, D)
        kc = torch.zeros(NB, B, Hkv, D, dtype=torch.bfloat16, device=dev)
        vc = torch.zeros_like(kc)
        slots = torch.randperm(NB * B, device=dev, generator=g)[:T]
        one = torch.ones(1, device=dev)

        def two():
            ops.rotary_embedding(pos, q, k, D, cache, True)
            ops.reshape_and_cache_flash(k3, v3, kc, vc, slots, "auto", one, one)

        def fused():
            ops.rotary_embedding_and_cache_flash(pos, q, k3, v3, D, cache, True, kc, vc, slots, "auto", one, one)

        nbytes = T * (2 * H * D * 2 + 2 * Hkv * D * 2 + 2 * Hkv * D * 2 + 2 * Hkv * D * 2)  # q r/w, k r/w, v read, cache writes
        med, best = timeit(two, iters)
        report("rotary_embedding + reshape_and_cache_flash, T=%d (2 launches)" % T, nbytes + T * 2 * Hkv * D * 2, med, best)
        med, best = timeit(fused, iters)
        report("rotary_embedding_and_cache_flash, T=%d (1 launch)" % T, nbytes, med, best)


if __name__ == "__main__":
    main()
