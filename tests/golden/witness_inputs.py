"""Deterministic inputs shared by tools/upstream_witness.py (run on the GPU box against UPSTREAM vLLM's compiled
kernels) and by the tests that replay its recorded outputs (tests/golden/fp8_upstream_*.pt).

Everything is generated on the CPU from fixed seeds or enumerated bit patterns, so both sides see identical bytes
without shipping the inputs.  Only torch is imported.
"""
from __future__ import annotations

import struct

import torch

QUANT_SCALES = (1.0, 0.5, 0.0123, 3.7, 1.0 / 3.0, 448.0)
DEQUANT_SCALES = (1.0, 0.5, 0.0123, 3.7)


def all_bf16_patterns() -> torch.Tensor:
    """Every bf16 bit pattern as a [1024, 64] bf16 tensor (NaNs and infinities included)."""
    bits = torch.arange(65536, dtype=torch.int32).to(torch.int16)
    return bits.view(torch.bfloat16).reshape(1024, 64)


def all_f16_patterns() -> torch.Tensor:
    bits = torch.arange(65536, dtype=torch.int32).to(torch.int16)
    return bits.view(torch.float16).reshape(1024, 64)


def _f32_bits(x: float) -> int:
    return struct.unpack("<I", struct.pack("<f", x))[0]


def tricky_f32() -> torch.Tensor:
    """fp32 inputs around every e4m3 rounding boundary: each e4m3 value, each midpoint between neighbours, one fp32
    ulp either side of both, +-448 .. +-480 (the saturating range), large values, signed zeros, the e4m3 subnormal
    range and fp32 subnormals.  Shape [N/64, 64] (padded with zeros)."""
    codes = torch.arange(256, dtype=torch.int32).to(torch.uint8).view(torch.float8_e4m3fn).to(torch.float32)
    vals = sorted(set(float(v) for v in codes.tolist() if v == v and v >= 0))
    pts = []
    for a, b in zip(vals, vals[1:]):
        pts += [a, (a + b) / 2, b]
    pts += [448.0, 456.0, 463.9, 464.0, 464.1, 472.0, 479.9, 480.0, 512.0, 6.0e4, 3.0e38, 1e-3, 2.0 ** -10, 2.0 ** -11,
            1.5 * 2.0 ** -10, 2.5 * 2.0 ** -10, 1e-40, 0.0]
    bits = set()
    for p in pts:
        b = _f32_bits(p)
        for d in (-2, -1, 0, 1, 2):
            if 0 <= b + d < 0x7F800000:
                bits.add(b + d)
    bits = sorted(bits)
    allb = bits + [b | 0x80000000 for b in bits]
    while len(allb) % 64:
        allb.append(0)
    t = torch.tensor(allb, dtype=torch.int64).to(torch.int32)  # wraps the sign bit
    return t.view(torch.float32).reshape(-1, 64)


def tricky_f32_scaled(scale: float) -> torch.Tensor:
    """The rounding-boundary values times `scale` (fp32), and their fp32 neighbours: x / scale lands on or next to
    an e4m3 rounding boundary, which tells an IEEE divide from a multiplication by the reciprocal."""
    t = tricky_f32().reshape(-1) * torch.tensor(scale, dtype=torch.float32)
    up = torch.nextafter(t, torch.full_like(t, float("inf")))
    dn = torch.nextafter(t, torch.full_like(t, float("-inf")))
    return torch.cat([t, up, dn]).reshape(-1, 64)


def special_f32() -> torch.Tensor:
    """inf / nan inputs (recorded, not asserted: the reference never defines them)."""
    t = torch.zeros(1, 64, dtype=torch.float32)
    t[0, 0], t[0, 1], t[0, 2] = float("inf"), float("-inf"), float("nan")
    return t


def all_bytes() -> torch.Tensor:
    """Every e4m3 code, [4, 64] uint8."""
    return torch.arange(256, dtype=torch.int32).to(torch.uint8).reshape(4, 64)


def cache_case(name: str):
    """Inputs of one reshape_and_cache / reshape_and_cache_flash / concat_and_cache_mla case."""
    cases = {
        # name: (T, Hkv, D, B, NB, k_scale, v_scale, seed)
        "v1_d64_b16": (42, 8, 64, 16, 32, 0.0123, 0.037, 11),
        "v1_d128_b32": (21, 4, 128, 32, 16, 0.5, 2.0, 12),
        "flash_d128_b16": (21, 8, 128, 16, 32, 0.0123, 0.037, 13),
        "flash_d64_b32": (33, 4, 64, 32, 16, 0.021, 0.3, 14),
    }
    T, Hkv, D, B, NB, ks, vs, seed = cases[name]
    g = torch.Generator().manual_seed(seed)
    key = (torch.randn(T, Hkv, D, generator=g) * 0.7).to(torch.bfloat16)
    value = (torch.randn(T, Hkv, D, generator=g) * 2.0).to(torch.bfloat16)
    # a few values that saturate and a few exact zeros
    key[0, 0, :4] = torch.tensor([1000.0, -1000.0, 0.0, -0.0], dtype=torch.bfloat16)
    value[1, 0, :4] = torch.tensor([1e4, -1e4, 0.0, 448.0 * vs], dtype=torch.bfloat16)
    slots = torch.randperm(NB * B, generator=g)[:T].to(torch.int64)
    slots[3] = -1  # padding token
    return dict(T=T, Hkv=Hkv, D=D, B=B, NB=NB, k_scale=ks, v_scale=vs, key=key, value=value, slot_mapping=slots)


CACHE_CASES_V1 = ("v1_d64_b16", "v1_d128_b32")
CACHE_CASES_FLASH = ("flash_d128_b16", "flash_d64_b32")


def mla_case():
    T, LORA, ROPE, B, NB, scale, seed = 19, 512, 64, 16, 8, 0.031, 21
    g = torch.Generator().manual_seed(seed)
    kv_c = (torch.randn(T, LORA, generator=g) * 1.3).to(torch.bfloat16)
    k_pe = (torch.randn(T, ROPE, generator=g) * 1.3).to(torch.bfloat16)
    slots = torch.randperm(NB * B, generator=g)[:T].to(torch.int64)
    slots[5] = -1
    return dict(T=T, LORA=LORA, ROPE=ROPE, B=B, NB=NB, scale=scale, kv_c=kv_c, k_pe=k_pe, slot_mapping=slots)


ATTN_CASES = {
    # name: (S, H, Hkv, D, B, max_len, NB, kv, k_scale, v_scale, seed)
    "attn_fp8_gqa4_scaled": (7, 16, 4, 128, 16, 600, 300, "fp8", 0.0123, 0.037, 31),
    "attn_fp8_gqa8_b32": (5, 16, 2, 128, 32, 900, 160, "fp8", 0.5, 2.0, 32),
    "attn_bf16_gqa4": (7, 16, 4, 128, 16, 600, 300, "auto", 1.0, 1.0, 33),
    "attn_fp8_mha_d64": (4, 8, 8, 64, 16, 300, 90, "fp8", 0.02, 0.01, 34),
}


def attn_case(name: str):
    """A decode batch over a v1/v2-layout cache; fp8 caches are produced by the oracle-independent route of casting
    uniform values in the e4m3 range to float8 with torch (bytes only: nothing here quantises with a scale)."""
    S, H, Hkv, D, B, max_len, NB, kv, ks, vs, seed = ATTN_CASES[name]
    g = torch.Generator().manual_seed(seed)
    fp8 = kv != "auto"
    x = 16 if fp8 else 8
    if fp8:
        # bytes whose dequantised magnitude is ~ D^-0.5: value / scale in e4m3
        kc = ((torch.rand(NB, Hkv, D // x, B, x, generator=g) * 2 - 1) * (D ** -0.5) / ks).to(torch.float8_e4m3fn).view(torch.uint8)
        vc = ((torch.rand(NB, Hkv, D, B, generator=g) * 2 - 1) * (D ** -0.5) / vs).to(torch.float8_e4m3fn).view(torch.uint8)
    else:
        kc = ((torch.rand(NB, Hkv, D // x, B, x, generator=g) * 2 - 1) * (D ** -0.5)).to(torch.bfloat16)
        vc = ((torch.rand(NB, Hkv, D, B, generator=g) * 2 - 1) * (D ** -0.5)).to(torch.bfloat16)
    q = ((torch.rand(S, H, D, generator=g) * 2 - 1) * (D ** -0.5)).to(torch.bfloat16)
    lens = torch.randint(1, max_len + 1, (S,), generator=g, dtype=torch.int32)
    lens[-1] = max_len
    nblk = (max_len + B - 1) // B
    bt = torch.randint(0, NB, (S, nblk), generator=g, dtype=torch.int32)
    return dict(S=S, H=H, Hkv=Hkv, D=D, B=B, max_len=max_len, NB=NB, kv=kv, k_scale=ks, v_scale=vs, scale=D ** -0.5,
                query=q, key_cache=kc, value_cache=vc, seq_lens=lens, block_tables=bt)
