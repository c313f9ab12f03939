"""Independent integer restatement of FP8-e4m3 (OCP "E4M3FN") encode / decode — TEST INFRASTRUCTURE.

Used only to cross-check the torch-based byte oracle (torch_oracle.fp8_e4m3_quant) with arithmetic that
shares no code with torch's cast: the encoder works on the fp32 bit pattern with explicit
round-to-nearest-even and saturate-to-finite, i.e. what the CUDA conversion
`__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3)` is specified to do and what the reference's
quantising call sites intend (csrc/cache_kernels.cu:249-252; the MetaX body is a stub,
csrc/quantization/fp8/metax/quant_utils.cuh:13-23).

Format: 1 sign, 4 exponent (bias 7), 3 mantissa bits; no infinities; 0x7F / 0xFF are NaN; largest finite
448 = 0x7E; subnormals m * 2^-9.
"""
from __future__ import annotations

import numpy as np


def decode(codes: np.ndarray) -> np.ndarray:
    """uint8 e4m3 -> float32 (exact)."""
    c = codes.astype(np.uint32)
    sign = np.where(c & 0x80, -1.0, 1.0).astype(np.float32)
    e = (c >> 3) & 0xF
    m = c & 0x7
    normal = (1.0 + m / 8.0) * np.exp2(e.astype(np.float64) - 7.0)
    sub = (m / 8.0) * 2.0 ** -6
    v = np.where(e == 0, sub, normal).astype(np.float32)
    v = np.where((e == 15) & (m == 7), np.float32(np.nan), v)
    return sign * v


def encode(x: np.ndarray) -> np.ndarray:
    """float32 -> uint8 e4m3, round-to-nearest-even, saturate to +-448, NaN -> 0x7F | sign."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    bits = x.view(np.uint32)
    sign = ((bits >> 24) & 0x80).astype(np.uint32)
    mag = bits & 0x7FFFFFFF
    exp = (mag >> 23).astype(np.int64) - 127  # unbiased fp32 exponent
    man = (mag & 0x7FFFFF).astype(np.int64) | 0x800000  # 24-bit significand with hidden one
    out = np.zeros(x.shape, dtype=np.uint32)

    isnan = mag > 0x7F800000
    # value = man * 2^(exp-23).  Target grid: multiples of 2^(max(exp,-6)-3).
    # shift = number of low bits of `man` dropped to land on that grid.
    qexp = np.maximum(exp, -6)
    shift = (qexp - 3) - (exp - 23)  # >= 20
    shift = np.minimum(shift, 40)  # everything below 2^-10 / 2 rounds to zero anyway
    keep = man >> shift
    rem = man & ((np.int64(1) << shift) - 1)
    half = np.int64(1) << (shift - 1)
    round_up = (rem > half) | ((rem == half) & ((keep & 1) == 1))
    keep = keep + round_up.astype(np.int64)  # in units of 2^(qexp-3); may reach 16 (carry)
    # rebuild the code: value = keep * 2^(qexp-3)
    #   subnormal range (qexp == -6): code = keep (keep <= 8; keep == 8 is the first normal, code 8)
    #   normal: keep in [8, 16]; keep == 16 -> exponent + 1, mantissa 0
    carry = keep >= 16
    e_field = np.where(carry, qexp + 1, qexp) + 7
    m_field = np.where(carry, 0, keep - 8)
    normal_code = (e_field << 3) | m_field
    code = np.where((qexp == -6) & (keep < 8), keep, normal_code)
    code = np.where(mag == 0, 0, code)
    # fp32 subnormals and anything tiny
    code = np.where(exp < -20, 0, code)
    # saturate (also +-inf)
    code = np.where((code > 0x7E) | (mag >= 0x7F800000), 0x7E, code)
    out = (code.astype(np.uint32) | sign)
    out = np.where(isnan, 0x7F | sign, out)
    return out.astype(np.uint8)
