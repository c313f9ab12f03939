"""FP8-e4m3 byte oracle: known-answer anchors and cross-checks between three independent encoders
(torch cast with pre-clamp, integer bit arithmetic in numpy, scalar C).  CPU only.

The reference never pins FP8 bytes (its conversion is a stub and its tests compare dequantised values with
rtol 0.1: tests/kernels/attention/test_cache.py:202-210), so these anchors define the byte contract."""
import numpy as np
import pytest
import torch

from oracle import c_oracle, e4m3_np
from oracle import torch_oracle as orc


def _codes():
    c = np.arange(256, dtype=np.uint8)
    return c[(c & 0x7F) != 0x7F]


def test_decode_table_known_answers():
    d = e4m3_np.decode(np.array([0x00, 0x01, 0x07, 0x08, 0x38, 0x40, 0x7E, 0x80, 0xFE], dtype=np.uint8))
    assert d.tolist() == [0.0, 2.0 ** -9, 7 * 2.0 ** -9, 2.0 ** -6, 1.0, 2.0, 448.0, -0.0, -448.0]
    t = torch.from_numpy(_codes()).view(torch.float8_e4m3fn).float().numpy()
    assert np.array_equal(e4m3_np.decode(_codes()), t)


def test_all_codes_roundtrip_in_every_encoder():
    codes = _codes()
    vals = e4m3_np.decode(codes)
    assert np.array_equal(e4m3_np.encode(vals), codes)
    assert np.array_equal(orc.fp8_e4m3_quant(torch.from_numpy(vals), 1.0).numpy(), codes)
    lib = c_oracle.lib()
    assert [lib.ref_f32_to_e4m3(float(v)) for v in vals] == codes.tolist()


def test_saturation_ties_and_zero():
    x = torch.tensor([448.0, 449.0, 464.0, 1e9, float("inf"), -1e9, 0.0, -0.0, 2.0 ** -10, 2.0 ** -10 * 1.0001,
                      17.0, 19.0, 21.0, 23.0, 0.0009765625 * 3])
    want = [0x7E, 0x7E, 0x7E, 0x7E, 0x7E, 0xFE, 0x00, 0x80, 0x00, 0x01, 0x58, 0x5A, 0x5A, 0x5C, 0x02]
    # 17 -> 16 (tie to even mantissa 0), 19 -> 20, 21 -> 20, 23 -> 24; 1.5 * 2^-9 -> 2 * 2^-9 (tie to even)
    assert orc.fp8_e4m3_quant(x, 1.0).tolist() == want
    assert e4m3_np.encode(x.numpy()).tolist() == want
    lib = c_oracle.lib()
    assert [lib.ref_f32_to_e4m3(float(v)) for v in x] == want


@pytest.mark.parametrize("scale", [1.0, 0.25, 0.0173, 3.7])
def test_three_encoders_agree_on_random_and_midpoint_inputs(scale):
    rng = np.random.default_rng(0)
    vals = np.sort(e4m3_np.decode(_codes()))
    mids = ((vals[:-1].astype(np.float64) + vals[1:]) / 2).astype(np.float32)
    x = np.concatenate([rng.uniform(-500, 500, 20000), rng.normal(0, 1, 20000), rng.normal(0, 0.01, 20000), mids,
                        np.nextafter(mids, np.float32(1e9)), np.nextafter(mids, np.float32(-1e9))]).astype(np.float32)
    xs = (x * np.float32(scale)).astype(np.float32)
    a = orc.fp8_e4m3_quant(torch.from_numpy(xs), scale).numpy()
    b = e4m3_np.encode((xs / np.float32(scale)).astype(np.float32))
    assert np.array_equal(a, b)
    dst = torch.empty(xs.shape, dtype=torch.uint8)
    c_oracle.convert_fp8(dst, torch.from_numpy(xs), scale)
    assert np.array_equal(dst.numpy(), a)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_dequant_then_quant_is_identity_for_power_of_two_scales(dtype):
    codes = torch.from_numpy(_codes())
    for scale in (1.0, 0.5, 4.0):
        f = orc.fp8_e4m3_dequant(codes, scale, dtype)
        assert torch.equal(orc.fp8_e4m3_quant(f, scale), codes)
