"""Parity of reshape_and_cache / reshape_and_cache_flash / convert_fp8 (CUDA) against the CPU oracle.

Grid and inputs follow the reference (tests/kernels/attention/test_cache.py:121-350,434-465); the bar is
stricter than the reference's (which compares dequantised fp8 values with rtol 0.1): every cache byte,
including FP8-e4m3 bytes, must equal the oracle's (BASELINE.json: "bit-exact for the slot_mapping cache
writes and FP8 cache bytes").
"""
import random

import pytest
import torch

from oracle import torch_oracle as orc

pytestmark = pytest.mark.gpu


def _inputs(T, H, D, dtype, seed, padded=0):
    g = torch.Generator().manual_seed(seed)
    qkv = torch.randn(T + padded, 3, H, D, generator=g, dtype=torch.float32).to(dtype)
    _, key, value = qkv.unbind(dim=1)  # strided views, as in the reference test
    return qkv, key, value


def _scales(key, value):
    # reference: tests/kernels/attention/test_cache.py:163-164
    return (key.amax() / 64.0).to(torch.float32), (value.amax() / 64.0).to(torch.float32)


@pytest.mark.parametrize("head_size", [64, 80, 128, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("num_blocks", [64, 1531])  # reference: 1024 / 10000
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
def test_reshape_and_cache(ops, head_size, block_size, num_blocks, dtype, kv_cache_dtype):
    if kv_cache_dtype == "fp8" and head_size % 16:
        pytest.skip()
    T, H = 42, 8
    random.seed(0)
    slots = random.sample(range(block_size * num_blocks), T)
    slots[5] = -1  # padding token: must be skipped (csrc/cache_kernels.cu:217-220)
    slot_mapping = torch.tensor(slots, dtype=torch.long)
    qkv, key, value = _inputs(T, H, head_size, dtype, 0)
    kc, vc = orc.make_kv_caches(num_blocks, block_size, H, head_size, kv_cache_dtype, dtype, seed=0)
    k_scale, v_scale = _scales(key, value)

    ref_k, ref_v = kc.clone(), vc.clone()
    orc.reshape_and_cache(key, value, ref_k, ref_v, slot_mapping, kv_cache_dtype, k_scale, v_scale)

    qkv_d = qkv.cuda()
    _, key_d, value_d = qkv_d.unbind(dim=1)
    kc_d, vc_d = kc.cuda(), vc.cuda()
    ops.reshape_and_cache(key_d, value_d, kc_d, vc_d, slot_mapping.cuda(), kv_cache_dtype, k_scale.cuda(),
                          v_scale.cuda())
    torch.cuda.synchronize()
    assert torch.equal(kc_d.cpu().view(torch.uint8), ref_k.view(torch.uint8))
    assert torch.equal(vc_d.cpu().view(torch.uint8), ref_v.view(torch.uint8))


@pytest.mark.parametrize("head_size", [64, 80, 128, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("num_blocks", [32, 777])  # reference: 512 / 5000
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("layout", ["NHD", "HND"])
def test_reshape_and_cache_flash(ops, head_size, block_size, num_blocks, dtype, kv_cache_dtype, layout):
    T, H = 21, 8
    random.seed(0)
    slots = random.sample(range(block_size * num_blocks), T)
    slots[3] = -1
    slot_mapping = torch.tensor(slots, dtype=torch.long)
    # key has more rows than slot_mapping: CUDA-graph padding (csrc/cache_kernels.cu:459-469)
    qkv, key, value = _inputs(T, H, head_size, dtype, 1, padded=3)
    k_scale, v_scale = _scales(key, value)
    cache_dtype = dtype if kv_cache_dtype == "auto" else torch.uint8
    g = torch.Generator().manual_seed(2)
    # storage [2, NB, B, H, D] (NHD) or [2, NB, H, B, D] viewed as NHD (HND), as the reference backend lays
    # it out (vllm_metax/v1/attention/backends/flash_attn.py:99-121)
    shape = (2, num_blocks, block_size, H, head_size) if layout == "NHD" else (2, num_blocks, H, block_size, head_size)
    store = torch.rand(shape, generator=g, dtype=torch.float32) - 0.5
    store = store.to(dtype) if kv_cache_dtype == "auto" else orc.fp8_e4m3_quant(store, 1.0)
    assert store.dtype == cache_dtype

    def views(st):
        v = st if layout == "NHD" else st.permute(0, 1, 3, 2, 4)
        return v[0], v[1]

    ref = store.clone()
    rk, rv = views(ref)
    orc.reshape_and_cache_flash(key, value, rk, rv, slot_mapping, kv_cache_dtype, k_scale, v_scale)

    dev = store.cuda()
    dk, dv = views(dev)
    qkv_d = qkv.cuda()
    _, key_d, value_d = qkv_d.unbind(dim=1)
    ops.reshape_and_cache_flash(key_d, value_d, dk, dv, slot_mapping.cuda(), kv_cache_dtype, k_scale.cuda(),
                                v_scale.cuda())
    torch.cuda.synchronize()
    assert torch.equal(dev.cpu().view(torch.uint8), ref.view(torch.uint8))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("scale", [1.0, 0.03125, 0.0173])
def test_convert_fp8_bytes_and_roundtrip(ops, dtype, scale):
    """Quantise: bytes equal the oracle's.  Dequantise: values equal the oracle's.  Round trip within the
    reference's tolerance (tests/kernels/attention/test_cache.py:442-465: U(-224,224), atol 1e-3, rtol 0.1)."""
    g = torch.Generator().manual_seed(0)
    x = ((torch.rand(257, 8, 64, 16, generator=g) * 2 - 1) * 224.0 * scale).to(dtype)
    # saturation, ties, zeros, tiny values
    x.view(-1)[:8] = torch.tensor([6e4, -6e4, 448.0, -448.0, 0.0, -0.0, 1e-9, 464.0], dtype=torch.float32).to(dtype) * scale
    xd = x.cuda()
    q = torch.empty_like(xd, dtype=torch.uint8)
    ops.convert_fp8(q, xd, scale, "fp8")
    ref_q = orc.fp8_e4m3_quant(x, scale)
    assert torch.equal(q.cpu(), ref_q)
    back = torch.empty_like(xd)
    ops.convert_fp8(back, q, scale, "fp8_e4m3")
    ref_back = orc.fp8_e4m3_dequant(ref_q, scale, dtype)
    assert torch.equal(back.cpu().view(torch.uint8), ref_back.view(torch.uint8))
    inside = x.float().abs() <= 448.0 * scale
    torch.testing.assert_close(back.cpu().float()[inside], x.float()[inside], atol=1e-3, rtol=0.1)


def test_all_e4m3_codes_roundtrip(ops):
    """Known-answer anchor: every non-NaN e4m3 code dequantises and re-quantises to itself."""
    codes = torch.arange(256, dtype=torch.uint8)
    codes = codes[(codes & 0x7F) != 0x7F]
    f = torch.empty(codes.numel(), dtype=torch.float32, device="cuda")
    ops.convert_fp8(f, codes.cuda(), 1.0, "fp8")
    assert torch.equal(f.cpu(), codes.view(torch.float8_e4m3fn).float())
    q = torch.empty(codes.numel(), dtype=torch.uint8, device="cuda")
    ops.convert_fp8(q, f, 1.0, "fp8")
    # -0.0 (0x80) keeps its sign bit
    assert torch.equal(q.cpu(), codes)


def test_cache_write_then_attend(ops):
    """End to end on the path: write K/V through reshape_and_cache (fp8), attend with paged_attention_v1,
    compare with the oracle run on the oracle-written cache."""
    S, H, Hkv, D, B, L = 4, 32, 8, 128, 16, 200
    g = torch.Generator().manual_seed(5)
    nb = 256
    kc = torch.zeros(nb, Hkv, D // 16, B, 16, dtype=torch.uint8)
    vc = torch.zeros(nb, Hkv, D, B, dtype=torch.uint8)
    bt = torch.randperm(nb, generator=g)[: S * ((L + B - 1) // B)].view(S, -1).to(torch.int32)
    key = torch.randn(S * L, Hkv, D, generator=g).to(torch.bfloat16)
    value = torch.randn(S * L, Hkv, D, generator=g).to(torch.bfloat16)
    slots = torch.empty(S * L, dtype=torch.long)
    for s in range(S):
        for t in range(L):
            slots[s * L + t] = int(bt[s, t // B]) * B + t % B
    ks, vs = torch.tensor(0.02), torch.tensor(0.03)
    rk, rv = kc.clone(), vc.clone()
    orc.reshape_and_cache(key, value, rk, rv, slots, "fp8", ks, vs)
    q = torch.randn(S, H, D, generator=g).to(torch.bfloat16)
    lens = torch.full((S,), L, dtype=torch.int32)
    ref = orc.paged_attention(q, rk, rv, bt, lens, D ** -0.5, None, "fp8", ks, vs)

    kc_d, vc_d = kc.cuda(), vc.cuda()
    ops.reshape_and_cache(key.cuda(), value.cuda(), kc_d, vc_d, slots.cuda(), "fp8", ks.cuda(), vs.cuda())
    assert torch.equal(kc_d.cpu(), rk) and torch.equal(vc_d.cpu(), rv)
    out = torch.empty_like(q, device="cuda")
    ops.paged_attention_v1(out, q.cuda(), kc_d, vc_d, Hkv, D ** -0.5, bt.cuda(), lens.cuda(), B, L, None, "fp8",
                           ks.cuda(), vs.cuda())
    # N(0,1) data through e4m3: the oracle rounds every dequantised element to bf16, the kernel folds the
    # scales; the reference tolerates 1e-2 for fp8 (tests/kernels/attention/test_attention.py:343-345)
    torch.testing.assert_close(out.cpu().float(), ref.float(), atol=5e-3, rtol=2e-2)
