"""Parity of rotary_embedding and the fused rotary_embedding_and_cache_flash (CUDA) against the CPU oracle.

Grid follows tests/kernels/core/test_pos_encoding.py:14-60 (NeoX / GPT-J, half / bf16 / float, head sizes incl. 80, 112,
120, partial rotary dim, flat / batched / padded-head tensor shapes, with and without key).  The oracle is the
PyTorch-native implementation that test compares with; the bar here is bit-exact (the reference uses
get_default_atol/rtol), because the kernel rounds where the native implementation rounds.
"""
import glob
import os
import random

import pytest
import torch

from oracle import torch_oracle as orc

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("shape_kind", ["batch", "flat", "padded"])
@pytest.mark.parametrize("seq_len", [11, 1024])
@pytest.mark.parametrize("head_size", [64, 80, 112, 120, 128, 256])
@pytest.mark.parametrize("rotary_dim", [None, 32])
@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16, torch.float])
@pytest.mark.parametrize("use_key", [True, False])
def test_rotary_embedding(ops, is_neox, shape_kind, seq_len, head_size, rotary_dim, dtype, use_key):
    if seq_len == 1024 and (head_size not in (80, 128) or shape_kind == "flat"):
        pytest.skip("large case on two head sizes only")
    batch, H = 5, 17
    R = head_size if rotary_dim is None else rotary_dim
    g = torch.Generator().manual_seed(0)
    cache = orc.compute_cos_sin_cache(R, 8192, dtype=dtype)
    positions = torch.randint(0, 8192, (batch, seq_len), generator=g)
    if shape_kind == "flat":
        q = torch.randn(batch, seq_len, H * head_size, generator=g).to(dtype)
    elif shape_kind == "batch":
        q = torch.randn(batch, seq_len, H, head_size, generator=g).to(dtype)
    else:  # sliced heads: head stride = head_size + 64 (test_pos_encoding.py:33-36,94-96)
        q = torch.randn(batch, seq_len, H, head_size + 64, generator=g).to(dtype)
    k = torch.randn(q.shape, generator=g).to(dtype) if use_key else None
    qd, kd = q.cuda(), (k.cuda() if use_key else None)
    if shape_kind == "padded":
        q, qd = q[..., :head_size], qd[..., :head_size]
        if use_key:
            k, kd = k[..., :head_size], kd[..., :head_size]
    ref_q, ref_k = orc.rotary_embedding(positions, q, k, head_size, cache, is_neox)
    ops.rotary_embedding(positions.cuda(), qd, kd, head_size, cache.cuda(), is_neox)
    assert torch.equal(qd.cpu(), ref_q)
    if use_key:
        assert torch.equal(kd.cpu(), ref_k)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "rotary_*.pt"))),
                         ids=lambda p: os.path.basename(p)[:-3])
def test_rotary_embedding_golden(ops, path):
    """Against upstream's RotaryEmbedding.forward_static output (tests/golden/make_golden.py)."""
    d = torch.load(path)
    q, k = d["query"].cuda(), d["key"].cuda()
    ops.rotary_embedding(d["positions"].cuda(), q, k, d["head_size"], d["cos_sin_cache"].cuda(), d["is_neox"])
    assert torch.equal(q.cpu(), d["ref_query"]) and torch.equal(k.cpu(), d["ref_key"])


@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("cfg", [  # tokens, H, Hkv, D, rot_dim, dtype
    (64, 32, 8, 128, 128, torch.bfloat16),
    (37, 8, 2, 80, 32, torch.float16),
    (200, 4, 4, 64, 64, torch.float32),
])
def test_rotary_embedding_and_cache_flash(ops, cfg, is_neox, layout, kv_cache_dtype):
    """The fused op = rotary_embedding then reshape_and_cache_flash, byte for byte (query, key and both caches)."""
    T, H, Hkv, D, R, dtype = cfg
    if kv_cache_dtype == "fp8" and D % 16:
        pytest.skip()
    B, NB, pad = 16, 40, 5
    g = torch.Generator().manual_seed(1)
    cache = orc.compute_cos_sin_cache(R, 4096, dtype=dtype)
    positions = torch.randint(0, 4096, (T + pad,), generator=g)
    qkv = torch.randn(T + pad, (H + 2 * Hkv) * D, generator=g).to(dtype)
    q, k, v = qkv.split([H * D, Hkv * D, Hkv * D], dim=-1)  # strided views of one projection output, as in the model
    k3, v3 = k.view(T + pad, Hkv, D), v.view(T + pad, Hkv, D)
    random.seed(2)
    slots = random.sample(range(B * NB), T)
    slots[3] = -1
    slot_mapping = torch.tensor(slots, dtype=torch.long)  # shorter than key: CUDA-graph padding
    cdt = torch.uint8 if kv_cache_dtype == "fp8" else dtype
    shape = (NB, Hkv, B, D) if layout == "HND" else (NB, B, Hkv, D)
    kc = torch.randint(0, 100, shape, generator=g).to(cdt)
    vc = torch.randint(0, 100, shape, generator=g).to(cdt)
    if layout == "HND":
        kc, vc = kc.permute(0, 2, 1, 3), vc.permute(0, 2, 1, 3)
    ks, vs = torch.tensor(0.05), torch.tensor(0.03)
    # oracle: the two reference ops back to back
    ref_q, ref_k = orc.rotary_embedding(positions, q, k, D, cache, is_neox)
    rk, rv = kc.clone(), vc.clone()
    orc.reshape_and_cache_flash(ref_k.reshape(T + pad, Hkv, D), v3, rk, rv, slot_mapping, kv_cache_dtype, ks, vs)

    qkv_d = qkv.cuda()
    qd, kd, vd = qkv_d.split([H * D, Hkv * D, Hkv * D], dim=-1)
    kcd, vcd = kc.cuda(), vc.cuda()
    assert kcd.stride() == kc.stride()
    ops.rotary_embedding_and_cache_flash(positions.cuda(), qd, kd.view(T + pad, Hkv, D), vd.view(T + pad, Hkv, D), D,
                                         cache.cuda(), is_neox, kcd, vcd, slot_mapping.cuda(), kv_cache_dtype,
                                         ks.cuda(), vs.cuda())
    assert torch.equal(qd.cpu(), ref_q) and torch.equal(kd.cpu(), ref_k)
    assert torch.equal(kcd.cpu().view(torch.uint8), rk.view(torch.uint8))
    assert torch.equal(vcd.cpu().view(torch.uint8), rv.view(torch.uint8))
    assert torch.equal(vd.cpu(), v)  # value is read-only

    # and it equals our two separate ops
    qkv_d2 = qkv.cuda()
    q2, k2, v2 = qkv_d2.split([H * D, Hkv * D, Hkv * D], dim=-1)
    kc2, vc2 = kc.cuda(), vc.cuda()
    ops.rotary_embedding(positions.cuda(), q2, k2, D, cache.cuda(), is_neox)
    ops.reshape_and_cache_flash(k2.view(T + pad, Hkv, D), v2.view(T + pad, Hkv, D), kc2, vc2, slot_mapping.cuda(),
                                kv_cache_dtype, ks.cuda(), vs.cuda())
    assert torch.equal(qkv_d2, qkv_d) and torch.equal(kc2, kcd) and torch.equal(vc2, vcd)
