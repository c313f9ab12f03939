"""Parity at the sizes BASELINE.json names and at the reference test's own bounds (VERDICT r1 "weak" #2, "missing" #7).

* one 131072-token sequence through paged_attention_v1 and _v2 (P = 256 partitions through the split reduce), bf16 and fp8;
* a Llama-3-70B-shaped case (H 64 / Hkv 8, ctx 8192) whose caches hold more than 2^31 ELEMENTS, with block ids on both
  sides of the 2^31 boundary, so the 64-bit offset arithmetic is compared with the oracle and not only timed;
* the reference grid once at the reference's own sequence-length bound: MAX_SEQ_LEN = max shared memory / 4 - 512 and
  NUM_BLOCKS = 4321 (tests/kernels/attention/test_attention.py:26-29), random non-unique block tables.

Inputs are generated on the device (seeded); the oracle runs on the CPU over a compact copy of exactly the cache blocks
each sequence references.
"""
import math

import pytest
import torch

from oracle import torch_oracle as orc

pytestmark = pytest.mark.gpu
ATOL = 2e-3  # BASELINE.json north_star: max-abs, bf16 output


def _rand_cache(shape, fp8, scale, gen, dev):
    r = (torch.rand(shape, device=dev, generator=gen) * 2 - 1) * scale
    return r.to(torch.float8_e4m3fn).view(torch.uint8) if fp8 else r.to(torch.bfloat16)


def _oracle_rows(q, kc, vc, bt, lens, scale, kvd, ks, vs, rows):
    """Oracle output of the sequences `rows`, each over a compact CPU copy of its own blocks."""
    outs = []
    B = vc.shape[3]
    for s in rows:
        L = int(lens[s])
        nb = max(1, (L + B - 1) // B)
        ids = bt[s, :nb].long()
        uniq, inv = torch.unique(ids, return_inverse=True)
        kcc, vcc = kc[uniq].cpu(), vc[uniq].cpu()
        outs.append(orc.paged_attention(q[s:s + 1].cpu(), kcc, vcc, inv.view(1, -1).to(torch.int32).cpu(),
                                        lens[s:s + 1].cpu(), scale, None, kvd, ks, vs))
    return torch.cat(outs)


def _call(ops, version, q, kc, vc, Hkv, scale, bt, lens, B, max_len, kvd, ks, vs):
    out = torch.full_like(q, float("nan"))
    dev = q.device
    kst = torch.tensor([ks], dtype=torch.float32, device=dev)
    vst = torch.tensor([vs], dtype=torch.float32, device=dev)
    if version == "v1":
        ops.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, lens, B, max_len, None, kvd, kst, vst)
    else:
        S, H, D = q.shape
        P = (max_len + 511) // 512
        es = torch.full((S, H, P), float("nan"), dtype=torch.float32, device=dev)
        ml = torch.full_like(es, float("nan"))
        tmp = torch.full((S, H, P, D), float("nan"), dtype=q.dtype, device=dev)
        ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, lens, B, max_len, None, kvd, kst, vst)
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("kvd,ks,vs", [("auto", 1.0, 1.0), ("fp8", 1.0, 1.0), ("fp8", 0.5, 2.0)])
@pytest.mark.parametrize("version", ["v1", "v2"])
def test_ctx_131072(ops, version, kvd, ks, vs):
    """BASELINE.json configs[4] at its full context: S = 3 (131072, 65537 and 1 tokens), Llama-3-8B head shape."""
    dev = torch.device("cuda")
    H, Hkv, D, B, L = 32, 8, 128, 16, 131072
    fp8 = kvd != "auto"
    x = 16 if fp8 else 8
    g = torch.Generator(device=dev).manual_seed(7)
    nblk = L // B
    NB = 2 * nblk + 64
    scale = D ** -0.5
    kc = _rand_cache((NB, Hkv, D // x, B, x), fp8, scale / ks, g, dev)
    vc = _rand_cache((NB, Hkv, D, B), fp8, scale / vs, g, dev)
    q = _rand_cache((3, H, D), False, scale, g, dev)
    perm = torch.randperm(NB, device=dev, generator=g).to(torch.int32)
    bt = torch.zeros(3, nblk, dtype=torch.int32, device=dev)
    bt[0] = perm[:nblk]
    bt[1] = perm[nblk:2 * nblk]
    bt[2, 0] = perm[2 * nblk]
    lens = torch.tensor([L, 65537, 1], dtype=torch.int32, device=dev)
    out = _call(ops, version, q, kc, vc, Hkv, scale, bt, lens, B, L, kvd, ks, vs)
    ref = _oracle_rows(q, kc, vc, bt, lens, scale, kvd, ks, vs, [0, 1, 2])
    err = (out.cpu().float() - ref.float()).abs().max().item()
    assert math.isfinite(err) and err <= ATOL, err


@pytest.mark.parametrize("kvd", ["auto", "fp8"])
@pytest.mark.parametrize("version", ["v1", "v2"])
def test_cache_beyond_2_31_elements(ops, version, kvd):
    """Llama-3-70B layer shape; NB * kv_block_stride = 135168 * 16384 > 2^31 elements.  Sequences draw their blocks
    from below and above the 2^31-element (and, for bf16, the 2^32-byte) boundary."""
    dev = torch.device("cuda")
    S, H, Hkv, D, B, L = 4, 64, 8, 128, 16, 8192
    fp8 = kvd != "auto"
    x = 16 if fp8 else 8
    NB = 135168  # 2^31 / 16384 = 131072 blocks
    nblk = L // B
    scale = D ** -0.5
    g = torch.Generator(device=dev).manual_seed(11)
    cdt = torch.uint8 if fp8 else torch.bfloat16
    kc = torch.empty((NB, Hkv, D // x, B, x), dtype=cdt, device=dev)
    vc = torch.empty((NB, Hkv, D, B), dtype=cdt, device=dev)
    assert kc.numel() > 2 ** 31
    # block ids: sequence 0 all above the boundary, 1 interleaved, 2 below (65536.. crosses 2^32 bytes for bf16), 3 mixed
    hi = 131072 + torch.randperm(NB - 131072, device=dev, generator=g)
    lo = torch.randperm(131072, device=dev, generator=g)
    bt = torch.empty(S, nblk, dtype=torch.int32, device=dev)
    bt[0] = hi[:nblk]
    bt[1, 0::2], bt[1, 1::2] = hi[nblk:nblk + nblk // 2], lo[:nblk // 2]
    bt[2] = lo[nblk:2 * nblk]
    bt[3, :nblk // 2], bt[3, nblk // 2:] = lo[2 * nblk:2 * nblk + nblk // 2], hi[2 * nblk:2 * nblk + nblk // 2]
    used = torch.unique(bt.flatten().long())
    kc[used] = _rand_cache((used.numel(), Hkv, D // x, B, x), fp8, scale, g, dev)
    vc[used] = _rand_cache((used.numel(), Hkv, D, B), fp8, scale, g, dev)
    q = _rand_cache((S, H, D), False, scale, g, dev)
    lens = torch.tensor([L, L - 5, 4097, L], dtype=torch.int32, device=dev)
    out = _call(ops, version, q, kc, vc, Hkv, scale, bt, lens, B, L, kvd, 1.0, 1.0)
    ref = _oracle_rows(q, kc, vc, bt, lens, scale, kvd, 1.0, 1.0, range(S))
    err = (out.cpu().float() - ref.float()).abs().max().item()
    assert math.isfinite(err) and err <= ATOL, err


def _reference_max_seq_len():
    """tests/kernels/attention/test_attention.py:23-26: FLOAT32_BYTES = 4; MAX_SEQ_LEN = get_max_shared_memory_bytes() //
    FLOAT32_BYTES - 512 (the opt-in per-block maximum: 232448 B on B200 -> 57600)."""
    smem = torch.cuda.get_device_properties(0).shared_memory_per_block_optin
    return smem // 4 - 512


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("num_heads", [(40, 40), (64, 8)])
@pytest.mark.parametrize("head_size,block_size,kvd", [(128, 16, "auto"), (128, 32, "fp8"), (80, 16, "auto")])
def test_reference_grid_at_reference_bound(ops, version, num_heads, head_size, block_size, kvd):
    """NUM_BLOCKS = 4321, 7 sequences of random length up to MAX_SEQ_LEN (last = max), random non-unique block tables
    (tests/kernels/attention/test_attention.py:29,157-185)."""
    dev = torch.device("cuda")
    H, Hkv = num_heads
    max_len = _reference_max_seq_len()
    assert max_len > 40000
    fp8 = kvd != "auto"
    x = 16 if fp8 else 8
    NB, S = 4321, 7
    scale = head_size ** -0.5
    g = torch.Generator(device=dev).manual_seed(0)
    kc = _rand_cache((NB, Hkv, head_size // x, block_size, x), fp8, scale, g, dev)
    vc = _rand_cache((NB, Hkv, head_size, block_size), fp8, scale, g, dev)
    q = _rand_cache((S, H, head_size), False, scale, g, dev)
    lens = torch.randint(1, max_len + 1, (S,), device=dev, generator=g, dtype=torch.int32)
    lens[-1] = max_len
    nblk = (max_len + block_size - 1) // block_size
    bt = torch.randint(0, NB, (S, nblk), device=dev, generator=g, dtype=torch.int32)
    out = _call(ops, version, q, kc, vc, Hkv, scale, bt, lens, block_size, max_len, kvd, 1.0, 1.0)
    ref = _oracle_rows(q, kc, vc, bt, lens, scale, kvd, 1.0, 1.0, range(S))
    err = (out.cpu().float() - ref.float()).abs().max().item()
    assert math.isfinite(err) and err <= (1e-3 if kvd == "auto" else ATOL), err


@pytest.mark.parametrize("kvd", ["auto", "fp8"])
def test_c2_batch64_every_sequence(ops, kvd):
    """BASELINE.json configs[1]/[2] at the benchmarked size (batch 64, ctx 4096, unique block tables over a 1.25x
    cache): every sequence, every head, against the oracle."""
    dev = torch.device("cuda")
    S, H, Hkv, D, B, L = 64, 32, 8, 128, 16, 4096
    fp8 = kvd != "auto"
    x = 16 if fp8 else 8
    nblk = L // B
    NB = int(S * nblk * 1.25)
    scale = D ** -0.5
    g = torch.Generator(device=dev).manual_seed(0)
    kc = _rand_cache((NB, Hkv, D // x, B, x), fp8, scale, g, dev)
    vc = _rand_cache((NB, Hkv, D, B), fp8, scale, g, dev)
    q = _rand_cache((S, H, D), False, scale, g, dev)
    bt = torch.randperm(NB, device=dev, generator=g)[:S * nblk].view(S, nblk).to(torch.int32)
    lens = torch.full((S,), L, dtype=torch.int32, device=dev)
    ref = _oracle_rows(q, kc, vc, bt, lens, scale, kvd, 1.0, 1.0, range(S))
    for version in ("v1", "v2"):
        out = _call(ops, version, q, kc, vc, Hkv, scale, bt, lens, B, L, kvd, 1.0, 1.0)
        err = (out.cpu().float() - ref.float()).abs().max().item()
        assert math.isfinite(err) and err <= ATOL, (version, err)
