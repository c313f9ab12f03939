"""Parity of the ops either side of the decode path: merge_attn_states, the LSE output of the decode
kernel, copy_blocks, swap_blocks, gather_and_maybe_dequant_cache, cp_gather_cache — CUDA vs the CPU oracle.

Grids follow the reference tests (tests/kernels/attention/test_merge_attn_states.py:50-53,
tests/kernels/attention/test_cache.py:19-40,54-118,352-432,704-862), scaled to sizes the oracle finishes in seconds.
Byte movers are bit-exact; merge_attn_states uses the reference's own tolerance (atol 1e-3, rtol 1e-2 bf16 / 1e-3).
"""
import glob
import os
import random

import pytest
import torch

from oracle import torch_oracle as orc

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ---- merge_attn_states --------------------------------------------------------------------------------------
def _merge_inputs(N, H, D, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    p_lse, s_lse = torch.randn(H, N, generator=g), torch.randn(H, N, generator=g)
    mp, ms = torch.rand(H, N, generator=g) < 0.1, torch.rand(H, N, generator=g) < 0.1
    both = mp & ms
    p_lse[mp & ~both] = float("inf")  # test_merge_attn_states.py:112-121
    s_lse[ms & ~both] = float("inf")
    return (torch.randn(N, H, D, generator=g).to(dtype), p_lse, torch.randn(N, H, D, generator=g).to(dtype), s_lse)


@pytest.mark.parametrize("num_tokens", [256, 613, 4096])
@pytest.mark.parametrize("num_heads", [4, 32, 48])
@pytest.mark.parametrize("head_size", [32, 48, 96, 128, 256])
@pytest.mark.parametrize("dtype", [torch.float32, torch.half, torch.bfloat16])
def test_merge_attn_states(ops, num_tokens, num_heads, head_size, dtype):
    if num_tokens == 4096 and num_heads == 48 and head_size == 256:
        pytest.skip("oracle size")
    p_out, p_lse, s_out, s_lse = _merge_inputs(num_tokens, num_heads, head_size, dtype, 0)
    ref_out, ref_lse = orc.merge_attn_states(p_out, p_lse, s_out, s_lse)
    out = torch.zeros(num_tokens, num_heads, head_size, dtype=dtype, device="cuda")
    lse = torch.zeros(num_heads, num_tokens, dtype=torch.float32, device="cuda")
    p_lse_d, s_lse_d = p_lse.cuda(), s_lse.cuda()
    ops.merge_attn_states(out, p_out.cuda(), p_lse_d, s_out.cuda(), s_lse_d, lse)
    rtol = 1e-2 if dtype == torch.bfloat16 else 1e-3  # test_merge_attn_states.py:222
    torch.testing.assert_close(out.cpu().float(), ref_out.float(), atol=1e-3, rtol=rtol)
    torch.testing.assert_close(lse.cpu(), ref_lse, atol=1e-3, rtol=rtol)
    assert torch.equal(p_lse_d.cpu(), p_lse) and torch.equal(s_lse_d.cpu(), s_lse)  # inputs are read-only
    # measured bar, tighter than the reference's: one rounding to the output dtype
    eps = {torch.float32: 2e-6, torch.half: 1e-3, torch.bfloat16: 8e-3}[dtype]
    assert (out.cpu().float() - ref_out.float()).abs().max() <= eps * max(1.0, float(ref_out.float().abs().max()))
    # without output_lse, and with strided (sliced) inputs
    out2 = torch.zeros_like(out)
    wide = torch.zeros(num_tokens, num_heads + 2, head_size, dtype=dtype, device="cuda")
    wide[:, 1:-1] = p_out.cuda()
    ops.merge_attn_states(out2, wide[:, 1:-1], p_lse_d, s_out.cuda(), s_lse_d)
    assert torch.equal(out2, out)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "merge_*.pt"))),
                         ids=lambda p: os.path.basename(p)[:-3])
def test_merge_attn_states_golden(ops, path):
    """Against the output of the reference's own merge_attn_states_torch (tests/golden/make_golden.py)."""
    d = torch.load(path)
    dtype = d["prefix_output"].dtype
    out = torch.empty_like(d["prefix_output"], device="cuda")
    lse = torch.empty_like(d["prefix_lse"], device="cuda")
    ops.merge_attn_states(out, d["prefix_output"].cuda(), d["prefix_lse"].cuda(), d["suffix_output"].cuda(),
                          d["suffix_lse"].cuda(), lse)
    rtol = 1e-2 if dtype == torch.bfloat16 else 1e-3
    torch.testing.assert_close(out.cpu().float(), d["ref_output"].float(), atol=1e-3, rtol=rtol)
    torch.testing.assert_close(lse.cpu(), d["ref_output_lse"], atol=1e-3, rtol=rtol)


def test_merge_attn_states_rejects_bad_head_size(ops):
    o = torch.zeros(4, 2, 36, dtype=torch.half, device="cuda")
    l = torch.zeros(2, 4, device="cuda")
    with pytest.raises(RuntimeError, match="pack_size"):
        ops.merge_attn_states(o, o.clone(), l, o.clone(), l)


# ---- decode kernel with LSE output, merged back ---------------------------------------------------------------
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("shape", [  # S, H, Hkv, D, B, L, dtype  -> streaming kernel / generic kernel
    (7, 32, 8, 128, 16, 1500, torch.bfloat16),
    (5, 8, 8, 128, 32, 700, torch.bfloat16),
    (3, 8, 2, 80, 16, 900, torch.float16),
    (3, 4, 4, 64, 8, 300, torch.float32),
])
def test_paged_attention_lse_and_split_merge(ops, shape, kv_cache_dtype):
    """lse[H, S] of paged_attention_v1_lse equals the oracle's log-sum-exp, and attending to two halves of every
    sequence then merging with merge_attn_states reproduces attention over the whole sequence (cascade / DCP use)."""
    S, H, Hkv, D, B, L, dtype = shape
    if kv_cache_dtype == "fp8" and D % 16:
        pytest.skip()
    NB = S * ((L + B - 1) // B) + 8
    d = orc.make_decode_inputs(S, H, Hkv, D, B, L, NB, dtype=dtype, kv_cache_dtype=kv_cache_dtype, seed=3,
                               unique_blocks=True)
    lens = d["seq_lens"].clone()
    lens[0] = L
    if S > 2:
        lens[1] = 0  # an empty sequence: out row 0, lse -inf
    ks = torch.tensor(0.5 if kv_cache_dtype == "fp8" else 1.0)
    vs = torch.tensor(2.0 if kv_cache_dtype == "fp8" else 1.0)
    args = (d["key_cache"], d["value_cache"], d["block_tables"])
    ref_lse = orc.paged_attention_lse(d["query"], *args, lens, d["scale"], None, kv_cache_dtype, ks, vs)
    ref_out = orc.paged_attention(d["query"], *args, lens, d["scale"], None, kv_cache_dtype, ks, vs)

    dev = "cuda"
    q, kc, vc, bt = d["query"].to(dev), d["key_cache"].to(dev), d["value_cache"].to(dev), d["block_tables"].to(dev)
    ksd, vsd = ks.to(dev), vs.to(dev)

    def run(tables, seq_lens):
        out = torch.full_like(q, float("nan"))
        lse = torch.full((H, S), float("nan"), dtype=torch.float32, device=dev)
        ops.paged_attention_v1_lse(out, lse, q, kc, vc, Hkv, d["scale"], tables, seq_lens.to(dev), B,
                                   int(seq_lens.max()), None, kv_cache_dtype, ksd, vsd)
        return out, lse

    out, lse = run(bt, lens)
    nz = lens > 0
    torch.testing.assert_close(lse.cpu()[:, nz], ref_lse[:, nz], atol=2e-3, rtol=1e-4)
    assert torch.isneginf(lse.cpu()[:, ~nz]).all() and (out.cpu()[~nz] == 0).all()
    assert (out.cpu().float()[nz] - ref_out.float()[nz]).abs().max() <= 2e-3

    cut = (lens // 2 // B * B).to(torch.int32)
    rest = lens - cut
    bt2 = torch.stack([torch.roll(d["block_tables"][i], -int(cut[i]) // B) for i in range(S)]).to(dev)
    o1, l1 = run(bt, cut)
    o2, l2 = run(bt2, rest)
    merged = torch.empty_like(q)
    mlse = torch.empty_like(l1)
    ops.merge_attn_states(merged, o1, l1, o2, l2, mlse)
    both_empty = (lens == 0)
    assert (merged.cpu().float()[~both_empty] - ref_out.float()[~both_empty]).abs().max() <= 2e-3
    torch.testing.assert_close(mlse.cpu()[:, ~both_empty], ref_lse[:, ~both_empty], atol=2e-3, rtol=1e-4)


# ---- copy_blocks ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("num_mappings", [1, 37])
@pytest.mark.parametrize("num_layers", [1, 3, 130])  # 130 > one launch's pointer table
@pytest.mark.parametrize("head_size", [64, 80, 128])
@pytest.mark.parametrize("block_size", [8, 16])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
def test_copy_blocks(ops, num_mappings, num_layers, head_size, block_size, dtype, kv_cache_dtype):
    if kv_cache_dtype == "fp8" and head_size % 16:
        pytest.skip()
    if num_layers == 130 and (head_size != 64 or dtype != torch.bfloat16):
        pytest.skip("one large-layer-count case per kv dtype is enough")
    num_blocks, H = 128, 2
    random.seed(0)
    src_blocks = random.sample(range(num_blocks), num_mappings)
    remaining = sorted(set(range(num_blocks)) - set(src_blocks))
    dst_blocks = random.sample(remaining, 2 * num_mappings)
    mapping = []
    for i in range(num_mappings):  # test_cache.py:71-81
        mapping += [(src_blocks[i], dst_blocks[2 * i]), (src_blocks[i], dst_blocks[2 * i + 1])]
    caches = [orc.make_kv_caches(num_blocks, block_size, H, head_size, kv_cache_dtype, dtype, seed=l)
              for l in range(num_layers)]
    kcs, vcs = [c[0] for c in caches], [c[1] for c in caches]
    kd, vd = [k.cuda() for k in kcs], [v.cuda() for v in vcs]
    m = torch.tensor(mapping, dtype=torch.int64).view(-1, 2)
    ops.copy_blocks(kd, vd, m.cuda())
    orc.copy_blocks(kcs, vcs, m)
    for a, b in zip(kd + vd, kcs + vcs):
        assert torch.equal(a.cpu().view(torch.uint8), b.view(torch.uint8))


def test_copy_blocks_golden(ops):
    d = torch.load(os.path.join(GOLDEN, "movers_copy_blocks_bf16.pt"))
    kd, vd = [k.cuda() for k in d["key_caches"]], [v.cuda() for v in d["value_caches"]]
    ops.copy_blocks(kd, vd, d["block_mapping"].cuda())
    for a, b in zip(kd + vd, d["ref_key_caches"] + d["ref_value_caches"]):
        assert torch.equal(a.cpu(), b)


# ---- swap_blocks ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("direction", [("cuda", "cpu"), ("cuda", "cuda"), ("cpu", "cuda")])
@pytest.mark.parametrize("num_mappings", [1, 64])
@pytest.mark.parametrize("head_size", [64, 120])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
def test_swap_blocks(ops, direction, num_mappings, head_size, dtype, kv_cache_dtype):
    if kv_cache_dtype == "fp8" and (head_size % 16 or "cpu" in direction):
        pytest.skip()  # test_cache.py:379-382
    num_blocks, block_size, H = 128, 16, 4
    random.seed(1)
    src_blocks = random.sample(range(num_blocks), num_mappings)
    if direction[0] == direction[1]:
        dst_blocks = random.sample(sorted(set(range(num_blocks)) - set(src_blocks)), num_mappings)
    else:
        dst_blocks = random.sample(range(num_blocks), num_mappings)
    if num_mappings >= 8 and direction[0] != direction[1]:  # a run consecutive on both sides travels as one copy
        src_blocks = [100, 101, 102, 103] + [b for b in src_blocks if b not in (100, 101, 102, 103)][:num_mappings - 4]
        dst_blocks = [20, 21, 22, 23] + [b for b in dst_blocks if b not in (20, 21, 22, 23)][:num_mappings - 4]
    mapping = torch.tensor(list(zip(src_blocks, dst_blocks)), dtype=torch.int64).view(-1, 2)
    for which in (0, 1):  # the key cache, then the value cache
        src = orc.make_kv_caches(num_blocks, block_size, H, head_size, kv_cache_dtype, dtype, seed=2)[which]
        dst = orc.make_kv_caches(num_blocks, block_size, H, head_size, kv_cache_dtype, dtype, seed=3)[which]
        ref = dst.clone()
        orc.swap_blocks(src, ref, mapping)
        src_t, dst_t = src.to(direction[0]), dst.to(direction[1])
        ops.swap_blocks(src_t, dst_t, mapping)
        torch.cuda.synchronize()
        assert torch.equal(dst_t.cpu().view(torch.uint8), ref.view(torch.uint8))
        assert torch.equal(src_t.cpu().view(torch.uint8), src.view(torch.uint8))


def test_swap_blocks_rejects_device_mapping(ops):
    c = torch.zeros(4, 16, 8, device="cuda")
    with pytest.raises(RuntimeError, match="block_mapping must be on CPU"):
        ops.swap_blocks(c, c.clone(), torch.zeros(1, 2, dtype=torch.int64, device="cuda"))


# ---- gathers -------------------------------------------------------------------------------------------------
def _gather_inputs(NB, B, entry, batch, max_len, dtype, kv_cache_dtype, seed):
    g = torch.Generator().manual_seed(seed)
    if kv_cache_dtype == "fp8":
        src = orc.fp8_e4m3_quant(torch.randn(NB, B, entry, generator=g), 0.1).view(NB, B, entry)
    else:
        src = torch.randn(NB, B, entry, generator=g).to(dtype)
    seq_len = torch.randint(0, max_len + 1, (batch,), generator=g)
    seq_len[batch // 2] = 0
    cu = torch.zeros(batch + 1, dtype=torch.int32)
    cu[1:] = seq_len.cumsum(0).to(torch.int32)
    table = torch.stack([torch.randperm(NB, generator=g) for _ in range(batch)]).to(torch.int32)
    return src, seq_len, cu, table


@pytest.mark.parametrize("entry_size", [576, 512 + 64 + 4, 96])  # MLA entry (test_cache.py:704-705); one not 16-divisible
@pytest.mark.parametrize("batch_size", [1, 8, 70, 200])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("with_starts", [False, True])
def test_gather_and_maybe_dequant_cache(ops, entry_size, batch_size, dtype, kv_cache_dtype, with_starts):
    NB, B, max_len = 256, 16, 300 if batch_size <= 8 else 60
    src, seq_len, cu, table = _gather_inputs(NB, B, entry_size, batch_size, max_len, dtype, kv_cache_dtype, 11)
    scale = torch.tensor(0.1, dtype=torch.float32)
    starts = None
    if with_starts:
        starts = torch.randint(0, NB * B - max_len - B, (batch_size,), generator=torch.Generator().manual_seed(4)).int()
    ref = orc.gather_cache(src, table, cu, batch_size, kv_cache_dtype, scale, starts, False, dtype)
    total = int(cu[-1])
    dst = torch.zeros(total + 3, entry_size, dtype=dtype, device="cuda")
    ops.gather_and_maybe_dequant_cache(src.cuda(), dst, table.cuda(), cu.cuda(), batch_size, kv_cache_dtype,
                                       scale.cuda(), starts.cuda() if with_starts else None)
    assert torch.equal(dst[:total].cpu().view(torch.uint8), ref.contiguous().view(torch.uint8))
    assert (dst[total:] == 0).all()  # rows past the last sequence are untouched


def test_gather_and_maybe_dequant_cache_golden(ops):
    d = torch.load(os.path.join(GOLDEN, "movers_gather_f32.pt"))
    dst = torch.zeros_like(d["ref_dst"], device="cuda")
    one = torch.ones(1, device="cuda")
    ops.gather_and_maybe_dequant_cache(d["src_cache"].cuda(), dst, d["block_table"].cuda(), d["cu_seq_lens"].cuda(),
                                       d["batch_size"], "auto", one, None)
    assert torch.equal(dst.cpu(), d["ref_dst"])
    dst2 = torch.zeros_like(dst)
    ops.cp_gather_cache(d["src_cache"].cuda(), dst2, d["block_table"].cuda(), d["cu_seq_lens"].cuda(), d["batch_size"])
    assert torch.equal(dst2.cpu(), d["ref_dst"])


@pytest.mark.parametrize("entry_size", [576, 656, 30])  # 656: the fp8 ds-MLA entry width upstream gathers as bytes
@pytest.mark.parametrize("batch_size", [1, 8, 130])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.uint8])
@pytest.mark.parametrize("with_starts", [False, True])
def test_cp_gather_cache(ops, entry_size, batch_size, dtype, with_starts):
    NB, B, max_len = 256, 16, 300 if batch_size <= 8 else 60
    src, seq_len, cu, table = _gather_inputs(NB, B, entry_size, batch_size, max_len,
                                             torch.float32 if dtype == torch.uint8 else dtype,
                                             "fp8" if dtype == torch.uint8 else "auto", 13)
    starts = None
    if with_starts:  # token-granular: need not be a multiple of the block size (csrc/cache_kernels.cu:777-778)
        starts = torch.randint(0, NB * B - max_len - B, (batch_size,), generator=torch.Generator().manual_seed(6)).int()
    ref = orc.gather_cache(src, table, cu, batch_size, "auto", None, starts, True)
    dst = torch.zeros(int(cu[-1]), entry_size, dtype=src.dtype, device="cuda")
    ops.cp_gather_cache(src.cuda(), dst, table.cuda(), cu.cuda(), batch_size, starts.cuda() if with_starts else None)
    assert torch.equal(dst.cpu().view(torch.uint8), ref.contiguous().view(torch.uint8))


def test_gather_rejects_mismatched_dtypes(ops):
    src = torch.zeros(4, 16, 32, device="cuda")
    dst = torch.zeros(8, 32, dtype=torch.half, device="cuda")
    t = torch.zeros(1, 4, dtype=torch.int32, device="cuda")
    cu = torch.tensor([0, 8], dtype=torch.int32, device="cuda")
    with pytest.raises(RuntimeError, match="same dtype"):
        ops.cp_gather_cache(src, dst, t, cu, 1)
    with pytest.raises(RuntimeError, match="must be int32"):
        ops.cp_gather_cache(src, src.new_zeros(8, 32), t.long(), cu, 1)


# ---- MLA cache ops ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kv_lora_rank,pe_dim", [(512, 64), (500, 60)])
@pytest.mark.parametrize("num_tokens", [1, 42, 300])
@pytest.mark.parametrize("block_size", [16, 64])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
def test_concat_and_cache_mla(ops, kv_lora_rank, pe_dim, num_tokens, block_size, dtype, kv_cache_dtype):
    """Grid of tests/kernels/attention/test_cache.py:494-572 (kv_lora_rank 512, rope dim 64) plus an unaligned entry."""
    NB, pad = 40, 4
    g = torch.Generator().manual_seed(3)
    random.seed(3)
    slots = random.sample(range(NB * block_size), num_tokens)
    if num_tokens > 2:
        slots[1] = -1
    slot_mapping = torch.tensor(slots, dtype=torch.long)
    kv_c = torch.randn(num_tokens + pad, kv_lora_rank, generator=g).to(dtype)
    wide = torch.randn(num_tokens + pad, pe_dim + 8, generator=g).to(dtype)
    k_pe = wide[:, :pe_dim]  # row stride != pe_dim, as the squeeze of a [T, 1, pe] slice gives in the model
    cdt = torch.uint8 if kv_cache_dtype == "fp8" else dtype
    cache = torch.randint(0, 100, (NB, block_size, kv_lora_rank + pe_dim), generator=g).to(cdt)
    scale = torch.tensor(0.1, dtype=torch.float32)
    ref = cache.clone()
    orc.concat_and_cache_mla(kv_c, k_pe, ref, slot_mapping, kv_cache_dtype, scale)
    cd = cache.cuda()
    ops.concat_and_cache_mla(kv_c.cuda(), wide.cuda()[:, :pe_dim], cd, slot_mapping.cuda(), kv_cache_dtype, scale.cuda())
    assert torch.equal(cd.cpu().view(torch.uint8), ref.view(torch.uint8))


@pytest.mark.parametrize("num_layers", [1, 5, 130])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.uint8])
def test_copy_blocks_mla(ops, num_layers, dtype):
    """tests/kernels/attention/test_cache.py:591-652: every source block copied onto two destinations, all layers."""
    NB, B, E, nmap = 64, 16, 576, 9
    random.seed(5)
    src_blocks = random.sample(range(NB), nmap)
    remaining = sorted(set(range(NB)) - set(src_blocks))
    dst_blocks = random.sample(remaining, 2 * nmap)
    mapping = []
    for i in range(nmap):
        mapping += [(src_blocks[i], dst_blocks[2 * i]), (src_blocks[i], dst_blocks[2 * i + 1])]
    m = torch.tensor(mapping, dtype=torch.int64).view(-1, 2)
    g = torch.Generator().manual_seed(6)
    caches = [torch.randint(0, 200, (NB, B, E), generator=g).to(dtype) for _ in range(num_layers)]
    dev = [c.cuda() for c in caches]
    ops.copy_blocks_mla(dev, m.cuda())
    for s_, d_ in mapping:  # the reference loop (:644-647)
        for c in caches:
            c[d_].copy_(c[s_])
    for a, b in zip(dev, caches):
        assert torch.equal(a.cpu(), b)
