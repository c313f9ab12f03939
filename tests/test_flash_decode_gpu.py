"""Parity of the flash-layout decode kernel (paged_attention_flash / flash_attn_with_kvcache) against the CPU oracle.

The oracle restates the reference's ref_paged_attn (tests/kernels/attention/test_triton_unified_attention.py:25-76),
pinned by tests/golden/flash_*.pt.  Tolerance: max-abs 2e-3 (BASELINE.json, bf16 output / fp32 accumulate).
"""
import glob
import os

import pytest
import torch

from oracle import torch_oracle as orc

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _to(d, dev="cuda"):
    return {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in d.items()}


@pytest.mark.parametrize("hnd", [False, True])
@pytest.mark.parametrize("shape", [  # S, H, Hkv, D, B, L
    (7, 32, 8, 128, 16, 1500),    # Llama-3-8B head shape
    (5, 8, 8, 128, 32, 700),      # MHA, block 32
    (4, 16, 1, 128, 16, 333),     # GQA 16 (two head tiles)
    (3, 8, 1, 128, 64, 1000),     # TP8 shard, block 64
    (150, 8, 2, 128, 16, 77),     # more (seq, kv head) pairs than CTAs, short sequences
    (2, 32, 8, 128, 128, 5000),   # block 128, long: split pairs merged in the kernel
])
@pytest.mark.parametrize("alibi", [False, True])
def test_paged_attention_flash(ops, shape, hnd, alibi):
    S, H, Hkv, D, B, L = shape
    if alibi and S > 7:
        pytest.skip()
    NB = S * ((L + B - 1) // B) + 5
    d = orc.make_flash_decode_inputs(S, H, Hkv, D, B, L, NB, seed=21, hnd=hnd, use_alibi=alibi)
    lens = d["seq_lens"].clone()
    if S > 3:
        lens[1] = 0      # empty sequence
        lens[2] = 1      # single token
        lens[3] = min(L, 2 * B)  # exactly two blocks
    ref = orc.paged_attention_flash(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], lens, d["scale"],
                                    d["alibi_slopes"])
    ref_lse = orc.paged_attention_flash_lse(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], lens,
                                            d["scale"], d["alibi_slopes"])
    g = _to(d)
    assert g["key_cache"].stride() == d["key_cache"].stride()  # .to() keeps the NHD / HND strides
    out = torch.full_like(g["query"], float("nan"))
    lse = torch.full((H, S), float("nan"), dtype=torch.float32, device="cuda")
    ops.paged_attention_flash(out, g["query"], g["key_cache"], g["value_cache"], d["scale"], g["block_tables"],
                              lens.cuda(), L, g["alibi_slopes"], lse)
    nz = lens > 0
    assert (out.cpu().float()[nz] - ref.float()[nz]).abs().max() <= 2e-3
    assert (out.cpu()[~nz] == 0).all() and torch.isneginf(lse.cpu()[:, ~nz]).all()
    torch.testing.assert_close(lse.cpu()[:, nz], ref_lse[:, nz], atol=2e-3, rtol=1e-4)
    # without the LSE output, twice: deterministic, and rows of empty sequences are left alone
    o1 = torch.zeros_like(out)
    o2 = torch.zeros_like(out)
    for o in (o1, o2):
        ops.paged_attention_flash(o, g["query"], g["key_cache"], g["value_cache"], d["scale"], g["block_tables"],
                                  lens.cuda(), L, g["alibi_slopes"])
    assert torch.equal(o1, o2) and torch.equal(o1[nz.cuda()], out[nz.cuda()])


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "flash_*.pt"))),
                         ids=lambda p: os.path.basename(p)[:-3])
def test_paged_attention_flash_golden(ops, path):
    d = torch.load(path)
    g = _to(d)
    out = torch.empty_like(g["query"])
    ops.paged_attention_flash(out, g["query"], g["key_cache"], g["value_cache"], d["scale"], g["block_tables"],
                              g["seq_lens"], d["max_seq_len"])
    assert (out.cpu().float() - d["ref_out"].float()).abs().max() <= 2e-3


def test_flash_attn_with_kvcache_wrapper(ops):
    """The call the reference backend makes for decode (vllm_metax/v1/attention/backends/flash_attn.py:742-757)."""
    from vllm_metax_b200.attention import flash_attn_with_kvcache

    S, H, Hkv, D, B, L = 6, 32, 8, 128, 16, 600
    d = orc.make_flash_decode_inputs(S, H, Hkv, D, B, L, S * ((L + B - 1) // B) + 3, seed=2)
    ref = orc.paged_attention_flash(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"], d["scale"])
    g = _to(d)
    out, lse = flash_attn_with_kvcache(q=g["query"].unsqueeze(1), k_cache=g["key_cache"], v_cache=g["value_cache"],
                                       block_table=g["block_tables"], cache_seqlens=g["seq_lens"],
                                       softmax_scale=d["scale"], causal=True, window_size=(-1, -1), alibi_slopes=None,
                                       softcap=0.0, return_softmax_lse=True)
    assert out.shape == (S, 1, H, D) and lse.shape == (S, H, 1)
    assert (out[:, 0].cpu().float() - ref.float()).abs().max() <= 2e-3
    ref_lse = orc.paged_attention_flash_lse(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"], d["scale"])
    torch.testing.assert_close(lse[:, :, 0].t().cpu(), ref_lse, atol=2e-3, rtol=1e-4)
    with pytest.raises(NotImplementedError):
        flash_attn_with_kvcache(q=g["query"].unsqueeze(1), k_cache=g["key_cache"], v_cache=g["value_cache"],
                                block_table=g["block_tables"], cache_seqlens=g["seq_lens"], window_size=(128, 0))


def test_flash_cache_written_by_reshape_and_cache_flash_then_decoded(ops):
    """The serving sequence: reshape_and_cache_flash writes this step's K/V, the flash decode kernel reads them."""
    S, H, Hkv, D, B, L = 8, 32, 8, 128, 16, 257
    d = orc.make_flash_decode_inputs(S, H, Hkv, D, B, L, S * ((L + B - 1) // B) + 3, seed=4)
    lens = d["seq_lens"]
    gen = torch.Generator().manual_seed(1)
    key = (torch.randn(S, Hkv, D, generator=gen) * 0.05).to(torch.bfloat16)
    value = (torch.randn(S, Hkv, D, generator=gen) * 0.05).to(torch.bfloat16)
    slots = torch.tensor([int(d["block_tables"][s, (int(lens[s]) - 1) // B]) * B + (int(lens[s]) - 1) % B
                          for s in range(S)], dtype=torch.long)
    one = torch.ones(1)
    rk, rv = d["key_cache"].clone(), d["value_cache"].clone()
    orc.reshape_and_cache_flash(key, value, rk, rv, slots, "auto", one, one)
    ref = orc.paged_attention_flash(d["query"], rk, rv, d["block_tables"], lens, d["scale"])
    g = _to(d)
    ops.reshape_and_cache_flash(key.cuda(), value.cuda(), g["key_cache"], g["value_cache"], slots.cuda(), "auto",
                                one.cuda(), one.cuda())
    assert torch.equal(g["key_cache"].cpu(), rk) and torch.equal(g["value_cache"].cpu(), rv)
    out = torch.empty_like(g["query"])
    ops.paged_attention_flash(out, g["query"], g["key_cache"], g["value_cache"], d["scale"], g["block_tables"],
                              g["seq_lens"], L)
    assert (out.cpu().float() - ref.float()).abs().max() <= 2e-3


def test_paged_attention_flash_rejects_unsupported(ops):
    q = torch.zeros(2, 8, 64, dtype=torch.bfloat16, device="cuda")
    kc = torch.zeros(4, 16, 2, 64, dtype=torch.bfloat16, device="cuda")
    bt = torch.zeros(2, 2, dtype=torch.int32, device="cuda")
    sl = torch.ones(2, dtype=torch.int32, device="cuda")
    with pytest.raises(RuntimeError, match="head_size is not 128"):
        ops.paged_attention_flash(torch.empty_like(q), q, kc, kc.clone(), 0.1, bt, sl, 16)
