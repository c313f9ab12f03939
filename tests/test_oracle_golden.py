"""The oracle against the golden vectors produced by the reference's own Python oracles
(tests/golden/make_golden.py).  CPU only: this is what pins the oracle before it is trusted."""
import glob
import os

import pytest
import torch

from oracle import c_oracle
from oracle import torch_oracle as orc

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ATTN = sorted(glob.glob(os.path.join(GOLDEN, "attn_*.pt")))
CACHE = sorted(glob.glob(os.path.join(GOLDEN, "cache_*.pt")))


def test_fixtures_present():
    assert len(ATTN) >= 6 and len(CACHE) >= 2


@pytest.mark.parametrize("path", ATTN, ids=[os.path.basename(p)[:-3] for p in ATTN])
def test_attention_oracle_matches_reference_oracle(path):
    d = torch.load(path)
    out = orc.paged_attention(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"],
                              d["scale"], d["alibi_slopes"], d["kv_cache_dtype"], d.get("k_scale", 1.0),
                              d.get("v_scale", 1.0))
    # same op sequence as the reference oracle; only the gather is vectorised
    torch.testing.assert_close(out.float(), d["ref_out"].float(), atol=2e-6, rtol=1e-6)


@pytest.mark.parametrize("path", ATTN, ids=[os.path.basename(p)[:-3] for p in ATTN])
def test_c_oracle_matches_reference_oracle(path):
    d = torch.load(path)
    out = c_oracle.paged_attention(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"],
                                   d["scale"], d["alibi_slopes"], d["kv_cache_dtype"], d.get("k_scale", 1.0),
                                   d.get("v_scale", 1.0))
    # the C port keeps fp32 probabilities (the reference oracle rounds them to the value dtype)
    torch.testing.assert_close(out.float(), d["ref_out"].float(), atol=1e-3, rtol=1e-2)


@pytest.mark.parametrize("path", CACHE, ids=[os.path.basename(p)[:-3] for p in CACHE])
@pytest.mark.parametrize("impl", ["torch", "c"])
def test_cache_writers_match_reference_loops(path, impl):
    d = torch.load(path)
    mod = orc if impl == "torch" else c_oracle
    one = torch.tensor(1.0)
    kc, vc = d["key_cache"].clone(), d["value_cache"].clone()
    mod.reshape_and_cache(d["key"], d["value"], kc, vc, d["slot_mapping"], "auto", one, one)
    assert torch.equal(kc, d["ref_key_cache"]) and torch.equal(vc, d["ref_value_cache"])
    k, v = d["nhd_k"].clone(), d["nhd_v"].clone()
    mod.reshape_and_cache_flash(d["key"], d["value"], k, v, d["slot_mapping"], "auto", one, one)
    assert torch.equal(k, d["ref_nhd_k"]) and torch.equal(v, d["ref_nhd_v"])
    # HND storage viewed as [blk, off, head, dim]
    k, v = d["hnd_k"].clone(), d["hnd_v"].clone()
    mod.reshape_and_cache_flash(d["key"], d["value"], k.permute(0, 2, 1, 3), v.permute(0, 2, 1, 3),
                                d["slot_mapping"], "auto", one, one)
    assert torch.equal(k, d["ref_hnd_k"]) and torch.equal(v, d["ref_hnd_v"])


MERGE = sorted(glob.glob(os.path.join(GOLDEN, "merge_*.pt")))


@pytest.mark.parametrize("path", MERGE, ids=[os.path.basename(p)[:-3] for p in MERGE])
def test_merge_attn_states_oracle_matches_reference_torch(path):
    """oracle.merge_attn_states against the output of the reference's merge_attn_states_torch
    (tests/kernels/attention/test_merge_attn_states.py:16-47, executed by make_golden.py)."""
    d = torch.load(path)
    assert len(MERGE) >= 3
    p_lse, s_lse = d["prefix_lse"].clone(), d["suffix_lse"].clone()
    out, lse = orc.merge_attn_states(d["prefix_output"], p_lse, d["suffix_output"], s_lse)
    assert torch.equal(p_lse, d["prefix_lse"]) and torch.equal(s_lse, d["suffix_lse"])  # inputs untouched
    assert torch.equal(lse, d["ref_output_lse"])
    # the reference function returns the fp32 result; the oracle rounds it once to the output dtype
    assert torch.equal(out, d["ref_output"].to(out.dtype))
    assert torch.isinf(d["prefix_lse"]).any() and torch.isinf(d["suffix_lse"]).any()  # the +inf path is exercised


def test_copy_blocks_oracle_matches_reference_loop():
    d = torch.load(os.path.join(GOLDEN, "movers_copy_blocks_bf16.pt"))
    kcs = [k.clone() for k in d["key_caches"]]
    vcs = [v.clone() for v in d["value_caches"]]
    orc.copy_blocks(kcs, vcs, d["block_mapping"])
    for a, b in zip(kcs + vcs, d["ref_key_caches"] + d["ref_value_caches"]):
        assert torch.equal(a, b)


def test_gather_cache_oracle_matches_reference_loop():
    d = torch.load(os.path.join(GOLDEN, "movers_gather_f32.pt"))
    for granular in (False, True):  # no seq_starts: both ops gather the same rows
        out = orc.gather_cache(d["src_cache"], d["block_table"], d["cu_seq_lens"], d["batch_size"],
                               token_granular_starts=granular)
        assert torch.equal(out, d["ref_dst"])


def test_gather_cache_oracle_seq_starts_semantics():
    """seq_starts: block-granular for gather_and_maybe_dequant_cache (csrc/cache_kernels.cu:660-662),
    token-granular for cp_gather_cache (:806-811)."""
    B, NB, E = 4, 6, 3
    src = torch.arange(NB * B * E, dtype=torch.float32).view(NB, B, E)
    table = torch.tensor([[5, 2, 0, 1]], dtype=torch.int32)
    cu = torch.tensor([0, 5], dtype=torch.int32)
    starts = torch.tensor([6], dtype=torch.int32)
    blockwise = orc.gather_cache(src, table, cu, 1, seq_starts=starts, token_granular_starts=False)
    exact = orc.gather_cache(src, table, cu, 1, seq_starts=starts, token_granular_starts=True)
    assert torch.equal(blockwise, torch.cat([src[2], src[0, :1]]))           # starts at token 4 = block index 1
    assert torch.equal(exact, torch.cat([src[2, 2:], src[0, :3]]))           # starts at token 6


def test_paged_attention_lse_merges_back_to_full_attention():
    """Splitting a sequence's keys in two, attending to each half and merging with merge_attn_states reproduces
    attention over the whole sequence (section 2.2 of arXiv 2501.01005) — ties the LSE oracle to the attention oracle."""
    S, H, Hkv, D, B, L, NB = 3, 8, 2, 64, 16, 90, 32
    d = orc.make_decode_inputs(S, H, Hkv, D, B, L, NB, dtype=torch.float32, seed=5, unique_blocks=True)
    lens = d["seq_lens"]
    full = orc.paged_attention_f64(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], lens, d["scale"])
    cut = (lens // 2 // B * B).to(torch.int32)  # block-aligned split point
    first = orc.paged_attention_f64(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], cut, d["scale"])
    lse1 = orc.paged_attention_lse(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], cut, d["scale"])
    bt2 = torch.stack([torch.roll(d["block_tables"][i], -int(cut[i]) // B) for i in range(S)])
    rest = lens - cut
    second = orc.paged_attention_f64(d["query"], d["key_cache"], d["value_cache"], bt2, rest, d["scale"])
    lse2 = orc.paged_attention_lse(d["query"], d["key_cache"], d["value_cache"], bt2, rest, d["scale"])
    first = torch.nan_to_num(first)  # an empty first half (cut == 0) has lse = -inf and weight 0
    out, lse = orc.merge_attn_states(first.float(), lse1, second.float(), lse2)
    full_lse = orc.paged_attention_lse(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], lens, d["scale"])
    torch.testing.assert_close(out.double(), full, atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(lse, full_lse, atol=1e-5, rtol=1e-5)


FLASH = sorted(glob.glob(os.path.join(GOLDEN, "flash_*.pt")))


@pytest.mark.parametrize("path", FLASH, ids=[os.path.basename(p)[:-3] for p in FLASH])
def test_flash_layout_oracle_matches_reference_ref_paged_attn(path):
    """oracle.paged_attention_flash against the reference's ref_paged_attn
    (tests/kernels/attention/test_triton_unified_attention.py:25-76, run by make_golden.py)."""
    assert len(FLASH) >= 2
    d = torch.load(path)
    out = orc.paged_attention_flash(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"],
                                    d["scale"])
    assert torch.equal(out, d["ref_out"])


def test_flash_and_v1_layout_oracles_agree():
    """The same keys/values stored in the v1/v2 layout and in the flash layout give the same attention (the two
    reference oracles differ only in where the scale is applied)."""
    S, H, Hkv, D, B, L, NB = 3, 8, 2, 128, 16, 100, 32
    d = orc.make_flash_decode_inputs(S, H, Hkv, D, B, L, NB, seed=9)
    x = 8
    kc = d["key_cache"].permute(0, 2, 3, 1).reshape(NB, Hkv, D // x, x, B).permute(0, 1, 2, 4, 3).contiguous()
    vc = d["value_cache"].permute(0, 2, 3, 1).contiguous()
    a = orc.paged_attention_flash(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"], d["scale"])
    b = orc.paged_attention(d["query"], kc, vc, d["block_tables"], d["seq_lens"], d["scale"])
    torch.testing.assert_close(a.float(), b.float(), atol=1e-3, rtol=1e-2)


ROTARY = sorted(glob.glob(os.path.join(GOLDEN, "rotary_*.pt")))


@pytest.mark.parametrize("path", ROTARY, ids=[os.path.basename(p)[:-3] for p in ROTARY])
def test_rotary_oracle_matches_upstream_native(path):
    """oracle.rotary_embedding against upstream vLLM's RotaryEmbedding.forward_static (the `forward_native` the
    reference's tests/kernels/core/test_pos_encoding.py compares the kernel with), run by make_golden.py."""
    assert len(ROTARY) >= 3
    d = torch.load(path)
    q, k = orc.rotary_embedding(d["positions"], d["query"], d["key"], d["head_size"], d["cos_sin_cache"], d["is_neox"])
    assert torch.equal(q, d["ref_query"]) and torch.equal(k, d["ref_key"])
    assert not torch.equal(q, d["query"])


MLA = sorted(p for p in glob.glob(os.path.join(GOLDEN, "mla_*.pt")) if "upstream" not in p)


@pytest.mark.parametrize("path", MLA, ids=[os.path.basename(p)[:-3] for p in MLA])
def test_mla_oracle_matches_reference_oracles(path):
    """oracle.mla_decode against the outputs of the reference's own `ref_mla` (test_mla_decode_cpu.py:13-40) and
    of the cutlass test's fp32 oracle (test_cutlass_mla_decode.py:133-154), tests/golden/make_golden_mla.py."""
    d = torch.load(path)
    out, lse = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
    torch.testing.assert_close(out, d["ref_out_f32"], atol=2e-6, rtol=1e-5)
    torch.testing.assert_close(lse, d["ref_lse_f32"], atol=5e-6, rtol=1e-6)
    # ref_mla runs SDPA in the cache dtype: agreement to that dtype's rounding of O(1) values
    tol = 4e-3 if d["query"].dtype == torch.bfloat16 else 5e-4
    assert (out - d["ref_out_sdpa_cpu"].float()).abs().max().item() <= tol


def test_mla_fixtures_present():
    assert len(MLA) >= 3
