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
