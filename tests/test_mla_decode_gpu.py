"""MLA decode attention (csrc/mla_decode.cu) against the oracle and the reference-generated goldens.
Template: tests/kernels/attention/test_cutlass_mla_decode.py (cos-diff < 1e-5 on out, < 1e-3 on lse) and
tests/kernels/attention/test_mla_decode_cpu.py (assert_close against ref_mla); the max-abs bound 2e-3 of
BASELINE.json's north_star is asserted on top."""
import glob
import os

import pytest
import torch

from oracle import torch_oracle as orc

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops():
    import vllm_metax_b200
    vllm_metax_b200.import_kernels()
    from vllm_metax_b200 import _custom_ops
    return _custom_ops


def cos_diff(x, y):
    x, y = x.double(), y.double()
    return 1 - 2 * (x * y).sum().item() / max((x * x + y * y).sum().item(), 1e-12)


def run(ops, d, num_kv_splits=-1, via="sm100", heads_pad=None):
    q = d["query"].to(DEV)
    B, H, _ = q.shape
    q_nope, q_pe = q[..., :512].contiguous(), q[..., 512:].contiguous()
    cache, table, lens = d["kv_cache"].to(DEV), d["block_tables"].to(DEV), d["seq_lens"].to(DEV)
    out = torch.full((B, H, 512), float("nan"), dtype=q.dtype, device=DEV)
    lse = torch.full((B, H), float("nan"), dtype=torch.float32, device=DEV)
    if via == "reference_schema":
        ops.cutlass_mla_decode(out, q_nope, q_pe, cache, lens, table, d["scale"])
        return out.cpu(), None
    ws = torch.empty(ops.sm100_cutlass_mla_get_workspace_size(table.shape[1] * cache.shape[1], B, 148, num_kv_splits),
                     dtype=torch.uint8, device=DEV)
    ops.sm100_cutlass_mla_decode(out, lse, q_nope, q_pe, cache, lens, table, ws, d["scale"], num_kv_splits)
    torch.cuda.synchronize()
    return out.cpu(), lse.cpu()


def check(out, lse, ref_out, ref_lse, dtype):
    assert torch.isfinite(out.float()).all()
    assert cos_diff(out.float(), ref_out) < 1e-5
    # the reference's own bars: cos-diff above, and assert_close at the dtype's default tolerances
    # (tests/kernels/attention/test_mla_decode_cpu.py:93); randn data gives O(1) outputs on short sequences, where
    # the 16-bit probabilities the tensor core consumes (as in the reference kernel) cost up to ~3e-3 absolute
    if dtype == torch.float16:
        torch.testing.assert_close(out.float(), ref_out, atol=1e-3, rtol=1e-3)
    else:
        torch.testing.assert_close(out.float(), ref_out, atol=4e-3, rtol=1.6e-2)
    if lse is not None:
        finite = torch.isfinite(ref_lse)
        assert torch.equal(torch.isfinite(lse), finite)   # empty sequences: -inf on both sides
        assert cos_diff(lse[finite], ref_lse[finite]) < 1e-3
        torch.testing.assert_close(lse[finite], ref_lse[finite], atol=2e-3, rtol=1e-4)


@pytest.mark.parametrize("path", sorted(p for p in glob.glob(os.path.join(GOLDEN, "mla_*.pt")) if "upstream" not in p),
                         ids=lambda p: os.path.basename(p)[:-3])
def test_goldens(ops, path):
    d = torch.load(path)
    out, lse = run(ops, d)
    check(out, lse, d["ref_out_f32"], d["ref_lse_f32"], d["query"].dtype)


@pytest.mark.parametrize("H", [16, 32, 64, 128, 8, 40, 96])
@pytest.mark.parametrize("page", [16, 64, 128, 256])
@pytest.mark.parametrize("L", [1, 17, 128, 129, 700, 2100])
def test_grid(ops, H, page, L):
    if H in (8, 40, 96) and (page != 64 or L not in (17, 700)):
        pytest.skip("odd head counts: two lengths on one page size")
    d = orc.make_mla_inputs(3, H, page, L, seed=H * 1000 + page + L)
    ref_out, ref_lse = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
    out, lse = run(ops, d)
    check(out, lse, ref_out, ref_lse, torch.bfloat16)


def test_small_magnitude_data_meets_the_2e_3_bar(ops):
    """north_star's max-abs 2e-3 on data scaled like the reference's attention tests (|x| <= D^-0.5 ... here std 0.2)"""
    d = orc.make_mla_inputs(4, 128, 64, 900, seed=11, std=0.2)
    d["scale"] = 1.0
    ref_out, ref_lse = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
    out, lse = run(ops, d)
    assert (out.float() - ref_out).abs().max().item() <= 2e-3
    torch.testing.assert_close(lse, ref_lse, atol=2e-3, rtol=1e-4)


@pytest.mark.parametrize("splits", [1, 2, 5, -1])
def test_token_splits(ops, splits):
    d = orc.make_mla_inputs(2, 32, 64, 3000, seed=splits + 50)
    ref_out, ref_lse = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
    out, lse = run(ops, d, num_kv_splits=splits)
    check(out, lse, ref_out, ref_lse, torch.bfloat16)


def test_reference_schema_f16_and_empty_sequence(ops):
    d = orc.make_mla_inputs(4, 128, 64, 500, dtype=torch.float16, seed=9, min_len=0)
    d["seq_lens"][1] = 0
    ref_out, _ = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
    out, _ = run(ops, d, via="reference_schema")
    check(out, None, ref_out, None, torch.float16)
    assert (out[1] == 0).all()


def test_large_logits_trigger_rereference(ops):
    """logits grow along the sequence by far more than 2^16: the lazy reference must be raised and O^T rescaled"""
    d = orc.make_mla_inputs(2, 16, 64, 1500, seed=77)
    ramp = torch.linspace(0.0, 6.0, 1500 + 200)
    cache = d["kv_cache"].float()
    flat_pos = torch.zeros(cache.shape[0], cache.shape[1])
    for s in range(2):
        for j, pg in enumerate(d["block_tables"][s].tolist()):
            flat_pos[pg] = ramp[j * 64:(j + 1) * 64] if (j + 1) * 64 <= ramp.numel() else 0
    d["kv_cache"] = (cache * (1 + flat_pos.unsqueeze(-1) * 8)).to(torch.bfloat16)
    d["scale"] = 0.5
    ref_out, ref_lse = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
    out, lse = run(ops, d, num_kv_splits=1)
    assert torch.isfinite(out.float()).all()
    assert cos_diff(out.float(), ref_out) < 1e-4
    torch.testing.assert_close(lse, ref_lse, atol=5e-2, rtol=1e-3)


def test_stale_rows_past_the_sequence_do_not_leak(ops):
    d = orc.make_mla_inputs(2, 16, 64, 70, seed=5, varlen=False)
    ref_out, ref_lse = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
    cache = d["kv_cache"].clone()
    for s in range(2):
        pg = int(d["block_tables"][s, 1])
        cache[pg, 6:] = float("nan")      # rows 70.. of the sequence (inside its last page)
    d["kv_cache"] = cache
    out, lse = run(ops, d)
    check(out, lse, ref_out, ref_lse, torch.bfloat16)


@pytest.mark.parametrize("page,splits", [(16, 4), (64, 1)])
def test_triton_decode_attention_fwd_call_shape(ops, page, splits):
    """the Triton MLA backend's call (vllm_metax/v1/attention/backends/mla/triton_mla.py:160-200) on our kernel"""
    from vllm_metax_b200.attention.mla_decode import decode_attention_fwd
    d = orc.make_mla_inputs(3, 16, page, 900, seed=21)
    ref_out, ref_lse = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
    q = d["query"].to(DEV)
    cache = d["kv_cache"].to(DEV).unsqueeze(2)
    o = torch.zeros(3, 16, 512, dtype=q.dtype, device=DEV)
    lse = torch.zeros(3, 16, dtype=q.dtype, device=DEV)
    attn_logits = torch.empty(3, 16, splits, 513, dtype=torch.float32, device=DEV)
    decode_attention_fwd(q, cache, cache[..., :512], o, lse, d["block_tables"].to(DEV), d["seq_lens"].to(DEV), attn_logits,
                         splits, d["scale"], page)
    check(o.cpu(), None, ref_out, None, torch.bfloat16)
    torch.testing.assert_close(lse.float().cpu(), ref_lse, atol=5e-2, rtol=1e-2)  # lse in the query dtype, as the caller allocates it


def _ragged(bs, H, page, lens, seed):
    d = orc.make_mla_inputs(bs, H, page, max(max(lens), 1), seed=seed, varlen=False)
    d["seq_lens"] = torch.tensor(lens, dtype=torch.int32)
    return d


@pytest.mark.parametrize("H", [16, 64, 128])
def test_flat_schedule_ragged_batch_with_empty_sequences(ops, H):
    """more items than SMs, lengths from 0 to a few tiles: the persistent schedule walks several items per CTA (or CTA
    pair), empty sequences get zeros / -inf, and whichever dealing mode the kernel picks the result is the oracle's"""
    import random
    rnd = random.Random(H)
    lens = [rnd.choice([0, 1, 17, 128, 129, 300, 700, 1100]) for _ in range(200)]
    lens[0], lens[7] = 0, 1100
    d = _ragged(200, H, 64, lens, seed=H + 1)
    ref_out, ref_lse = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
    out, lse = run(ops, d)
    check(out, lse, ref_out, ref_lse, torch.bfloat16)
    empty = [i for i, n in enumerate(lens) if n == 0]
    assert (out[empty] == 0).all() and torch.isinf(lse[empty]).all()


@pytest.mark.parametrize("H", [16, 128])
def test_flat_schedule_one_long_sequence_among_short_ones(ops, H):
    """one 20k-token sequence among 1-tile ones: range mode, the long item is cut across many CTAs (partial slots +
    mla_merge_flat_kernel), the short ones are whole items inside a CTA's range"""
    lens = [20000] + [100] * 30 + [0] + [257] * 8
    d = _ragged(len(lens), H, 64, lens, seed=H + 9)
    ref_out, ref_lse = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
    out, lse = run(ops, d)
    check(out, lse, ref_out, ref_lse, torch.bfloat16)


@pytest.mark.parametrize("H", [16, 32, 64, 128])
@pytest.mark.parametrize("L", [100, 1500, 5000])
def test_fp8_e4m3_inputs(ops, H, L):
    """the float8_e4m3fn form of the reference op (tests/kernels/attention/test_cutlass_mla_decode.py:53-58,91-101):
    q and cache quantised to e4m3 (descale 1), bf16 output; the reference's bar for it is cos-diff < 1e-4 (out) and
    < 1e-3 (lse) against the fp32 oracle on the dequantised values"""
    d = orc.make_mla_inputs(3, H, 64, L, seed=H + L)
    q8 = d["query"].to(torch.float8_e4m3fn)
    c8 = d["kv_cache"].to(torch.float8_e4m3fn)
    ref_out, ref_lse = orc.mla_decode(q8.float(), c8.float(), d["scale"], d["block_tables"], d["seq_lens"])
    qd, cd = q8.to(DEV), c8.to(DEV)
    out = torch.full((3, H, 512), float("nan"), dtype=torch.bfloat16, device=DEV)
    lse = torch.full((3, H), float("nan"), dtype=torch.float32, device=DEV)
    ws = torch.empty(ops.sm100_cutlass_mla_get_workspace_size(d["block_tables"].shape[1] * 64, 3, 148, -1), dtype=torch.uint8,
                     device=DEV)
    ops.sm100_cutlass_mla_decode(out, lse, qd[..., :512].contiguous(), qd[..., 512:].contiguous(), cd, d["seq_lens"].to(DEV),
                                 d["block_tables"].to(DEV), ws, d["scale"], -1)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
    assert cos_diff(out.float().cpu(), ref_out) < 1e-5        # two-term probabilities: far inside the reference's 1e-4
    assert cos_diff(lse.cpu(), ref_lse) < 1e-3
    torch.testing.assert_close(out.float().cpu(), ref_out, atol=2e-2, rtol=2e-2)
    torch.testing.assert_close(lse.cpu(), ref_lse, atol=2e-3, rtol=1e-4)
