"""Parity of paged_attention_v1 / paged_attention_v2 (CUDA, sm_100a) against the CPU oracle.

Grid, input distributions and oracle follow the reference's test_paged_attention
(reference: tests/kernels/attention/test_attention.py:29-47,121-345); the bar is BASELINE.json's
max-abs <= 2e-3 for bf16 output (the reference's own tolerance is atol 1e-3 for "auto", 1e-2 for fp8).
"""
import math

import pytest
import torch

from oracle import torch_oracle as orc

pytestmark = pytest.mark.gpu

PARTITION_SIZE = 512
NUM_BLOCKS = 613  # the reference uses 4321; the table entries stay random and non-unique
ATOL_AUTO = 1e-3   # reference: tests/kernels/attention/test_attention.py:342
ATOL_BAR = 2e-3    # BASELINE.json north_star


def _run(ops, version, d, dev="cuda", via_abi=False):
    t = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
    q = t["query"]
    out = torch.empty_like(q)
    ks = torch.tensor(d.get("k_scale", 1.0), dtype=torch.float32, device=dev)
    vs = torch.tensor(d.get("v_scale", 1.0), dtype=torch.float32, device=dev)
    S, H, D = q.shape
    if version == "v1":
        ops.paged_attention_v1(out, q, t["key_cache"], t["value_cache"], t["num_kv_heads"], t["scale"],
                               t["block_tables"], t["seq_lens"], t["block_size"], t["max_seq_len"],
                               t["alibi_slopes"], t["kv_cache_dtype"], ks, vs)
        extra = None
    else:
        P = (t["max_seq_len"] + PARTITION_SIZE - 1) // PARTITION_SIZE
        tmp = torch.full((S, H, P, D), float("nan"), dtype=q.dtype, device=dev)
        es = torch.full((S, H, P), float("nan"), dtype=torch.float32, device=dev)
        ml = torch.full_like(es, float("nan"))
        ops.paged_attention_v2(out, es, ml, tmp, q, t["key_cache"], t["value_cache"], t["num_kv_heads"], t["scale"],
                               t["block_tables"], t["seq_lens"], t["block_size"], t["max_seq_len"],
                               t["alibi_slopes"], t["kv_cache_dtype"], ks, vs)
        extra = (es.cpu(), ml.cpu(), tmp.cpu())
    torch.cuda.synchronize()
    return out.cpu(), extra


def _oracle(d):
    return orc.paged_attention(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"],
                               d["scale"], d["alibi_slopes"], d["kv_cache_dtype"], d.get("k_scale", 1.0),
                               d.get("v_scale", 1.0))


def _check(out, ref, atol, rtol=1e-5):
    err = (out.float() - ref.float()).abs().max().item()
    assert math.isfinite(err)
    torch.testing.assert_close(out.float(), ref.float(), atol=atol, rtol=rtol)
    return err


# ---- the reference grid ------------------------------------------------------------------------
@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("num_heads", [(40, 40), (64, 8)])
@pytest.mark.parametrize("head_size", [32, 80, 128, 256])
@pytest.mark.parametrize("use_alibi", [False, True])
@pytest.mark.parametrize("block_size", [16, 32])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
def test_reference_grid(ops, version, num_heads, head_size, use_alibi, block_size, kv_cache_dtype):
    if kv_cache_dtype == "fp8" and head_size % 16:
        pytest.skip("fp8 cache needs head_size % 16 == 0 (reference test skips too)")
    H, Hkv = num_heads
    d = orc.make_decode_inputs(7, H, Hkv, head_size, block_size, max_seq_len=1571, num_blocks=NUM_BLOCKS,
                               kv_cache_dtype=kv_cache_dtype, seed=0, use_alibi=use_alibi)
    out, _ = _run(ops, version, d)
    _check(out, _oracle(d), ATOL_AUTO if kv_cache_dtype == "auto" else ATOL_BAR)


# ---- named configurations (BASELINE.json configs, at sizes the oracle finishes in seconds) -----------
@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("G,Hkv", [(1, 32), (4, 8), (8, 8), (2, 4), (16, 2), (5, 2)])
@pytest.mark.parametrize("block_size", [16, 32])
def test_stream_variant_shapes(ops, version, kv_cache_dtype, G, Hkv, block_size):
    """D=128 bf16 shapes served by the streaming kernels: TMA ring + mma.sync (GQA 1..16), TMA ring + tcgen05 (fp8)."""
    from vllm_metax_b200 import _native

    d = orc.make_decode_inputs(5, G * Hkv, Hkv, 128, block_size, max_seq_len=1300, num_blocks=2048,
                               kv_cache_dtype=kv_cache_dtype, seed=3)
    out, _ = _run(ops, version, d)
    assert _native.last_attention_variant() == _native.VARIANT_STREAM_MMA
    _check(out, _oracle(d), ATOL_AUTO if kv_cache_dtype == "auto" else ATOL_BAR)
    if kv_cache_dtype == "fp8" and block_size == 16 and G <= 8:  # shapes the opt-in tcgen05 kernel serves too
        out2, _ = _run_umma(ops, version, d)
        _check(out2, _oracle(d), ATOL_BAR)


def _run_umma(ops, version, d):
    """Run with the tcgen05 fp8 kernel forced (it is opt-in: the mma.sync kernel is the default)."""
    from vllm_metax_b200 import _native

    _native.set_attention_variant(_native.VARIANT_UMMA_FP8)
    try:
        r = _run(ops, version, d)
        assert _native.last_attention_variant() == _native.VARIANT_UMMA_FP8
    finally:
        _native.set_attention_variant(_native.VARIANT_AUTO)
    return r


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("L", [1, 15, 16, 17, 31, 32, 33, 511, 512, 513, 1024, 1025, 4096])
def test_seq_len_edges(ops, version, kv_cache_dtype, L):
    """L = 1 exposes the +1e-6 normaliser; k*B and k*B +- 1; partition boundaries."""
    d = orc.make_decode_inputs(3, 32, 8, 128, 16, max_seq_len=L, num_blocks=1024, kv_cache_dtype=kv_cache_dtype,
                               seed=L, fixed_len=True)
    out, _ = _run(ops, version, d)
    _check(out, _oracle(d), ATOL_AUTO if kv_cache_dtype == "auto" else ATOL_BAR)


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("variant", ["generic", "stream"])
def test_stale_cache_tail_is_ignored(ops, version, variant):
    """Slots past seq_len in the last block may hold NaN/Inf from earlier use
    (reference zeroes them: csrc/attention/attention_kernels.cuh:409-419)."""
    from vllm_metax_b200 import _native

    d = orc.make_decode_inputs(4, 32, 8, 128, 16, max_seq_len=777, num_blocks=512, seed=5, unique_blocks=True)
    ref = _oracle(d)
    kc, vc = d["key_cache"].clone(), d["value_cache"].clone()
    for s in range(4):
        L = int(d["seq_lens"][s])
        blk = int(d["block_tables"][s, (L - 1) // 16])
        off = L % 16
        if off:
            kc[blk, :, :, off:, :] = float("nan")
            vc[blk, :, :, off:] = float("inf")
    d2 = dict(d, key_cache=kc, value_cache=vc)
    _native.set_attention_variant(_native.VARIANT_GENERIC if variant == "generic" else _native.VARIANT_STREAM_MMA)
    try:
        out, _ = _run(ops, version, d2)
    finally:
        _native.set_attention_variant(_native.VARIANT_AUTO)
    _check(out, ref, ATOL_AUTO)


@pytest.mark.parametrize("version", ["v1", "v2"])
def test_generic_matches_stream(ops, version):
    from vllm_metax_b200 import _native

    d = orc.make_decode_inputs(6, 32, 8, 128, 16, max_seq_len=2111, num_blocks=2048, seed=11, use_alibi=True)
    outs = []
    for v in (_native.VARIANT_GENERIC, _native.VARIANT_STREAM_MMA):
        _native.set_attention_variant(v)
        try:
            outs.append(_run(ops, version, d)[0])
            assert _native.last_attention_variant() == v
        finally:
            _native.set_attention_variant(_native.VARIANT_AUTO)
    _check(outs[0], outs[1], 5e-4)
    _check(outs[1], _oracle(d), ATOL_AUTO)


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32])
@pytest.mark.parametrize("version", ["v1", "v2"])
def test_other_dtypes(ops, version, dtype):
    """half and float queries/caches (reference dispatch: csrc/quantization/fp8/metax/quant_utils.cuh:29-39)."""
    d = orc.make_decode_inputs(4, 16, 4, 64, 16, max_seq_len=700, num_blocks=512, dtype=dtype, seed=2)
    out, _ = _run(ops, version, d)
    _check(out, _oracle(d), ATOL_AUTO)


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("head_size", [64, 96, 112, 120, 192])
@pytest.mark.parametrize("block_size", [8, 16])
def test_remaining_head_and_block_sizes(ops, version, head_size, block_size):
    d = orc.make_decode_inputs(3, 8, 2, head_size, block_size, max_seq_len=600, num_blocks=512, seed=4)
    out, _ = _run(ops, version, d)
    _check(out, _oracle(d), ATOL_AUTO)


@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
def test_strided_query_and_nonunit_scales(ops, kv_cache_dtype):
    """q is a slice of a fused qkv tensor (q_stride > H*D, csrc/attention/attention_kernels.cuh:166-168);
    fp8 with k_scale, v_scale != 1."""
    d = orc.make_decode_inputs(5, 32, 8, 128, 16, max_seq_len=900, num_blocks=1024, kv_cache_dtype=kv_cache_dtype,
                               seed=9)
    if kv_cache_dtype == "fp8":
        d["k_scale"], d["v_scale"] = 0.0123, 0.0457
    ref = _oracle(d)
    S, H, D = d["query"].shape
    qkv = torch.zeros(S, 3 * H * D, dtype=d["query"].dtype)
    qkv[:, : H * D] = d["query"].reshape(S, H * D)
    dev = dict(d)
    dev["query"] = qkv.cuda()[:, : H * D].view(S, H, D)
    assert dev["query"].stride(0) == 3 * H * D
    out, _ = _run(ops, "v1", dev)
    # fp8: the kernel folds the scales (no per-element rounding to bf16), the oracle dequantises first
    _check(out, ref, ATOL_AUTO if kv_cache_dtype == "auto" else ATOL_BAR, rtol=2e-2 if kv_cache_dtype == "fp8" else 1e-5)
    out2, _ = _run(ops, "v2", dev)
    _check(out2, ref, ATOL_AUTO if kv_cache_dtype == "auto" else ATOL_BAR, rtol=2e-2 if kv_cache_dtype == "fp8" else 1e-5)


def test_v2_scratch_has_reference_semantics(ops):
    """exp_sums / max_logits / tmp_out hold what the reference v2 kernel leaves there
    (csrc/attention/attention_kernels.cuh:337-346,472-484): per 512-token partition the max logit,
    sum exp(logit - max) and the partition-normalised output."""
    d = orc.make_decode_inputs(3, 8, 2, 128, 16, max_seq_len=1400, num_blocks=512, seed=21)
    out, (es, ml, tmp) = _run(ops, "v2", d)
    q = d["query"].float()
    for s in range(3):
        L = int(d["seq_lens"][s])
        k, v = orc.gather_kv(d["key_cache"], d["value_cache"], d["block_tables"][s].tolist(), L)
        k = torch.repeat_interleave(k.float(), 4, dim=1)
        v = torch.repeat_interleave(v.float(), 4, dim=1)
        logits = d["scale"] * torch.einsum("hd,khd->hk", q[s], k)
        for p in range((L + 511) // 512):
            lg = logits[:, p * 512:(p + 1) * 512]
            m = lg.max(-1).values
            e = torch.exp(lg - m[:, None])
            torch.testing.assert_close(ml[s, :, p], m, atol=1e-4, rtol=1e-4)
            torch.testing.assert_close(es[s, :, p], e.sum(-1), atol=1e-2, rtol=2e-3)
            o = torch.einsum("hk,khd->hd", e, v[p * 512:(p + 1) * 512]) / (e.sum(-1, keepdim=True) + 1e-6)
            torch.testing.assert_close(tmp[s, :, p].float(), o, atol=1e-3, rtol=1e-2)


def test_peaky_softmax_randn(ops):
    """N(0,1) q/k/v: logits ~ N(0,1), outputs O(0.1): checks the online-softmax rescaling."""
    g = torch.Generator().manual_seed(0)
    d = orc.make_decode_inputs(4, 32, 8, 128, 16, max_seq_len=3000, num_blocks=2048, seed=13)
    d["query"] = (torch.randn(d["query"].shape, generator=g) * 3).to(torch.bfloat16)
    d["key_cache"] = torch.randn(d["key_cache"].shape, generator=g).to(torch.bfloat16)
    d["value_cache"] = torch.randn(d["value_cache"].shape, generator=g).to(torch.bfloat16)
    ref64 = orc.paged_attention_f64(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"],
                                    d["scale"])
    for version in ("v1", "v2"):
        out, _ = _run(ops, version, d)
        torch.testing.assert_close(out.double(), ref64, atol=2e-3, rtol=1.6e-2)


def test_via_c_abi_matches_torch_op(ops):
    """Call the C ABI directly (raw pointers through ctypes), as a non-torch host would."""
    from vllm_metax_b200 import _native

    d = orc.make_decode_inputs(6, 32, 8, 128, 16, max_seq_len=2500, num_blocks=2048, seed=17)
    ref_out, _ = _run(ops, "v1", d)
    t = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
    out = torch.empty_like(t["query"])
    one = torch.ones(1, dtype=torch.float32, device="cuda")
    p = _native.attn_params(out, t["query"], t["key_cache"], t["value_cache"], 8, d["scale"], t["block_tables"],
                            t["seq_lens"], 16, d["max_seq_len"], None, "auto", one, one)
    ws = torch.zeros(max(16, _native.workspace_bytes(p)), dtype=torch.uint8, device="cuda")  # zero-filled once
    p.workspace, p.workspace_bytes = ws.data_ptr(), ws.numel()
    _native.paged_attention_v1(p, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert torch.equal(out.cpu(), ref_out)


def test_errors(ops):
    d = orc.make_decode_inputs(2, 8, 2, 128, 16, max_seq_len=64, num_blocks=64, seed=1)
    t = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
    out = torch.empty_like(t["query"])
    one = torch.ones(1, dtype=torch.float32, device="cuda")
    args = [out, t["query"], t["key_cache"], t["value_cache"], 2, d["scale"], t["block_tables"], t["seq_lens"], 16, 64,
            None, "auto", one, one]
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        ops.paged_attention_v1(*args[:11], "fp8_e5m2", one, one)
    with pytest.raises(RuntimeError, match="Unsupported block size"):
        bad = list(args)
        bad[8] = 64
        bad[3] = t["value_cache"].reshape(-1, 2, 128, 64)[:16]
        bad[2] = t["key_cache"].reshape(-1, 2, 16, 64, 8)[:16]
        ops.paged_attention_v1(*bad)
    with pytest.raises(RuntimeError, match="block-sparse"):
        ops.paged_attention_v1(*args, 0, 16, 8, 64, 0)


def test_v1_workspace_is_reusable_across_shapes(ops):
    """The shim's v1 scratch is shared by calls of different shapes on a stream: the arrival counters of the
    fused split-KV reduce must not land on a previous call's partials."""
    shapes = [(7, 40, 40, 1571), (3, 8, 2, 4000), (9, 64, 8, 700), (2, 16, 16, 2600), (7, 40, 40, 1571)]
    for i, (S, H, Hkv, L) in enumerate(shapes):
        d = orc.make_decode_inputs(S, H, Hkv, 128, 16, max_seq_len=L, num_blocks=NUM_BLOCKS, seed=i, fixed_len=True)
        out, _ = _run(ops, "v1", d)
        _check(out, _oracle(d), ATOL_AUTO)


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("H,Hkv,block_size,kv_cache_dtype",
                         [(40, 40, 32, "auto"), (64, 8, 32, "auto"), (40, 40, 16, "auto"), (64, 8, 16, "fp8")])
def test_streaming_kernel_is_deterministic(ops, version, H, Hkv, block_size, kv_cache_dtype):
    """Regression: a ring stage was handed back to the TMA producer while the consumer's last ldmatrix reads
    (bank-conflicted V rows at block_size 32, bf16) were still queued; 60 launches must be bit-identical."""
    d = orc.make_decode_inputs(7, H, Hkv, 128, block_size, max_seq_len=1571, num_blocks=NUM_BLOCKS,
                               kv_cache_dtype=kv_cache_dtype, seed=0)
    first, _ = _run(ops, version, d)
    _check(first, _oracle(d), ATOL_AUTO if kv_cache_dtype == "auto" else ATOL_BAR)
    for _ in range(60):
        again, _ = _run(ops, version, d)
        assert torch.equal(again, first)


# ---- tcgen05 fp8 kernel: cases aimed at its own machinery -------------------------------------------------
@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("G,Hkv", [(4, 8), (8, 1), (1, 4), (3, 2)])
@pytest.mark.parametrize("k_scale,v_scale", [(1.0, 1.0), (0.037, 0.21), (2.5, 0.5)])
def test_umma_fp8_scales_and_dynamic_range(ops, version, G, Hkv, k_scale, v_scale):
    """Non-unit scales; k_scale 2.5 on full-range e4m3 keys gives logits spread over tens of units, so the classes
    re-reference (rescale their TMEM accumulators) many times."""
    from vllm_metax_b200 import _native

    d = orc.make_decode_inputs(4, G * Hkv, Hkv, 128, 16, max_seq_len=2500, num_blocks=1024, kv_cache_dtype="fp8", seed=21)
    g = torch.Generator().manual_seed(5)
    d["key_cache"] = torch.randint(0, 0x7f, d["key_cache"].shape, generator=g, dtype=torch.uint8) | (
        torch.randint(0, 2, d["key_cache"].shape, generator=g, dtype=torch.uint8) << 7)
    d["key_cache"][(d["key_cache"] & 0x7f) > 0x50] &= 0xcf  # keep |k| <= 8: logits stay finite in every implementation
    d["query"] = (torch.randn(d["query"].shape, generator=g) * 0.7).to(torch.bfloat16)
    d["k_scale"], d["v_scale"] = k_scale, v_scale
    out, _ = _run_umma(ops, version, d)
    _check(out, _oracle(d), ATOL_BAR)


@pytest.mark.parametrize("version", ["v1", "v2"])
def test_umma_fp8_rising_logits_alibi(ops, version):
    """ALiBi with large slopes makes the logits grow monotonically along the sequence: worst case for the lazy reference."""
    from vllm_metax_b200 import _native

    d = orc.make_decode_inputs(3, 16, 4, 128, 16, max_seq_len=1777, num_blocks=512, kv_cache_dtype="fp8", seed=8, use_alibi=True)
    d["alibi_slopes"] = torch.linspace(0.01, 0.6, 16)
    out, _ = _run_umma(ops, version, d)
    _check(out, _oracle(d), ATOL_BAR)


@pytest.mark.parametrize("version", ["v1", "v2"])
def test_umma_fp8_stale_tail_nan_bytes(ops, version):
    """e4m3 NaN bytes (0x7f / 0xff) past seq_len in the last block, and in the ring's second block slot, must not leak."""
    from vllm_metax_b200 import _native

    d = orc.make_decode_inputs(5, 32, 8, 128, 16, max_seq_len=333, num_blocks=256, kv_cache_dtype="fp8", seed=9, unique_blocks=True)
    ref = _oracle(d)
    kc, vc = d["key_cache"].clone(), d["value_cache"].clone()
    for s in range(5):
        L = int(d["seq_lens"][s])
        blk = int(d["block_tables"][s, (L - 1) // 16])
        off = L % 16
        if off:
            kc[blk, :, :, off:, :] = 0x7f
            vc[blk, :, :, off:] = 0xff
    out, _ = _run_umma(ops, version, dict(d, key_cache=kc, value_cache=vc))
    _check(out, ref, ATOL_BAR)


def test_umma_fp8_deterministic(ops):
    d = orc.make_decode_inputs(64, 32, 8, 128, 16, max_seq_len=1500, num_blocks=8192, kv_cache_dtype="fp8", seed=4, unique_blocks=True)
    a, _ = _run_umma(ops, "v1", d)
    for _ in range(5):
        b, _ = _run_umma(ops, "v1", d)
        assert torch.equal(a, b)


@pytest.mark.parametrize("min_chunk", [2, 8])
@pytest.mark.parametrize("shape", [(64, 32, 8, 16, "auto", 2100), (40, 16, 16, 32, "fp8", 3000), (9, 64, 8, 16, "auto", 9000),
                                   (200, 8, 1, 16, "fp8", 700)])
def test_flat_schedule_dynamic_tail_region(ops, min_chunk, shape):
    """The opt-in dynamic tail of the flat schedule (b200_set_stream_dynamic; off by default because it measured slower):
    chunks claimed through a device counter, partial slots indexed by range.  Same results as the oracle, counters and
    claim counter left at zero (a second launch on the same workspace agrees), v1 and the flash layout."""
    from vllm_metax_b200 import _native

    S, H, Hkv, B, kvd, L = shape
    d = orc.make_decode_inputs(S, H, Hkv, 128, B, max_seq_len=L, num_blocks=max(2048, S * ((L + B - 1) // B)),
                               kv_cache_dtype=kvd, seed=S + L, unique_blocks=False)
    ref = _oracle(d)
    _native.set_stream_dynamic(min_chunk)
    try:
        for _ in range(3):
            out, _ = _run(ops, "v1", d)
            _check(out, ref, ATOL_AUTO if kvd == "auto" else ATOL_BAR)
    finally:
        _native.set_stream_dynamic(-1)
    out, _ = _run(ops, "v1", d)  # the static schedule on the same workspace afterwards
    _check(out, ref, ATOL_AUTO if kvd == "auto" else ATOL_BAR)
