"""The drop-in boundary under the dispatcher's own checks (VERDICT r1 "missing" #5, "weak" #11).

* torch.library.opcheck on the thirteen registered reference ops with the reference's argument tuples
  (tests/kernels/attention/test_attention.py:217-222,261-267; tests/kernels/attention/test_cache.py:100,177-180,
  301-304,415,553-557,635-639; tests/kernels/core/test_opcheck.py:12-15), test utilities as in tests/kernels/utils.py:27-38;
* CUDA-graph capture + replay of paged_attention_v1 / _v2 (bf16 and fp8) on shapes that split (seq, kv head) pairs across
  CTAs, so the arrival counters of the fused reduce live inside the graph; inputs change between replays;
* workspace rules: reserve_workspace, the error on first use under capture, the reset entry point;
* empty sequences without an LSE.
"""
import pytest
import torch

from oracle import torch_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda"
DEFAULT_UTILS = ("test_schema", "test_autograd_registration", "test_faketensor")          # tests/kernels/utils.py:27-31
ALL_UTILS = DEFAULT_UTILS + ("test_aot_dispatch_dynamic",)                                 # tests/kernels/utils.py:33-38


def _allclose_via_double(a, b, rtol=1e-05, atol=1e-08, equal_nan=False):
    """torch.allclose has no float8 kernels; the reference's opcheck wrapper swaps in a comparison in double for the
    duration of the check (tests/kernels/utils.py:1022-1044,1211)."""
    return bool(torch.all(torch.isclose(a.double(), b.double(), rtol=rtol, atol=atol, equal_nan=equal_nan)).item())


def _opcheck(op, args, utils=ALL_UTILS):
    import unittest.mock

    with unittest.mock.patch("torch.allclose", new=_allclose_via_double):
        torch.library.opcheck(op, args, test_utils=utils, raise_exception=True)


def _decode_args(kvd="auto", S=3, H=8, Hkv=2, D=32, B=16, L=300):
    d = orc.make_decode_inputs(S, H, Hkv, D, B, max_seq_len=L, num_blocks=128, kv_cache_dtype=kvd, seed=0, device=DEV)
    one = torch.ones(1, dtype=torch.float32, device=DEV)
    return d, one


@pytest.mark.parametrize("kvd", ["auto", "fp8"])
def test_opcheck_paged_attention_v1(ops, kvd):
    d, one = _decode_args(kvd)
    out = torch.empty_like(d["query"])
    _opcheck(torch.ops._C.paged_attention_v1,
             (out, d["query"], d["key_cache"], d["value_cache"], 2, d["scale"], d["block_tables"], d["seq_lens"], 16, 300, None,
              kvd, one, one, 0, 0, 0, 64, 0))


@pytest.mark.parametrize("kvd", ["auto", "fp8"])
def test_opcheck_paged_attention_v2(ops, kvd):
    d, one = _decode_args(kvd, L=1300)
    q = d["query"]
    S, H, D = q.shape
    P = (1300 + 511) // 512
    out = torch.empty_like(q)
    es = torch.empty(S, H, P, dtype=torch.float32, device=DEV)
    ml, tmp = torch.empty_like(es), torch.empty(S, H, P, D, dtype=q.dtype, device=DEV)
    _opcheck(torch.ops._C.paged_attention_v2,
             (out, es, ml, tmp, q, d["key_cache"], d["value_cache"], 2, d["scale"], d["block_tables"], d["seq_lens"], 16, 1300,
              None, kvd, one, one, 0, 0, 0, 64, 0))


@pytest.mark.parametrize("kvd", ["auto", "fp8"])
@pytest.mark.parametrize("flash", [False, True])
def test_opcheck_reshape_and_cache(ops, kvd, flash):
    T, Hkv, D, B, NB = 42, 8, 64, 8, 64
    g = torch.Generator().manual_seed(0)
    key = torch.randn(T, Hkv, D, generator=g).to(torch.bfloat16).to(DEV)
    value = torch.randn(T, Hkv, D, generator=g).to(torch.bfloat16).to(DEV)
    slots = torch.randperm(NB * B, generator=g)[:T].to(DEV)
    cdt = torch.bfloat16 if kvd == "auto" else torch.uint8
    x = 16 // torch.tensor([], dtype=cdt).element_size()
    scale = torch.tensor([0.05], dtype=torch.float32, device=DEV)
    if flash:
        kc = torch.zeros(NB, B, Hkv, D, dtype=cdt, device=DEV)
        vc = torch.zeros_like(kc)
        _opcheck(torch.ops._C_cache_ops.reshape_and_cache_flash, (key, value, kc, vc, slots, kvd, scale, scale))
    else:
        kc = torch.zeros(NB, Hkv, D // x, B, x, dtype=cdt, device=DEV)
        vc = torch.zeros(NB, Hkv, D, B, dtype=cdt, device=DEV)
        _opcheck(torch.ops._C_cache_ops.reshape_and_cache, (key, value, kc, vc, slots, kvd, scale, scale))


def test_opcheck_convert_fp8(ops):
    """tests/kernels/core/test_opcheck.py:12-15 (a float8_e4m3fn destination)."""
    data = torch.randn((256, 256), dtype=torch.float32, device=DEV)
    result = torch.empty_like(data, dtype=torch.float8_e4m3fn)
    _opcheck(torch.ops._C_cache_ops.convert_fp8, (result, data, 1.0, "fp8"))
    back = torch.empty_like(data)
    _opcheck(torch.ops._C_cache_ops.convert_fp8, (back, result.view(torch.uint8), 1.0, "fp8"))


def test_opcheck_merge_attn_states_and_rotary(ops):
    N, H, D = 64, 8, 128
    p_out = torch.randn(N, H, D, dtype=torch.bfloat16, device=DEV)
    s_out = torch.randn(N, H, D, dtype=torch.bfloat16, device=DEV)
    p_lse, s_lse = torch.randn(H, N, device=DEV), torch.randn(H, N, device=DEV)
    out, lse = torch.empty_like(p_out), torch.empty(H, N, device=DEV)
    _opcheck(torch.ops._C.merge_attn_states, (out, lse, p_out, p_lse, s_out, s_lse), DEFAULT_UTILS)
    _opcheck(torch.ops._C.merge_attn_states, (out, None, p_out, p_lse, s_out, s_lse), DEFAULT_UTILS)
    # rotary_embedding (tests/kernels/core/test_rotary_embedding.py: opcheck with positions, query, key, head_size, cache, is_neox)
    T, hs = 11, 64
    pos = torch.randint(0, 100, (T,), device=DEV)
    q = torch.randn(T, 4 * hs, dtype=torch.bfloat16, device=DEV)
    k = torch.randn(T, 2 * hs, dtype=torch.bfloat16, device=DEV)
    cache = orc.compute_cos_sin_cache(hs, 128, dtype=torch.bfloat16).to(DEV)
    _opcheck(torch.ops._C.rotary_embedding, (pos, q, k, hs, cache, True), DEFAULT_UTILS)
    _opcheck(torch.ops._C.rotary_embedding, (pos, q, None, hs, cache, False), DEFAULT_UTILS)


def test_opcheck_cache_movers(ops):
    NB, B, Hkv, D = 32, 16, 4, 64
    kcs = [torch.randn(NB, Hkv, D // 8, B, 8, dtype=torch.bfloat16, device=DEV) for _ in range(2)]
    vcs = [torch.randn(NB, Hkv, D, B, dtype=torch.bfloat16, device=DEV) for _ in range(2)]
    mapping = torch.tensor([[0, 5], [1, 6], [2, 7]], dtype=torch.int64, device=DEV)
    _opcheck(torch.ops._C_cache_ops.copy_blocks, (kcs, vcs, mapping), DEFAULT_UTILS)      # test_cache.py:100-103
    mla = [torch.randn(NB, B, 576, dtype=torch.bfloat16, device=DEV) for _ in range(2)]
    _opcheck(torch.ops._C_cache_ops.copy_blocks_mla, (mla, mapping), DEFAULT_UTILS)       # test_cache.py:635-639
    src, dst = kcs[0], torch.zeros_like(kcs[0])
    _opcheck(torch.ops._C_cache_ops.swap_blocks, (src, dst, mapping.cpu()))               # test_cache.py:415-420
    # concat_and_cache_mla (test_cache.py:553-557)
    T = 13
    kv_c = torch.randn(T, 512, dtype=torch.bfloat16, device=DEV)
    k_pe = torch.randn(T, 64, dtype=torch.bfloat16, device=DEV)
    slots = torch.randperm(NB * B)[:T].to(DEV)
    scale = torch.tensor([0.1], dtype=torch.float32, device=DEV)
    for kvd, cdt in (("auto", torch.bfloat16), ("fp8", torch.uint8)):
        cache = torch.zeros(NB, B, 576, dtype=cdt, device=DEV)
        _opcheck(torch.ops._C_cache_ops.concat_and_cache_mla, (kv_c, k_pe, cache, slots, kvd, scale), DEFAULT_UTILS)
    # gather_and_maybe_dequant_cache / cp_gather_cache (test_cache.py:787-789, 861)
    seq_lens = torch.tensor([40, 17, 64], dtype=torch.int32)
    cu = torch.zeros(4, dtype=torch.int32)
    cu[1:] = torch.cumsum(seq_lens, 0)
    bt = torch.randperm(NB)[:12].view(3, 4).to(torch.int32).to(DEV)
    src_cache = torch.randn(NB, B, 576, dtype=torch.bfloat16, device=DEV)
    dst = torch.zeros(int(cu[-1]), 576, dtype=torch.bfloat16, device=DEV)
    _opcheck(torch.ops._C_cache_ops.gather_and_maybe_dequant_cache, (src_cache, dst, bt, cu.to(DEV), 3, "auto", scale, None),
             DEFAULT_UTILS)
    _opcheck(torch.ops._C_cache_ops.cp_gather_cache, (src_cache, dst, bt, cu.to(DEV), 3, None), DEFAULT_UTILS)


# ---- CUDA graphs --------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("kvd", ["auto", "fp8"])
def test_cuda_graph_capture_and_replay(ops, version, kvd):
    """S * Hkv = 10 pairs of ~4000 tokens on 148 CTAs: every pair is split across CTAs (v1: merged through the device
    arrival counters; v2: 8 partitions through the reduce kernel).  Captured once, replayed three times with new
    queries, sequence lengths and block tables written into the captured buffers; each replay against the oracle."""
    S, H, Hkv, D, B, L = 5, 16, 2, 128, 16, 4000
    ds = [orc.make_decode_inputs(S, H, Hkv, D, B, max_seq_len=L, num_blocks=1400, kv_cache_dtype=kvd, seed=40 + i,
                                 unique_blocks=True) for i in range(3)]
    if kvd == "fp8":
        for d in ds:
            d["k_scale"], d["v_scale"] = 0.5, 2.0
    kc, vc = ds[0]["key_cache"].to(DEV), ds[0]["value_cache"].to(DEV)
    q = ds[0]["query"].to(DEV).clone()
    bt, lens = ds[0]["block_tables"].to(DEV).clone(), ds[0]["seq_lens"].to(DEV).clone()
    out = torch.zeros_like(q)
    ks = torch.tensor([ds[0].get("k_scale", 1.0)], dtype=torch.float32, device=DEV)
    vs = torch.tensor([ds[0].get("v_scale", 1.0)], dtype=torch.float32, device=DEV)
    P = (L + 511) // 512
    es = torch.empty(S, H, P, dtype=torch.float32, device=DEV)
    ml, tmp = torch.empty_like(es), torch.empty(S, H, P, D, dtype=q.dtype, device=DEV)

    def call():
        if version == "v1":
            ops.paged_attention_v1(out, q, kc, vc, Hkv, ds[0]["scale"], bt, lens, B, L, None, kvd, ks, vs)
        else:
            ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, ds[0]["scale"], bt, lens, B, L, None, kvd, ks, vs)

    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        ops.reserve_workspace(q, L)  # the warm-up call a plugin makes: sizes the side stream's workspace outside capture
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            call()
    for i in range(3):
        d = ds[i]
        q.copy_(d["query"].to(DEV))
        bt.copy_(d["block_tables"].to(DEV))
        lens.copy_(d["seq_lens"].to(DEV))
        out.fill_(float("nan"))
        torch.cuda.synchronize()
        graph.replay()
        torch.cuda.synchronize()
        # same cache for the three replays (the block tables of every seed index into it)
        ref = orc.paged_attention(d["query"], ds[0]["key_cache"], ds[0]["value_cache"], d["block_tables"], d["seq_lens"],
                                  d["scale"], None, kvd, ds[0].get("k_scale", 1.0), ds[0].get("v_scale", 1.0))
        err = (out.cpu().float() - ref.float()).abs().max().item()
        assert err <= (1e-3 if kvd == "auto" else 2e-3), (i, err)


def test_first_use_under_capture_raises_and_reserve_fixes_it(ops):
    d, one = _decode_args("auto", S=2, H=8, Hkv=2, D=128, B=16, L=200)
    out = torch.empty_like(d["query"])
    args = (out, d["query"], d["key_cache"], d["value_cache"], 2, d["scale"], d["block_tables"], d["seq_lens"], 16, 200, None,
            "auto", one, one)
    fresh = torch.cuda.Stream()  # no workspace has been created for this stream
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(fresh):
        torch.cuda.synchronize()
        g.capture_begin()
        try:
            with pytest.raises(RuntimeError, match="reserve_workspace"):
                ops.paged_attention_v1(*args)
        finally:
            g.capture_end()
    torch.cuda.synchronize()
    fresh2 = torch.cuda.Stream()
    with torch.cuda.stream(fresh2):
        ops.reserve_workspace(d["query"], 200, True)
        torch.cuda.synchronize()
        g2 = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g2, stream=fresh2):
            ops.paged_attention_v1(*args)
    g2.replay()
    torch.cuda.synchronize()
    ref = orc.paged_attention(d["query"].cpu(), d["key_cache"].cpu(), d["value_cache"].cpu(), d["block_tables"].cpu(),
                              d["seq_lens"].cpu(), d["scale"])
    torch.testing.assert_close(out.cpu().float(), ref.float(), atol=1e-3, rtol=1e-5)


def test_workspace_reset_entry_point(ops):
    """A workspace recycled from an unknown state (counters scribbled over) works again after the reset entry point."""
    from vllm_metax_b200 import _native

    d = orc.make_decode_inputs(3, 16, 2, 128, 16, max_seq_len=3000, num_blocks=700, seed=5, unique_blocks=True, fixed_len=True)
    t = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
    out = torch.empty_like(t["query"])
    one = torch.ones(1, dtype=torch.float32, device=DEV)
    p = _native.attn_params(out, t["query"], t["key_cache"], t["value_cache"], 2, d["scale"], t["block_tables"], t["seq_lens"], 16,
                            3000, None, "auto", one, one)
    ws = torch.full((max(16, _native.workspace_bytes(p)),), 0x5a, dtype=torch.uint8, device=DEV)  # garbage, not zeros
    p.workspace, p.workspace_bytes = ws.data_ptr(), ws.numel()
    stream = torch.cuda.current_stream().cuda_stream
    assert _native.load().b200_paged_attention_workspace_reset(ws.data_ptr(), ws.numel(), stream) == 0
    _native.paged_attention_v1(p, stream)
    torch.cuda.synchronize()
    ref = orc.paged_attention(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"], d["scale"])
    torch.testing.assert_close(out.cpu().float(), ref.float(), atol=1e-3, rtol=1e-5)
    assert _native.load().b200_paged_attention_workspace_reset(None, 0, stream) == _native.ERR_WORKSPACE


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("shape", [(32, 8, 128, "auto"), (32, 8, 128, "fp8"), (8, 2, 80, "auto")])
def test_empty_sequences_without_lse(ops, version, shape):
    """seq_len == 0 rows (CUDA-graph padding) are written as zeros, as the reference kernels leave them
    (acc = 0 times 1/(0 + 1e-6), csrc/attention/attention_kernels.cuh:331) — with no LSE asked for."""
    H, Hkv, D, kvd = shape
    d = orc.make_decode_inputs(6, H, Hkv, D, 16, max_seq_len=700, num_blocks=400, kv_cache_dtype=kvd, seed=3)
    d["seq_lens"][1] = 0
    d["seq_lens"][4] = 0
    t = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
    out = torch.full_like(t["query"], float("nan"))
    one = torch.ones(1, dtype=torch.float32, device=DEV)
    if version == "v1":
        ops.paged_attention_v1(out, t["query"], t["key_cache"], t["value_cache"], Hkv, d["scale"], t["block_tables"], t["seq_lens"],
                               16, 700, None, kvd, one, one)
    else:
        P = 2
        es = torch.empty(6, H, P, dtype=torch.float32, device=DEV)
        ml, tmp = torch.empty_like(es), torch.empty(6, H, P, D, dtype=out.dtype, device=DEV)
        ops.paged_attention_v2(out, es, ml, tmp, t["query"], t["key_cache"], t["value_cache"], Hkv, d["scale"], t["block_tables"],
                               t["seq_lens"], 16, 700, None, kvd, one, one)
    torch.cuda.synchronize()
    o = out.cpu()
    assert torch.equal(o[1], torch.zeros_like(o[1])) and torch.equal(o[4], torch.zeros_like(o[4]))
    keep = [0, 2, 3, 5]
    dd = dict(d)
    ref = orc.paged_attention(d["query"][keep], d["key_cache"], d["value_cache"], d["block_tables"][keep], d["seq_lens"][keep],
                              d["scale"], None, kvd)
    torch.testing.assert_close(o[keep].float(), ref.float(), atol=2e-3, rtol=1e-5)
    # every sequence empty: max_seq_len 0
    out.fill_(float("nan"))
    zl = torch.zeros_like(t["seq_lens"])
    if version == "v1":
        ops.paged_attention_v1(out, t["query"], t["key_cache"], t["value_cache"], Hkv, d["scale"], t["block_tables"], zl, 16, 0,
                               None, kvd, one, one)
        torch.cuda.synchronize()
        assert torch.equal(out.cpu(), torch.zeros_like(o))


# ---- MLA decode ops (csrc/torch_bindings.cpp:287-290; upstream 0.22 sm100_cutlass_mla_decode) ----
def _mla_args(H=16, bs=3, page=64, L=500):
    d = orc.make_mla_inputs(bs, H, page, L, seed=4)
    q = d["query"].to(DEV)
    out = torch.empty(bs, H, 512, dtype=q.dtype, device=DEV)
    lse = torch.empty(bs, H, dtype=torch.float32, device=DEV)
    return d, out, lse, q[..., :512].contiguous(), q[..., 512:].contiguous(), d["kv_cache"].to(DEV), d["seq_lens"].to(DEV), \
        d["block_tables"].to(DEV)


@pytest.mark.parametrize("H", [16, 128])
def test_opcheck_mla_decode(ops, H):
    d, out, lse, qn, qp, cache, lens, table = _mla_args(H)
    _opcheck(torch.ops._C.cutlass_mla_decode, (out, qn, qp, cache, lens, table, d["scale"]))
    n = torch.ops._C.sm100_cutlass_mla_get_workspace_size(table.shape[1] * cache.shape[1], out.shape[0], 148, -1)
    ws = torch.empty(max(n, 16), dtype=torch.uint8, device=DEV)
    # upstream's schema (kept string-identical) declares the scratch `Tensor workspace` immutable although the op's whole
    # purpose for it is to be written (the stock kernel writes its split-KV partials there too): test_schema would flag
    # exactly that, so it is left out for this one op
    _opcheck(torch.ops._C.sm100_cutlass_mla_decode, (out, lse, qn, qp, cache, lens, table, ws, d["scale"], -1),
             utils=("test_autograd_registration", "test_faketensor", "test_aot_dispatch_dynamic"))


@pytest.mark.parametrize("H", [16, 128])
def test_mla_decode_cuda_graph_capture_and_replay(ops, H):
    """the flat schedule (persistent grid, partial slots, merge kernel) inside a CUDA graph; queries change between replays"""
    d, out, lse, qn, qp, cache, lens, table = _mla_args(H, bs=5, L=1500)
    n = torch.ops._C.sm100_cutlass_mla_get_workspace_size(table.shape[1] * cache.shape[1], out.shape[0], 148, -1)
    ws = torch.empty(max(n, 16), dtype=torch.uint8, device=DEV)
    ops.sm100_cutlass_mla_decode(out, lse, qn, qp, cache, lens, table, ws, d["scale"], -1)   # warm-up outside capture
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        ops.sm100_cutlass_mla_decode(out, lse, qn, qp, cache, lens, table, ws, d["scale"], -1)
    for step in range(3):
        q = torch.randn(d["query"].shape, generator=torch.Generator().manual_seed(100 + step)).to(torch.bfloat16)
        qn.copy_(q[..., :512]); qp.copy_(q[..., 512:])
        out.fill_(float("nan"))
        graph.replay()
        torch.cuda.synchronize()
        ref, ref_lse = orc.mla_decode(q, d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
        torch.testing.assert_close(out.float().cpu(), ref, atol=4e-3, rtol=1.6e-2)
        torch.testing.assert_close(lse.cpu(), ref_lse, atol=2e-3, rtol=1e-4)
