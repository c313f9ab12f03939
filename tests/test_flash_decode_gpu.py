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


def test_decode_step_is_cuda_graph_capturable(ops):
    """The C-ABI calls only enqueue work (no sync, no allocation, no host read of device memory), so a decode step —
    cache write, v1-layout and flash-layout attention, LSE, merge — can be captured once and replayed with new
    inputs in the same buffers (the reference's ops are used under CUDA graphs the same way)."""
    S, H, Hkv, D, B, L = 6, 32, 8, 128, 16, 900
    d = orc.make_flash_decode_inputs(S, H, Hkv, D, B, L, S * ((L + B - 1) // B) + 3, seed=8)
    g = _to(d)
    v1 = orc.make_decode_inputs(S, H, Hkv, D, B, L, S * ((L + B - 1) // B) + 3, seed=8, unique_blocks=True)
    v1g = _to(v1)
    one = torch.ones(1, device="cuda")
    key = torch.zeros(S, Hkv, D, dtype=torch.bfloat16, device="cuda")
    value = torch.zeros_like(key)
    slots = torch.zeros(S, dtype=torch.long, device="cuda")
    q = g["query"].clone()
    out_f, out_1, merged = torch.empty_like(q), torch.empty_like(q), torch.empty_like(q)
    lse_f = torch.empty(H, S, dtype=torch.float32, device="cuda")
    lse_1 = torch.empty_like(lse_f)

    def step():
        ops.reshape_and_cache_flash(key, value, g["key_cache"], g["value_cache"], slots, "auto", one, one)
        ops.paged_attention_flash(out_f, q, g["key_cache"], g["value_cache"], d["scale"], g["block_tables"],
                                  g["seq_lens"], L, None, lse_f)
        ops.paged_attention_v1_lse(out_1, lse_1, q, v1g["key_cache"], v1g["value_cache"], Hkv, v1["scale"],
                                   v1g["block_tables"], v1g["seq_lens"], B, L, None, "auto", one, one)
        ops.merge_attn_states(merged, out_f, lse_f, out_1, lse_1)

    stream = torch.cuda.Stream()
    stream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(stream):
        step()  # warm-up on the capture stream: the shim allocates its per-stream scratch here
    torch.cuda.current_stream().wait_stream(stream)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=stream):
        step()
    lens = d["seq_lens"]
    gen = torch.Generator().manual_seed(5)
    for it in range(3):  # new query / new token each replay, same buffers
        qn = ((torch.rand(S, H, D, generator=gen) * 2 - 1) * d["scale"]).to(torch.bfloat16)
        kn = (torch.randn(S, Hkv, D, generator=gen) * 0.05).to(torch.bfloat16)
        vn = (torch.randn(S, Hkv, D, generator=gen) * 0.05).to(torch.bfloat16)
        sl = torch.tensor([int(d["block_tables"][s, (int(lens[s]) - 1) // B]) * B + (int(lens[s]) - 1) % B
                           for s in range(S)], dtype=torch.long)
        q.copy_(qn); key.copy_(kn); value.copy_(vn); slots.copy_(sl)
        graph.replay()
        torch.cuda.synchronize()
        rk, rv = g["key_cache"].cpu(), g["value_cache"].cpu()  # the cache after this replay's write
        ref_f = orc.paged_attention_flash(qn, rk, rv, d["block_tables"], lens, d["scale"])
        ref_1 = orc.paged_attention(qn, v1["key_cache"], v1["value_cache"], v1["block_tables"], v1["seq_lens"], v1["scale"])
        assert torch.equal(rk[sl // B, sl % B], kn) and torch.equal(rv[sl // B, sl % B], vn)
        assert (out_f.cpu().float() - ref_f.float()).abs().max() <= 2e-3
        assert (out_1.cpu().float() - ref_1.float()).abs().max() <= 2e-3
        l_f = orc.paged_attention_flash_lse(qn, rk, rv, d["block_tables"], lens, d["scale"])
        l_1 = orc.paged_attention_lse(qn, v1["key_cache"], v1["value_cache"], v1["block_tables"], v1["seq_lens"], v1["scale"])
        ref_m, _ = orc.merge_attn_states(ref_f, l_f, ref_1, l_1)
        assert (merged.cpu().float() - ref_m.float()).abs().max() <= 4e-3


@pytest.mark.parametrize("layout", ["NHD", "HND"])
def test_decode_impl_forward_and_split_merge(ops, layout):
    """B200FlashDecodeImpl.forward = the decode branch of the reference's FlashAttentionImpl.forward
    (vllm_metax/v1/attention/backends/flash_attn.py:661-762): cache write from slot_mapping, then decode attention."""
    from vllm_metax_b200.attention import B200FlashDecodeBackend, B200FlashDecodeImpl, DecodeMetadata

    S, H, Hkv, D, B, L = 9, 32, 8, 128, 16, 500
    NB = S * ((L + B - 1) // B) + 3
    d = orc.make_flash_decode_inputs(S, H, Hkv, D, B, L, NB, seed=12, hnd=(layout == "HND"))
    lens = d["seq_lens"]
    shape = B200FlashDecodeBackend.get_kv_cache_shape(NB, B, Hkv, D)
    order = B200FlashDecodeBackend.get_kv_cache_stride_order(layout)
    # allocate in storage order, view in logical order — as the engine does with the stride order
    storage = torch.empty([shape[i] for i in order], dtype=torch.bfloat16, device="cuda")
    inv = [order.index(i) for i in range(5)]
    kv_cache = storage.permute(*inv)
    assert tuple(kv_cache.shape) == shape
    kv_cache[0].copy_(d["key_cache"].cuda())
    kv_cache[1].copy_(d["value_cache"].cuda())
    gen = torch.Generator().manual_seed(3)
    pad = 3  # key/value padded past slot_mapping (CUDA-graph padding, flash_attn.py:671-677)
    key = (torch.randn(S + pad, Hkv, D, generator=gen) * 0.05).to(torch.bfloat16)
    value = (torch.randn(S + pad, Hkv, D, generator=gen) * 0.05).to(torch.bfloat16)
    slots = torch.tensor([int(d["block_tables"][s, (int(lens[s]) - 1) // B]) * B + (int(lens[s]) - 1) % B
                          for s in range(S)], dtype=torch.long)
    one = torch.ones(1)
    rk, rv = d["key_cache"].clone(), d["value_cache"].clone()
    orc.reshape_and_cache_flash(key, value, rk, rv, slots, "auto", one, one)
    ref = orc.paged_attention_flash(d["query"], rk, rv, d["block_tables"], lens, d["scale"])
    impl = B200FlashDecodeImpl(H, D, d["scale"], Hkv)
    md = DecodeMetadata(num_actual_tokens=S, num_decode_tokens=S, slot_mapping=slots.cuda(),
                        decode_block_table=d["block_tables"].cuda(), decode_seq_lens=lens.cuda(), max_seq_len=L)
    q = torch.cat([d["query"], torch.zeros(pad, H, D, dtype=torch.bfloat16)]).cuda()
    output = torch.zeros_like(q)
    impl.forward(q, key.cuda(), value.cuda(), kv_cache, md, output)
    assert torch.equal(kv_cache[0].cpu(), rk) and torch.equal(kv_cache[1].cpu(), rv)
    assert (output[:S].cpu().float() - ref.float()).abs().max() <= 2e-3
    assert (output[S:] == 0).all()
    # split at a block boundary + merge_attn_states == attention over the whole sequence
    split = (lens // 3 // B * B).to(torch.int32)
    merged = torch.empty(S, H, D, dtype=torch.bfloat16, device="cuda")
    mlse = torch.empty(H, S, dtype=torch.float32, device="cuda")
    impl.forward_decode_split(q[:S], kv_cache, d["block_tables"].cuda(), lens.cuda(), split.cuda(), L, merged, mlse)
    assert (merged.cpu().float() - ref.float()).abs().max() <= 2e-3
    ref_lse = orc.paged_attention_flash_lse(d["query"], rk, rv, d["block_tables"], lens, d["scale"])
    torch.testing.assert_close(mlse.cpu(), ref_lse, atol=2e-3, rtol=1e-4)
