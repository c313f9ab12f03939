"""The v1-engine backend (vllm_metax_b200/v1_backend.py) under the reference's backend test template
(tests/v1/attention/test_attention_backends.py:285-460): BatchSpec -> CommonAttentionMetadata -> builder.build() ->
Impl.forward over a prepopulated, block-permuted paged cache, against per-sequence scaled_dot_product_attention with the
template's tolerances (rtol 1e-2, atol 5e-3).  CPU part: the class hierarchy, the metadata builder and the selection
rule (no kernel is called)."""
import pytest
import torch

vb = pytest.importorskip("vllm_metax_b200.v1_backend")  # needs upstream vLLM's interface module

from v1_backend_common import BATCH_SPECS, BatchSpec, MockAttentionLayer, create_common_attn_metadata, stand_in_config  # noqa: E402


def _spec(block_size=16, num_kv_heads=8, head_size=128):
    from vllm.v1.kv_cache_interface import FullAttentionSpec

    return FullAttentionSpec(block_size=block_size, num_kv_heads=num_kv_heads, head_size=head_size, dtype=torch.bfloat16)


def test_class_hierarchy_and_backend_contract():
    from vllm.v1.attention.backend import AttentionBackend, AttentionImpl, AttentionMetadataBuilder

    B = vb.B200PagedAttentionBackend
    assert issubclass(B, AttentionBackend) and issubclass(B.get_impl_cls(), AttentionImpl)
    assert issubclass(B.get_builder_cls(), AttentionMetadataBuilder)
    assert B.get_name() == "B200_PAGED_ATTN"
    assert B.get_kv_cache_shape(100, 16, 8, 128) == (2, 100, 16, 8, 128)  # flash_attn.py:101-110
    with pytest.raises(ValueError, match="multiple of 16"):
        B.get_kv_cache_shape(100, 8, 8, 128)
    assert B.get_kv_cache_stride_order(layout="NHD") == (0, 1, 2, 3, 4) and B.get_kv_cache_stride_order(layout="HND") == (0, 1, 3, 2, 4)
    assert B.supports_head_size(128) and not B.supports_head_size(64)
    assert B.supports_dtype(torch.bfloat16) and not B.supports_dtype(torch.float32)
    assert B.supports_block_size(16) and B.supports_block_size(64) and not B.supports_block_size(8)
    assert B.full_cls_name() == ("vllm_metax_b200.v1_backend", "B200PagedAttentionBackend")
    Impl = B.get_impl_cls()
    with pytest.raises(NotImplementedError):
        Impl(32, 128, 0.1, 8, sliding_window=128)
    with pytest.raises(NotImplementedError):
        Impl(32, 128, 0.1, 8, kv_cache_dtype="fp8")
    with pytest.raises(ValueError, match="Head size"):
        Impl(32, 64, 0.1, 8)


def test_selection_rule():
    assert vb.get_attn_backend_cls(None, 128, torch.bfloat16, "auto", 16, True, False, False, False) == vb.BACKEND_PATH
    assert vb.get_attn_backend_cls("B200_PAGED_ATTN", 128, torch.bfloat16, None, 32) == vb.BACKEND_PATH
    for bad in (dict(head_size=64), dict(dtype=torch.float16), dict(kv_cache_dtype="fp8"), dict(block_size=8), dict(use_mla=True),
                dict(has_sink=True)):
        kw = dict(selected_backend=None, head_size=128, dtype=torch.bfloat16, kv_cache_dtype="auto", block_size=16)
        kw.update(bad)
        assert vb.get_attn_backend_cls(**kw) is None  # the platform goes on to its own backends
        kw["selected_backend"] = "B200_PAGED_ATTN"
        with pytest.raises(ValueError, match="does not serve"):
            vb.get_attn_backend_cls(**kw)
    assert vb.get_attn_backend_cls("FLASH_ATTN", 128, torch.bfloat16, "auto", 16) is None
    # the path resolves to the class
    import importlib

    mod, name = vb.BACKEND_PATH.rsplit(".", 1)
    assert getattr(importlib.import_module(mod), name) is vb.B200PagedAttentionBackend


@pytest.mark.parametrize("name", sorted(BATCH_SPECS))
def test_metadata_builder_split(name):
    """decode-first split as flash_attn.py:307-379 computes it; per-token lengths of the prefill part."""
    spec = BATCH_SPECS[name]
    cfg = stand_in_config()
    builder = vb.B200AttentionMetadataBuilder(_spec(), ["layer"], cfg, torch.device("cpu"))
    assert builder.num_heads_q == 32 and builder.num_heads_kv == 8 and builder.reorder_batch_threshold == 1
    common = create_common_attn_metadata(spec, 16, torch.device("cpu"))
    m = builder.build(0, common)
    nd = sum(1 for q in spec.query_lens if q == 1)
    assert m.num_decodes == nd and m.num_prefills == spec.batch_size - nd
    assert m.num_decode_tokens == nd and m.num_prefill_tokens == spec.compute_num_tokens() - nd
    assert m.num_actual_tokens == spec.compute_num_tokens() and m.max_seq_len == max(spec.seq_lens)
    if nd:
        assert m.decode_seq_lens.tolist() == spec.seq_lens[:nd] and m.decode_block_table.shape[0] == nd
        assert m.decode_query_start_loc.tolist() == list(range(nd + 1))
    if m.num_prefills:
        assert m.prefill_seq_lens.tolist() == spec.seq_lens[nd:]
        assert m.prefill_query_start_loc[0] == 0 and m.prefill_query_start_loc[-1] == m.num_prefill_tokens
        want_len, want_req = [], []
        for r, (s, q) in enumerate(zip(spec.seq_lens[nd:], spec.query_lens[nd:])):
            want_len += [s - q + i + 1 for i in range(q)]
            want_req += [r] * q
        assert m.prefill_token_seq_lens.tolist() == want_len and m.prefill_token_req.tolist() == want_req


def test_builder_reads_per_rank_head_counts():
    """TP shards by KV head through the builder's config reads (flash_attn.py:238-241): 64q/8kv at TP 8 -> 8q/1kv."""
    cfg = stand_in_config(num_heads=64, num_kv_heads=8, tp=8)
    builder = vb.B200AttentionMetadataBuilder(_spec(num_kv_heads=1), ["layer"], cfg, torch.device("cpu"))
    assert (builder.num_heads_q, builder.num_heads_kv) == (8, 1)


# ---- the template on the GPU ---------------------------------------------------------------------------------------
def _prepopulate(k_contexts, v_contexts, block_size, num_kv_heads, head_size, device, num_blocks, common, layout):
    """tests/v1/attention/test_attention_backends.py:73-173: contexts written block-aligned from block 1 (block 0 is the
    null block), blocks randomly permuted, block table and slot mapping made consistent with the permutation."""
    seq_lens = common._seq_lens_cpu
    qsl = common.query_start_loc_cpu
    query_lens = qsl[1:] - qsl[:-1]
    context_lens = common._num_computed_tokens_cpu
    kv = torch.empty(2, num_blocks, block_size, num_kv_heads, head_size, dtype=torch.bfloat16, device=device)
    flat = kv.view(2, -1, num_kv_heads, head_size)
    start_block = 1
    for i, (kc, vc) in enumerate(zip(k_contexts, v_contexts)):
        start = start_block * block_size
        flat[0, start:start + kc.shape[0]] = kc
        flat[1, start:start + vc.shape[0]] = vc
        start_block += -(-int(seq_lens[i]) // block_size)
    blocks_end = start_block
    perm = torch.randperm(blocks_end - 1) + 1
    inv = torch.zeros(blocks_end, dtype=torch.long, device=device)
    inv[1:] = (torch.argsort(perm) + 1).to(device)
    kv[:, 1:blocks_end] = kv[:, perm.to(device)]
    start_block = 1
    for i in range(len(k_contexts)):
        n = -(-int(seq_lens[i]) // block_size)
        common.block_table_tensor[i, :n] = inv[start_block:start_block + n].to(torch.int32)
        start_block += n
    for i in range(len(k_contexts)):
        offs = torch.arange(int(query_lens[i])) + int(context_lens[i])
        common.slot_mapping[int(qsl[i]):int(qsl[i + 1])] = (
            common.block_table_tensor[i, (offs // block_size).to(device)].long() * block_size + (offs % block_size).to(device))
    if layout == "HND":  # same logical shape, HND storage (flash_attn.py:112-123)
        kv = kv.permute(0, 1, 3, 2, 4).contiguous().permute(0, 1, 3, 2, 4)
    return kv


@pytest.mark.gpu
@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("name", sorted(BATCH_SPECS))
def test_backend_correctness(ops, name, layout):
    spec = BATCH_SPECS[name]
    device = torch.device("cuda:0")
    H, Hkv, D, B = 32, 8, 128, 16
    scale = 1.0 / D ** 0.5
    torch.manual_seed(0)
    qs, ks, vs, refs, kctx, vctx = [], [], [], [], [], []
    for s_len, q_len in zip(spec.seq_lens, spec.query_lens):
        ctx = s_len - q_len
        q = torch.randn(q_len, H, D, dtype=torch.bfloat16, device=device)
        k = torch.randn(s_len, Hkv, D, dtype=torch.bfloat16, device=device)
        v = torch.randn(s_len, Hkv, D, dtype=torch.bfloat16, device=device)
        mask = torch.full((q_len, s_len), float("-inf"), device=device, dtype=torch.bfloat16)
        for i in range(q_len):
            mask[i, :ctx + i + 1] = 0.0
        o = torch.nn.functional.scaled_dot_product_attention(
            q.unsqueeze(0).transpose(1, 2), k.unsqueeze(0).transpose(1, 2).repeat_interleave(H // Hkv, dim=1),
            v.unsqueeze(0).transpose(1, 2).repeat_interleave(H // Hkv, dim=1), attn_mask=mask, scale=scale)
        refs.append(o.transpose(1, 2).squeeze(0))
        qs.append(q); ks.append(k[ctx:]); vs.append(v[ctx:]); kctx.append(k[:ctx]); vctx.append(v[:ctx])
    query, key, value, ref = torch.cat(qs), torch.cat(ks), torch.cat(vs), torch.cat(refs)
    common = create_common_attn_metadata(spec, B, device)
    kv_cache = _prepopulate(kctx, vctx, B, Hkv, D, device, 8192, common, layout)

    backend = vb.B200PagedAttentionBackend
    builder = backend.get_builder_cls()(_spec(), ["placeholder"], stand_in_config(), device)
    meta = builder.build(common_prefix_len=0, common_attn_metadata=common)
    impl = backend.get_impl_cls()(num_heads=H, head_size=D, scale=scale, num_kv_heads=Hkv, alibi_slopes=None,
                                  sliding_window=None, kv_cache_dtype="auto")
    out = torch.empty_like(query)
    res = impl.forward(MockAttentionLayer(device), query, key, value, kv_cache, meta, output=out)
    torch.cuda.synchronize()
    assert res.shape == ref.shape and res.dtype == ref.dtype and torch.isfinite(res).all()
    torch.testing.assert_close(res, ref, rtol=1e-2, atol=5e-3)  # test_attention_backends.py:441-442


@pytest.mark.gpu
def test_prefill_can_be_refused(ops):
    device = torch.device("cuda:0")
    spec = BatchSpec(seq_lens=[40], query_lens=[8])
    common = create_common_attn_metadata(spec, 16, device, max_block_idx=8)
    builder = vb.B200AttentionMetadataBuilder(_spec(), ["l"], stand_in_config(), device)
    meta = builder.build(0, common)
    impl = vb.B200AttentionImpl(32, 128, 0.1, 8, prefill_as_decode=False)
    q = torch.zeros(8, 32, 128, dtype=torch.bfloat16, device=device)
    kv = torch.zeros(2, 16, 16, 8, 128, dtype=torch.bfloat16, device=device)
    with pytest.raises(NotImplementedError, match="prefill"):
        impl.forward(MockAttentionLayer(device), q, None, None, kv, meta, output=torch.empty_like(q))
