"""Generate the golden vectors that pin the oracle to the reference (run in the BUILD container only).

    python tests/golden/make_golden.py        # needs /root/reference; writes tests/golden/*.pt

What is executed is the reference's OWN Python oracle code, unmodified: the source text of
`ref_masked_attention` and `ref_single_query_cached_kv_attention` is cut out of
/root/reference/tests/kernels/attention/test_attention.py (lines 50-118) with `ast` and exec'd with only
`torch` / `Optional` in scope.  (The module itself cannot be imported here: it pulls in vllm._C, xformers
and a CUDA platform.)  The cache-writer goldens execute the scatter loops that the reference keeps inline
in its tests (tests/kernels/attention/test_cache.py:190-200 for the v1/v2 layout, :325-337 for the flash
layouts); they are transcribed below next to their line numbers because they are not separate functions.

The fixtures hold inputs and outputs, so the GPU box (which has no /root/reference) can replay them.
FP8 cases use caches dequantised by oracle.torch_oracle.fp8_e4m3_dequant before the reference oracle runs,
as the reference test dequantises with convert_fp8 before calling it (test_attention.py:303-319).
"""
from __future__ import annotations

import ast
import os
import random
import sys
from typing import Optional

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF_TEST = "/root/reference/tests/kernels/attention/test_attention.py"

from oracle import torch_oracle as orc  # noqa: E402


def load_reference_oracle():
    src = open(REF_TEST).read()
    tree = ast.parse(src)
    wanted = {"ref_masked_attention", "ref_single_query_cached_kv_attention"}
    ns = {"torch": torch, "Optional": Optional}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in wanted:
            code = compile(ast.Module(body=[node], type_ignores=[]), REF_TEST, "exec")
            exec(code, ns)
    assert wanted <= set(ns), "reference oracle functions not found"
    return ns["ref_single_query_cached_kv_attention"]


ATTN_CASES = [
    # name, S, H, Hkv, D, B, max_len, NB, dtype, kv, alibi
    ("attn_bf16_gqa4", 3, 8, 2, 128, 16, 150, 40, torch.bfloat16, "auto", False),
    ("attn_bf16_mha_alibi", 3, 4, 4, 64, 16, 97, 24, torch.bfloat16, "auto", True),
    ("attn_bf16_gqa8_b32", 2, 8, 1, 128, 32, 130, 12, torch.bfloat16, "auto", False),
    ("attn_fp8_gqa4", 3, 8, 2, 128, 16, 120, 32, torch.bfloat16, "fp8", False),
    ("attn_f32_d80", 2, 4, 2, 80, 16, 70, 16, torch.float32, "auto", True),
    ("attn_f16_d32_b8", 2, 4, 2, 32, 8, 50, 16, torch.float16, "auto", False),
]


def make_attention(ref_fn):
    for idx, (name, S, H, Hkv, D, B, L, NB, dtype, kv, alibi) in enumerate(ATTN_CASES):
        d = orc.make_decode_inputs(S, H, Hkv, D, B, L, NB, dtype=dtype, kv_cache_dtype=kv, seed=100 + idx,
                                   use_alibi=alibi)
        kc, vc = d["key_cache"], d["value_cache"]
        if kv != "auto":
            d["k_scale"], d["v_scale"] = 0.5, 2.0  # exact powers of two: dequant stays exact in bf16
            kc = orc.fp8_e4m3_dequant(kc, d["k_scale"], dtype)
            vc = orc.fp8_e4m3_dequant(vc, d["v_scale"], dtype)
        out = torch.empty_like(d["query"])
        ref_fn(out, d["query"], H // Hkv, kc, vc, d["block_tables"], d["seq_lens"], d["scale"], d["alibi_slopes"])
        d["ref_out"] = out
        torch.save(d, os.path.join(HERE, name + ".pt"))
        print("wrote", name, tuple(out.shape), float(out.float().abs().max()))


CACHE_CASES = [
    # name, T, H, D, B, NB, dtype
    ("cache_bf16_d64_b16", 11, 2, 64, 16, 8, torch.bfloat16),
    ("cache_f32_d80_b8", 7, 2, 80, 8, 8, torch.float32),
]


def make_cache():
    for name, T, H, D, B, NB, dtype in CACHE_CASES:
        rnd = random.Random(7)
        g = torch.Generator().manual_seed(11)
        slots = rnd.sample(range(B * NB), T)
        slot_mapping = torch.tensor(slots, dtype=torch.long)
        key = torch.randn(T, H, D, generator=g).to(dtype)
        value = torch.randn(T, H, D, generator=g).to(dtype)
        x = 16 // torch.tensor([], dtype=dtype).element_size()
        key_cache = torch.randn(NB, H, D // x, B, x, generator=g).to(dtype)
        value_cache = torch.randn(NB, H, D, B, generator=g).to(dtype)
        nhd_k = torch.randn(NB, B, H, D, generator=g).to(dtype)
        nhd_v = torch.randn(NB, B, H, D, generator=g).to(dtype)
        hnd_k = torch.randn(NB, H, B, D, generator=g).to(dtype)
        hnd_v = torch.randn(NB, H, B, D, generator=g).to(dtype)
        fix = dict(key=key, value=value, slot_mapping=slot_mapping, key_cache=key_cache.clone(),
                   value_cache=value_cache.clone(), nhd_k=nhd_k.clone(), nhd_v=nhd_v.clone(), hnd_k=hnd_k.clone(),
                   hnd_v=hnd_v.clone())
        # reference: tests/kernels/attention/test_cache.py:190-200
        reshaped_key = key.reshape(T, *key_cache[0, :, :, 0, :].shape)
        block_indices = torch.div(slot_mapping, B, rounding_mode="floor").tolist()
        block_offsets = (slot_mapping % B).tolist()
        for i in range(T):
            key_cache[block_indices[i], :, :, block_offsets[i], :] = reshaped_key[i]
            value_cache[block_indices[i], :, :, block_offsets[i]] = value[i]
        # reference: tests/kernels/attention/test_cache.py:325-337 (NHD, then HND)
        for i in range(T):
            nhd_k[block_indices[i], block_offsets[i], :, :] = key[i]
            nhd_v[block_indices[i], block_offsets[i], :, :] = value[i]
            hnd_k[block_indices[i], :, block_offsets[i], :] = key[i]
            hnd_v[block_indices[i], :, block_offsets[i], :] = value[i]
        fix.update(ref_key_cache=key_cache, ref_value_cache=value_cache, ref_nhd_k=nhd_k, ref_nhd_v=nhd_v,
                   ref_hnd_k=hnd_k, ref_hnd_v=hnd_v)
        torch.save(fix, os.path.join(HERE, name + ".pt"))
        print("wrote", name)


REF_MERGE_TEST = "/root/reference/tests/kernels/attention/test_merge_attn_states.py"


def load_reference_merge():
    """merge_attn_states_torch, cut out of the reference test file unmodified (test_merge_attn_states.py:16-47)."""
    tree = ast.parse(open(REF_MERGE_TEST).read())
    ns = {"torch": torch, "Optional": Optional}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == "merge_attn_states_torch":
            exec(compile(ast.Module(body=[node], type_ignores=[]), REF_MERGE_TEST, "exec"), ns)
    return ns["merge_attn_states_torch"]


MERGE_CASES = [
    # name, tokens, heads, head_size, dtype  (sizes from the reference grid, test_merge_attn_states.py:50-53, scaled down)
    ("merge_bf16_h8_d128", 37, 8, 128, torch.bfloat16),
    ("merge_f16_h4_d96", 29, 4, 96, torch.float16),
    ("merge_f32_h4_d48", 21, 4, 48, torch.float32),
]


def make_merge(ref_fn):
    for idx, (name, N, H, D, dtype) in enumerate(MERGE_CASES):
        g = torch.Generator().manual_seed(300 + idx)
        # inputs as the reference test draws them (:104-136): randn lse with ~10 % +inf on either side, never both
        p_lse = torch.randn(H, N, generator=g)
        s_lse = torch.randn(H, N, generator=g)
        mp = torch.rand(H, N, generator=g) < 0.1
        ms = torch.rand(H, N, generator=g) < 0.1
        both = mp & ms
        p_lse[mp & ~both] = float("inf")
        s_lse[ms & ~both] = float("inf")
        p_out = torch.randn(N, H, D, generator=g).to(dtype)
        s_out = torch.randn(N, H, D, generator=g).to(dtype)
        out0 = torch.zeros(N, H, D, dtype=dtype)
        lse0 = torch.zeros(H, N)
        ref_out, ref_lse = ref_fn(out0, p_out, p_lse.clone(), s_out, s_lse.clone(), lse0)  # the function mutates its lse args
        fix = dict(prefix_output=p_out, prefix_lse=p_lse, suffix_output=s_out, suffix_lse=s_lse,
                   ref_output=ref_out, ref_output_lse=ref_lse)
        torch.save(fix, os.path.join(HERE, name + ".pt"))
        print("wrote", name, ref_out.dtype, tuple(ref_out.shape))


REF_UNIFIED_TEST = "/root/reference/tests/kernels/attention/test_triton_unified_attention.py"


def make_flash():
    """Flash-layout decode goldens: the reference's ref_paged_attn (test_triton_unified_attention.py:25-76),
    cut out unmodified, run with query_lens = 1."""
    tree = ast.parse(open(REF_UNIFIED_TEST).read())
    ns = {"torch": torch, "Optional": Optional}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == "ref_paged_attn":
            exec(compile(ast.Module(body=[node], type_ignores=[]), REF_UNIFIED_TEST, "exec"), ns)
    ref_fn = ns["ref_paged_attn"]
    cases = [  # name, S, H, Hkv, D, B, L, NB, hnd
        ("flash_bf16_gqa4_nhd", 3, 8, 2, 128, 16, 150, 40, False),
        ("flash_bf16_gqa8_hnd_b32", 2, 8, 1, 128, 32, 130, 12, True),
    ]
    for idx, (name, S, H, Hkv, D, B, L, NB, hnd) in enumerate(cases):
        d = orc.make_flash_decode_inputs(S, H, Hkv, D, B, L, NB, seed=200 + idx, hnd=hnd)
        lens = d["seq_lens"].tolist()
        # ref_paged_attn scales its query argument in place (:46) and views the cache (:51): hand it copies
        out = ref_fn(d["query"].clone(), d["key_cache"].contiguous(), d["value_cache"].contiguous(), [1] * S, lens,
                     d["block_tables"], d["scale"])
        d["ref_out"] = out
        torch.save(d, os.path.join(HERE, name + ".pt"))
        print("wrote", name, tuple(out.shape))


def make_rotary():
    """Rotary goldens from upstream vLLM's own PyTorch-native implementation — the oracle of the reference's kernel
    test (tests/kernels/core/test_pos_encoding.py:79-81 `rope.forward_native`).  The class cannot be instantiated
    without an engine config, so the static method that forward_native calls is used directly; the installed
    upstream (0.22) differs from the plugin's pinned version only in packaging, not in this arithmetic."""
    from vllm.model_executor.layers.rotary_embedding.base import RotaryEmbedding

    cases = [  # name, tokens, H, Hkv, head_size, rot_dim, neox, dtype
        ("rotary_bf16_neox_d128", 9, 8, 2, 128, 128, True, torch.bfloat16),
        ("rotary_f16_gptj_d80_r32", 7, 4, 4, 80, 32, False, torch.float16),
        ("rotary_f32_neox_d64", 5, 4, 2, 64, 64, True, torch.float32),
    ]
    for idx, (name, n, H, Hkv, D, R, neox, dtype) in enumerate(cases):
        g = torch.Generator().manual_seed(400 + idx)
        cache = orc.compute_cos_sin_cache(R, 512, dtype=dtype)
        positions = torch.randint(0, 512, (n,), generator=g)
        q = torch.randn(n, H * D, generator=g).to(dtype)
        k = torch.randn(n, Hkv * D, generator=g).to(dtype)
        rq, rk = RotaryEmbedding.forward_static(positions, q.clone(), k.clone(), D, R, cache, neox)
        torch.save(dict(positions=positions, query=q, key=k, head_size=D, cos_sin_cache=cache, is_neox=neox,
                        ref_query=rq, ref_key=rk), os.path.join(HERE, name + ".pt"))
        print("wrote", name)


def make_movers():
    """copy_blocks / gather goldens: the reference tests' own expectation loops, transcribed next to their lines."""
    rnd = random.Random(5)
    g = torch.Generator().manual_seed(17)
    NB, B, H, D, layers, nmap = 24, 16, 2, 64, 3, 5
    dtype = torch.bfloat16
    x = 8
    key_caches = [torch.randn(NB, H, D // x, B, x, generator=g).to(dtype) for _ in range(layers)]
    value_caches = [torch.randn(NB, H, D, B, generator=g).to(dtype) for _ in range(layers)]
    # tests/kernels/attention/test_cache.py:71-81: every source block is mapped to two destinations
    src_blocks = rnd.sample(range(NB), nmap)
    remaining = sorted(set(range(NB)) - set(src_blocks))
    dst_blocks = rnd.sample(remaining, 2 * nmap)
    mapping = []
    for i in range(nmap):
        mapping.append((src_blocks[i], dst_blocks[2 * i]))
        mapping.append((src_blocks[i], dst_blocks[2 * i + 1]))
    ref_k = [k.clone() for k in key_caches]
    ref_v = [v.clone() for v in value_caches]
    for src, dst in mapping:  # :106-111
        for c in ref_k:
            c[dst].copy_(c[src])
        for c in ref_v:
            c[dst].copy_(c[src])
    torch.save(dict(key_caches=key_caches, value_caches=value_caches,
                    block_mapping=torch.tensor(mapping, dtype=torch.int64).view(-1, 2), ref_key_caches=ref_k,
                    ref_value_caches=ref_v), os.path.join(HERE, "movers_copy_blocks_bf16.pt"))
    print("wrote movers_copy_blocks_bf16")

    # gather (tests/kernels/attention/test_cache.py:720-777): MLA-shaped entries, float32, auto
    NB, B, entry, batch, max_len = 20, 16, 576, 5, 100
    src_cache = torch.randn(NB, B, entry, generator=g)
    seq_len = torch.randint(0, max_len + 1, (batch,), generator=g)
    seq_len[1] = 0  # the reference's loop skips empty sequences (:750-751)
    cu = torch.zeros(batch + 1, dtype=torch.int32)
    cu[1:] = seq_len.cumsum(0).to(torch.int32)
    tot_blocks = (seq_len + B - 1) // B
    block_table = torch.stack([torch.randperm(NB, generator=g) for _ in range(batch)]).to(torch.int32)
    expected = []
    for b in range(batch):  # :748-776
        s_ = int(seq_len[b])
        if s_ == 0:
            continue
        tot = int(tot_blocks[b])
        blocks = block_table[b, :tot].tolist()
        rows = [src_cache[blocks[i]] for i in range(tot - 1)]
        remaining_tokens = s_ - (tot - 1) * B
        rows.append(src_cache[blocks[-1], :remaining_tokens, :])
        expected.append(torch.cat(rows, dim=0))
    torch.save(dict(src_cache=src_cache, block_table=block_table, cu_seq_lens=cu, batch_size=batch,
                    ref_dst=torch.cat(expected, dim=0)), os.path.join(HERE, "movers_gather_f32.pt"))
    print("wrote movers_gather_f32")


if __name__ == "__main__":
    if not os.path.exists(REF_TEST):
        sys.exit("the reference is not mounted at /root/reference: goldens can only be regenerated in the build container")
    make_attention(load_reference_oracle())
    make_cache()
    make_merge(load_reference_merge())
    make_movers()
    make_flash()
    make_rotary()
