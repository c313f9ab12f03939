"""Reduced kernel grid for compute-sanitizer (SURVEY.md §5 "race detection"):

    compute-sanitizer --tool memcheck  python tools/sanitizer_cases.py
    compute-sanitizer --tool racecheck python tools/sanitizer_cases.py
    compute-sanitizer --tool initcheck python tools/sanitizer_cases.py     (optional)

Every hot kernel once or twice at small sizes: the streaming decode kernel (block 16 / 32, bf16 / fp8, flat schedule,
v2 items, flash layout), the generic kernel, the split reduce, cache writers, movers, merge_attn_states, rotary, MLA decode
(16 / 128 heads, token splits), DCP gather / fold with emulated peers.  Results are checked against the oracle so a
sanitizer-clean run is also a correct one."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import torch_oracle as orc  # noqa: E402
import vllm_metax_b200  # noqa: E402

vllm_metax_b200.import_kernels()
from vllm_metax_b200 import _custom_ops as ops  # noqa: E402

DEV = "cuda:0"


def attention_case(S, H, Hkv, D, B, L, kvd, v2):
    d = orc.make_decode_inputs(S, H, Hkv, D, B, L, 64 + S * ((L + B - 1) // B), kv_cache_dtype=kvd, seed=L + B, unique_blocks=True)
    ks = torch.tensor(d.get("k_scale", 1.0)).to(DEV)
    vs = torch.tensor(d.get("v_scale", 1.0)).to(DEV)
    ref = orc.paged_attention(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"], d["scale"],
                              None, kvd, d.get("k_scale", 1.0), d.get("v_scale", 1.0))
    q, kc, vc = d["query"].to(DEV), d["key_cache"].to(DEV), d["value_cache"].to(DEV)
    bt, lens = d["block_tables"].to(DEV), d["seq_lens"].to(DEV)
    out = torch.empty_like(q)
    if v2:
        P = (L + 511) // 512
        tmp = torch.empty(S, H, P, D, dtype=q.dtype, device=DEV)
        es = torch.empty(S, H, P, dtype=torch.float32, device=DEV)
        ml = torch.empty_like(es)
        ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, d["scale"], bt, lens, B, L, None, kvd, ks, vs)
    else:
        ops.paged_attention_v1(out, q, kc, vc, Hkv, d["scale"], bt, lens, B, L, None, kvd, ks, vs)
    torch.cuda.synchronize()
    err = (out.cpu().float() - ref.float()).abs().max().item()
    assert err <= 2e-3, (S, H, Hkv, D, B, L, kvd, v2, err)
    return err


def mla_case(bs, H, page, L, splits):
    d = orc.make_mla_inputs(bs, H, page, L, seed=H + L)
    ref, ref_lse = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
    q = d["query"].to(DEV)
    out = torch.empty(bs, H, 512, dtype=q.dtype, device=DEV)
    lse = torch.empty(bs, H, dtype=torch.float32, device=DEV)
    ws = torch.empty(max(16, ops.sm100_cutlass_mla_get_workspace_size(d["block_tables"].shape[1] * page, bs, 148, splits)),
                     dtype=torch.uint8, device=DEV)
    ops.sm100_cutlass_mla_decode(out, lse, q[..., :512].contiguous(), q[..., 512:].contiguous(), d["kv_cache"].to(DEV),
                                 d["seq_lens"].to(DEV), d["block_tables"].to(DEV), ws, d["scale"], splits)
    torch.cuda.synchronize()
    err = (out.cpu().float() - ref).abs().max().item()
    assert err <= 8e-3, (bs, H, page, L, splits, err)
    return err


def main():
    print("stream bf16 B16 flat:", attention_case(3, 8, 2, 128, 16, 700, "auto", False), flush=True)
    print("stream fp8  B16 flat:", attention_case(3, 8, 2, 128, 16, 700, "fp8", False), flush=True)
    print("stream bf16 B32 flat:", attention_case(2, 8, 1, 128, 32, 600, "auto", False), flush=True)
    print("stream bf16 B16 v2  :", attention_case(2, 8, 2, 128, 16, 1300, "auto", True), flush=True)
    print("stream fp8  B16 v2  :", attention_case(2, 8, 2, 128, 16, 1300, "fp8", True), flush=True)
    print("generic f16-like D64:", attention_case(2, 4, 2, 64, 16, 200, "auto", False), flush=True)
    print("mla H16 page16      :", mla_case(2, 16, 16, 300, 1), flush=True)
    print("mla H128 page64     :", mla_case(2, 128, 64, 400, 1), flush=True)
    print("mla H32 splits      :", mla_case(2, 32, 64, 1500, 3), flush=True)
    # flash layout + merge + rotary + movers through smoke()'s sequence
    import __graft_entry__ as g
    g.smoke()
    print("sanitizer cases done", flush=True)


if __name__ == "__main__":
    main()
