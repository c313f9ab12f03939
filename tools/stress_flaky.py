"""debug: repeat one config many times; report run-to-run differences (races) and error vs oracle."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import torch_oracle as orc
from vllm_metax_b200 import _custom_ops as ops, _native

def run(t, version):
    q = t["query"]
    out = torch.empty_like(q)
    one = torch.ones(1, dtype=torch.float32, device="cuda")
    S, H, D = q.shape
    if version == "v1":
        ops.paged_attention_v1(out, q, t["key_cache"], t["value_cache"], t["num_kv_heads"], t["scale"], t["block_tables"],
                               t["seq_lens"], t["block_size"], t["max_seq_len"], t["alibi_slopes"], t["kv_cache_dtype"], one, one)
    else:
        P = (t["max_seq_len"] + 511) // 512
        tmp = torch.empty(S, H, P, D, dtype=q.dtype, device="cuda"); es = torch.empty(S, H, P, dtype=torch.float32, device="cuda"); ml = torch.empty_like(es)
        ops.paged_attention_v2(out, es, ml, tmp, q, t["key_cache"], t["value_cache"], t["num_kv_heads"], t["scale"], t["block_tables"],
                               t["seq_lens"], t["block_size"], t["max_seq_len"], t["alibi_slopes"], t["kv_cache_dtype"], one, one)
    torch.cuda.synchronize()
    return out

CFGS = [(40, 40, 32, "auto"), (40, 40, 16, "auto"), (64, 8, 32, "auto"), (64, 8, 16, "fp8"), (64, 8, 32, "fp8")]
if len(sys.argv) > 1:
    CFGS = [CFGS[int(a)] for a in sys.argv[1:]]
for (H, Hkv, B, kvd) in CFGS:
    d = orc.make_decode_inputs(7, H, Hkv, 128, B, max_seq_len=1571, num_blocks=613, kv_cache_dtype=kvd, seed=0)
    ref = orc.paged_attention(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"], d["scale"], None, kvd, 1.0, 1.0).float()
    ref2 = orc.paged_attention(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"], d["scale"], None, kvd, 1.0, 1.0).float()
    ref64 = orc.paged_attention_f64(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"], d["scale"], None, kvd, 1.0, 1.0)
    print("cfg", H, Hkv, B, kvd, "oracle deterministic:", torch.equal(ref, ref2), "oracle vs f64 max err %.3e" % (ref.double() - ref64).abs().max())
    t = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
    for version in ("v1", "v2"):
        base = run(t, version)
        ndiff = 0; worst = 0.0
        for i in range(200):
            o = run(t, version)
            if not torch.equal(o, base):
                ndiff += 1
                idx = (o != base).nonzero()
                if ndiff <= 3:
                    print("   run %d differs at %d elements, e.g. %s: %g vs %g" % (i, idx.shape[0], idx[0].tolist(), float(o[tuple(idx[0])]), float(base[tuple(idx[0])])))
            worst = max(worst, (o.cpu().float() - ref).abs().max().item())
        e64 = (base.cpu().double() - ref64).abs()
        print("  %s: %d/200 runs differ from the first; worst err vs oracle %.3e; first-run err vs f64 %.3e at %s" % (version, ndiff, worst, e64.max(), (e64 == e64.max()).nonzero()[0].tolist()))
