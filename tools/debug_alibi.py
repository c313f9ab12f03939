"""debug: stream vs generic vs oracle with ALiBi, per (seq, head) error."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import torch_oracle as orc
from vllm_metax_b200 import _custom_ops as ops, _native

def run(d, variant):
    t = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
    out = torch.empty_like(t["query"])
    one = torch.ones(1, dtype=torch.float32, device="cuda")
    _native.set_attention_variant(variant)
    ops.paged_attention_v1(out, t["query"], t["key_cache"], t["value_cache"], t["num_kv_heads"], t["scale"],
                           t["block_tables"], t["seq_lens"], t["block_size"], t["max_seq_len"], t["alibi_slopes"],
                           t["kv_cache_dtype"], one, one)
    torch.cuda.synchronize()
    _native.set_attention_variant(0)
    return out.cpu().float()

for (H, Hkv, S, L) in [(40, 40, 7, 1571), (8, 8, 1, 64), (8, 8, 1, 32), (8, 2, 2, 300), (64, 8, 7, 1571)]:
    d = orc.make_decode_inputs(S, H, Hkv, 128, 16, max_seq_len=L, num_blocks=613, seed=0, use_alibi=True)
    ref = orc.paged_attention(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"],
                              d["scale"], d["alibi_slopes"], "auto", 1.0, 1.0).float()
    g = run(d, 1); s = run(d, 2)
    eg = (g - ref).abs().amax(-1); es = (s - ref).abs().amax(-1)
    print("H=%d Hkv=%d S=%d L=%d lens=%s generic max %.2e stream max %.2e" % (H, Hkv, S, L, d["seq_lens"].tolist(), eg.max(), es.max()))
    bad = (es > 1e-3).nonzero().tolist()
    print("  bad (seq, head):", bad[:40], "of", es.numel())
    print("  slopes of bad heads:", [round(float(d["alibi_slopes"][h]), 3) for _, h in bad[:40]])
    print("  slopes of good heads:", [round(float(d["alibi_slopes"][h]), 3) for h in range(H) if all(b[1] != h for b in bad)][:40])
