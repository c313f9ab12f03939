"""MLA decode goldens (run in the BUILD container only; needs /root/reference):

    python tests/golden/make_golden_mla.py      # writes tests/golden/mla_*.pt

Executes the reference's own oracles, cut out of its test files with `ast` and run unmodified:
  * `ref_mla` (tests/kernels/attention/test_mla_decode_cpu.py:13-40), the CPU paged oracle;
  * the nested `scaled_dot_product_attention` of tests/kernels/attention/test_cutlass_mla_decode.py:133-154
    (fp32 logits / logsumexp / softmax), which also pins the LSE.
The fixtures hold inputs and both outputs so the GPU box can replay them without /root/reference.
"""
from __future__ import annotations

import ast
import math
import os
import sys

import torch
import torch.nn.functional as F
from torch import Tensor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF_CPU = "/root/reference/tests/kernels/attention/test_mla_decode_cpu.py"
REF_CUTLASS = "/root/reference/tests/kernels/attention/test_cutlass_mla_decode.py"

from oracle import torch_oracle as orc  # noqa: E402


def find_function(path, name):
    for node in ast.walk(ast.parse(open(path).read())):
        if isinstance(node, ast.FunctionDef) and node.name == name:
            return node
    raise KeyError(name)


def load(path, name, ns):
    exec(compile(ast.Module(body=[find_function(path, name)], type_ignores=[]), path, "exec"), ns)
    return ns[name]


CASES = [
    # name, bs, H, page, max_len, dtype, varlen
    ("mla_bf16_h16_p16", 4, 16, 16, 300, torch.bfloat16, True),
    ("mla_bf16_h128_p64", 2, 128, 64, 700, torch.bfloat16, True),
    ("mla_f16_h32_p128", 3, 32, 128, 257, torch.float16, False),
]


def main():
    ref_mla = load(REF_CPU, "ref_mla", {"torch": torch, "F": F, "Tensor": Tensor})
    for idx, (name, bs, H, page, max_len, dtype, varlen) in enumerate(CASES):
        d = orc.make_mla_inputs(bs, H, page, max_len, dtype=dtype, seed=100 + idx, varlen=varlen)
        q, cache, table, lens, scale = d["query"], d["kv_cache"], d["block_tables"], d["seq_lens"], d["scale"]
        out_cpu = ref_mla(torch.empty(bs, H, 512, dtype=dtype), q, cache, scale, table.long(), lens)
        # the cutlass test's oracle works per sequence on gathered fp32 rows with h_kv = 1 and scale = 1/sqrt(576)
        sdpa = load(REF_CUTLASS, "scaled_dot_product_attention", {"torch": torch, "math": math, "h_q": H, "h_kv": 1})
        out_sd = torch.empty(bs, H, 512)
        lse_sd = torch.empty(bs, H)
        for i in range(bs):
            L = int(lens[i])
            kv = cache[table[i].long()].reshape(-1, 576)[:L]
            o, l = sdpa(q[i].unsqueeze(1), kv.unsqueeze(0), kv[:, :512].unsqueeze(0), is_causal=True)
            out_sd[i], lse_sd[i] = o[:, 0], l[:, 0]
        torch.save(dict(d, ref_out_sdpa_cpu=out_cpu, ref_out_f32=out_sd, ref_lse_f32=lse_sd),
                   os.path.join(HERE, name + ".pt"))
        mine, mine_lse = orc.mla_decode(q, cache, scale, table, lens)
        print("wrote %s: restatement vs cutlass-test oracle out %.2e lse %.2e; vs ref_mla (%s SDPA) %.2e" % (
            name, (mine - out_sd).abs().max(), (mine_lse - lse_sd).abs().max(), str(dtype).split(".")[-1],
            (mine - out_cpu.float()).abs().max()))


if __name__ == "__main__":
    main()
