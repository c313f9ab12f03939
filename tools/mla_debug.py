"""Stage-by-stage check of the MLA decode kernel on a GPU box (development tool, not a test):
raw S^T of the first tile (debug_logits) against q.k^T, then the output against the oracle, for a few tiny shapes.

    python tools/mla_debug.py
"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import torch_oracle as orc  # noqa: E402
from vllm_metax_b200 import _native  # noqa: E402

lib = _native.load()
CTA_TIMES = None
DEV = "cuda:0"


def run(d, splits=1, debug=True):
    q = d["query"].to(DEV)
    B, H, _ = q.shape
    qn, qp = q[..., :512].contiguous(), q[..., 512:].contiguous()
    cache, table, lens = d["kv_cache"].to(DEV), d["block_tables"].to(DEV), d["seq_lens"].to(DEV)
    fp8 = bool(os.environ.get("B200_MLA_DEBUG_FP8"))
    if fp8:
        qn, qp, cache = qn.to(torch.float8_e4m3fn), qp.to(torch.float8_e4m3fn), cache.to(torch.float8_e4m3fn)
    out = torch.full((B, H, 512), float("nan"), dtype=q.dtype, device=DEV)
    lse = torch.full((B, H), float("nan"), dtype=torch.float32, device=DEV)
    dbg = torch.zeros(128 * 64 + 3 * 64 * 4 * 2 + 3 * 2 * 4096, dtype=torch.float32, device=DEV)
    p = _native.MlaDecodeParams()
    p.out, p.lse, p.q_nope, p.q_pe, p.kv_cache = out.data_ptr(), lse.data_ptr(), qn.data_ptr(), qp.data_ptr(), cache.data_ptr()
    p.seq_lens, p.page_table = lens.data_ptr(), table.data_ptr()
    p.out_seq_stride, p.out_head_stride = out.stride(0), out.stride(1)
    p.q_nope_seq_stride, p.q_nope_head_stride = qn.stride(0), qn.stride(1)
    p.q_pe_seq_stride, p.q_pe_head_stride = qp.stride(0), qp.stride(1)
    p.block_stride, p.entry_stride = cache.stride(0), cache.stride(1)
    p.page_table_stride, p.lse_stride, p.num_blocks = table.stride(0), lse.stride(0), cache.shape[0]
    p.scale, p.num_seqs, p.num_heads, p.page_size, p.max_pages = d["scale"], B, H, cache.shape[1], table.shape[1]
    p.dtype = _native.DTYPE_BF16 if q.dtype == torch.bfloat16 else _native.DTYPE_F16
    p.num_kv_splits = splits
    p.kv_dtype = 1 if fp8 else 0
    need = lib.b200_mla_decode_workspace_bytes(B, H, table.shape[1] * cache.shape[1], splits)
    ws = torch.empty(max(need, 16), dtype=torch.uint8, device=DEV)
    p.workspace, p.workspace_bytes = ws.data_ptr(), need
    p.debug_logits = dbg.data_ptr() if debug else None
    rc = lib.b200_mla_decode(C.byref(p), None)
    assert rc == 0, lib.b200_last_error().decode()
    torch.cuda.synchronize()
    global CTA_TIMES
    CTA_TIMES = dbg[128 * 64 + 3 * 64 * 4 * 2:].cpu().view(torch.int64).view(-1, 3)
    stamps = dbg[128 * 64:128 * 64 + 3 * 64 * 4 * 2].cpu().view(torch.int64).view(3, 64, 4)
    return out.cpu(), lse.cpu(), dbg[:128 * 64].view(128, 64).cpu(), stamps


def main():
    for (B, H, page, L, splits) in [(1, 16, 64, 128, 1), (1, 16, 16, 128, 1), (1, 64, 64, 128, 1), (1, 16, 64, 100, 1),
                                    (2, 16, 64, 300, 1), (1, 128, 64, 700, 1), (2, 32, 128, 1500, 3)]:
        d = orc.make_mla_inputs(B, H, page, L, seed=3, varlen=False)
        ref_out, ref_lse = orc.mla_decode(d["query"], d["kv_cache"], d["scale"], d["block_tables"], d["seq_lens"])
        out, lse, dbg, _ = run(d, splits)
        kv = d["kv_cache"][d["block_tables"][0].long()].reshape(-1, 576)[:min(L, 128)].float()
        s_ref = kv @ d["query"][0, :min(H, 64)].float().t()                      # [tokens, heads]
        n = s_ref.shape[0]
        s_err = (dbg[:n, :s_ref.shape[1]] - s_ref).abs().max().item()
        o_err = (out.float() - ref_out).abs().max().item()
        l_err = (lse - ref_lse).abs().max().item()
        print("B%d H%d page%d L%d splits%d: S^T err %.3e (|S| max %.1f)  out err %.3e (|out| max %.2f)  lse err %.3e" % (
            B, H, page, L, splits, s_err, s_ref.abs().max().item(), o_err, ref_out.abs().max().item(), l_err), flush=True)
        if o_err > 1e-2:
            bad = (out.float() - ref_out).abs()
            print("   worst (seq, head, dv):", [int(v) for v in torch.unravel_index(bad.argmax(), bad.shape)],
                  " per-dv-block err:", [round(bad[..., j * 128:(j + 1) * 128].max().item(), 4) for j in range(4)],
                  " nan count:", int(torch.isnan(out.float()).sum()))


def timeline(B, H, L, page=64):
    """per-tile clock64 stamps of CTA 0 (seq 0, group 0, split 0) while the whole batch runs"""
    d = orc.make_mla_inputs(B, H, page, L, seed=3, varlen=False)
    for _ in range(3):
        out, lse, dbg, st = run(d, 0 if os.environ.get("B200_MLA_STAMP_CTA") else 1)   # a stamped CTA index means the flat schedule
    t0 = int(st[2, 0, 0])
    print("timeline B%d H%d L%d (cycles from the first producer stamp)" % (B, H, L))
    print(" tile | MMA: S-start S-issued PV-start PV-issued | softmax: S-ready pass1 P-free P-written | producer0: K-start K-end V-start V-end")
    print(" MMA lane of CTA 0 over %d tiles: cycles in full-barrier waits %d, S^T mma issue %d, O^T mma issue %d, commits %d" % (
        int(st[1, 63, 0]), int(st[0, 63, 0]), int(st[0, 63, 1]), int(st[0, 63, 2]), int(st[0, 63, 3])))
    for i in range(min(63, L // 128)):
        r = lambda a: " ".join("%7d" % (int(v) - t0 if int(v) else -1) for v in a)
        print(" %4d | %s | %s | %s" % (i, r(st[0, i]), r(st[1, i]), r(st[2, i])))


def cta_times(B, H, L, page=64):
    """start / end (globaltimer ns) and SM of every CTA of the decode kernel (B200_MLA_CTA_TIMES=1)"""
    os.environ["B200_MLA_CTA_TIMES"] = "1"
    d = orc.make_mla_inputs(B, H, page, L, seed=3, varlen=False)
    for _ in range(3):
        run(d, int(os.environ.get("B200_MLA_DEBUG_SPLITS", "0")))
    t = CTA_TIMES[CTA_TIMES[:, 1] > 0]
    t0 = int(t[:, 0].min())
    dur = (t[:, 1] - t[:, 0]).float() / 1e3
    print("cta times B%d H%d L%d: %d CTAs, kernel span %.1f us; start spread %.1f us; duration min/median/max %.1f/%.1f/%.1f us" % (
        B, H, L, t.shape[0], (int(t[:, 1].max()) - t0) / 1e3, (int(t[:, 0].max()) - t0) / 1e3, dur.min(), dur.median(), dur.max()))
    print(" per CTA (us):", " ".join("%d" % round(float(x)) for x in dur))
    order = torch.argsort(dur, descending=True)[:8]
    print(" slowest:", [(int(i), round(float(dur[i]), 1), "sm%d" % int(t[i, 2]), "start+%.1f" % ((int(t[i, 0]) - t0) / 1e3)) for i in order])


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--cta-times":
        cta_times(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]))
    elif len(sys.argv) > 1 and sys.argv[1] == "--timeline":
        if len(sys.argv) > 4:
            timeline(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]))
        else:
            timeline(296, 16, 2048)
            timeline(148, 128, 2048)
    else:
        main()
