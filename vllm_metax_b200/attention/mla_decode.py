"""MLA decode through the call shapes the reference's MLA backends use.

* `decode_attention_fwd` — the Triton split-K MLA decode's signature
  (vllm_metax/attention/ops/triton_decode_attention.py:673-721), as called by the Triton MLA backend
  (vllm_metax/v1/attention/backends/mla/triton_mla.py:188-200): q [B, H, 576], k_buffer = the cache with a unit head
  dim [NB, page, 1, 576], v_buffer = its first 512 columns, o [B, H, 512], lse [B, H], req_to_token = block table,
  attn_logits = split-K scratch [B, H, num_kv_splits, 513] fp32 (used as this kernel's workspace: same layout).
* `cutlass_mla_decode` — the reference op's wrapper, re-exported from _custom_ops.

Only the MLA shape is served (one shared latent head, 576 / 512 wide, page size a power of two >= 16, no logit cap);
everything else raises NotImplementedError — there is no fallback path.
"""
from __future__ import annotations

import torch

from .. import _custom_ops as ops

cutlass_mla_decode = ops.cutlass_mla_decode


def decode_attention_fwd(q, k_buffer, v_buffer, o, lse, req_to_token, b_seq_len, attn_logits, num_kv_splits, sm_scale,
                         page_size=1, logit_cap=0.0):
    assert num_kv_splits == attn_logits.shape[2]  # the reference's own check (:687)
    if logit_cap != 0.0:
        raise NotImplementedError("logit_cap is not implemented for the B200 MLA decode kernel")
    if k_buffer.dim() == 4:
        if k_buffer.shape[2] != 1:
            raise NotImplementedError("only the MLA shape (one shared KV head) is served")
        cache = k_buffer.squeeze(2)
    else:
        cache = k_buffer
    if cache.shape[-1] != 576 or o.shape[-1] != 512 or q.shape[-1] != 576:
        raise NotImplementedError("only head sizes 576 (qk) / 512 (v) are served")
    if v_buffer.data_ptr() != k_buffer.data_ptr():
        raise NotImplementedError("v_buffer must be the first 512 columns of k_buffer (the MLA latent)")
    if cache.shape[1] != page_size or page_size < 16 or page_size & (page_size - 1):
        raise NotImplementedError("page size must be a power of two >= 16 and match the cache")
    table = req_to_token if req_to_token.dim() == 2 else req_to_token.squeeze(-1)
    if table.dtype != torch.int32:
        table = table.to(torch.int32)
    lens = b_seq_len if b_seq_len.dtype == torch.int32 else b_seq_len.to(torch.int32)
    lse32 = lse if lse.dtype == torch.float32 else torch.empty(lse.shape, dtype=torch.float32, device=lse.device)
    workspace = attn_logits.view(torch.uint8).view(-1) if attn_logits.is_contiguous() else attn_logits.new_empty(0).view(torch.uint8)
    ops.sm100_cutlass_mla_decode(o, lse32, q[..., :512], q[..., 512:], cache, lens, table, workspace, sm_scale,
                                 num_kv_splits)
    if lse32 is not lse:
        lse.copy_(lse32)
    return o, lse
