"""Decode context parallel (DCP): the one place where the decode path has a real exchange step.

For long contexts the reference can shard the TOKENS of every sequence across the ranks of a DCP group instead of
(or on top of) sharding heads: token t of a sequence lives on rank t % N, stored densely in that rank's own paged cache
(reference: vllm_metax/v1/attention/backends/flash_attn.py:421-431).
A decode step then is (flash_attn.py:794-872, `_forward_with_dcp`):

  1. all-gather the query over the group along the head dimension: every rank attends with ALL heads of the group,
  2. attention over the local token shard, returning the partial output and its log-sum-exp,
  3. exchange: every rank needs, for ITS heads, the partial (output, lse) of all ranks (upstream's `cp_lse_ag_out_rs`:
     all-gather of the lse + reduce-scatter of the corrected outputs),
  4. combine the N partials with the merge_attn_states rule.

Here step 2 is `paged_attention_flash(..., lse)` (or `paged_attention_v1_lse`), step 3 is ONE all-to-all of the packed
(output, lse) head slices — each rank receives exactly the N partials of its own heads, N x less than an all-gather —
and step 4 folds them with `merge_attn_states`.  `torch.distributed` (NCCL over NVLink/NVSwitch) is plumbing; the
attention and merge kernels are the product.  The attention / merge callables are parameters so that the host logic can
be exercised on CPU (gloo) in the tests.
"""
from __future__ import annotations

from typing import Callable

import torch
import torch.distributed as dist


def dcp_local_seq_lens(seq_lens: torch.Tensor, world_size: int, rank: int) -> torch.Tensor:
    """Tokens of each sequence that live on `rank` when token t is stored on rank t % world_size: L // N plus one for
    the ranks below L % N.  (The reference computes `L // N + (rank <= (L - 1) % N)` at flash_attn.py:425-429, which is
    the same count except when L is a multiple of N, where it is one too many on every rank; the ownership count is
    used here.)"""
    return seq_lens // world_size + (rank < seq_lens % world_size).to(seq_lens.dtype)


def dcp_token_owner(position: int, world_size: int) -> tuple[int, int]:
    """(rank, local position) of global token `position`."""
    return position % world_size, position // world_size


def _cuda_attend(query, key_cache, value_cache, scale, block_table, seq_lens, max_seq_len):
    from .. import _custom_ops as ops

    S, H, _ = query.shape
    out = torch.empty_like(query)
    lse = torch.empty((H, S), dtype=torch.float32, device=query.device)
    ops.paged_attention_flash(out, query, key_cache, value_cache, scale, block_table, seq_lens, max_seq_len, None, lse)
    return out, lse


def _cuda_merge(prefix_out, prefix_lse, suffix_out, suffix_lse):
    from .. import _custom_ops as ops

    out = torch.empty_like(prefix_out)
    lse = torch.empty_like(prefix_lse)
    ops.merge_attn_states(out, prefix_out, prefix_lse, suffix_out, suffix_lse, lse)
    return out, lse


def forward_decode_dcp(
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    block_table: torch.Tensor,
    local_seq_lens: torch.Tensor,
    max_local_seq_len: int,
    scale: float,
    group: dist.ProcessGroup | None = None,
    attend: Callable = _cuda_attend,
    merge: Callable = _cuda_merge,
) -> tuple[torch.Tensor, torch.Tensor]:
    """One DCP decode step.  query [S, H_local, D] (this rank's heads); key_cache / value_cache: this rank's token shard
    in the flash layout; block_table / local_seq_lens describe the shard.  Returns (out [S, H_local, D], lse [H_local, S])
    of attention over the WHOLE sequences for this rank's heads."""
    world = dist.get_world_size(group)
    S, Hl, D = query.shape
    # 1. all heads of the group on every rank: [S, N * H_local, D], rank-major along heads
    gathered = torch.empty((world * S, Hl, D), dtype=query.dtype, device=query.device)
    dist.all_gather_into_tensor(gathered, query.contiguous(), group=group)
    q_all = gathered.view(world, S, Hl, D).permute(1, 0, 2, 3).reshape(S, world * Hl, D).contiguous()
    # 2. partial attention over the local tokens
    out_all, lse_all = attend(q_all, key_cache, value_cache, scale, block_table, local_seq_lens, max_local_seq_len)
    # a rank that holds no token of a sequence contributes weight 0: its row must be finite (0) with lse = -inf
    # 3. exchange: head slice j of (out, lse) goes to rank j.  Packed as float32 [N, S, H_local, D + 1] so that one
    #    all-to-all carries both (the lse rides in the last column).
    send = torch.empty((world, S, Hl, D + 1), dtype=torch.float32, device=query.device)
    send[..., :D] = out_all.view(S, world, Hl, D).permute(1, 0, 2, 3)
    send[..., D] = lse_all.view(world, Hl, S).permute(0, 2, 1)
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv.view(world * S, Hl, D + 1), send.view(world * S, Hl, D + 1), group=group)
    # 4. fold the N partials of this rank's heads (merge_attn_states is associative on (out, lse) pairs)
    #    in float32, rounded once at the end.  Rank 0 holds token 0 of every non-empty sequence, so the running
    #    pair never starts from two empty partials.
    out = recv[0, ..., :D].contiguous()
    lse = recv[0, ..., D].t().contiguous()
    for r in range(1, world):
        out, lse = merge(out, lse, recv[r, ..., :D].contiguous(), recv[r, ..., D].t().contiguous())
    return out.to(query.dtype), lse


class _DeviceBytes:
    """A raw device allocation seen through __cuda_array_interface__ (torch.as_tensor wraps it without copying)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


class _Plan:
    __slots__ = ("q_all", "out_all", "lse_all", "gather", "fold")


def _align256(n: int) -> int:
    return (n + 255) // 256 * 256


class DcpExchange:
    """DCP decode step in three launches, no NCCL call on the data path: query gather by peer stores
    (b200_dcp_gather_query), attention, exchange + combine over peer loads (b200_dcp_exchange_fold).

    Steps 3 and 4 of `forward_decode_dcp` — pack, all-to-all, N-1 merge launches — become ONE launch: the attention
    kernel writes its partial (out, lse) for all heads straight into a buffer that the peers have mapped through CUDA
    IPC (NVLink / NVSwitch peer access), and the fold kernel signals the peers, waits for their flags and pulls the N
    slices of this rank's heads, combining them in fp32 on the fly.

    The step number lives in DEVICE memory (one word per rank, advanced by the gather launch, read by the fold launch),
    so no launch argument changes from step to step and `forward_decode` can be captured into a CUDA graph and replayed
    — every rank of the group must run (or replay) the same number of steps.  One set of buffers is enough: a rank's
    gather of step k+1 returns only after every peer has published its gather flag k+1, which a peer does after its
    fold of step k has finished reading this rank's partial, so the attention launch of step k+1 overwrites nothing
    that is still being read; and a peer pushes its step-k+1 query into this rank's gathered query only after its own
    fold of step k, which waited for this rank's fold flag k, i.e. for this rank's attention launch of step k.
    One instance per (group, shape bound).  Same node only (CUDA IPC).
    """

    def __init__(self, max_seqs: int, num_heads_total: int, head_size: int, dtype: torch.dtype,
                 group: dist.ProcessGroup | None = None):
        from .. import _native as native

        self._native = native
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > native.DCP_MAX_WORLD:
            raise ValueError("DCP group of %d ranks (max %d)" % (self.world, native.DCP_MAX_WORLD))
        if num_heads_total % self.world != 0:
            raise ValueError("num_heads_total must be a multiple of the group size")
        self.max_seqs, self.num_heads_total, self.head_size, self.dtype = max_seqs, num_heads_total, head_size, dtype
        self._esize = torch.empty((), dtype=dtype).element_size()
        self._out_bytes = _align256(max_seqs * num_heads_total * head_size * self._esize)
        self._lse_bytes = _align256(num_heads_total * max_seqs * 4)
        self._q_bytes = self._out_bytes  # the gathered query has the shape of the partial output
        self._data = self._out_bytes + self._lse_bytes + self._q_bytes
        # (partial out, partial lse, gathered query) + 256 bytes of control words:
        # fold flags [16] at +0, gather flags [16] at +64, the gather kernel's CTA counter at +128, the step at +132
        total = self._data + 256
        self._device = torch.device("cuda", torch.cuda.current_device())
        self._ptr, handle = native.peer_alloc(total)
        handles: list = [None] * self.world
        dist.all_gather_object(handles, handle, group=group)
        self._bases = [self._ptr if r == self.rank else native.peer_open(handles[r]) for r in range(self.world)]
        self._local = torch.as_tensor(_DeviceBytes(self._ptr, total), device=self._device)
        self._plans: dict[int, _Plan] = {}
        dist.barrier(group=group)  # every flag array is zeroed and mapped before the first signal

    def close(self) -> None:
        if self._bases is None:
            return
        torch.cuda.synchronize(self._device)
        dist.barrier(group=self.group)  # nobody is still reading this rank's buffers
        for r, b in enumerate(self._bases):
            if r != self.rank:
                self._native.peer_close(b)
        self._native.peer_free(self._ptr)
        self._bases = None
        self._local = None
        self._plans = {}

    def step(self) -> int:
        """Steps completed on this rank (reads the device word: synchronises the current stream)."""
        torch.cuda.current_stream(self._device).synchronize()
        return int(self._local[self._data + 132:self._data + 136].view(torch.int32).item()) & 0xFFFFFFFF

    def forward_decode(self, query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                       block_table: torch.Tensor, local_seq_lens: torch.Tensor, max_local_seq_len: int,
                       scale: float) -> tuple[torch.Tensor, torch.Tensor]:
        """Same contract as `forward_decode_dcp`: (out [S, H_local, D], lse [H_local, S]).  CUDA-graph capturable."""
        from .. import _custom_ops as ops

        S, Hl, D = query.shape
        H = Hl * self.world
        if H != self.num_heads_total or D != self.head_size or query.dtype != self.dtype or S > self.max_seqs:
            raise ValueError("query %s %s does not fit the exchange buffers (max_seqs %d, heads %d, head_size %d, %s)"
                             % (tuple(query.shape), query.dtype, self.max_seqs, self.num_heads_total, self.head_size,
                                self.dtype))
        q_all = self._gather(query)
        # 2. local attention, partial (out, lse) of all heads written into the peer-mapped buffer
        plan = self._plan(S, Hl, D)
        ops.paged_attention_flash(plan.out_all, q_all, key_cache, value_cache, scale, block_table, local_seq_lens,
                                  max_local_seq_len, None, plan.lse_all)
        return self._fold(S, Hl, D, query.device)

    def _plan(self, S: int, Hl: int, D: int) -> _Plan:
        """Per batch size: the views of the exchange buffer and the launch parameters (built once; a decode step then
        only fills in the query / output pointers)."""
        plan = self._plans.get(S)
        if plan is None:
            native, H, ctl = self._native, Hl * self.world, self._data
            lse_off = self._out_bytes
            q_off = lse_off + self._lse_bytes
            plan = _Plan()
            plan.q_all = self._local[q_off:q_off + S * H * D * self._esize].view(self.dtype).view(S, H, D)
            plan.out_all = self._local[:S * H * D * self._esize].view(self.dtype).view(S, H, D)
            plan.lse_all = self._local[lse_off:lse_off + H * S * 4].view(torch.float32).view(H, S)
            step = self._ptr + ctl + 132
            plan.gather = native.dcp_gather_params([b + q_off for b in self._bases], [b + ctl + 64 for b in self._bases],
                                                   self._ptr + ctl + 128, self.rank, S, Hl, D, self.dtype, step)
            plan.fold = native.dcp_fold_params(list(self._bases), [b + lse_off for b in self._bases],
                                               [b + ctl for b in self._bases], self.rank, S, Hl, D, self.dtype, step)
            self._plans[S] = plan
        return plan

    def _gather(self, query: torch.Tensor) -> torch.Tensor:
        """Step 1 (starts a new step: the launch advances the device-side step word): every rank pushes its heads into
        every rank's gathered query and waits for the others' pushes.  Returns this rank's gathered query [S, H, D]
        (a view of the exchange buffer)."""
        S, Hl, D = query.shape
        plan = self._plan(S, Hl, D)
        if query.stride(1) != D or query.stride(2) != 1:
            query = query.contiguous()
        self._native.dcp_gather_launch(plan.gather, query, 0, torch.cuda.current_stream(query.device).cuda_stream)
        return plan.q_all

    def _fold(self, S: int, Hl: int, D: int, device) -> tuple[torch.Tensor, torch.Tensor]:
        """Steps 3 + 4 of the current step: signal, wait, pull this rank's head slices from every peer and fold them."""
        plan = self._plan(S, Hl, D)
        out = torch.empty((S, Hl, D), dtype=self.dtype, device=device)
        lse = torch.empty((Hl, S), dtype=torch.float32, device=device)
        self._native.dcp_fold_launch(plan.fold, out, lse, 0, torch.cuda.current_stream(device).cuda_stream)
        return out, lse
