"""v1-engine attention backend on the B200 decode kernels.

The reference's serving path is a backend behind upstream vLLM's `AttentionBackend` / `AttentionMetadataBuilder` /
`AttentionImpl` interfaces (vllm_metax/v1/attention/backends/flash_attn.py:55-132 backend, :200-526 builder, :532-792 impl),
chosen by the platform hook (vllm_metax/platform.py:259-400).  This module is that triple for the flash cache layout
`[2, num_blocks, block_size, num_kv_heads, head_size]`, subclassing upstream's ABCs (upstream vLLM is the plugin's pinned,
un-vendored dependency; the pure-Python interface module imports without a GPU):

  B200PagedAttentionBackend   names, cache shape / stride order, supported head sizes, dtypes and block sizes
  B200AttentionMetadata       the reference's FlashAttentionMetadata fields for the decode / prefill split (:160-198)
  B200AttentionMetadataBuilder.build()   decode-first split of the batch as the reference does it (:286-379):
                              `split_decodes_and_prefills`, per-part query_start_loc / seq_lens / block-table slices
  B200AttentionImpl.forward() `reshape_and_cache_flash` of the step's K/V (slot_mapping decides the token count, :661-687),
                              then the flash-layout decode kernel for the decode tokens (:742-762 in the reference calls
                              flash_attn_with_kvcache of the external library)
  get_attn_backend_cls()      the selection rule in the shape of MacaPlatform.get_attn_backend_cls

Prefill tokens: this repository's scope is the decode path; there is no prefill kernel here.  So that the backend is
nevertheless usable on mixed batches (and testable with the reference's backend test template,
tests/v1/attention/test_attention_backends.py), a prefill token is served as what it is mathematically — a decode query
over the first `context + i + 1` tokens of its sequence, read back from the paged cache the step has just written.  That
re-reads the context once per query token: correct (bit-for-bit the decode kernel), not fast; `prefill_as_decode=False`
makes forward() raise instead.

Load order: import this module (or vllm) BEFORE anything that loads the kernels of this package — see the note at
the imports.  No engine, scheduler or model-runner code is here or needed: the builder consumes upstream's CommonAttentionMetadata.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import ClassVar

import torch

from vllm.v1.attention.backend import (  # upstream interfaces (pure Python)
    AttentionBackend,
    AttentionCGSupport,
    AttentionImpl,
    AttentionMetadataBuilder,
    AttentionType,
    CommonAttentionMetadata,
    MultipleOf,
)
from vllm.v1.attention.backends.utils import split_decodes_and_prefills

# Upstream first, this repository's kernels second: on a CUDA box the imports above load upstream's compiled
# `vllm._C`, which DEFINES the torch.ops._C schemas; our shim then only overrides the CUDA kernels of the identical ones
# (csrc/torch_shim.cpp register_op).  The other order aborts inside upstream's TORCH_LIBRARY definition — which is why
# this module sits at the top of the package: importing it must not run a package __init__ that loads the kernels.
from . import _custom_ops as ops  # noqa: E402

BACKEND_PATH = "vllm_metax_b200.v1_backend.B200PagedAttentionBackend"


class B200PagedAttentionBackend(AttentionBackend):
    """reference: MacaFlashAttentionBackend (flash_attn.py:55-132)."""

    accept_output_buffer: bool = True
    supported_dtypes: ClassVar[list[torch.dtype]] = [torch.bfloat16]
    supported_kv_cache_dtypes: ClassVar[list[str]] = ["auto", "bfloat16"]

    @staticmethod
    def get_name() -> str:
        return "B200_PAGED_ATTN"

    @staticmethod
    def get_impl_cls() -> type["B200AttentionImpl"]:
        return B200AttentionImpl

    @staticmethod
    def get_metadata_cls() -> type["B200AttentionMetadata"]:
        return B200AttentionMetadata

    @staticmethod
    def get_builder_cls() -> type["B200AttentionMetadataBuilder"]:
        return B200AttentionMetadataBuilder

    @classmethod
    def get_supported_head_sizes(cls) -> list[int]:
        return [128]  # the streaming kernel; the reference lists what its external library serves (:63-65)

    @staticmethod
    def get_supported_kernel_block_sizes() -> list:
        return [MultipleOf(16)]  # flash_attn.py:67-69

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int,
                           cache_dtype_str: str = "auto") -> tuple[int, ...]:
        if block_size % 16 != 0:  # flash_attn.py:108-109
            raise ValueError("Block size must be a multiple of 16.")
        return (2, num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def get_kv_cache_stride_order(include_num_layers_dimension: bool = False, layout: str | None = None) -> tuple[int, ...]:
        """NHD storage, or HND (the kernel takes both through the cache's strides; flash_attn.py:112-123)."""
        if layout is None:
            try:
                from vllm.v1.attention.backends.utils import get_kv_cache_layout

                layout = get_kv_cache_layout()
            except Exception:
                layout = "NHD"
        if layout == "NHD":
            order = (0, 1, 2, 3, 4)
        elif layout == "HND":
            order = (0, 1, 3, 2, 4)
        else:
            raise ValueError("Unknown cache layout format %s." % layout)
        if include_num_layers_dimension:
            order = (0,) + tuple(o + 1 for o in order)
        return order

    @classmethod
    def supports_attn_type(cls, attn_type: str) -> bool:
        return attn_type == AttentionType.DECODER


@dataclass
class B200AttentionMetadata:
    """reference: FlashAttentionMetadata (flash_attn.py:135-198), the fields of the decode / prefill split."""
    num_actual_tokens: int  # number of tokens excluding padding
    max_query_len: int
    query_start_loc: torch.Tensor
    max_seq_len: int
    seq_lens: torch.Tensor
    block_table: torch.Tensor
    slot_mapping: torch.Tensor

    num_decodes: int
    num_decode_tokens: int
    decode_query_start_loc: torch.Tensor | None
    decode_max_seq_len: int
    decode_seq_lens: torch.Tensor | None
    decode_block_table: torch.Tensor | None

    num_prefills: int
    num_prefill_tokens: int
    prefill_query_start_loc: torch.Tensor | None
    prefill_max_seq_len: int
    prefill_seq_lens: torch.Tensor | None
    prefill_block_table: torch.Tensor | None

    # prefill tokens served as decode queries (module docstring): per-token sequence length and request index
    prefill_token_seq_lens: torch.Tensor | None = None
    prefill_token_req: torch.Tensor | None = None

    use_cascade: bool = False
    common_prefix_len: int = 0
    causal: bool = True


class B200AttentionMetadataBuilder(AttentionMetadataBuilder[B200AttentionMetadata]):
    """reference: FlashAttentionMetadataBuilder (flash_attn.py:200-526).  The runner is asked to put decode requests
    (one query token) first — `reorder_batch_threshold = 1` — so that the two parts are contiguous slices."""

    # the decode kernel launches are capturable for any batch of single-token requests (tests/test_boundary_gpu.py)
    _cudagraph_support: ClassVar[AttentionCGSupport] = AttentionCGSupport.UNIFORM_SINGLE_TOKEN_DECODE
    reorder_batch_threshold: int = 1

    def __init__(self, kv_cache_spec, layer_names: list[str], vllm_config, device: torch.device):
        super().__init__(kv_cache_spec, layer_names, vllm_config, device)
        self.model_config = vllm_config.model_config
        self.parallel_config = vllm_config.parallel_config
        self.cache_config = vllm_config.cache_config
        # per-rank head counts: the KV-head sharding of the path happens here, the kernel never sees TP
        # (flash_attn.py:238-241)
        self.num_heads_q = self.model_config.get_num_attention_heads(self.parallel_config)
        self.num_heads_kv = self.model_config.get_num_kv_heads(self.parallel_config)
        self.headdim = self.model_config.get_head_size()
        self.block_size = kv_cache_spec.block_size
        self.kv_cache_dtype = kv_cache_spec.dtype

    def build(self, common_prefix_len: int, common_attn_metadata: CommonAttentionMetadata,
              fast_build: bool = False) -> B200AttentionMetadata:
        m = common_attn_metadata
        num_reqs, num_actual_tokens = m.num_reqs, m.num_actual_tokens
        num_decodes, num_prefills, num_decode_tokens, num_prefill_tokens = split_decodes_and_prefills(
            m, decode_threshold=self.reorder_batch_threshold)
        assert num_decode_tokens + num_prefill_tokens == num_actual_tokens  # flash_attn.py:313-316
        assert num_decodes + num_prefills == num_reqs
        if common_prefix_len > 0:
            # cascade attention (shared-prefix pass + merge_attn_states) is the engine's optimisation; not built here
            common_prefix_len = 0
        qsl_cpu = m.query_start_loc_cpu
        d = dict(decode_query_start_loc=None, decode_max_seq_len=0, decode_seq_lens=None, decode_block_table=None,
                 prefill_query_start_loc=None, prefill_max_seq_len=0, prefill_seq_lens=None, prefill_block_table=None,
                 prefill_token_seq_lens=None, prefill_token_req=None)
        if num_decodes > 0:  # flash_attn.py:343-360
            d["decode_max_seq_len"] = m.max_seq_len  # an upper bound is all the kernel needs (no host read of seq_lens)
            d["decode_query_start_loc"] = m.query_start_loc[: num_decodes + 1]
            d["decode_seq_lens"] = m.seq_lens[:num_decodes]
            d["decode_block_table"] = m.block_table_tensor[:num_decodes]
        if num_prefills > 0:  # flash_attn.py:362-379
            d["prefill_max_seq_len"] = m.max_seq_len
            d["prefill_query_start_loc"] = m.query_start_loc[num_decodes: num_reqs + 1] - m.query_start_loc[num_decodes]
            d["prefill_seq_lens"] = m.seq_lens[num_decodes:num_reqs]
            d["prefill_block_table"] = m.block_table_tensor[num_decodes:num_reqs]
            # token t of prefill request r (i-th of its q_r query tokens) sees context_r + i + 1 tokens
            qlens_cpu = (qsl_cpu[num_decodes + 1: num_reqs + 1] - qsl_cpu[num_decodes:num_reqs]).to(torch.int64)
            req_cpu = torch.repeat_interleave(torch.arange(num_prefills), qlens_cpu)
            start_cpu = (qsl_cpu[num_decodes:num_reqs] - qsl_cpu[num_decodes]).to(torch.int64)
            pos_cpu = torch.arange(num_prefill_tokens) - start_cpu[req_cpu]
            dev = m.seq_lens.device
            req = req_cpu.to(dev, non_blocking=True)
            ahead = (pos_cpu + 1 - qlens_cpu[req_cpu]).to(torch.int32).to(dev, non_blocking=True)  # i + 1 - q_r  (<= 0)
            d["prefill_token_req"] = req
            d["prefill_token_seq_lens"] = d["prefill_seq_lens"].index_select(0, req) + ahead
        return B200AttentionMetadata(
            num_actual_tokens=num_actual_tokens, max_query_len=m.max_query_len, query_start_loc=m.query_start_loc,
            max_seq_len=m.max_seq_len, seq_lens=m.seq_lens, block_table=m.block_table_tensor, slot_mapping=m.slot_mapping,
            num_decodes=num_decodes, num_decode_tokens=num_decode_tokens, num_prefills=num_prefills,
            num_prefill_tokens=num_prefill_tokens, use_cascade=False, common_prefix_len=common_prefix_len,
            causal=m.causal, **d)

    def use_cascade_attention(self, *args, **kwargs) -> bool:
        return False


class B200AttentionImpl(AttentionImpl[B200AttentionMetadata]):
    """reference: FlashAttentionImpl (flash_attn.py:532-792)."""

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: int | None = None,
                 alibi_slopes: list[float] | None = None, sliding_window: int | None = None, kv_cache_dtype: str = "auto",
                 logits_soft_cap: float | None = None, attn_type: str = AttentionType.DECODER,
                 kv_sharing_target_layer_name: str | None = None, sinks: torch.Tensor | None = None,
                 prefill_as_decode: bool = True) -> None:
        if attn_type != AttentionType.DECODER:
            raise NotImplementedError("B200AttentionImpl serves decoder self-attention only")  # flash_attn.py:590-595 likewise
        if sliding_window is not None or logits_soft_cap or sinks is not None or kv_sharing_target_layer_name is not None:
            raise NotImplementedError("sliding window / logits soft cap / sinks / KV sharing are not served by the decode kernel")
        if kv_cache_dtype not in ("auto", "bfloat16"):
            # the reference's flash path has no fp8 cache either (vllm_metax/attention/utils/fa_utils.py:26-30)
            raise NotImplementedError("the flash-layout decode kernel takes a bf16 cache")
        if head_size not in B200PagedAttentionBackend.get_supported_head_sizes():
            raise ValueError("Head size %d is not supported by B200PagedAttention. Supported head sizes are: %s." % (
                head_size, B200PagedAttentionBackend.get_supported_head_sizes()))
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        self.kv_cache_dtype = "auto"
        self.prefill_as_decode = prefill_as_decode
        self._alibi_list = alibi_slopes
        self._alibi: torch.Tensor | None = None
        self._one: torch.Tensor | None = None

    def _consts(self, device):
        if self._one is None or self._one.device != device:
            self._one = torch.ones(1, dtype=torch.float32, device=device)
            self._alibi = (torch.tensor(self._alibi_list, dtype=torch.float32, device=device)
                           if self._alibi_list is not None else None)
        return self._one, self._alibi

    def forward(self, layer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, kv_cache: torch.Tensor,
                attn_metadata: B200AttentionMetadata, output: torch.Tensor | None = None,
                output_scale: torch.Tensor | None = None, output_block_scale: torch.Tensor | None = None) -> torch.Tensor:
        """query [num_tokens, num_heads, head_size]; key / value [num_tokens, num_kv_heads, head_size];
        kv_cache [2, num_blocks, block_size, num_kv_heads, head_size]; output like query (flash_attn.py:600-640)."""
        assert output is not None, "Output tensor must be provided."
        if output_scale is not None or output_block_scale is not None:
            raise NotImplementedError("fused output quantization is not supported")  # flash_attn.py:636-640
        if attn_metadata is None:  # profiling run (flash_attn.py:642-644)
            return output.fill_(0)
        one, alibi = self._consts(query.device)
        m = attn_metadata
        n = m.num_actual_tokens
        q = query.view(-1, self.num_heads, self.head_size)
        out = output.view(-1, self.num_heads, self.head_size)
        key_cache, value_cache = kv_cache.unbind(0)
        if key is not None and value is not None:
            # key / value may be padded past slot_mapping; the op takes the token count from slot_mapping (:661-687)
            k_scale = getattr(layer, "_k_scale", one)
            v_scale = getattr(layer, "_v_scale", one)
            ops.reshape_and_cache_flash(key.view(-1, self.num_kv_heads, self.head_size),
                                        value.view(-1, self.num_kv_heads, self.head_size), key_cache, value_cache,
                                        m.slot_mapping, self.kv_cache_dtype, k_scale.reshape(1).float(), v_scale.reshape(1).float())
        nd = m.num_decode_tokens
        if nd > 0:
            ops.paged_attention_flash(out[:nd], q[:nd], key_cache, value_cache, self.scale, m.decode_block_table,
                                      m.decode_seq_lens, m.decode_max_seq_len, alibi)
        if m.num_prefill_tokens > 0:
            if not self.prefill_as_decode:
                raise NotImplementedError("prefill tokens in the batch: this backend serves the decode path "
                                          "(prefill_as_decode=True serves them as decode queries over the cache)")
            # every prefill token as a decode query over the first context + i + 1 tokens of its sequence; its own
            # key / value were written to the cache above
            table = m.prefill_block_table.index_select(0, m.prefill_token_req)
            ops.paged_attention_flash(out[nd:n], q[nd:n], key_cache, value_cache, self.scale, table,
                                      m.prefill_token_seq_lens, m.prefill_max_seq_len, alibi)
        return output


def get_attn_backend_cls(selected_backend, head_size: int, dtype: torch.dtype, kv_cache_dtype: str | None, block_size: int,
                         use_v1: bool = True, use_mla: bool = False, has_sink: bool = False,
                         use_sparse: bool = False) -> str | None:
    """Selection rule in the shape of MacaPlatform.get_attn_backend_cls (vllm_metax/platform.py:259-400): the qualified
    class name when this backend serves the configuration, else None so that the platform goes on to its own choices
    (this repository replaces one path and nothing else).  `selected_backend` may name the backend explicitly
    ("B200_PAGED_ATTN"); an explicit selection of an unsupported configuration raises, as the reference does for its
    backends (platform.py:372-389)."""
    reasons = []
    if use_mla or use_sparse:
        reasons.append("MLA / sparse attention")
    if not use_v1:
        reasons.append("the V0 engine")
    if has_sink:
        reasons.append("attention sinks")
    if head_size not in B200PagedAttentionBackend.get_supported_head_sizes():
        reasons.append("head size %d" % head_size)
    if dtype not in B200PagedAttentionBackend.supported_dtypes:
        reasons.append("dtype %s" % dtype)
    if kv_cache_dtype not in (None, "auto", "bfloat16"):
        reasons.append("kv cache dtype %s" % kv_cache_dtype)
    if block_size is not None and block_size % 16 != 0:
        reasons.append("block size %d" % block_size)
    explicit = selected_backend is not None and str(getattr(selected_backend, "name", selected_backend)) == \
        B200PagedAttentionBackend.get_name()
    if reasons:
        if explicit:
            raise ValueError("B200_PAGED_ATTN does not serve: " + ", ".join(reasons))
        return None
    if selected_backend is not None and not explicit:
        return None  # another backend was asked for by name
    return BACKEND_PATH
