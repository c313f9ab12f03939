"""Host-side callers of the decode hot path (the layer the reference's attention backends sit in)."""
from .paged_attn import PagedAttention, PinnedDecodeInputs, shard_kv_heads  # noqa: F401
from .flash_decode import flash_attn_with_kvcache  # noqa: F401,E402
from .backend import B200FlashDecodeBackend, B200FlashDecodeImpl, DecodeMetadata  # noqa: F401,E402
from .dcp import DcpExchange, dcp_local_seq_lens, dcp_token_owner, forward_decode_dcp  # noqa: F401,E402
