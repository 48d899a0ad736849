"""b200attn: B200-native (sm_100a) paged decode attention behind the sglang
``AttentionBackend`` plug-in surface.  Importing the package loads
``libb200attn.so`` and registers ``torch.ops.b200attn.*``; it raises if the
library has not been built -- there is no CPU fallback.
"""
from . import capi

capi.lib()

from . import ops  # noqa: E402  (registers torch.ops.b200attn.*)
from .backend import B200AttnBackend, B200MultiStepDraftBackend, ExtendMetadata, ForwardMetadata, rope_store_kv, store_kv  # noqa: E402
from .interfaces import (  # noqa: E402
    AttentionBackend,
    AttentionType,
    AttnLayer,
    ForwardBatch,
    ForwardMode,
    MHATokenToKVPool,
    ReqToTokenPool,
    SWAKVPool,
)

__all__ = [
    "capi",
    "ops",
    "B200AttnBackend",
    "B200MultiStepDraftBackend",
    "ForwardMetadata",
    "ExtendMetadata",
    "store_kv",
    "rope_store_kv",
    "AttentionBackend",
    "AttnLayer",
    "AttentionType",
    "ForwardBatch",
    "ForwardMode",
    "MHATokenToKVPool",
    "ReqToTokenPool",
    "SWAKVPool",
]
