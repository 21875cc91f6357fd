"""Shared pieces of the decode path (reference: common/__init__.py)."""
from .data_structures import (
    BlockRange,
    InferenceState,
    KVCacheBlock,
    ModelShardConfig,
    SessionConfig,
    WorkerInfo,
    WorkerRole,
    WorkerState,
    compute_prefix_hash,
    estimate_kv_cache_size,
)
from .serialization import StreamingTensorBuffer, TensorSerializer, deserialize_tensor, serialize_tensor

__all__ = [
    "BlockRange",
    "WorkerInfo",
    "WorkerRole",
    "WorkerState",
    "InferenceState",
    "KVCacheBlock",
    "SessionConfig",
    "ModelShardConfig",
    "TensorSerializer",
    "StreamingTensorBuffer",
    "serialize_tensor",
    "deserialize_tensor",
    "compute_prefix_hash",
    "estimate_kv_cache_size",
]
