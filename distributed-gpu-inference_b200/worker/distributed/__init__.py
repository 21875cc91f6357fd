"""Layer-range inference components (reference: worker/distributed/__init__.py)."""
from .kv_cache import CacheBlock, DistributedKVCacheManager, KVCachePool, PagedKVCache
from .model_shard import ModelShard, ShardedModelLoader, get_layer_range_for_worker
from .servicer import InferenceServicer

__all__ = [
    "ModelShard",
    "ShardedModelLoader",
    "get_layer_range_for_worker",
    "DistributedKVCacheManager",
    "PagedKVCache",
    "KVCachePool",
    "CacheBlock",
    "InferenceServicer",
]
