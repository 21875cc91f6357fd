"""Plain records of the reference's ``common/data_structures.py`` that the decode path does not use itself but
callers may construct: the per-session cursor, the wire-level description of a KV block, the layer→worker map
(reference: common/data_structures.py:123-180, 257-290).  Field names and defaults are the reference's."""
from __future__ import annotations

import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple


@dataclass
class InferenceState:
    """Per-session cursor handed between stages (reference: common/data_structures.py:123-143)."""

    session_id: str
    position: int = 0
    kv_cache_keys: List[str] = field(default_factory=list)
    hidden_states_data: Optional[bytes] = None
    hidden_states_shape: Optional[Tuple[int, ...]] = None
    hidden_states_dtype: str = "float16"
    input_tokens: int = 0
    output_tokens: int = 0
    created_at: float = field(default_factory=time.time)
    updated_at: float = field(default_factory=time.time)

    def update_position(self, new_tokens: int) -> None:
        self.position += new_tokens
        self.updated_at = time.time()


@dataclass
class KVCacheBlock:
    """Wire-level description of one KV block (reference: common/data_structures.py:146-180)."""

    block_id: str
    layer_idx: int
    block_size: int = 16
    keys_data: Optional[bytes] = None
    values_data: Optional[bytes] = None
    num_heads: int = 0
    head_dim: int = 0
    ref_count: int = 1
    prefix_hash: str = ""
    location: str = "gpu"

    @property
    def is_shared(self) -> bool:
        return self.ref_count > 1

    def increment_ref(self) -> None:
        self.ref_count += 1

    def decrement_ref(self) -> int:
        # Floors at zero, as the reference does (tests/test_common_data_structures.py:57-65).
        self.ref_count = max(0, self.ref_count - 1)
        return self.ref_count


@dataclass
class ModelShardConfig:
    """Layer → worker map (reference: common/data_structures.py:257-290)."""

    model_id: str
    total_layers: int
    shard_mapping: Dict[str, "BlockRange"] = field(default_factory=dict)
    hidden_size: int = 0
    num_attention_heads: int = 0
    num_key_value_heads: int = 0
    intermediate_size: int = 0
    vocab_size: int = 0
    memory_per_layer_gb: float = 0.0
    kv_cache_per_token_bytes: int = 0

    def get_worker_for_layer(self, layer_idx: int) -> Optional[str]:
        for worker_id, rng in self.shard_mapping.items():
            if layer_idx in rng:
                return worker_id
        return None

    def get_inference_route(self) -> List[Tuple[str, "BlockRange"]]:
        return sorted(self.shard_mapping.items(), key=lambda kv: kv[1].start)
