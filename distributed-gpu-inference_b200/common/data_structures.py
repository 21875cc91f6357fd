"""Plain data types shared by the layer-range decode path.

Mirrors the subset of the reference's ``common/data_structures.py`` that the hot path touches
(reference: common/data_structures.py:13-309).  Field names, defaults and method behaviour are kept so
callers written against the reference keep working; everything the control plane alone uses is left out.
"""
from __future__ import annotations

import hashlib
import time
from dataclasses import dataclass, field
from enum import Enum
from typing import Any, Dict, List, Optional


class WorkerRole(Enum):
    """Which phase a worker serves (reference: common/data_structures.py:13-17)."""

    PREFILL = "prefill"
    DECODE = "decode"
    HYBRID = "hybrid"


class WorkerState(Enum):
    """Liveness of a worker (reference: common/data_structures.py:20-26)."""

    OFFLINE = 0
    JOINING = 1
    ONLINE = 2
    BUSY = 3
    ERROR = 4


@dataclass
class BlockRange:
    """Half-open range of decoder layers ``[start, end)`` (reference: common/data_structures.py:29-47)."""

    start: int
    end: int

    @property
    def length(self) -> int:
        return self.end - self.start

    def __contains__(self, layer_idx: int) -> bool:
        return self.start <= layer_idx < self.end

    def to_dict(self) -> Dict[str, int]:
        return {"start": self.start, "end": self.end}

    @classmethod
    def from_dict(cls, data: Dict[str, int]) -> "BlockRange":
        return cls(start=data["start"], end=data["end"])


@dataclass
class WorkerInfo:
    """One stage of a route (reference: common/data_structures.py:50-120).

    Only the fields the session/pipeline code reads are carried; ``rank`` is new and names the
    ``torch.distributed`` rank that owns the stage when the hop is NCCL instead of HTTP.
    """

    worker_id: str
    blocks: Optional[BlockRange] = None
    role: WorkerRole = WorkerRole.HYBRID
    state: WorkerState = WorkerState.OFFLINE
    gpu_name: str = ""
    gpu_memory_gb: float = 0.0
    gpu_memory_used_gb: float = 0.0
    throughput_tokens_per_sec: float = 0.0
    latency_ms: float = 0.0
    reliability_score: float = 1.0
    peer_address: str = ""
    api_endpoint: str = ""
    cache_tokens_available: int = 0
    cache_tokens_used: int = 0
    model_id: str = ""
    supported_models: List[str] = field(default_factory=list)
    last_heartbeat: float = field(default_factory=time.time)
    rank: int = -1

    @property
    def cache_utilization(self) -> float:
        return self.cache_tokens_used / self.cache_tokens_available if self.cache_tokens_available else 0.0

    @property
    def gpu_utilization(self) -> float:
        return self.gpu_memory_used_gb / self.gpu_memory_gb if self.gpu_memory_gb else 0.0

    def is_healthy(self, timeout_seconds: float = 60.0) -> bool:
        if self.state in (WorkerState.OFFLINE, WorkerState.ERROR):
            return False
        return (time.time() - self.last_heartbeat) < timeout_seconds

    def to_dict(self) -> Dict[str, Any]:
        out = dict(self.__dict__)
        out["blocks"] = self.blocks.to_dict() if self.blocks else None
        out["role"] = self.role.value
        out["state"] = self.state.value
        out.pop("rank", None)
        return out


@dataclass
class SessionConfig:
    """Knobs of a distributed session (reference: common/data_structures.py:231-254)."""

    model_name: str
    max_length: int = 4096
    temperature: float = 0.7
    top_p: float = 0.9
    top_k: int = 50
    use_cache: bool = True
    stream: bool = False
    connect_timeout: float = 30.0
    request_timeout: float = 120.0
    max_retries: int = 3
    use_speculative_decoding: bool = False
    speculative_depth: int = 5


def compute_prefix_hash(token_ids: List[int]) -> str:
    """First 16 hex digits of sha256 over ``bytes(token_ids)`` (reference: common/data_structures.py:293-296).

    ``bytes()`` of a list only accepts values 0..255, so — exactly like the reference — this raises
    ``ValueError`` for larger token ids.  ``compute_prefix_hash_wide`` below is the documented extension.
    """
    return hashlib.sha256(bytes(token_ids)).hexdigest()[:16]


def compute_prefix_hash_wide(token_ids: List[int]) -> str:
    """Extension for real vocabularies: little-endian uint32 per token, same digest and truncation.

    No reference counterpart (SURVEY.md §9 defect 7); equals ``compute_prefix_hash`` only for empty input.
    """
    buf = b"".join(int(t).to_bytes(4, "little", signed=False) for t in token_ids)
    return hashlib.sha256(buf).hexdigest()[:16]


def estimate_kv_cache_size(
    num_layers: int,
    num_heads: int,
    head_dim: int,
    seq_length: int,
    batch_size: int = 1,
    dtype_bytes: int = 2,
) -> int:
    """Bytes of K+V for a batch (reference: common/data_structures.py:299-309)."""
    return 2 * num_layers * batch_size * seq_length * num_heads * head_dim * dtype_bytes


# Records the decode path itself never touches (per-session cursor, wire-level block description, layer→worker
# map) live in records.py; they stay importable from here, as in the reference.
from .records import InferenceState, KVCacheBlock, ModelShardConfig  # noqa: E402,F401
