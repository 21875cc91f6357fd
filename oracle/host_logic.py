"""Oracle (test infrastructure): pure-Python restatement of the reference's host-side logic on the hot path.

No torch, no product imports.  Each function cites what it restates; oracle/make_golden.py pins every one of
them against the reference's own code.
"""
from __future__ import annotations

import hashlib
import struct
from collections import OrderedDict
from typing import Dict, List, Optional, Tuple


def layer_range(total_layers: int, num_workers: int, worker_idx: int) -> Tuple[int, int]:
    """Restates get_layer_range_for_worker (reference: worker/distributed/model_shard.py:372-394)."""
    per = total_layers // num_workers
    rem = total_layers % num_workers
    start = worker_idx * per + (worker_idx if worker_idx < rem else rem)
    end = start + per + (1 if worker_idx < rem else 0)
    return start, end


def prefix_hash(token_ids: List[int]) -> str:
    """Restates compute_prefix_hash (reference: common/data_structures.py:293-296, kv_cache.py:384-387)."""
    return hashlib.sha256(bytes(token_ids)).hexdigest()[:16]


DTYPE_IDS = {"float16": 0, "float32": 1, "bfloat16": 2, "int8": 3, "int32": 4, "int64": 5}


def stream_header(shape: Tuple[int, ...], dtype_str: str) -> bytes:
    """Restates StreamingTensorBuffer.write_header (reference: common/serialization.py:218-228)."""
    out = struct.pack("I", len(shape))
    for dim in shape:
        out += struct.pack("Q", dim)
    return out + struct.pack("B", DTYPE_IDS.get(dtype_str, 0))


def blocks_for(seq_len: int, block_size: int) -> int:
    """Ceil-div block count of KVCachePool.allocate_sequence (reference: worker/distributed/kv_cache.py:298-299)."""
    return (seq_len + block_size - 1) // block_size


def slot_mapping(block_table: List[int], positions: List[int], block_size: int) -> List[int]:
    """Device-side addressing derived in SURVEY.md §8(a): slot = table[t // bs] * bs + t % bs."""
    return [block_table[t // block_size] * block_size + t % block_size for t in positions]


class RefAllocator:
    """Id-level restatement of PagedKVCache's allocator (reference: worker/distributed/kv_cache.py:153-238).

    Tracks only what decides WHICH block id comes next: the free list (popped from the end, appended on free),
    ref-counts, and the LRU order used by eviction.  ``touch`` models get_block's move-to-end.
    """

    def __init__(self, max_blocks: int):
        self.free: List[str] = [f"block_{i}" for i in range(max_blocks)]
        self.ref: Dict[str, int] = {}
        self.lru: "OrderedDict[str, None]" = OrderedDict()
        self.stats = {"allocations": 0, "evictions": 0, "hits": 0, "misses": 0}

    def allocate(self) -> Optional[str]:
        if not self.free and not self._evict():
            return None
        bid = self.free.pop()
        self.ref[bid] = 1
        self.lru[bid] = None
        self.stats["allocations"] += 1
        return bid

    def add_ref(self, bid: str) -> None:
        self.ref[bid] += 1

    def release(self, bid: str) -> None:
        if bid not in self.ref:
            return
        self.ref[bid] = max(0, self.ref[bid] - 1)
        if self.ref[bid] == 0:
            del self.ref[bid]
            self.lru.pop(bid, None)
            self.free.append(bid)

    def touch(self, bid: Optional[str]) -> bool:
        if bid in self.ref:
            self.lru.move_to_end(bid)
            self.stats["hits"] += 1
            return True
        self.stats["misses"] += 1
        return False

    def _evict(self) -> bool:
        for bid in self.lru:
            if self.ref.get(bid) == 1:
                self.release(bid)
                self.stats["evictions"] += 1
                return True
        return False


def run_allocator_schedule(max_blocks: int, schedule: List[Tuple[str, int]]) -> Dict[str, object]:
    """Replay a scripted schedule on RefAllocator.

    ``schedule`` entries: ("alloc", _), ("free", k) frees the k-th allocated id, ("ref", k) adds a reference to
    it, ("touch", k) looks it up.  Returns the sequence of ids handed out (None = exhausted), the final free
    list and the stats — the quantities the golden fixture stores.
    """
    alloc = RefAllocator(max_blocks)
    handed: List[Optional[str]] = []
    for op, arg in schedule:
        if op == "alloc":
            handed.append(alloc.allocate())
        elif op == "free":
            if handed[arg] is not None:
                alloc.release(handed[arg])
        elif op == "ref":
            if handed[arg] is not None and handed[arg] in alloc.ref:
                alloc.add_ref(handed[arg])
        elif op == "touch":
            alloc.touch(handed[arg])
        else:
            raise ValueError(op)
    return {
        "handed": handed,
        "free": list(alloc.free),
        "stats": dict(alloc.stats, total_blocks=len(alloc.ref), free_blocks=len(alloc.free)),
    }
