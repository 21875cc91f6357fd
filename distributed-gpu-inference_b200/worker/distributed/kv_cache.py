"""Paged KV-cache allocator with the reference's semantics, laid out for the CUDA kernels.

Drop-in for the reference's ``worker/distributed/kv_cache.py``: same classes, constructor arguments, return
values, ``get_stats()`` keys and — bit for bit — the same block-id sequence for any alloc/free/evict schedule
(pinned by tests/golden/allocator_*.json which were produced by the reference itself).

What is kept exactly
  * pool layout ``[max_blocks, num_heads, block_size, head_dim]`` for K and for V (reference :130-144)
  * ids ``block_{slot}``, free list initialised ascending and popped from the END (reference :177)
  * ``free_block``: ref-count floor 0, slot zeroed, id appended to the END of the free list (reference :196-216)
  * LRU eviction of the oldest ``ref_count == 1`` entry (reference :229-238)
  * ``KVCachePool.allocate_sequence``: ceil-div block count, layer-major (reference :283-313)
  * tier keys ``"{prefix_hash}:{layer}"``, CPU LRU capacity formula, pickle payload (reference :326-555)

Deliberate deviations (each is listed in DESIGN.md)
  1. pools are created on any device, not only CUDA (#1) — the reference leaves CPU caches unusable
  2. ``KVCachePool`` / ``DistributedKVCacheManager`` accept ``dtype=`` (#2); default stays fp16
  3. blocks handed out earlier in the same ``allocate_sequence`` call, and blocks pinned with ``pin()``,
     are skipped by eviction (#8); a failing ``allocate_sequence`` also returns the partial layer (#9)
  4. freeing a block drops its ``_prefix_index`` entry (#11)

Added for the device path (no reference counterpart): ``slot_of``, ``pin``/``unpin``, ``block_table`` and the
single backing allocation in ``KVCachePool`` whose per-layer views are what the C-ABI kernels receive.
"""
from __future__ import annotations

import asyncio
import hashlib
import io
import logging
import pickle
import time
from collections import OrderedDict
from dataclasses import dataclass, field  # noqa: F401
from enum import Enum
from typing import Any, Dict, Iterable, List, Optional, Sequence, Tuple

import torch

logger = logging.getLogger(__name__)


class CacheLocation(Enum):
    GPU = "gpu"
    CPU = "cpu"
    REDIS = "redis"
    REMOTE = "remote"


class CacheBlock:
    """One fixed-size page of K and V for one layer (reference: worker/distributed/kv_cache.py:34-76).

    Same constructor, fields and defaults as the reference's dataclass.  ``keys`` / ``values`` are views into the
    owning pool, shape ``[num_heads, block_size, head_dim]``; for blocks handed out by ``PagedKVCache`` the views are
    made on first access (a 4k-token sequence of a 32-layer model is 8192 blocks — building 16384 tensor views
    nobody reads was most of the cost of ``allocate_sequence``), the device path only needs ``slot``.
    """

    __slots__ = ("block_id", "block_size", "_keys", "_values", "layer_idx", "num_tokens", "ref_count", "prefix_hash",
                 "location", "last_access", "created_at", "_owner", "_slot")

    def __init__(self, block_id: str, block_size: int = 16, keys: Optional[torch.Tensor] = None,
                 values: Optional[torch.Tensor] = None, layer_idx: int = 0, num_tokens: int = 0, ref_count: int = 1,
                 prefix_hash: str = "", location: CacheLocation = CacheLocation.GPU,
                 last_access: Optional[float] = None, created_at: Optional[float] = None):
        now = time.time() if last_access is None or created_at is None else 0.0
        self.block_id = block_id
        self.block_size = block_size
        self._keys = keys
        self._values = values
        self.layer_idx = layer_idx
        self.num_tokens = num_tokens
        self.ref_count = ref_count
        self.prefix_hash = prefix_hash
        self.location = location
        self.last_access = now if last_access is None else last_access
        self.created_at = now if created_at is None else created_at
        self._owner = None  # the PagedKVCache whose pool rows back this block, if any
        self._slot = -1

    @property
    def keys(self) -> Optional[torch.Tensor]:
        if self._keys is None and self._owner is not None:
            self._keys = self._owner._key_pool[self._slot]
        return self._keys

    @keys.setter
    def keys(self, t: Optional[torch.Tensor]) -> None:
        self._keys = t

    @property
    def values(self) -> Optional[torch.Tensor]:
        if self._values is None and self._owner is not None:
            self._values = self._owner._value_pool[self._slot]
        return self._values

    @values.setter
    def values(self, t: Optional[torch.Tensor]) -> None:
        self._values = t

    def __repr__(self) -> str:
        return (f"CacheBlock(block_id={self.block_id!r}, block_size={self.block_size}, layer_idx={self.layer_idx}, "
                f"num_tokens={self.num_tokens}, ref_count={self.ref_count}, prefix_hash={self.prefix_hash!r}, "
                f"location={self.location})")

    @property
    def is_full(self) -> bool:
        return self.num_tokens >= self.block_size

    @property
    def is_shared(self) -> bool:
        return self.ref_count > 1

    @property
    def slot(self) -> int:
        """Pool row this block occupies (the integer after ``block_``)."""
        if self._slot < 0:
            self._slot = int(self.block_id.rsplit("_", 1)[1])
        return self._slot

    def add_ref(self) -> None:
        self.ref_count += 1

    def remove_ref(self) -> int:
        self.ref_count = max(0, self.ref_count - 1)
        return self.ref_count

    def touch(self) -> None:
        self.last_access = time.time()


class PagedKVCache:
    """Block allocator over one K pool and one V pool (reference: worker/distributed/kv_cache.py:79-247)."""

    def __init__(
        self,
        num_layers: int,
        num_heads: int,
        head_dim: int,
        block_size: int = 16,
        max_blocks: int = 1000,
        device: str = "cuda",
        dtype: torch.dtype = torch.float16,
        *,
        _storage: Optional[Tuple[torch.Tensor, torch.Tensor]] = None,
    ):
        self.num_layers = num_layers
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.block_size = block_size
        self.max_blocks = max_blocks
        self.device = device
        self.dtype = dtype

        self._blocks: Dict[str, CacheBlock] = {}
        self._free_blocks: List[str] = []
        self._key_pool: Optional[torch.Tensor] = None
        self._value_pool: Optional[torch.Tensor] = None
        self._block_to_slot: Dict[str, int] = {}
        self._lru_queue: "OrderedDict[str, float]" = OrderedDict()
        self._pinned: Dict[str, int] = {}  # block id -> number of pins (live sequences that use the block)
        self._stats = {"allocations": 0, "evictions": 0, "hits": 0, "misses": 0}
        self._on_free = None  # optional callback(block_id), used by the tier manager to drop index entries

        self._init_memory_pool(_storage)

    # -- pool ---------------------------------------------------------------------------------------
    def _init_memory_pool(self, storage: Optional[Tuple[torch.Tensor, torch.Tensor]] = None) -> None:
        shape = (self.max_blocks, self.num_heads, self.block_size, self.head_dim)
        if storage is not None:
            k, v = storage
            if tuple(k.shape) != shape or tuple(v.shape) != shape or not (k.is_contiguous() and v.is_contiguous()):
                raise ValueError(f"backing storage must be two contiguous tensors of shape {shape}")
            self._key_pool, self._value_pool = k, v
        else:
            self._key_pool = torch.zeros(shape, dtype=self.dtype, device=self.device)
            self._value_pool = torch.zeros(shape, dtype=self.dtype, device=self.device)
        self._free_blocks = [f"block_{i}" for i in range(self.max_blocks)]
        self._slot_of_id = {bid: i for i, bid in enumerate(self._free_blocks)}

    def _get_pool_memory_gb(self) -> float:
        if self._key_pool is None:
            return 0.0
        return 2 * self._key_pool.numel() * self._key_pool.element_size() / (1024 ** 3)

    @property
    def key_pool(self) -> torch.Tensor:
        return self._key_pool

    @property
    def value_pool(self) -> torch.Tensor:
        return self._value_pool

    # -- allocation ---------------------------------------------------------------------------------
    def allocate_block(self, layer_idx: int, prefix_hash: str = "") -> Optional[CacheBlock]:
        if not self._free_blocks and not self._evict_lru():
            logger.warning("No free blocks and eviction failed")
            return None
        block_id = self._free_blocks.pop()
        now = time.time()
        blk = CacheBlock(block_id, self.block_size, None, None, layer_idx, 0, 1, prefix_hash, CacheLocation.GPU, now, now)
        blk._owner = self
        slot = blk._slot = self._slot_of_id[block_id]
        self._blocks[block_id] = blk
        self._block_to_slot[block_id] = slot
        self._lru_queue[block_id] = now
        self._stats["allocations"] += 1
        return blk

    def allocate_blocks(self, layer_idx: int, n: int, prefix_hash: str = "") -> Optional[List[CacheBlock]]:
        """``n`` × ``allocate_block`` with the same observable result (ids in pop order, LRU order, statistics) when
        the free list already holds ``n`` ids; returns None — having changed nothing — when it does not, so that the
        caller falls back to the block-by-block path with its eviction."""
        if n > len(self._free_blocks):
            return None
        if n <= 0:
            return []
        ids = self._free_blocks[-n:]
        ids.reverse()
        del self._free_blocks[-n:]
        now = time.time()
        slot_of, bs, gpu = self._slot_of_id, self.block_size, CacheLocation.GPU
        out = []
        for block_id in ids:
            blk = CacheBlock(block_id, bs, None, None, layer_idx, 0, 1, prefix_hash, gpu, now, now)
            blk._owner = self
            blk._slot = slot_of[block_id]
            out.append(blk)
        self._blocks.update(zip(ids, out))
        self._block_to_slot.update((i, slot_of[i]) for i in ids)
        self._lru_queue.update((i, now) for i in ids)
        self._stats["allocations"] += n
        return out

    def free_block(self, block_id: str) -> None:
        blk = self._blocks.get(block_id)
        if blk is None:
            return
        if blk.remove_ref() != 0:
            return
        slot = self._block_to_slot.get(block_id)
        if slot is not None:
            self._key_pool[slot].zero_()
            self._value_pool[slot].zero_()
        del self._blocks[block_id]
        self._block_to_slot.pop(block_id, None)
        self._lru_queue.pop(block_id, None)
        self._pinned.pop(block_id, None)
        self._free_blocks.append(block_id)
        if self._on_free is not None:
            self._on_free(block_id)

    def free_blocks(self, block_ids: Iterable[str]) -> None:
        """``free_block`` for many ids with the same observable result, but the pool rows that reach ref-count 0
        are zeroed with one ``index_fill_`` per pool instead of two launches per block."""
        dead: List[int] = []
        blocks, slots, lru, pinned, free = self._blocks, self._block_to_slot, self._lru_queue, self._pinned, self._free_blocks
        on_free = self._on_free
        for block_id in block_ids:
            blk = blocks.get(block_id)
            if blk is None:
                continue
            if blk.ref_count > 1:  # still shared: only drop this holder's reference
                blk.ref_count -= 1
                continue
            blk.ref_count = 0
            slot = slots.pop(block_id, None)
            if slot is not None:
                dead.append(slot)
            del blocks[block_id]
            lru.pop(block_id, None)
            pinned.pop(block_id, None)
            free.append(block_id)
            if on_free is not None:
                on_free(block_id)
        if dead:
            idx = torch.tensor(dead, dtype=torch.int64, device=self._key_pool.device)
            self._key_pool.index_fill_(0, idx, 0)
            self._value_pool.index_fill_(0, idx, 0)

    def get_block(self, block_id: Optional[str]) -> Optional[CacheBlock]:
        blk = self._blocks.get(block_id)
        if blk is None:
            self._stats["misses"] += 1
            return None
        blk.touch()
        self._lru_queue.move_to_end(block_id)
        self._stats["hits"] += 1
        return blk

    def _evict_lru(self) -> bool:
        for block_id in self._lru_queue:
            blk = self._blocks.get(block_id)
            if blk is not None and blk.ref_count == 1 and block_id not in self._pinned:
                self.free_block(block_id)
                self._stats["evictions"] += 1
                return True
        return False

    # -- additions for the device path --------------------------------------------------------------
    def slot_of(self, block_id: str) -> int:
        return self._block_to_slot[block_id]

    def pin(self, block_ids: Iterable[str]) -> None:
        """Protect blocks of a live sequence from LRU eviction.  Pins are counted: a block shared
        by several live sequences is pinned once per sequence and stays protected until each has called ``unpin``.
        Freeing a sequence lifts no pin of a block that survives the free — whoever pinned must unpin."""
        pinned = self._pinned
        for b in block_ids:
            if b in self._blocks:
                pinned[b] = pinned.get(b, 0) + 1

    def unpin(self, block_ids: Iterable[str]) -> None:
        pinned = self._pinned
        for b in block_ids:
            n = pinned.get(b, 0)
            if n > 1:
                pinned[b] = n - 1
            elif n == 1:
                del pinned[b]

    def get_stats(self) -> Dict[str, Any]:
        return {
            **self._stats,
            "total_blocks": len(self._blocks),
            "free_blocks": len(self._free_blocks),
            "memory_gb": self._get_pool_memory_gb(),
        }


class KVCachePool:
    """One ``PagedKVCache`` per layer (reference: worker/distributed/kv_cache.py:250-323).

    All layers share one backing tensor ``[num_layers, 2, max_blocks, heads, block_size, head_dim]`` so the pools
    of a shard are a single HBM allocation; layer ``l`` uses ``backing[l, 0]`` for K and ``backing[l, 1]`` for V.
    """

    def __init__(
        self,
        num_layers: int,
        num_heads: int,
        head_dim: int,
        block_size: int = 16,
        max_blocks_per_layer: int = 100,
        device: str = "cuda",
        dtype: torch.dtype = torch.float16,
        backing: Optional[torch.Tensor] = None,
    ):
        self.num_layers = num_layers
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.block_size = block_size
        self.max_blocks_per_layer = max_blocks_per_layer
        self.device = device
        self.dtype = dtype
        shape = (num_layers, 2, max_blocks_per_layer, num_heads, block_size, head_dim)
        if backing is None:
            backing = torch.zeros(shape, dtype=dtype, device=device)
        elif tuple(backing.shape) != shape or backing.dtype != dtype or not backing.is_contiguous():
            # e.g. a symmetric-memory allocation whose peers can be written over NVLink (dgi_b200.kv_transfer)
            raise ValueError(f"backing must be a contiguous {dtype} tensor of shape {shape}")
        self.backing = backing
        self._layer_caches: List[PagedKVCache] = [
            PagedKVCache(
                num_layers=1,
                num_heads=num_heads,
                head_dim=head_dim,
                block_size=block_size,
                max_blocks=max_blocks_per_layer,
                device=device,
                dtype=dtype,
                _storage=(self.backing[l, 0], self.backing[l, 1]),
            )
            for l in range(num_layers)
        ]

    def allocate_sequence(self, seq_len: int, prefix_hash: str = "",
                          shared_prefix: Optional[List[List[CacheBlock]]] = None) -> List[List[CacheBlock]]:
        """Blocks for ``seq_len`` tokens in every layer (reference :283-313).

        ``shared_prefix`` (addition): per layer, blocks of an existing sequence that hold the
        same leading tokens.  They are reused with ``add_ref`` — the copy-on-write sharing ``CacheBlock.ref_count``
        was designed for (reference :63-72) — and only the remainder is freshly allocated, so the sequences' block
        tables start with identical slots.  Shared blocks (ref_count > 1) are never evicted (reference :232-234).
        """
        bs = self._layer_caches[0].block_size
        n_blocks = (seq_len + bs - 1) // bs
        out: List[List[CacheBlock]] = []
        held: List[List[str]] = []  # per layer: ids pinned by THIS call (never a shared block its owner pinned)
        for layer_idx, cache in enumerate(self._layer_caches):
            mine: List[CacheBlock] = []
            if shared_prefix is not None:
                for blk in shared_prefix[layer_idx][:n_blocks]:
                    blk.add_ref()
                    mine.append(blk)
            fresh = cache.allocate_blocks(layer_idx, n_blocks - len(mine), prefix_hash)
            held.append([])
            if fresh is not None:  # enough free ids: no eviction can happen, so nothing needs protecting either
                mine.extend(fresh)
                out.append(mine)
                continue
            for _ in range(n_blocks - len(mine)):
                blk = cache.allocate_block(layer_idx, prefix_hash)
                if blk is None:
                    for li, ids in enumerate(held):
                        self._layer_caches[li].unpin(ids)
                    self._free_sequence_blocks(out + [mine])
                    raise RuntimeError(f"Failed to allocate KV-Cache for layer {layer_idx}")
                cache.pin((blk.block_id,))  # not evictable by the rest of this call
                held[layer_idx].append(blk.block_id)
                mine.append(blk)
            out.append(mine)
        for layer_idx, ids in enumerate(held):
            self._layer_caches[layer_idx].unpin(ids)
        return out

    def _free_sequence_blocks(self, blocks: List[List[CacheBlock]]) -> None:
        for layer_idx, layer_blocks in enumerate(blocks):
            cache = self._layer_caches[layer_idx]
            # pins belong to whoever set them (a block that another holder still references keeps them); a block
            # this call releases for good loses its pin entry together with everything else (free_blocks)
            cache.free_blocks([blk.block_id for blk in layer_blocks])

    def free_sequence(self, blocks: List[List[CacheBlock]]) -> None:
        """Public spelling of the reference's private ``_free_sequence_blocks``."""
        self._free_sequence_blocks(blocks)

    def get_total_memory_gb(self) -> float:
        return sum(c._get_pool_memory_gb() for c in self._layer_caches)

    # -- additions for the device path --------------------------------------------------------------
    def layer(self, layer_idx: int) -> PagedKVCache:
        return self._layer_caches[layer_idx]

    @staticmethod
    def block_table(layer_blocks: Sequence[CacheBlock]) -> List[int]:
        """Pool slots of a sequence's blocks in allocation order — one row of the device block table."""
        return [b.slot for b in layer_blocks]


class DistributedKVCacheManager:
    """GPU → CPU → Redis tiers keyed by prefix (reference: worker/distributed/kv_cache.py:326-555)."""

    def __init__(
        self,
        num_layers: int,
        num_heads: int,
        head_dim: int,
        gpu_cache_blocks: int = 500,
        cpu_cache_gb: float = 16.0,
        redis_client=None,
        block_size: int = 16,
        device: str = "cuda",
        dtype: torch.dtype = torch.float16,
    ):
        self.num_layers = num_layers
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.block_size = block_size
        self.gpu_cache = KVCachePool(
            num_layers=num_layers,
            num_heads=num_heads,
            head_dim=head_dim,
            block_size=block_size,
            max_blocks_per_layer=gpu_cache_blocks,
            device=device,
            dtype=dtype,
        )
        self.cpu_cache: "OrderedDict[str, Tuple[torch.Tensor, torch.Tensor]]" = OrderedDict()
        # capacity in items; the 2-byte element size is hard-wired as in the reference (:366-368)
        self.cpu_cache_max_items = int(cpu_cache_gb * 1024 * 1024 * 1024 / (2 * num_heads * block_size * head_dim * 2))
        self.redis = redis_client
        self._prefix_index: Dict[str, str] = {}
        self._stats = {"l1_hits": 0, "l2_hits": 0, "l3_hits": 0, "l4_hits": 0, "misses": 0}
        for layer_idx, cache in enumerate(self.gpu_cache._layer_caches):
            cache._on_free = self._make_index_dropper(layer_idx)

    def _make_index_dropper(self, layer_idx: int):
        suffix = f":{layer_idx}"

        def drop(block_id: str) -> None:
            stale = [k for k, v in self._prefix_index.items() if v == block_id and k.endswith(suffix)]
            for k in stale:
                del self._prefix_index[k]

        return drop

    def compute_prefix_hash(self, token_ids: List[int]) -> str:
        return hashlib.sha256(bytes(token_ids)).hexdigest()[:16]

    async def get_or_compute(self, prefix_hash: str, layer_idx: int, compute_fn) -> Tuple[torch.Tensor, torch.Tensor]:
        cache_key = f"{prefix_hash}:{layer_idx}"
        layer_cache = self.gpu_cache._layer_caches[layer_idx]

        blk = layer_cache.get_block(self._prefix_index.get(cache_key))
        if blk is not None and blk.keys is not None:
            self._stats["l1_hits"] += 1
            return blk.keys, blk.values

        if cache_key in self.cpu_cache:
            keys, values = self.cpu_cache[cache_key]
            self.cpu_cache.move_to_end(cache_key)
            self._stats["l2_hits"] += 1
            await self._promote_to_gpu(cache_key, keys, values, layer_idx, prefix_hash)
            dev = self.gpu_cache._layer_caches[0].device
            return keys.to(dev), values.to(dev)

        if self.redis:
            blob = await self._get_from_redis(cache_key)
            if blob:
                keys, values = self._deserialize_kv(blob)
                self._stats["l3_hits"] += 1
                self._add_to_cpu_cache(cache_key, keys, values)
                await self._promote_to_gpu(cache_key, keys, values, layer_idx, prefix_hash)
                return keys, values

        self._stats["misses"] += 1
        keys, values = await compute_fn()
        await self._store_kv(cache_key, keys, values, layer_idx, prefix_hash)
        return keys, values

    async def _promote_to_gpu(self, cache_key, keys, values, layer_idx, prefix_hash) -> None:
        blk = self.gpu_cache._layer_caches[layer_idx].allocate_block(layer_idx, prefix_hash)
        if blk is not None and blk.keys is not None:
            blk.keys.copy_(keys)
            blk.values.copy_(values)
            self._prefix_index[cache_key] = blk.block_id

    def _add_to_cpu_cache(self, cache_key, keys, values) -> None:
        while len(self.cpu_cache) >= self.cpu_cache_max_items:
            self.cpu_cache.popitem(last=False)
        self.cpu_cache[cache_key] = (keys.cpu(), values.cpu())

    async def _get_from_redis(self, cache_key: str) -> Optional[bytes]:
        if self.redis is None:
            return None
        try:
            return await self.redis.get(f"kv:{cache_key}")
        except Exception as exc:  # noqa: BLE001 - same swallow-and-log contract as the reference
            logger.warning(f"Redis get error: {exc}")
            return None

    async def _store_kv(self, cache_key, keys, values, layer_idx, prefix_hash) -> None:
        await self._promote_to_gpu(cache_key, keys, values, layer_idx, prefix_hash)
        self._add_to_cpu_cache(cache_key, keys, values)
        if self.redis:
            asyncio.create_task(self._write_to_redis(cache_key, keys, values))

    async def _write_to_redis(self, cache_key, keys, values, ttl: int = 3600) -> None:
        if self.redis is None:
            return
        try:
            await self.redis.setex(f"kv:{cache_key}", ttl, self._serialize_kv(keys, values))
        except Exception as exc:  # noqa: BLE001
            logger.warning(f"Redis write error: {exc}")

    def _serialize_kv(self, keys: torch.Tensor, values: torch.Tensor) -> bytes:
        sink = io.BytesIO()
        pickle.dump({"keys": keys.cpu().numpy(), "values": values.cpu().numpy()}, sink)
        return sink.getvalue()

    def _deserialize_kv(self, data: bytes) -> Tuple[torch.Tensor, torch.Tensor]:
        kv = pickle.load(io.BytesIO(data))
        return torch.from_numpy(kv["keys"]), torch.from_numpy(kv["values"])

    def get_stats(self) -> Dict[str, Any]:
        total = sum(self._stats.values())
        denom = max(1, total)
        return {
            **self._stats,
            "total_requests": total,
            "l1_hit_rate": self._stats["l1_hits"] / denom,
            "l2_hit_rate": self._stats["l2_hits"] / denom,
            "l3_hit_rate": self._stats["l3_hits"] / denom,
            "gpu_memory_gb": self.gpu_cache.get_total_memory_gb(),
            "cpu_cache_items": len(self.cpu_cache),
        }


# ---------------------------------------------------------------------------------------------------------
# Prefix reuse per block.  The reference keeps ONE block per (prefix hash, layer) in
# ``_prefix_index`` and hashes ``bytes(token_ids)``, which fails for ids >= 256 (reference :384-387, :462).
# ---------------------------------------------------------------------------------------------------------
def chained_block_hashes(token_ids: Sequence[int], block_size: int) -> List[str]:
    """One hash per FULL block of tokens, each chained on the previous one, so equal hashes mean equal prefixes
    up to and including that block.  Tokens are encoded as little-endian uint32 (any vocabulary size)."""
    out: List[str] = []
    prev = b""
    for i in range(len(token_ids) // block_size):
        chunk = token_ids[i * block_size:(i + 1) * block_size]
        h = hashlib.sha256(prev + b"".join(int(t).to_bytes(4, "little") for t in chunk)).digest()
        out.append(h.hex()[:16])
        prev = h
    return out


class RadixPrefixIndex:
    """Chained block hash → the blocks (one per layer) that hold that block of tokens.

    ``match(token_ids)`` returns how many leading full blocks are already cached and the per-layer block lists to
    pass as ``KVCachePool.allocate_sequence(shared_prefix=...)``; ``insert`` registers a sequence's full blocks.
    Entries are held by object identity, so a block that was freed and whose id has since been handed out again is
    never mistaken for the cached one; such entries are dropped lazily on lookup.

    With ``hold=True`` the index is a prefix CACHE that outlives the sequences: it takes its own reference on every
    block it registers (``CacheBlock.add_ref``, the sharing the reference declared in worker/distributed/kv_cache.py:
    63-72 and never used).  A block referenced only by the index has ``ref_count == 1`` and is therefore exactly what
    the allocator's LRU eviction may reclaim when the pool runs dry (reference :225-238); blocks that live sequences
    still use have ``ref_count > 1`` and stay.  ``match`` touches the blocks it returns (``get_block``), so recently
    used prefixes are evicted last.
    """

    def __init__(self, pool: "KVCachePool", hold: bool = False):
        self.pool = pool
        self.hold = hold
        self._by_hash: Dict[str, List[CacheBlock]] = {}  # hash -> the block of every layer
        self.stats = {"lookups": 0, "hit_blocks": 0, "inserted_blocks": 0}

    def _alive(self, blks: List[CacheBlock]) -> bool:
        return all(self.pool.layer(l)._blocks.get(b.block_id) is b for l, b in enumerate(blks))

    def insert(self, token_ids: Sequence[int], blocks: List[List[CacheBlock]]) -> int:
        hashes = chained_block_hashes(token_ids, self.pool.block_size)
        for i, h in enumerate(hashes):
            if i >= len(blocks[0]):
                break
            known = self._by_hash.get(h)
            if known is not None and self._alive(known):
                continue
            entry = [layer_blocks[i] for layer_blocks in blocks]
            self._by_hash[h] = entry
            self.stats["inserted_blocks"] += 1
            if self.hold:
                for b in entry:
                    b.add_ref()
        return len(hashes)

    def match(self, token_ids: Sequence[int]) -> Tuple[int, Optional[List[List[CacheBlock]]]]:
        hashes = chained_block_hashes(token_ids, self.pool.block_size)
        per_layer: List[List[CacheBlock]] = [[] for _ in range(self.pool.num_layers)]
        n = 0
        self.stats["lookups"] += 1
        for h in hashes:
            blks = self._by_hash.get(h)
            if blks is None:
                break
            if not self._alive(blks):
                # freed (or evicted) since it was indexed.  Hashes are chained, so whatever was registered BEHIND this
                # block can never be reached again: give those blocks back now instead of waiting for the LRU.
                for dead in hashes[n:]:
                    self._forget(dead)
                break
            for l, b in enumerate(blks):
                per_layer[l].append(b)
            n += 1
        # LRU touch (and hit statistics), last block first: the head of a chain ends up most recently used, so under
        # pressure eviction shortens cached prefixes from the tail and what remains is still a usable prefix
        for i in range(n - 1, -1, -1):
            for l in range(self.pool.num_layers):
                self.pool.layer(l).get_block(per_layer[l][i].block_id)
        self.stats["hit_blocks"] += n
        return n, (per_layer if n else None)

    def _forget(self, h: str) -> None:
        blks = self._by_hash.pop(h, None)
        if blks is not None and self.hold:
            for l, b in enumerate(blks):
                if self.pool.layer(l)._blocks.get(b.block_id) is b:
                    self.pool.layer(l).free_blocks([b.block_id])

    def drop_all(self) -> None:
        """Forget every entry, giving back the references a holding index took."""
        for h in list(self._by_hash):
            self._forget(h)
