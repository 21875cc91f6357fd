"""KV-cache migration glue: the reference's ``KVCacheMigrator`` surface with a real data mover behind it.

Reference: server/app/services/pd_scheduler.py:404-479.  There ``_do_migrate`` is ``await asyncio.sleep(0.05)``
followed by the index update; the surface is kept — ``KVCacheMigrator(scheduler)``, ``await migrate(kv_cache_key,
source_worker, target_worker) -> bool``, de-duplication of a migration that is already in flight, the update of
``scheduler._kv_cache_locations[key]`` and ``scheduler._stats["migrations"]``, ``False`` on failure — and the sleep
is replaced by the hand-off of ``dgi_b200.kv_transfer`` (pack → NVLink → unpack, or the fused peer push).

The scheduler is duck-typed (anything with ``_kv_cache_locations`` and ``_stats``), so the reference's
``PrefillDecodeScheduler`` can be passed as is.  What a migrator needs to know about a cached sequence — which worker
holds it and under which block table — lives in ``SequenceDirectory``.
"""
from __future__ import annotations

import asyncio
import inspect
import logging
import time
from dataclasses import dataclass, field
from typing import Any, Awaitable, Callable, Dict, List, Optional, Union

logger = logging.getLogger(__name__)

# mover(kv_cache_key, source_worker, target_worker) -> bytes moved (or None); may be sync or async
Mover = Callable[[str, str, str], Union[Optional[int], Awaitable[Optional[int]]]]


@dataclass
class CachedSequence:
    """Where a sequence's KV lives: per worker, the block table (pool slots, allocation order) and the CacheBlocks."""

    n_tokens: int
    tables: Dict[str, List[int]] = field(default_factory=dict)    # worker id -> block table
    blocks: Dict[str, Any] = field(default_factory=dict)          # worker id -> [[CacheBlock]] per layer


class SequenceDirectory:
    """kv_cache_key -> CachedSequence; the data-plane companion of the scheduler's ``_kv_cache_locations``.

    With a ``scheduler`` (the reference's ``PrefillDecodeScheduler`` or anything with ``_kv_cache_locations``) the
    directory feeds that index: the first worker a sequence is registered on becomes its location (what
    ``transition_to_decode`` would record, server/app/services/pd_scheduler.py:231), and dropping the copy the index
    points at moves the index to a remaining copy or removes the entry — so ``_assign_decode`` (:274-323) always
    sees where the blocks really are."""

    def __init__(self, scheduler=None):
        self._seqs: Dict[str, CachedSequence] = {}
        self.scheduler = scheduler

    def put(self, key: str, worker: str, n_tokens: int, blocks) -> CachedSequence:
        from .worker.distributed.kv_cache import KVCachePool

        seq = self._seqs.setdefault(key, CachedSequence(n_tokens=n_tokens))
        seq.n_tokens = n_tokens
        seq.tables[worker] = KVCachePool.block_table(blocks[0])
        seq.blocks[worker] = blocks
        if self.scheduler is not None:
            self.scheduler._kv_cache_locations.setdefault(key, worker)
        return seq

    def get(self, key: str) -> Optional[CachedSequence]:
        return self._seqs.get(key)

    def drop(self, key: str, worker: str) -> None:
        seq = self._seqs.get(key)
        if seq is not None:
            seq.tables.pop(worker, None)
            seq.blocks.pop(worker, None)
            if self.scheduler is not None and self.scheduler._kv_cache_locations.get(key) == worker:
                if seq.tables:
                    self.scheduler._kv_cache_locations[key] = next(iter(seq.tables))
                else:
                    del self.scheduler._kv_cache_locations[key]
            if not seq.tables:
                del self._seqs[key]


class KVCacheMigrator:
    def __init__(self, scheduler, mover: Optional[Mover] = None):
        self.scheduler = scheduler
        self.mover = mover
        self._pending_migrations: Dict[str, asyncio.Task] = {}
        self.stats = {"bytes_moved": 0, "seconds": 0.0, "migrations": 0, "failures": 0}

    async def migrate(self, kv_cache_key: str, source_worker: str, target_worker: str) -> bool:
        migration_id = f"{kv_cache_key}:{source_worker}:{target_worker}"
        running = self._pending_migrations.get(migration_id)
        if running is not None:  # the same move is already under way: wait for it instead of starting another
            await running
            return True
        task = asyncio.create_task(self._do_migrate(kv_cache_key, source_worker, target_worker))
        self._pending_migrations[migration_id] = task
        try:
            return await task
        finally:
            self._pending_migrations.pop(migration_id, None)

    async def _do_migrate(self, kv_cache_key: str, source_worker: str, target_worker: str) -> bool:
        try:
            t0 = time.perf_counter()
            moved = None
            if self.mover is not None:
                moved = self.mover(kv_cache_key, source_worker, target_worker)
                if inspect.isawaitable(moved):
                    moved = await moved
            self.scheduler._kv_cache_locations[kv_cache_key] = target_worker
            self.scheduler._stats["migrations"] = self.scheduler._stats.get("migrations", 0) + 1
            self.stats["migrations"] += 1
            self.stats["bytes_moved"] += int(moved or 0)
            self.stats["seconds"] += time.perf_counter() - t0
            return True
        except Exception as exc:  # noqa: BLE001 - the reference reports failure as False
            logger.error(f"KV-Cache migration failed: {exc}")
            self.stats["failures"] += 1
            return False


def same_device_mover(directory: SequenceDirectory, pools: Dict[str, Any], chunks: int = 1) -> Mover:
    """Mover for workers whose pools live on one GPU (tests, hybrid workers): pack → unpack between two
    ``KVCachePool``s, destination blocks taken from the target's own allocator."""
    import torch

    from .kv_transfer import KVTransferEndpoint

    endpoints = {w: KVTransferEndpoint(p, chunks=chunks) for w, p in pools.items()}

    def move(key: str, src: str, dst: str) -> int:
        seq = directory.get(key)
        if seq is None or src not in seq.tables:
            raise KeyError(f"{key} is not cached on {src}")
        dst_blocks = pools[dst].allocate_sequence(seq.n_tokens)
        directory.put(key, dst, seq.n_tokens, dst_blocks)
        dev = endpoints[src].device
        st = torch.tensor(seq.tables[src], dtype=torch.int32, device=dev)
        dt = torch.tensor(seq.tables[dst], dtype=torch.int32, device=dev)
        endpoints[src].copy_to(endpoints[dst], st, dt)
        return endpoints[src].payload_bytes(len(seq.tables[src]))

    return move
