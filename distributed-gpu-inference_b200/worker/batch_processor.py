"""Continuous batcher with the reference's surface (reference: worker/batch_processor.py).

Engine-agnostic request coalescing: a priority heap of pending requests, optional grouping by the hash of the
system messages, a batch fired either when ``max_batch_size`` requests wait or ``max_wait_ms`` after the first one,
results delivered through per-request futures.  The engine protocol is the reference's (:302-336): the first of
``batch_inference_async(params_list)``, ``batch_inference(params_list)`` (run in the default executor),
``inference_async(params)``, ``inference(params)``; an ``Exception`` instance in the returned list fails only that
request (:237-242).  ``dgi_b200.engine.PagedDecodeEngine`` and ``dgi_b200.worker.engines.PagedLLMEngine`` provide
``batch_inference``, so a batch really runs as one batch on the GPU (in the reference no engine implements it and
every batch degrades to the serial fallback).
"""
from __future__ import annotations

import asyncio
import hashlib
import heapq
import itertools
import logging
import time
from collections import defaultdict
from dataclasses import dataclass, field
from enum import Enum
from typing import Any, Dict, List, Optional

logger = logging.getLogger(__name__)


class RequestPriority(Enum):
    HIGH = 0
    NORMAL = 1
    LOW = 2


@dataclass(order=True)
class PendingRequest:
    priority: int
    timestamp: float
    job_id: str = field(compare=False)
    params: Dict[str, Any] = field(compare=False)
    future: asyncio.Future = field(compare=False)
    prefix_hash: str = field(compare=False, default="")

    @classmethod
    def create(cls, job_id: str, params: Dict[str, Any], priority: RequestPriority = RequestPriority.NORMAL,
               prefix_hash: str = "") -> "PendingRequest":
        return cls(priority.value, time.time(), job_id, params, asyncio.get_event_loop().create_future(), prefix_hash)


class ContinuousBatcher:
    def __init__(self, engine, max_batch_size: int = 32, max_wait_ms: float = 50, enable_prefix_grouping: bool = True,
                 max_queue_size: int = 1000):
        self.engine = engine
        self.max_batch_size = max_batch_size
        self.max_wait_ms = max_wait_ms
        self.enable_prefix_grouping = enable_prefix_grouping
        self.max_queue_size = max_queue_size
        self._pending: List[PendingRequest] = []
        self._pending_by_prefix: Dict[str, List[PendingRequest]] = defaultdict(list)
        self._batch_task: Optional[asyncio.Task] = None
        self._lock = asyncio.Lock()
        self._stats = {"total_requests": 0, "total_batches": 0, "avg_batch_size": 0.0, "avg_wait_time_ms": 0.0}
        self._running = False

    async def start(self) -> None:
        self._running = True

    async def stop(self) -> None:
        self._running = False
        async with self._lock:
            for req in self._pending:
                if not req.future.done():
                    req.future.cancel()
            self._pending.clear()
            self._pending_by_prefix.clear()
        if self._batch_task is not None:
            self._batch_task.cancel()
            try:
                await self._batch_task
            except asyncio.CancelledError:
                pass

    async def submit(self, job_id: str, params: Dict[str, Any], priority: RequestPriority = RequestPriority.NORMAL,
                     timeout: float = 120.0) -> Dict[str, Any]:
        if not self._running:
            raise RuntimeError("Batcher is not running")
        if len(self._pending) >= self.max_queue_size:
            raise RuntimeError(f"Queue full (max={self.max_queue_size})")
        prefix = self._compute_prefix_hash(params) if self.enable_prefix_grouping else ""
        req = PendingRequest.create(job_id, params, priority, prefix)
        async with self._lock:
            heapq.heappush(self._pending, req)
            if prefix:
                self._pending_by_prefix[prefix].append(req)
            self._stats["total_requests"] += 1
            if len(self._pending) >= self.max_batch_size:
                asyncio.create_task(self._process_batch())
            elif self._batch_task is None or self._batch_task.done():
                self._batch_task = asyncio.create_task(self._wait_and_process())
        try:
            return await asyncio.wait_for(req.future, timeout=timeout)
        except asyncio.TimeoutError:
            async with self._lock:
                if req in self._pending:
                    self._pending.remove(req)
                    heapq.heapify(self._pending)
                if prefix and req in self._pending_by_prefix.get(prefix, []):
                    self._pending_by_prefix[prefix].remove(req)
            raise

    async def _wait_and_process(self) -> None:
        await asyncio.sleep(self.max_wait_ms / 1000)
        await self._process_batch()

    async def _process_batch(self) -> None:
        async with self._lock:
            if not self._pending:
                return
            t0 = time.time()
            batch = self._select_batch_with_prefix_grouping() if self.enable_prefix_grouping else self._select_batch_simple()
            if not batch:
                return
            taken = set(map(id, batch))
            self._pending = [r for r in self._pending if id(r) not in taken]
            heapq.heapify(self._pending)
            for r in batch:
                if r.prefix_hash:
                    grp = self._pending_by_prefix.get(r.prefix_hash, [])
                    if r in grp:
                        grp.remove(r)
                    if not grp:
                        self._pending_by_prefix.pop(r.prefix_hash, None)
        try:
            results = await self._execute_batch(batch)
            for req, res in zip(batch, results):
                if req.future.done():
                    continue
                if isinstance(res, Exception):
                    req.future.set_exception(res)
                else:
                    req.future.set_result(res)
        except Exception as exc:  # noqa: BLE001 - a failing batch fails all its requests
            logger.error(f"Batch processing error: {exc}")
            for req in batch:
                if not req.future.done():
                    req.future.set_exception(exc)
        n = self._stats["total_batches"] + 1
        ms = (time.time() - t0) * 1000
        self._stats["avg_batch_size"] += (len(batch) - self._stats["avg_batch_size"]) / n
        self._stats["avg_wait_time_ms"] += (ms - self._stats["avg_wait_time_ms"]) / n
        self._stats["total_batches"] = n

    def _select_batch_simple(self) -> List[PendingRequest]:
        return heapq.nsmallest(self.max_batch_size, self._pending)

    def _select_batch_with_prefix_grouping(self) -> List[PendingRequest]:
        batch: List[PendingRequest] = []
        for _, group in sorted(self._pending_by_prefix.items(), key=lambda kv: len(kv[1]), reverse=True):
            room = self.max_batch_size - len(batch)
            if room <= 0:
                break
            batch.extend(group[:room])
        room = self.max_batch_size - len(batch)
        if room > 0:
            batch.extend(itertools.islice((r for r in self._pending if not r.prefix_hash), room))
        return batch

    async def _execute_batch(self, batch: List[PendingRequest]) -> List[Any]:
        params_list = [r.params for r in batch]
        if hasattr(self.engine, "batch_inference_async"):
            return await self.engine.batch_inference_async(params_list)
        loop = asyncio.get_event_loop()
        if hasattr(self.engine, "batch_inference"):
            return await loop.run_in_executor(None, self.engine.batch_inference, params_list)
        out: List[Any] = []
        for p in params_list:
            try:
                if hasattr(self.engine, "inference_async"):
                    out.append(await self.engine.inference_async(p))
                else:
                    out.append(await loop.run_in_executor(None, self.engine.inference, p))
            except Exception as exc:  # noqa: BLE001
                out.append(exc)
        return out

    def _compute_prefix_hash(self, params: Dict[str, Any]) -> str:
        system = [m.get("content", "") for m in params.get("messages", []) if m.get("role") == "system"]
        return hashlib.sha256("".join(system).encode()).hexdigest()[:16] if system else ""

    def get_stats(self) -> Dict[str, Any]:
        return {**self._stats, "queue_size": len(self._pending), "prefix_groups": len(self._pending_by_prefix)}


class AdaptiveBatcher(ContinuousBatcher):
    """Batcher that steers its batch size toward a latency target (reference: worker/batch_processor.py:368-433).

    Starts at half of ``max_batch_size``; once ten batch latencies are known, the mean of the last ten moves the
    working size by ×0.8 (above 1.2 × target) or ×1.2 (below 0.8 × target) within ``[min_batch_size,
    max_batch_size]``.  At most the last 100 latencies are kept.
    """

    def __init__(self, engine, min_batch_size: int = 1, max_batch_size: int = 64, target_latency_ms: float = 100,
                 **kwargs):
        super().__init__(engine, max_batch_size=max_batch_size, **kwargs)
        self.min_batch_size = min_batch_size
        self.target_latency_ms = target_latency_ms
        self._current_batch_size = max_batch_size // 2
        self._latency_history: List[float] = []
        self._max_history = 100

    async def _process_batch(self) -> None:
        t0 = time.time()
        ceiling, self.max_batch_size = self.max_batch_size, self._current_batch_size
        try:
            await super()._process_batch()
        finally:
            self.max_batch_size = ceiling
        self._latency_history.append((time.time() - t0) * 1000)
        del self._latency_history[:-self._max_history]
        self._adapt_batch_size()

    def _adapt_batch_size(self) -> None:
        recent = self._latency_history[-10:]
        if len(self._latency_history) < 10:
            return
        mean = sum(recent) / 10
        if mean > 1.2 * self.target_latency_ms:
            self._current_batch_size = max(self.min_batch_size, int(self._current_batch_size * 0.8))
        elif mean < 0.8 * self.target_latency_ms:
            self._current_batch_size = min(self.max_batch_size, int(self._current_batch_size * 1.2))
        logger.debug(f"Adaptive batch size: {self._current_batch_size} (avg latency: {mean:.1f}ms)")
