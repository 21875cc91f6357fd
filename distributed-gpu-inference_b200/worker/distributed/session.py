"""Client side of the layer pipeline with the reference's session surface, transport made pluggable.

Reference: worker/distributed/session.py.  Kept for callers:

* ``WorkerSession(worker_info)`` with ``connect / forward(hidden_states, position, kv_cache_keys) / close``,
  attributes ``session_id, state, position, history, next_session`` (reference :58-195);
* ``DistributedInferenceSession(config, route)`` with ``setup / step(inputs, kv_cache_keys) / close``, the
  ``position`` property that fans out to the hop sessions, ``get_stats()`` keys ``total_tokens, total_steps,
  total_latency_ms, retries, avg_latency_ms, tokens_per_second``, the ``ValueError`` when ``max_length`` would be
  exceeded, and the per-hop retry loop with linear back-off (reference :198-395);
* ``SessionManager`` (reference :398-455) — without the self-deadlock of the original, whose ``create_session``
  holds the lock while ``_cleanup_expired_sessions`` re-acquires it through ``close_session``.

What changed: the reference's hop is an aiohttp POST of base64 JSON (:125-160).  Inside one box the hop is either
an in-process call into a stage (``InProcessTransport``) or NCCL point-to-point (``dgi_b200.pipeline``), so
``WorkerSession`` delegates to a ``transport`` object instead of owning an HTTP client.  ``setup`` still builds
its hops from the module-global name ``WorkerSession`` — the seam the reference's tests patch.
"""
from __future__ import annotations

import asyncio
import logging
import time
import uuid
from dataclasses import dataclass, field
from enum import Enum
from typing import Any, Callable, Dict, List, Optional, Tuple

from ...common.data_structures import SessionConfig, WorkerInfo
from ...common.serialization import deserialize_tensor, serialize_tensor

try:  # only the HTTP hop needs it; the NVLink paths never touch the network
    import aiohttp
except ImportError:  # pragma: no cover
    aiohttp = None

logger = logging.getLogger(__name__)


class SessionState(Enum):
    INITIALIZING = "initializing"
    READY = "ready"
    ACTIVE = "active"
    ERROR = "error"
    CLOSED = "closed"


class InProcessTransport:
    """Hop that calls a stage function living in this process: ``fn(hidden_states, position) -> hidden_states``."""

    def __init__(self, fn: Callable[[Any, int], Any]):
        self.fn = fn

    async def health(self) -> bool:
        return True

    async def forward(self, session_id: str, hidden_states: Any, position: int, kv_cache_keys: List[str]):
        return self.fn(hidden_states, position), kv_cache_keys

    async def close(self, session_id: str) -> None:
        return None


class HttpTransport:
    """The reference's own hop (worker/distributed/session.py:78-184): JSON over HTTP to ``api_endpoint`` —
    ``GET /health``, ``POST /inference/forward`` with the serialized hidden states, ``POST /inference/close``.
    Kept wire-compatible so a route may mix stock workers and workers of this package; it is what a
    ``WorkerSession`` falls back to when no in-process / NCCL transport is registered for its worker id."""

    def __init__(self, session: "WorkerSession", timeout: float = 30.0):
        if aiohttp is None:
            raise ConnectionError("aiohttp is not installed")
        self._ws = session
        self._http = aiohttp.ClientSession(timeout=aiohttp.ClientTimeout(total=timeout))

    @property
    def _base(self) -> str:
        return self._ws.worker_info.api_endpoint

    async def health(self) -> bool:
        async with self._http.get(f"{self._base}/health") as response:
            if response.status != 200:
                raise ConnectionError(f"Worker health check failed: {response.status}")
        return True

    async def forward(self, session_id: str, hidden_states: Any, position: int, kv_cache_keys: List[str]):
        info = self._ws.worker_info
        payload = {
            "session_id": session_id,
            "input": serialize_tensor(hidden_states),
            "position": position,
            "kv_cache_keys": kv_cache_keys,
            "blocks": info.blocks.to_dict() if info.blocks else None,
        }
        nxt = self._ws.next_session
        if nxt is not None:
            payload["next_worker"] = {"address": nxt.worker_info.api_endpoint, "session_id": nxt.session_id}
        async with self._http.post(f"{self._base}/inference/forward", json=payload) as response:
            if response.status != 200:
                raise RuntimeError(f"Forward failed: {await response.text()}")
            result = await response.json()
        return deserialize_tensor(result["output"]), result.get("kv_cache_keys", [])

    async def close(self, session_id: str) -> None:
        try:
            async with self._http.post(f"{self._base}/inference/close", json={"session_id": session_id}):
                pass
        finally:
            await self._http.close()


def _run_coroutine_in_new_thread(coro):
    """Run ``coro`` to completion on a fresh event loop in a helper thread (for sync code called from a running loop)."""
    import threading
    from concurrent.futures import Future

    done: Future = Future()

    def runner() -> None:
        try:
            done.set_result(asyncio.run(coro))
        except BaseException as exc:  # noqa: BLE001
            done.set_exception(exc)

    threading.Thread(target=runner, daemon=True).start()
    return done.result()


# worker_id -> transport; lets `WorkerSession(worker_info=...)` (the only call the reference's setup makes) find
# its hop without a network address
TRANSPORTS: Dict[str, Any] = {}


@dataclass
class WorkerSession:
    worker_info: WorkerInfo
    session_id: str = field(default_factory=lambda: str(uuid.uuid4()))
    state: SessionState = SessionState.INITIALIZING
    position: int = 0
    history: Optional[Any] = None
    next_session: Optional["WorkerSession"] = None
    transport: Optional[Any] = None

    async def connect(self, timeout: float = 30.0) -> None:
        if self.transport is None:
            self.transport = TRANSPORTS.get(self.worker_info.worker_id)
        try:
            if self.transport is None and self.worker_info.api_endpoint:
                self.transport = HttpTransport(self, timeout)
            if self.transport is None:
                raise ConnectionError(f"no transport registered for worker {self.worker_info.worker_id}")
            if not await asyncio.wait_for(self.transport.health(), timeout):
                raise ConnectionError("Worker health check failed")
        except Exception as exc:
            self.state = SessionState.ERROR
            raise ConnectionError(f"Failed to connect to worker: {exc}")
        self.state = SessionState.READY

    async def forward(self, hidden_states: Any, position: int, kv_cache_keys: List[str] = None) -> Tuple[Any, List[str]]:
        if self.state not in (SessionState.READY, SessionState.ACTIVE):
            raise RuntimeError(f"Session not ready: {self.state}")
        self.state = SessionState.ACTIVE
        try:
            out, keys = await self.transport.forward(self.session_id, hidden_states, position, kv_cache_keys or [])
        except Exception as exc:
            self.state = SessionState.ERROR
            raise RuntimeError(f"Forward error: {exc}")
        self.position = position + (hidden_states.shape[1] if hasattr(hidden_states, "shape") else 1)
        return out, keys

    async def close(self) -> None:
        if self.transport is not None:
            try:
                await self.transport.close(self.session_id)
            except Exception as exc:  # noqa: BLE001
                logger.warning(f"Error closing worker session: {exc}")
            if isinstance(self.transport, HttpTransport):
                self.transport = None  # its client session is gone; a later connect() opens a new one
        self.state = SessionState.CLOSED

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        """Synchronous ``with`` support (reference :186-195): close on a private loop, from a helper thread when
        the caller is itself running inside an event loop."""
        try:
            asyncio.get_running_loop()
        except RuntimeError:
            asyncio.run(self.close())
        else:
            _run_coroutine_in_new_thread(self.close())


class DistributedInferenceSession:
    def __init__(self, config: SessionConfig, route: List[WorkerInfo]):
        self.config = config
        self.route = route
        self.session_id = str(uuid.uuid4())
        self.state = SessionState.INITIALIZING
        self._worker_sessions: List[WorkerSession] = []
        self._position = 0
        self._max_length = config.max_length
        self._stats = {"total_tokens": 0, "total_steps": 0, "total_latency_ms": 0, "retries": 0}

    @property
    def position(self) -> int:
        return self._position

    @position.setter
    def position(self, value: int) -> None:
        self._position = value
        for s in self._worker_sessions:
            s.position = value

    async def setup(self) -> None:
        try:
            for info in self.route:
                hop = WorkerSession(worker_info=info)  # module-global name on purpose (patch seam)
                await hop.connect(timeout=self.config.connect_timeout)
                self._worker_sessions.append(hop)
            for a, b in zip(self._worker_sessions, self._worker_sessions[1:]):
                a.next_session = b
            self.state = SessionState.READY
        except Exception:
            self.state = SessionState.ERROR
            for hop in self._worker_sessions:
                await hop.close()
            self._worker_sessions.clear()
            raise

    async def step(self, inputs: Any, kv_cache_keys: List[str] = None) -> Any:
        if self.state not in (SessionState.READY, SessionState.ACTIVE):
            raise RuntimeError(f"Session not ready: {self.state}")
        self.state = SessionState.ACTIVE
        t0 = time.time()
        n_new = inputs.shape[1] if hasattr(inputs, "shape") else 1
        if self._position + n_new > self._max_length:
            raise ValueError(f"Maximum length exceeded: {self._position} + {n_new} > {self._max_length}")

        hidden, keys = inputs, (kv_cache_keys or [])
        for idx, hop in enumerate(self._worker_sessions):
            for attempt in range(self.config.max_retries):
                try:
                    hidden, keys = await hop.forward(hidden, position=self._position, kv_cache_keys=keys)
                    break
                except Exception as exc:
                    logger.warning(f"Worker {hop.worker_info.worker_id} failed "
                                   f"(attempt {attempt + 1}/{self.config.max_retries}): {exc}")
                    self._stats["retries"] += 1
                    if attempt + 1 == self.config.max_retries:
                        await self._handle_failure(idx, exc)
                        hidden, keys = await hop.forward(hidden, position=self._position, kv_cache_keys=keys)
                    else:
                        if hop.state == SessionState.ERROR:
                            hop.state = SessionState.READY  # the next attempt re-enters forward()
                        await asyncio.sleep(0.5 * (attempt + 1))

        self._position += n_new
        self._stats["total_tokens"] += n_new
        self._stats["total_steps"] += 1
        self._stats["total_latency_ms"] += (time.time() - t0) * 1000
        return hidden

    async def _handle_failure(self, failed_idx: int, error: Exception) -> None:
        hop = self._worker_sessions[failed_idx]
        await hop.close()
        raise RuntimeError(f"Worker failure recovery not implemented. Failed worker: {hop.worker_info.worker_id}")

    async def close(self) -> None:
        for hop in self._worker_sessions:
            try:
                await hop.close()
            except Exception as exc:  # noqa: BLE001
                logger.warning(f"Error closing session: {exc}")
        self._worker_sessions.clear()
        self.state = SessionState.CLOSED

    def get_stats(self) -> Dict[str, Any]:
        stats = dict(self._stats)
        if stats["total_steps"] > 0:
            stats["avg_latency_ms"] = stats["total_latency_ms"] / stats["total_steps"]
            ms = stats["total_latency_ms"]
            stats["tokens_per_second"] = stats["total_tokens"] / (ms / 1000) if ms > 0 else 0
        return stats

    async def __aenter__(self):
        await self.setup()
        return self

    async def __aexit__(self, *exc):
        await self.close()


class SessionManager:
    def __init__(self, max_sessions: int = 100):
        self.max_sessions = max_sessions
        self._sessions: Dict[str, DistributedInferenceSession] = {}
        self._lock = asyncio.Lock()

    async def create_session(self, config: SessionConfig, route: List[WorkerInfo]) -> DistributedInferenceSession:
        async with self._lock:
            if len(self._sessions) >= self.max_sessions:
                await self._cleanup_expired_sessions()
            if len(self._sessions) >= self.max_sessions:
                raise RuntimeError(f"Maximum sessions reached: {self.max_sessions}")
            sess = DistributedInferenceSession(config, route)
            await sess.setup()
            self._sessions[sess.session_id] = sess
            return sess

    async def get_session(self, session_id: str) -> Optional[DistributedInferenceSession]:
        return self._sessions.get(session_id)

    async def close_session(self, session_id: str) -> None:
        async with self._lock:
            await self._close_unlocked(session_id)

    async def _close_unlocked(self, session_id: str) -> None:
        sess = self._sessions.pop(session_id, None)
        if sess is not None:
            await sess.close()

    async def _cleanup_expired_sessions(self) -> None:
        # called with the lock held: must not go through close_session()
        dead = [sid for sid, s in self._sessions.items() if s.state in (SessionState.CLOSED, SessionState.ERROR)]
        for sid in dead:
            await self._close_unlocked(sid)

    async def close_all(self) -> None:
        async with self._lock:
            for sess in list(self._sessions.values()):
                await sess.close()
            self._sessions.clear()
