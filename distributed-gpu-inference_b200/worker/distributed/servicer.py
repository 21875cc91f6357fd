"""Worker-side servicer with the reference's method surface, over paged per-session KV.

Reference: ``InferenceServicer`` (worker/distributed/grpc_server.py:36-422) — the object a worker's gRPC / HTTP front
end calls: ``CreateSession / Forward / StreamInference / TransferKVCache / CloseSession / HealthCheck``, taking
protobuf-like request objects and returning plain dicts, with the shard and the cache manager injected through the
constructor (:43-53).  Only the servicer is mirrored here; the transports around it (gRPC server, aiohttp routes,
:399-560) are out of scope.

What is different, and why:

* The reference's ``_forward`` (:351-374) calls ``model_shard.forward(h, position_ids=arange(pos, pos+T)[None],
  use_cache=True)`` with NO ``past_key_values`` and throws the returned KV away, so decoding through the servicer
  has no cache at all.  Here a session owns blocks of the injected manager's GPU pool (``kv_cache.gpu_cache``, the
  ``KVCachePool`` the reference's manager also builds, :360-363): on the first call the session's sequences get
  their blocks, every call appends the new tokens' K/V through the slot mapping and attends over the block table
  (the CUDA kernels behind ``PagedLlamaDecoderLayer``), and ``CloseSession`` frees the blocks (the reference's
  ``TODO``, :277-280).  A repeated call for a position that was already served (the client's retry loop,
  worker/distributed/session.py:303-329) overwrites the same slots, so retries are idempotent.
* A one-token call replays a CUDA graph of the shard's decode step (``decode_graphs``, on by default): the session's
  first such call builds a ``PagedDecodeEngine`` that borrows the session's blocks and captures the step; later calls
  copy the input into the graph's buffer, set the position the request names and replay.  Measured on B200 (8B shape,
  32 layers, batch 64, context 2k, profiles/r02_servicer_bench.json): 6.44 ms per call (5.58 ms with
  ``fuse_projections``) against 7.97 ms with eager launches and 6.43 ms for the bare graph.  Multi-token (prompt) calls launch eagerly with one RoPE table per call.
* ``HealthCheck`` and ``_get_cache_capacity`` use the pool's real block size and free-block count instead of the
  hard-coded 16 and a stats key that does not exist (:311,394); ``total_tokens`` is counted (:61), so
  ``throughput_tokens_per_sec`` is live.
* ``TransferKVCache`` keeps the reference's behaviour for one-block messages (``shape = [heads, block_size,
  head_dim]`` stored under ``"{prefix_key}:{layer_idx}"`` through ``_store_kv``, :190-235); a request whose
  ``prefix_key`` names a session carries whole sequences (``shape = [n_blocks, heads, block_size, head_dim]`` per
  layer, as ``kv_transfer.WireHandoff.export`` emits) and lands in that session's blocks on the GPU.

Requests may be attribute objects (as in the reference) or dicts; tensors may arrive as ``TensorSerializer`` bytes
(``input`` + ``shape`` + ``dtype``, the wire form) or, inside one process, as ``torch.Tensor`` objects.
"""
from __future__ import annotations

import logging
import time
import uuid
from typing import Any, AsyncIterator, Dict, List, Optional

import torch

from ...common.serialization import TensorSerializer
from ...paged_layer import PagedBatchState, PagedLayerKV, rope_tables
from .kv_cache import DistributedKVCacheManager, KVCachePool

logger = logging.getLogger(__name__)

DEFAULT_MAX_LENGTH = 2048


def _get(request: Any, name: str, default: Any = None) -> Any:
    if isinstance(request, dict):
        return request.get(name, default)
    return getattr(request, name, default)


class _SessionKV:
    """Blocks, block table and length of one session's sequences in the servicer's pool."""

    def __init__(self, pool: KVCachePool, batch: int, max_length: int, device: torch.device):
        self.pool = pool
        self.batch = batch
        self.max_length = max_length
        self.blocks = []
        self.pinned = False
        try:
            for _ in range(batch):
                self.blocks.append(pool.allocate_sequence(max_length))
        except Exception:
            self.free()
            raise
        for per_layer in self.blocks:
            for li, layer_blocks in enumerate(per_layer):
                pool.layer(li).pin(b.block_id for b in layer_blocks)
        self.pinned = True
        rows = []
        for per_layer in self.blocks:
            row = KVCachePool.block_table(per_layer[0])
            if any(KVCachePool.block_table(lb) != row for lb in per_layer):
                self.free()
                raise RuntimeError("layers disagree on block slots; per-layer tables are not supported here")
            rows.append(row)
        self.tables = torch.tensor(rows, dtype=torch.int32, device=device)

    def free(self) -> None:
        for per_layer in self.blocks:
            if self.pinned:
                for li, layer_blocks in enumerate(per_layer):
                    self.pool.layer(li).unpin(b.block_id for b in layer_blocks)
            self.pool.free_sequence(per_layer)
        self.blocks = []
        self.pinned = False


class InferenceServicer:
    """See the module docstring.  Constructor and public methods as in the reference."""

    def __init__(
        self,
        worker_id: str,
        model_shard,
        kv_cache_manager: DistributedKVCacheManager,
        max_sessions: int = 100,
    ):
        self.worker_id = worker_id
        self.model_shard = model_shard
        self.kv_cache = kv_cache_manager
        self.max_sessions = max_sessions
        self._sessions: Dict[str, Dict[str, Any]] = {}
        self._stats = {"total_requests": 0, "total_tokens": 0, "total_latency_ms": 0, "errors": 0}
        # attention recipe of multi-token (prompt) calls: "sdpa", or "eager" = the reference layer's exact O(T^2) recipe
        self.prefill_mode = "sdpa"
        # one-token calls replay a CUDA graph of the shard's decode step, captured per session on its first such call
        # (the ~290 launches of a 32-layer step cost more on the host than on the GPU); False: eager launches
        self.decode_graphs = True
        # True: the decode engine lays q|k|v and gate|up out as one K-major GEMM each and fuses the residual adds into
        # the norms (PagedDecodeEngine(fuse_projections=True): same dot products, same roundings, fewer launches);
        # it rewrites the shard's weight storage in place, so it is the owner's choice
        self.fuse_projections = False

    # ---- helpers ---------------------------------------------------------------------------------------------
    def _device(self) -> str:
        return str(self.model_shard.device) if hasattr(self.model_shard, "device") else "cuda"

    def _pool(self) -> KVCachePool:
        pool = getattr(self.kv_cache, "gpu_cache", None)
        if not isinstance(pool, KVCachePool):
            raise RuntimeError("the injected cache manager has no GPU pool (gpu_cache)")
        if self.model_shard is not None and pool.num_layers < self.model_shard.get_layer_count():
            raise RuntimeError(f"the cache manager holds {pool.num_layers} layers, the shard owns "
                               f"{self.model_shard.get_layer_count()}")
        return pool

    def _tensor_of(self, payload: Any, request: Any) -> torch.Tensor:
        if isinstance(payload, torch.Tensor):
            return payload.to(self._device())
        return TensorSerializer.deserialize(payload, tuple(_get(request, "shape")), _get(request, "dtype"),
                                            device=self._device())

    def _create_session(self, session_id: str) -> Dict[str, Any]:
        session = {"session_id": session_id, "position": 0, "created_at": time.time()}
        self._sessions[session_id] = session
        return session

    def _get_or_create_session(self, session_id: str, request: Any) -> Dict[str, Any]:
        if session_id not in self._sessions:
            return self._create_session(session_id)
        return self._sessions[session_id]

    def _release_kv(self, session: Dict[str, Any]) -> None:
        eng = session.pop("_engine", None)
        if eng is not None:
            eng.release()  # the blocks are the session's: dropped, not freed
        kv = session.pop("_kv", None)
        if kv is not None:
            kv.free()

    def _decode_engine(self, session: Dict[str, Any], kv: _SessionKV, position: int):
        """The session's decode-step graph: a ``PagedDecodeEngine`` over this shard that borrows the session's blocks
        (``adopt_sequences``) and stops at the shard's output — hidden states, or the normed rows on the last shard,
        which is what ``ModelShard.forward`` returns."""
        eng = session.get("_engine")
        if eng is None:
            from ...engine import PagedDecodeEngine

            eng = PagedDecodeEngine([self.model_shard], kv.batch, kv.max_length, block_size=kv.pool.block_size,
                                    kv_pools=[kv.pool], external_head=True, fuse_projections=self.fuse_projections)
            eng.adopt_sequences([[per_layer] for per_layer in kv.blocks], [position] * kv.batch, table_rows=kv.tables)
            session["_engine"] = eng
        return eng

    def _session_kv(self, session: Dict[str, Any], batch: int, need_len: int) -> _SessionKV:
        kv: Optional[_SessionKV] = session.get("_kv")
        if kv is not None and kv.batch != batch:
            raise ValueError(f"session {session['session_id']} was opened for batch {kv.batch}, got batch {batch}")
        if kv is None:
            max_length = int(session.get("max_length") or DEFAULT_MAX_LENGTH)
            kv = _SessionKV(self._pool(), batch, max(max_length, need_len), torch.device(self._device()))
            session["_kv"] = kv
        if need_len > kv.max_length:
            raise ValueError(f"Maximum length exceeded: {need_len} > {kv.max_length}")
        return kv

    def _get_cache_capacity(self) -> int:
        """Tokens that still fit the GPU pool: free blocks (of the fullest layer) x the pool's block size."""
        pool = getattr(self.kv_cache, "gpu_cache", None)
        if not isinstance(pool, KVCachePool):
            return 0
        free = min(pool.layer(li).get_stats()["free_blocks"] for li in range(pool.num_layers))
        return free * pool.block_size

    # ---- the hot path ----------------------------------------------------------------------------------------
    @torch.no_grad()
    def forward_sync(self, session: Dict[str, Any], hidden_states: torch.Tensor, position: int,
                     kv_cache_keys: list) -> tuple:
        """``_forward`` without the coroutine wrapper (nothing in it awaits)."""
        if self.model_shard is None:
            return hidden_states, kv_cache_keys  # reference :359-361: echo
        B, T = hidden_states.shape[0], hidden_states.shape[1]
        dev = hidden_states.device
        held = int(session.get("position", 0))
        if position < 0 or position > held:
            # a retry (position <= what the session holds) rewrites the same slots; a gap would attend over slots
            # nobody wrote
            raise ValueError(f"position {position} is beyond the {held} tokens session {session['session_id']} holds")
        kv = self._session_kv(session, B, position + T)
        pool = kv.pool
        bs = pool.block_size
        keys = list(kv_cache_keys) if kv_cache_keys else [f"session:{session['session_id']}"]
        if T == 1 and self.decode_graphs and hidden_states.is_cuda and hasattr(self.model_shard, "layers"):
            eng = self._decode_engine(session, kv, position)
            # the step runs at the position the request names (a retry repeats one, a hand-off may have moved it)
            eng.context_lens[:B].fill_(position)
            (eng.ids_buf if eng.is_first else eng.h_in)[:B].copy_(hidden_states)
            eng.stage_step()
            output = (eng.hn_out[:B].unsqueeze(1) if eng.is_last else eng.h_out[:B]).clone()
            session["position"] = max(held, position + 1)
            self._stats["total_tokens"] += B
            return output, keys
        t = torch.arange(position, position + T, device=dev)
        slots = kv.tables.to(torch.int64)[:, t // bs] * bs + (t % bs)[None, :]
        state = PagedBatchState(
            block_tables=kv.tables,
            context_lens=torch.full((B,), position + T, dtype=torch.int32, device=dev),
            slot_mapping=slots.reshape(-1).contiguous(),
            max_context_len=position + T,
            prefill_mode=session.get("prefill_mode", self.prefill_mode),
            past_len=position if T > 1 else 0,
        )
        handles = [PagedLayerKV(pool.layer(li).key_pool, pool.layer(li).value_pool, state)
                   for li in range(self.model_shard.get_layer_count())]
        position_ids = t[None, :].expand(B, T)  # reference :365-369: arange(position, position + T)
        layers = getattr(self.model_shard, "layers", None)
        layer0 = layers[0] if layers is not None and len(layers) else None
        if getattr(layer0, "inv_freq", None) is not None:  # RoPE tables once per call, not once per layer
            state.cos, state.sin = rope_tables(position_ids, layer0.inv_freq.to(dev), getattr(self.model_shard, "dtype", None)
                                               or hidden_states.dtype, float(getattr(layer0, "rope_attention_scaling", 1.0)))
        output, _ = self.model_shard.forward(hidden_states, position_ids=position_ids, past_key_values=handles,
                                             use_cache=True)
        session["position"] = max(held, position + T)
        self._stats["total_tokens"] += B * T
        return output, keys

    async def _forward(self, session: Dict[str, Any], hidden_states: torch.Tensor, position: int,
                       kv_cache_keys: list) -> tuple:
        return self.forward_sync(session, hidden_states, position, kv_cache_keys)

    async def _forward_to_next(self, next_address: str, next_session_id: str, hidden_states, position: int,
                               kv_cache_keys: list) -> None:
        """Server-to-server forwarding is a transport matter (reference :376-388 is a stub as well)."""
        logger.debug("forwarding to %s (session %s) is left to the transport", next_address, next_session_id)

    # ---- RPC surface -----------------------------------------------------------------------------------------
    async def Forward(self, request, context=None) -> Dict[str, Any]:
        try:
            start = time.time()
            session_id = _get(request, "session_id")
            session = self._sessions.get(session_id) or self._create_session(session_id)
            x = self._tensor_of(_get(request, "input"), request)
            output, kv_keys = await self._forward(session, x, int(_get(request, "position", 0)),
                                                  list(_get(request, "kv_cache_keys", None) or []))
            if isinstance(_get(request, "input"), torch.Tensor):
                out_payload, out_shape, out_dtype = output, tuple(output.shape), str(output.dtype).replace("torch.", "")
            else:
                out_payload, out_shape, out_dtype = TensorSerializer.serialize(output)
            latency_ms = (time.time() - start) * 1000
            self._stats["total_requests"] += 1
            self._stats["total_latency_ms"] += latency_ms
            return {"output": out_payload, "shape": list(out_shape), "dtype": out_dtype, "updated_kv_keys": kv_keys,
                    "success": True, "error_message": "", "latency_ms": int(latency_ms)}
        except Exception as exc:  # noqa: BLE001 - the RPC reports failure in its response (reference :183-188)
            logger.error("Forward error: %s", exc)
            self._stats["errors"] += 1
            return {"success": False, "error_message": str(exc)}

    async def StreamInference(self, request_iterator: AsyncIterator, context=None) -> AsyncIterator:
        session_id = None
        async for request in request_iterator:
            try:
                start = time.time()
                session_id = _get(request, "session_id")
                session = self._get_or_create_session(session_id, request)
                x = self._tensor_of(_get(request, "hidden_states"), request)
                output, kv_keys = await self._forward(session, x, int(_get(request, "position", 0)),
                                                      list(_get(request, "kv_cache_keys", None) or []))
                if isinstance(_get(request, "hidden_states"), torch.Tensor):
                    payload, shape, dtype = output, tuple(output.shape), str(output.dtype).replace("torch.", "")
                else:
                    payload, shape, dtype = TensorSerializer.serialize(output)
                if _get(request, "next_worker_address"):
                    await self._forward_to_next(_get(request, "next_worker_address"), _get(request, "next_session_id"),
                                                output, int(_get(request, "position", 0)), kv_keys)
                latency_ms = (time.time() - start) * 1000
                self._stats["total_requests"] += 1
                self._stats["total_latency_ms"] += latency_ms
                yield {"session_id": session_id, "step_id": _get(request, "step_id"), "hidden_states": payload,
                       "shape": list(shape), "dtype": dtype, "updated_kv_keys": kv_keys, "latency_ms": int(latency_ms),
                       "tokens_processed": x.shape[1] if hasattr(x, "shape") else 1, "success": True,
                       "error_message": ""}
            except Exception as exc:  # noqa: BLE001
                logger.error("StreamInference error: %s", exc)
                self._stats["errors"] += 1
                yield {"session_id": session_id or "", "success": False, "error_message": str(exc)}

    async def TransferKVCache(self, request, context=None) -> Dict[str, Any]:
        try:
            start = time.time()
            prefix_key = _get(request, "prefix_key")
            layers = list(_get(request, "layers") or [])
            session = self._sessions.get(prefix_key)
            if session is not None:
                return self._import_into_session(session, request, layers)
            total = 0
            for layer in layers:  # the reference's unit: one block per layer per prefix key
                shape, dtype = tuple(_get(layer, "shape")), _get(layer, "dtype")
                keys = TensorSerializer.deserialize(_get(layer, "keys"), shape, dtype)
                values = TensorSerializer.deserialize(_get(layer, "values"), shape, dtype)
                idx = _get(layer, "layer_idx")
                await self.kv_cache._store_kv(f"{prefix_key}:{idx}", keys, values, idx, prefix_key)
                total += len(_get(layer, "keys")) + len(_get(layer, "values"))
            return {"success": True, "error_message": "", "bytes_transferred": total,
                    "latency_ms": int((time.time() - start) * 1000)}
        except Exception as exc:  # noqa: BLE001
            logger.error("TransferKVCache error: %s", exc)
            return {"success": False, "error_message": str(exc)}

    def _import_into_session(self, session: Dict[str, Any], request, layers) -> Dict[str, Any]:
        """Whole sequences for a session (the prefill -> decode hand-off): one request per sequence of the session's
        batch, ``shape = [n_blocks, heads, block_size, head_dim]``; ``num_tokens`` (optional) is the context length
        the blocks hold."""
        from ...kv_transfer import WireHandoff

        if not layers:
            raise ValueError("empty KVCacheRequest")
        shape = list(_get(layers[0], "shape"))
        if len(shape) != 4:
            raise ValueError("a session hand-off carries [n_blocks, heads, block_size, head_dim] per layer")
        n_blocks = shape[0]
        pool = self._pool()
        seq = int(_get(request, "sequence_index", 0) or 0)
        batch = int(_get(request, "batch", 1) or 1)
        num_tokens = int(_get(request, "num_tokens", 0) or n_blocks * pool.block_size)
        if num_tokens > n_blocks * pool.block_size:
            raise ValueError(f"{n_blocks} blocks cannot hold {num_tokens} tokens")
        kv = self._session_kv(session, batch, num_tokens)
        if not 0 <= seq < kv.batch:
            raise ValueError(f"sequence_index {seq} outside the session's batch of {kv.batch}")
        if n_blocks > kv.tables.shape[1]:
            raise ValueError(f"{n_blocks} blocks do not fit the session's {kv.tables.shape[1]} blocks per sequence")
        start_layer = int(_get(request, "start_layer", 0) or 0)
        wire = session.get("_wire")
        if wire is None or wire.start_layer != start_layer:
            wire = WireHandoff(pool, start_layer=start_layer)
            session["_wire"] = wire
        msg = {"prefix_key": _get(request, "prefix_key"), "start_layer": start_layer,
               "end_layer": _get(request, "end_layer"),
               "layers": [{k: _get(l, k) for k in ("layer_idx", "keys", "values", "shape", "dtype")} for l in layers]}
        resp = wire.import_([msg], kv.tables[seq, :n_blocks].contiguous())
        if resp.get("success"):
            session["position"] = max(session.get("position", 0), num_tokens)
        return resp

    async def CreateSession(self, request, context=None) -> Dict[str, Any]:
        try:
            if len(self._sessions) >= self.max_sessions:
                return {"success": False, "error_message": f"Max sessions reached: {self.max_sessions}"}
            session_id = str(uuid.uuid4())
            self._sessions[session_id] = {
                "session_id": session_id,
                "model_name": _get(request, "model_name"),
                "max_length": _get(request, "max_length"),
                "start_layer": _get(request, "start_layer"),
                "end_layer": _get(request, "end_layer"),
                "position": 0,
                "created_at": time.time(),
            }
            return {"session_id": session_id, "success": True, "error_message": "",
                    "cache_tokens_available": self._get_cache_capacity()}
        except Exception as exc:  # noqa: BLE001
            logger.error("CreateSession error: %s", exc)
            return {"success": False, "error_message": str(exc)}

    async def CloseSession(self, request, context=None) -> Dict[str, Any]:
        try:
            session = self._sessions.pop(_get(request, "session_id"), None)
            if session:
                self._release_kv(session)
            return {"success": True, "error_message": ""}
        except Exception as exc:  # noqa: BLE001
            return {"success": False, "error_message": str(exc)}

    async def HealthCheck(self, request=None, context=None) -> Dict[str, Any]:
        try:
            response = {"healthy": True, "worker_id": self.worker_id, "status": "online",
                        "active_sessions": len(self._sessions)}
            if _get(request, "include_stats", False):
                dev = torch.device(self._device())
                if dev.type == "cuda" and torch.cuda.is_available():
                    response["gpu_memory_used_gb"] = torch.cuda.memory_allocated(dev) / (1024 ** 3)
                    response["gpu_memory_total_gb"] = torch.cuda.get_device_properties(dev).total_memory / (1024 ** 3)
                pool = getattr(self.kv_cache, "gpu_cache", None)
                if isinstance(pool, KVCachePool):
                    used = max(pool.layer(li).get_stats()["total_blocks"] for li in range(pool.num_layers))
                    response["cache_tokens_used"] = used * pool.block_size
                response["cache_tokens_available"] = self._get_cache_capacity()
                if self._stats["total_requests"] > 0:
                    response["avg_latency_ms"] = self._stats["total_latency_ms"] / self._stats["total_requests"]
                    response["throughput_tokens_per_sec"] = (
                        self._stats["total_tokens"] / (self._stats["total_latency_ms"] / 1000)
                        if self._stats["total_latency_ms"] > 0 else 0)
            return response
        except Exception as exc:  # noqa: BLE001
            return {"healthy": False, "status": "error", "error_message": str(exc)}

    def get_stats(self) -> Dict[str, Any]:
        return {**self._stats, "active_sessions": len(self._sessions)}


class ServicerTransport:
    """Hop of a ``WorkerSession`` that calls a servicer living in this process — what the reference's HTTP route
    ``POST /inference/forward`` does around ``InferenceServicer._forward`` (worker/distributed/grpc_server.py:
    470-520), minus the wire.  Register it in ``session.TRANSPORTS[worker_id]``."""

    def __init__(self, servicer: InferenceServicer, max_length: Optional[int] = None, wire: bool = False):
        self.servicer = servicer
        self.max_length = max_length
        self.wire = wire  # True: tensors cross as TensorSerializer bytes, exactly like the RPC messages

    async def health(self) -> bool:
        return bool((await self.servicer.HealthCheck({}))["healthy"])

    async def forward(self, session_id: str, hidden_states: Any, position: int, kv_cache_keys: List[str]):
        sv = self.servicer
        if session_id not in sv._sessions:
            sv._create_session(session_id)["max_length"] = self.max_length
        req: Dict[str, Any] = {"session_id": session_id, "position": position, "kv_cache_keys": list(kv_cache_keys or [])}
        if self.wire:
            data, shape, dtype = TensorSerializer.serialize(hidden_states)
            req.update(input=data, shape=list(shape), dtype=dtype)
        else:
            req["input"] = hidden_states
        resp = await sv.Forward(req)
        if not resp.get("success"):
            raise RuntimeError(f"Forward failed: {resp.get('error_message')}")
        out = resp["output"]
        if not isinstance(out, torch.Tensor):
            out = TensorSerializer.deserialize(out, tuple(resp["shape"]), resp["dtype"], device=sv._device())
        return out, resp.get("updated_kv_keys", [])

    async def close(self, session_id: str) -> None:
        await self.servicer.CloseSession({"session_id": session_id})
