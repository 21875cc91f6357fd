"""Prefill → decode KV hand-off inside one box: pack kernel → NCCL send/recv over NVLink → unpack kernel.

Replaces the body of ``InferenceServicer.TransferKVCache`` (reference: worker/distributed/grpc_server.py:190-235),
which receives ONE block per layer per prefix key as host bytes and copies it into a freshly allocated GPU block
(``DistributedKVCacheManager._promote_to_gpu``, worker/distributed/kv_cache.py:447-462), and supplies the sender
the reference never wrote (``KVCacheMigrator._do_migrate`` is a ``sleep(0.05)``, server/app/services/
pd_scheduler.py:462-468).  Here a whole sequence moves at once:

  prefill rank:  one launch gathers the sequence's blocks of ALL owned layers into a contiguous buffer
                 ``[L, 2, n_blocks, Hkv, bs, D]`` (b200_kv_pack_layers), then ``dist.isend``;
  decode rank:   ``dist.irecv`` into a staging buffer, then one launch scatters it into the slots its OWN allocator
                 handed out (b200_kv_unpack_layers).  Source and destination block ids are independent.

Overlap comes from pipelining SEQUENCES, not layers: ``isend`` / ``irecv`` only enqueue work on the communicator's
stream, so with two staging buffers pack(i+1) runs while send(i) is on the wire and unpack(i) while recv(i+1) is
(tools/pd_bench.py).  Measured on 2 B200 (profiles/r02_pd_bench_2gpu.jsonl, 8 x 512 MiB): 611 GB/s per pair end to
end, against 680 GB/s for NCCL's bare send/recv of the same bytes — the ceiling of this path, reached up to the
pipeline's fill and drain (first pack + last unpack) — and 698 GB/s for the fused peer push below, which is the
product path inside a box; this NCCL path is what remains when peer mapping is not available.  Cutting one
sequence into ``chunks`` layer groups only adds launches and NCCL operations (484 GB/s with 2..8 groups), so the
default is one group.  The payload is bit-exact (no fp16 detour as in the host serializer,
common/serialization.py:73-76).
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from . import ops
from .worker.distributed.kv_cache import KVCachePool


class KVTransferEndpoint:
    """Pack / unpack side of the hand-off for one ``KVCachePool`` (all layers a rank owns)."""

    def __init__(self, pool: KVCachePool, chunks: int = 1):
        self.pool = pool
        L = pool.num_layers
        self.chunks = max(1, min(chunks, L))
        self.pool_shape = tuple(pool.layer(0).key_pool.shape)
        self.dtype = pool.layer(0).key_pool.dtype
        self.device = pool.layer(0).key_pool.device
        bounds = [round(i * L / self.chunks) for i in range(self.chunks + 1)]
        self.layer_groups = [(bounds[i], bounds[i + 1]) for i in range(self.chunks) if bounds[i + 1] > bounds[i]]
        self._k_ptrs = [ops.pool_pointer_array([pool.layer(l).key_pool for l in range(a, b)]) for a, b in self.layer_groups]
        self._v_ptrs = [ops.pool_pointer_array([pool.layer(l).value_pool for l in range(a, b)]) for a, b in self.layer_groups]

    def staging(self, n_blocks: int) -> List[torch.Tensor]:
        """One contiguous buffer per layer group, ``[l_group, 2, n_blocks, Hkv, bs, D]``."""
        _, Hkv, bs, D = self.pool_shape
        return [torch.empty((b - a, 2, n_blocks, Hkv, bs, D), dtype=self.dtype, device=self.device)
                for a, b in self.layer_groups]

    def payload_bytes(self, n_blocks: int) -> int:
        _, Hkv, bs, D = self.pool_shape
        return self.pool.num_layers * 2 * n_blocks * Hkv * bs * D * torch.empty((), dtype=self.dtype).element_size()

    # ---- sender ------------------------------------------------------------------------------------------
    def send(self, block_table: torch.Tensor, dst: int, group=None, bufs: Optional[List[torch.Tensor]] = None) -> List[dist.Work]:
        """Pack the blocks in ``block_table`` (int32, device) group by group and ship each as soon as it is packed."""
        bufs = bufs if bufs is not None else self.staging(block_table.shape[0])
        works = []
        for (a, b), kp, vp, buf in zip(self.layer_groups, self._k_ptrs, self._v_ptrs, bufs):
            ops.kv_pack_layers(kp, vp, self.pool_shape, self.dtype, block_table, out=buf)
            works.append(dist.isend(buf, dst=dst, group=group))
        return works

    # ---- receiver ----------------------------------------------------------------------------------------
    def recv(self, block_table: torch.Tensor, src: int, group=None, bufs: Optional[List[torch.Tensor]] = None) -> None:
        """Receive group by group and scatter each into the destination slots of ``block_table``."""
        bufs = bufs if bufs is not None else self.staging(block_table.shape[0])
        works = [dist.irecv(buf, src=src, group=group) for buf in bufs]
        for w, kp, vp, buf in zip(works, self._k_ptrs, self._v_ptrs, bufs):
            w.wait()
            ops.kv_unpack_layers(buf, kp, vp, self.pool_shape, block_table)

    # ---- same-device reference path (tests, single-GPU hand-off between two pools) ---------------------------
    def copy_to(self, other: "KVTransferEndpoint", src_table: torch.Tensor, dst_table: torch.Tensor) -> None:
        for (kp, vp, okp, ovp) in zip(self._k_ptrs, self._v_ptrs, other._k_ptrs, other._v_ptrs):
            buf = ops.kv_pack_layers(kp, vp, self.pool_shape, self.dtype, src_table)
            ops.kv_unpack_layers(buf, okp, ovp, other.pool_shape, dst_table)


class HostSwap:
    """Host tier for whole sequences: the L2 ("CPU memory") level of the reference's cache hierarchy
    (``DistributedKVCacheManager._add_to_cpu_cache`` / ``_promote_to_gpu``, worker/distributed/kv_cache.py:447-504),
    which moves one block of one layer per ``.cpu()`` / ``copy_`` call.

    Here a sequence's blocks of ALL layers leave the GPU with one pack launch and one device→host copy into pinned
    memory, and come back — into whatever slots the allocator hands out then — with one host→device copy and one
    unpack launch.  Bit-exact (no fp16 detour).  ``swap_out`` returns immediately after queueing the copy; the
    returned handle's ``ready`` event says when the blocks may be freed or overwritten.
    """

    class Handle:
        __slots__ = ("host", "n_blocks", "ready", "_staging")

        def __init__(self, host, n_blocks, ready, staging):
            self.host, self.n_blocks, self.ready, self._staging = host, n_blocks, ready, staging

    def __init__(self, pool: KVCachePool):
        self.ep = KVTransferEndpoint(pool, chunks=1)

    def nbytes(self, n_blocks: int) -> int:
        return self.ep.payload_bytes(n_blocks)

    def swap_out(self, block_table: torch.Tensor) -> "HostSwap.Handle":
        n = block_table.shape[0]
        buf = self.ep.staging(n)[0]
        ops.kv_pack_layers(self.ep._k_ptrs[0], self.ep._v_ptrs[0], self.ep.pool_shape, self.ep.dtype, block_table, out=buf)
        host = torch.empty(buf.shape, dtype=buf.dtype, pin_memory=True)
        host.copy_(buf, non_blocking=True)
        ready = torch.cuda.Event()
        ready.record()
        return HostSwap.Handle(host, n, ready, buf)  # the staging tensor lives until the copy has been waited for

    def swap_in(self, handle: "HostSwap.Handle", block_table: torch.Tensor) -> None:
        if block_table.shape[0] != handle.n_blocks:
            raise ValueError("block table does not match the swapped-out sequence")
        handle.ready.synchronize()
        handle._staging = None
        dev = handle.host.to(block_table.device, non_blocking=True)
        ops.kv_unpack_layers(dev, self.ep._k_ptrs[0], self.ep._v_ptrs[0], self.ep.pool_shape, block_table)


class WireHandoff:
    """Cross-box form of the hand-off (no NVLink between sender and receiver): the messages of the reference's
    ``TransferKVCache`` RPC, produced and consumed on the GPU.

    Reference: ``KVCacheRequest{prefix_key, start_layer, end_layer, layers[]}`` with ``KVCacheLayer{layer_idx, keys,
    values, shape, dtype}`` (proto/inference.proto:110-127), where ``keys`` / ``values`` are
    ``TensorSerializer.serialize`` bytes (common/serialization.py:55-104: bfloat16 travels as float16) and the
    receiver deserializes per layer and copies into a pool slot (worker/distributed/grpc_server.py:190-235).

    Here the sender makes the bytes of ALL layers and blocks with one launch (b200_kv_pack_layers_wire for bfloat16
    pools — the narrowing runs on the GPU; b200_kv_pack_layers for float16 / float32 pools, whose wire bytes are
    their own) and one device→host copy into pinned memory; the receiver does one host→device copy and one launch
    that widens and scatters into the slots of its own allocator.  The message dictionaries are field-for-field
    the proto's, so they can be handed to a gRPC / JSON layer unchanged; ``per_block=True`` emits one request per
    block with ``shape = [heads, block_size, head_dim]`` — the unit a stock reference receiver stores.
    """

    def __init__(self, pool: KVCachePool, start_layer: int = 0):
        self.ep = KVTransferEndpoint(pool, chunks=1)
        self.start_layer = start_layer
        self.num_layers = pool.num_layers
        self.dtype = self.ep.dtype
        self.wire_dtype = torch.float16 if self.dtype == torch.bfloat16 else self.dtype
        self.dtype_str = {torch.bfloat16: "bfloat16", torch.float16: "float16", torch.float32: "float32"}[self.dtype]

    def _wire_tensor(self, block_table: torch.Tensor) -> torch.Tensor:
        kp, vp = self.ep._k_ptrs[0], self.ep._v_ptrs[0]
        if self.dtype == torch.bfloat16:
            return ops.kv_pack_layers_wire(kp, vp, self.ep.pool_shape, block_table)
        return ops.kv_pack_layers(kp, vp, self.ep.pool_shape, self.dtype, block_table)

    def export(self, prefix_key: str, block_table: torch.Tensor, per_block: bool = False) -> List[dict]:
        """``KVCacheRequest`` dictionaries for the blocks in ``block_table`` (int32, device).  One request holding
        ``[n_blocks, heads, block_size, head_dim]`` per layer, or (``per_block``) one request per block keyed
        ``"{prefix_key}:{j}"`` with the reference's one-block shape."""
        n = block_table.shape[0]
        wire = self._wire_tensor(block_table)                      # [L, 2, n, Hkv, bs, D] on the device
        host = torch.empty(wire.shape, dtype=wire.dtype, pin_memory=True)
        host.copy_(wire, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        raw = host.view(torch.uint8).numpy()                        # zero-copy view of the pinned buffer
        _, Hkv, bs, D = self.ep.pool_shape

        def layer_msgs(j0: int, j1: int, shape: List[int]) -> List[dict]:
            return [{"layer_idx": self.start_layer + l, "keys": raw[l, 0, j0:j1].tobytes(),
                     "values": raw[l, 1, j0:j1].tobytes(), "shape": shape, "dtype": self.dtype_str}
                    for l in range(self.num_layers)]

        head = {"start_layer": self.start_layer, "end_layer": self.start_layer + self.num_layers}
        if per_block:
            return [{"prefix_key": f"{prefix_key}:{j}", **head, "layers": layer_msgs(j, j + 1, [Hkv, bs, D])}
                    for j in range(n)]
        return [{"prefix_key": prefix_key, **head, "layers": layer_msgs(0, n, [n, Hkv, bs, D])}]

    # ---- StreamingTensorBuffer framing (reference: common/serialization.py:209-265) --------------------------
    def export_stream(self, block_table: torch.Tensor, chunk_size: int = 1024 * 1024):
        """The same payload as ``export`` as ONE chunked tensor stream: first the ``StreamingTensorBuffer`` header
        (``[ndim u32][dims u64 ...][dtype id u8]`` for shape ``[L, 2, n_blocks, heads, block_size, head_dim]``), then
        ``chunk_size``-byte frames (1 MiB by default, the reference's chunk size) — what a stock
        ``StreamingTensorBuffer`` receiver consumes with ``read_header`` / ``write_chunk`` / ``finalize``.

        The wire tensor is made by one launch; the device→host copy runs layer by layer on a side stream and a
        layer's frames are yielded as soon as its copy has landed in the pinned buffer, so the consumer (a socket
        writer) overlaps with the remaining copies.  Frames are zero-copy ``memoryview`` slices of that pinned
        buffer, valid until the generator is closed; a frame never straddles more than the bytes in flight."""
        from .common.serialization import StreamingTensorBuffer

        wire = self._wire_tensor(block_table)                      # [L, 2, n, Hkv, bs, D] on the device
        host = torch.empty(wire.shape, dtype=wire.dtype, pin_memory=True)
        side = torch.cuda.Stream(device=wire.device)
        side.wait_stream(torch.cuda.current_stream())
        done = []
        with torch.cuda.stream(side):
            for l in range(self.num_layers):
                host[l].copy_(wire[l], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(side)
                done.append(ev)
        wire.record_stream(side)
        yield StreamingTensorBuffer(chunk_size).write_header(tuple(wire.shape), self.dtype_str)
        raw = memoryview(host.view(torch.uint8).numpy()).cast("B")
        per_layer = len(raw) // self.num_layers
        sent, landed = 0, 0
        for l in range(self.num_layers):
            done[l].synchronize()
            landed += per_layer
            last = l == self.num_layers - 1
            while sent + chunk_size <= landed or (last and sent < landed):
                end = min(sent + chunk_size, landed)
                yield raw[sent:end]
                sent = end

    def import_stream(self, frames, block_table: torch.Tensor) -> dict:
        """Receiver of ``export_stream``: header frame, then data frames of any size, written straight into a pinned
        buffer; one host→device copy and one launch scatter them into the slots of ``block_table``.  Returns the
        reference's ``KVCacheResponse`` fields."""
        import time

        from .common.serialization import StreamingTensorBuffer

        t0 = time.time()
        try:
            it = iter(frames)
            meta = StreamingTensorBuffer().read_header(bytes(next(it)))
            n = block_table.shape[0]
            _, Hkv, bs, D = self.ep.pool_shape
            want = (self.num_layers, 2, n, Hkv, bs, D)
            if tuple(meta["shape"]) != want:
                raise ValueError(f"stream carries shape {tuple(meta['shape'])}, expected {want}")
            if meta["dtype"] != self.dtype_str:
                raise TypeError(f"stream carries {meta['dtype']}, the pool holds {self.dtype_str}")
            host = torch.empty(want, dtype=self.wire_dtype, pin_memory=True)
            raw = memoryview(host.view(torch.uint8).numpy()).cast("B")
            off = 0
            for frame in it:
                size = len(frame)
                if off + size > len(raw):
                    raise ValueError("stream is longer than its header says")
                raw[off:off + size] = frame
                off += size
            if off != len(raw):
                raise ValueError(f"stream ended after {off} of {len(raw)} bytes")
            dev = host.to(block_table.device, non_blocking=True)
            kp, vp = self.ep._k_ptrs[0], self.ep._v_ptrs[0]
            if self.dtype == torch.bfloat16:
                ops.kv_unpack_layers_wire(dev, kp, vp, self.ep.pool_shape, block_table)
            else:
                ops.kv_unpack_layers(dev, kp, vp, self.ep.pool_shape, block_table)
            torch.cuda.current_stream().synchronize()
        except Exception as exc:  # noqa: BLE001 - the RPC reports failure in its response
            return {"success": False, "error_message": str(exc)}
        return {"success": True, "error_message": "", "bytes_transferred": off,
                "latency_ms": int((time.time() - t0) * 1000)}

    def import_(self, requests: List[dict], block_table: torch.Tensor) -> dict:
        """Store the blocks carried by ``requests`` (as produced by ``export``, either form, in order) into the
        slots of ``block_table``; returns the reference's ``KVCacheResponse`` fields."""
        import time

        import numpy as np

        t0 = time.time()
        n = block_table.shape[0]
        _, Hkv, bs, D = self.ep.pool_shape
        es = torch.empty((), dtype=self.wire_dtype).element_size()
        blk = Hkv * bs * D * es
        host = torch.empty((self.num_layers, 2, n, Hkv, bs, D), dtype=self.wire_dtype, pin_memory=True)
        raw = host.view(torch.uint8).numpy().reshape(self.num_layers, 2, n * blk)
        total, j = 0, 0
        try:
            for req in requests:
                cnt = None
                for msg in req["layers"]:
                    if msg["dtype"] != self.dtype_str:
                        raise TypeError(f"request carries {msg['dtype']}, the pool holds {self.dtype_str}")
                    l = msg["layer_idx"] - self.start_layer
                    if not 0 <= l < self.num_layers:
                        raise ValueError(f"layer {msg['layer_idx']} is not owned here")
                    for kv, key in ((0, "keys"), (1, "values")):
                        data = np.frombuffer(msg[key], dtype=np.uint8)
                        if data.size % blk or j * blk + data.size > n * blk:
                            raise ValueError("payload does not match the block geometry / block table")
                        raw[l, kv, j * blk:j * blk + data.size] = data
                        total += data.size
                    cnt = data.size // blk
                j += cnt or 0
            if j != n:
                raise ValueError(f"requests carry {j} blocks, the block table has {n}")
            dev = host.to(block_table.device, non_blocking=True)
            kp, vp = self.ep._k_ptrs[0], self.ep._v_ptrs[0]
            if self.dtype == torch.bfloat16:
                ops.kv_unpack_layers_wire(dev, kp, vp, self.ep.pool_shape, block_table)
            else:
                ops.kv_unpack_layers(dev, kp, vp, self.ep.pool_shape, block_table)
            torch.cuda.current_stream().synchronize()  # the pinned staging buffer is released on return
        except Exception as exc:  # noqa: BLE001 - the RPC reports failure in its response
            return {"success": False, "error_message": str(exc)}
        return {"success": True, "error_message": "", "bytes_transferred": total,
                "latency_ms": int((time.time() - t0) * 1000)}


class PeerPushEndpoint:
    """Hand-off without staging: the prefill GPU writes a sequence's blocks straight into the decode GPU's pools.

    Every rank allocates the backing tensor of its ``KVCachePool`` from torch's symmetric memory and exchanges the
    mappings once (``rendezvous``); afterwards a rank holds a device pointer to every peer's pools.  ``push`` is
    then ONE kernel (b200_kv_move_blocks_layers) whose loads hit local HBM and whose 16-byte stores travel over
    NVLink / NVSwitch into the blocks the receiver's own allocator handed out — the gather (pack), the transfer and
    the scatter (unpack) of the NCCL path fused into one pass, with no packed buffer on either side.

    Protocol per sequence: the receiver allocates blocks and tells the sender its block table (a few hundred
    int32); the sender pushes and then signals completion (``barrier`` on the symmetric-memory handle orders the
    two streams).
    """

    def __init__(self, num_layers: int, num_heads: int, head_dim: int, block_size: int, max_blocks_per_layer: int,
                 device: str, dtype: torch.dtype, group=None):
        import torch.distributed._symmetric_memory as symm_mem

        self.group = group if group is not None else dist.group.WORLD
        shape = (num_layers, 2, max_blocks_per_layer, num_heads, block_size, head_dim)
        try:  # needed by older torch releases, a deprecated no-op in newer ones
            symm_mem.enable_symm_mem_for_group(self.group.group_name)
        except Exception:  # noqa: BLE001
            pass
        backing = symm_mem.empty(shape, dtype=dtype, device=device)
        backing.zero_()
        self.handle = symm_mem.rendezvous(backing, self.group.group_name)
        self.pool = KVCachePool(num_layers, num_heads, head_dim, block_size, max_blocks_per_layer, device=device,
                                dtype=dtype, backing=backing)
        self.pool_shape = (max_blocks_per_layer, num_heads, block_size, head_dim)
        self.dtype = dtype
        self.device = torch.device(device)
        es = backing.element_size()
        plane = max_blocks_per_layer * num_heads * block_size * head_dim * es  # bytes of one K or V pool
        self._ptrs = {}
        for peer, base in enumerate(self.handle.buffer_ptrs):
            k = torch.tensor([base + (2 * l) * plane for l in range(num_layers)], dtype=torch.int64).to(self.device)
            v = torch.tensor([base + (2 * l + 1) * plane for l in range(num_layers)], dtype=torch.int64).to(self.device)
            self._ptrs[peer] = (k, v)
        self.rank = self.handle.rank

    def payload_bytes(self, n_blocks: int) -> int:
        _, Hkv, bs, D = self.pool_shape
        return self.pool.num_layers * 2 * n_blocks * Hkv * bs * D * torch.empty((), dtype=self.dtype).element_size()

    def push(self, src_table: torch.Tensor, dst_table: torch.Tensor, dst_rank: int) -> None:
        """Enqueue the copy of local blocks ``src_table`` into blocks ``dst_table`` of rank ``dst_rank``'s pools."""
        sk, sv = self._ptrs[self.rank]
        dk, dv = self._ptrs[dst_rank]
        ops.kv_move_blocks_layers(dk, dv, self.pool_shape, sk, sv, self.pool_shape, self.dtype, src_table, dst_table)

    def barrier(self) -> None:
        """Order the pushes issued so far on every rank before whatever the ranks enqueue next."""
        self.handle.barrier()
