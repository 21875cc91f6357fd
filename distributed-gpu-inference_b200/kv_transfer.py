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

To overlap the copy engines with NVLink the layers can be moved in ``chunks`` groups: pack(c+1) runs while
send(c) is on the wire, and unpack(c) while recv(c+1) is.  The payload is bit-exact (no fp16 detour as in the host
serializer, common/serialization.py:73-76).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

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
