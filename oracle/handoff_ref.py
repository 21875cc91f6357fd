"""Oracle (test infrastructure): the reference's prefill→decode KV hand-off as it is done TODAY, on the host.

Reference: the sender turns every block of every layer into bytes with ``TensorSerializer.serialize``
(common/serialization.py:55-104: bfloat16 → float16 first, ``.cpu().numpy().tobytes()``); the receiver's
``InferenceServicer.TransferKVCache`` (worker/distributed/grpc_server.py:190-235) runs
``TensorSerializer.deserialize`` per layer for keys and for values (common/serialization.py:106-160:
``np.frombuffer(...).reshape(shape)``, ``.copy()``, ``torch.from_numpy``, float16 → bfloat16) and stores the block in
a pool slot (worker/distributed/kv_cache.py:447-462, a whole-block ``copy_``).

Used (a) to state what bytes the GPU wire kernels must produce and what values arrive — including the loss for
bfloat16 values outside float16's range — and (b) as the timed CPU baseline of the hand-off in bench.py.
Pinned by tests/test_oracle_golden.py against tests/golden/serializer.json (outputs of the reference's serializer).
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np
import torch

_NP = {"float16": np.float16, "float32": np.float32, "bfloat16": np.float16, "int8": np.int8, "int32": np.int32,
       "int64": np.int64}
_NAME = {torch.float16: "float16", torch.float32: "float32", torch.bfloat16: "bfloat16", torch.int8: "int8",
         torch.int32: "int32", torch.int64: "int64"}


def wire_bytes(t: torch.Tensor) -> Tuple[bytes, Tuple[int, ...], str]:
    """common/serialization.py:70-88 — the bytes one tensor becomes on the wire."""
    if t.dtype == torch.bfloat16:
        arr = t.to(torch.float16).cpu().numpy()
    else:
        arr = t.cpu().numpy()
    return arr.tobytes(), tuple(arr.shape), _NAME[t.dtype]


def from_wire(data: bytes, shape: Tuple[int, ...], dtype_str: str) -> torch.Tensor:
    """common/serialization.py:142-158 — what the receiver rebuilds (host tensor)."""
    arr = np.frombuffer(data, dtype=_NP[dtype_str]).reshape(shape)
    t = torch.from_numpy(arr.copy())
    return t.to(torch.bfloat16) if dtype_str == "bfloat16" else t


def handoff_sequence(src_k: List[torch.Tensor], src_v: List[torch.Tensor], src_table: List[int],
                     dst_k: List[torch.Tensor], dst_v: List[torch.Tensor], dst_table: List[int], serializer=None) -> int:
    """Move one sequence's blocks, layer by layer and block by block, the way the reference does; pools are per
    layer ``[max_blocks, heads, block_size, head_dim]``.  Returns the bytes that crossed the wire.
    ``serializer``: the reference's own ``TensorSerializer`` class (oracle/_ref bytecode) to do the two conversions
    with instead of the restatements above."""
    to_wire = serializer.serialize if serializer is not None else wire_bytes
    back = serializer.deserialize if serializer is not None else from_wire
    total = 0
    for kp, vp, dk, dv in zip(src_k, src_v, dst_k, dst_v):
        for s, d in zip(src_table, dst_table):
            kb, shape, name = to_wire(kp[s])
            vb, _, _ = to_wire(vp[s])
            dk[d].copy_(back(kb, shape, name))
            dv[d].copy_(back(vb, shape, name))
            total += len(kb) + len(vb)
    return total
