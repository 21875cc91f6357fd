"""Host wire format for tensors: the reference's prefill→decode / inter-stage envelope.

Same observable behaviour as the reference's ``common/serialization.py``:

* ``TensorSerializer.serialize`` → ``(bytes, shape, dtype_str)``, C-order, native endian
  (reference: common/serialization.py:55-104).  bf16 goes through fp16 on the wire and is widened back on
  receipt (reference :73-76 and :143-155) — lossy outside the fp16 range; the device-side hand-off in
  ``dgi_b200.kv_transfer`` moves bf16 bit-exactly instead and never touches this path.
* ``serialize_tensor`` / ``deserialize_tensor``: base64 dict envelope (reference :163-206).
* ``StreamingTensorBuffer``: ``[ndim u32][dim u64]*[dtype_id u8]`` header + fixed-size chunks (reference :209-265).

lz4 / zstd are used when importable and silently skipped otherwise, as in the reference.
"""
from __future__ import annotations

import base64
import io
import struct
from typing import Any, Dict, Iterator, Optional, Tuple

import numpy as np

try:  # torch is optional for the wire format, as in the reference
    import torch

    HAS_TORCH = True
except ImportError:  # pragma: no cover
    HAS_TORCH = False

# Numeric ids used in the streaming header (reference: common/serialization.py:20-27).
DTYPE_TO_ID: Dict[str, int] = {"float16": 0, "float32": 1, "bfloat16": 2, "int8": 3, "int32": 4, "int64": 5}
ID_TO_DTYPE: Dict[int, str] = {i: n for n, i in DTYPE_TO_ID.items()}

DTYPE_TO_NUMPY = {
    "float16": np.float16,
    "float32": np.float32,
    "int8": np.int8,
    "int32": np.int32,
    "int64": np.int64,
}

if HAS_TORCH:
    DTYPE_TO_TORCH = {name: getattr(torch, name) for name in DTYPE_TO_ID}
    TORCH_TO_DTYPE = {t: n for n, t in DTYPE_TO_TORCH.items()}


def _maybe_compress(buf: bytes, how: str) -> bytes:
    if how == "lz4":
        try:
            import lz4.frame

            return lz4.frame.compress(buf)
        except ImportError:
            return buf
    if how == "zstd":
        try:
            import zstandard

            return zstandard.ZstdCompressor().compress(buf)
        except ImportError:
            return buf
    return buf


def _maybe_decompress(buf: bytes, how: str) -> bytes:
    if how == "lz4":
        try:
            import lz4.frame

            return lz4.frame.decompress(buf)
        except ImportError:
            return buf
    if how == "zstd":
        try:
            import zstandard

            return zstandard.ZstdDecompressor().decompress(buf)
        except ImportError:
            return buf
    return buf


class TensorSerializer:
    """Tensor ↔ bytes (reference: common/serialization.py:52-160)."""

    @staticmethod
    def serialize(data: Any, compression: str = "none") -> Tuple[bytes, Tuple[int, ...], str]:
        if HAS_TORCH and isinstance(data, torch.Tensor):
            if data.dtype == torch.bfloat16:
                arr = data.to(torch.float16).cpu().numpy()
                dtype_str = "bfloat16"
            else:
                arr = data.cpu().numpy()
                dtype_str = TORCH_TO_DTYPE.get(data.dtype, "float32")
        elif isinstance(data, np.ndarray):
            arr = data
            dtype_str = str(arr.dtype)
        else:
            raise TypeError(f"Unsupported type: {type(data)}")
        return _maybe_compress(arr.tobytes(), compression), arr.shape, dtype_str

    @staticmethod
    def deserialize(
        data_bytes: bytes,
        shape: Tuple[int, ...],
        dtype_str: str,
        compression: str = "none",
        device: str = "cpu",
    ) -> Any:
        raw = _maybe_decompress(data_bytes, compression)
        np_dtype = np.float16 if dtype_str == "bfloat16" else DTYPE_TO_NUMPY.get(dtype_str, np.float32)
        arr = np.frombuffer(raw, dtype=np_dtype).reshape(shape)
        if not HAS_TORCH or device == "numpy":
            return arr
        out = torch.from_numpy(arr.copy())
        if dtype_str == "bfloat16":
            out = out.to(torch.bfloat16)
        if device.startswith("cuda") and torch.cuda.is_available():
            out = out.to(device)
        return out


def serialize_tensor(data: Any, compression: str = "none") -> Dict[str, Any]:
    """Dict envelope with base64 payload (reference: common/serialization.py:163-187)."""
    payload, shape, dtype_str = TensorSerializer.serialize(data, compression)
    return {
        "data": base64.b64encode(payload).decode("ascii"),
        "shape": list(shape),
        "dtype": dtype_str,
        "compression": compression,
    }


def deserialize_tensor(serialized: Dict[str, Any], device: str = "cpu") -> Any:
    """Inverse of ``serialize_tensor`` (reference: common/serialization.py:190-206)."""
    return TensorSerializer.deserialize(
        base64.b64decode(serialized["data"]),
        tuple(serialized["shape"]),
        serialized["dtype"],
        serialized.get("compression", "none"),
        device,
    )


class StreamingTensorBuffer:
    """Chunked transfer of one tensor (reference: common/serialization.py:209-265)."""

    def __init__(self, chunk_size: int = 1024 * 1024):
        self.chunk_size = chunk_size
        self.buffer = io.BytesIO()
        self.metadata: Optional[Dict[str, Any]] = None

    def write_header(self, shape: Tuple[int, ...], dtype_str: str) -> bytes:
        # native-endian, unpadded pieces: u32 ndim, u64 per dim, u8 dtype id (unknown dtype → 0)
        parts = [struct.pack("I", len(shape))]
        parts += [struct.pack("Q", d) for d in shape]
        parts.append(struct.pack("B", DTYPE_TO_ID.get(dtype_str, 0)))
        self.metadata = {"shape": shape, "dtype": dtype_str}
        return b"".join(parts)

    def read_header(self, header_bytes: bytes) -> Dict[str, Any]:
        (ndim,) = struct.unpack_from("I", header_bytes, 0)
        dims = [struct.unpack_from("Q", header_bytes, 4 + 8 * i)[0] for i in range(ndim)]
        (dtype_id,) = struct.unpack_from("B", header_bytes, 4 + 8 * ndim)
        return {"shape": tuple(dims), "dtype": ID_TO_DTYPE.get(dtype_id, "float32")}

    def iter_chunks(self, data_bytes: bytes) -> Iterator[bytes]:
        for off in range(0, len(data_bytes), self.chunk_size):
            yield data_bytes[off : off + self.chunk_size]

    def write_chunk(self, chunk: bytes) -> None:
        self.buffer.write(chunk)

    def finalize(self, device: str = "cpu") -> Any:
        if self.metadata is None:
            raise ValueError("Header not received")
        return TensorSerializer.deserialize(
            self.buffer.getvalue(), self.metadata["shape"], self.metadata["dtype"], device=device
        )
