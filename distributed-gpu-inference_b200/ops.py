"""Torch-tensor front end of the C ABI (include/b200_kv.h).

Each function validates shapes on the host, passes raw device pointers plus torch's current CUDA stream to
libb200_kv.so and returns immediately (the work is queued on that stream).  No function here computes
anything itself and none falls back to torch: a missing library or a non-CUDA tensor is an error.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from . import _lib

DTYPE_ID = {torch.float16: 0, torch.float32: 1, torch.bfloat16: 2}

ATTN_FORCE_GENERIC = 1
ATTN_ROUND_SCORES = 2
ATTN_WS_COUNTERS_ZERO = 4  # shared-prefix calls: the caller vouches for the zeroed counter header of its workspace

# bench.py instrumentation: when set to a list, every attention call appends a (start, end) CUDA-event pair
# recorded on the launching stream around the C-ABI call.
ATTN_EVENTS = None


def _dtype_id(t: torch.Tensor) -> int:
    try:
        return DTYPE_ID[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported KV dtype {t.dtype}; expected float16, bfloat16 or float32") from None


def _need_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if not t.is_cuda:
            raise _lib.B200LibraryError(
                "paged-KV ops run only on CUDA tensors (sm_100a kernels); there is no CPU fallback"
            )


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _pool_dims(pool: torch.Tensor):
    if pool.dim() != 4 or not pool.is_contiguous():
        raise ValueError("pool must be a contiguous [max_blocks, Hkv, block_size, D] tensor")
    return pool.shape


def _row_stride(x: torch.Tensor, heads: int, D: int) -> int:
    """x is [n, heads, D] with heads*D contiguous inside a row; returns the row stride in elements."""
    if x.dim() != 3 or x.shape[1] != heads or x.shape[2] != D:
        raise ValueError(f"expected [n, {heads}, {D}], got {tuple(x.shape)}")
    if x.shape[0] > 0 and (x.stride(2) != 1 or x.stride(1) != D):
        raise ValueError("heads and head_dim must be contiguous inside a row")
    return x.stride(0) if x.shape[0] > 1 else max(x.stride(0), heads * D)


def kv_append(
    k_new: torch.Tensor,
    v_new: torch.Tensor,
    k_pool: torch.Tensor,
    v_pool: torch.Tensor,
    slot_mapping: torch.Tensor,
) -> None:
    """Scatter ``k_new/v_new [n_tok, Hkv, D]`` into the pools at ``slot_mapping`` (int64, negative = skip)."""
    _need_cuda(k_new, v_new, k_pool, v_pool, slot_mapping)
    max_blocks, Hkv, bs, D = _pool_dims(k_pool)
    if v_pool.shape != k_pool.shape or v_pool.dtype != k_pool.dtype or not v_pool.is_contiguous():
        raise ValueError("k_pool and v_pool must match")
    if k_new.dtype != k_pool.dtype or v_new.dtype != k_pool.dtype:
        raise TypeError("new K/V dtype must equal the pool dtype")
    if slot_mapping.dtype != torch.int64 or slot_mapping.dim() != 1 or not slot_mapping.is_contiguous():
        raise ValueError("slot_mapping must be a contiguous int64 vector")
    n_tok = k_new.shape[0]
    if v_new.shape[0] != n_tok or slot_mapping.shape[0] != n_tok:
        raise ValueError("k_new, v_new and slot_mapping disagree on the token count")
    ks, vs = _row_stride(k_new, Hkv, D), _row_stride(v_new, Hkv, D)
    lib = _lib.load()
    rc = lib.b200_kv_append(
        k_new.data_ptr(), v_new.data_ptr(), ks, vs, k_pool.data_ptr(), v_pool.data_ptr(), slot_mapping.data_ptr(),
        n_tok, Hkv, D, bs, max_blocks, _dtype_id(k_pool), _stream(),
    )
    _lib.check(rc, "b200_kv_append")


def add_rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float, delta: Optional[torch.Tensor] = None,
                want_residual: bool = False):
    """``h = x + delta`` (if given) then ``weight * rmsnorm(h)``; returns ``out`` or ``(out, h)``.

    ``x`` / ``delta``: ``[..., hidden]`` with a contiguous last dim and uniform row stride; outputs are contiguous.
    """
    _need_cuda(x, weight)
    hidden = x.shape[-1]
    x2 = x.reshape(-1, hidden) if x.is_contiguous() else x.view(-1, hidden)
    rows = x2.shape[0]
    out = torch.empty((rows, hidden), dtype=x.dtype, device=x.device)
    d2 = None
    resid = None
    if delta is not None:
        d2 = delta.reshape(-1, hidden) if delta.is_contiguous() else delta.view(-1, hidden)
        if want_residual:
            resid = torch.empty((rows, hidden), dtype=x.dtype, device=x.device)
    rc = _lib.load().b200_add_rmsnorm(
        out.data_ptr(), resid.data_ptr() if resid is not None else 0, x2.data_ptr(), d2.data_ptr() if d2 is not None else 0,
        weight.data_ptr(), float(eps), rows, hidden, x2.stride(0) if rows > 1 else hidden,
        (d2.stride(0) if rows > 1 else hidden) if d2 is not None else 0, _dtype_id(x), _stream())
    _lib.check(rc, "b200_add_rmsnorm")
    out = out.view(x.shape)
    if want_residual:
        return out, (resid.view(x.shape) if resid is not None else x)
    return out


def rope_kv_append(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor,
                   k_pool: torch.Tensor, v_pool: torch.Tensor, slot_mapping: torch.Tensor,
                   rotate_k_in_place: bool = False) -> None:
    """Rotate ``q [n, Hq, D]`` in place, rotate ``k [n, Hkv, D]`` and scatter it and ``v`` into the pools.

    ``cos`` / ``sin``: contiguous ``[n, D]`` tables in the storage dtype.  ``rotate_k_in_place``: also overwrite
    ``k`` with its rotated value (prompt pass, where the new tokens attend to each other densely).
    """
    _need_cuda(q, k, v, cos, sin, k_pool, v_pool, slot_mapping)
    max_blocks, Hkv, bs, D = _pool_dims(k_pool)
    n, Hq = q.shape[0], q.shape[1]
    if cos.shape != (n, D) or sin.shape != (n, D) or not cos.is_contiguous() or not sin.is_contiguous():
        raise ValueError("cos / sin must be contiguous [n_tok, D]")
    if slot_mapping.dtype != torch.int64 or slot_mapping.shape != (n,) or not slot_mapping.is_contiguous():
        raise ValueError("slot_mapping must be a contiguous int64 [n_tok]")
    if not (q.dtype == k.dtype == v.dtype == cos.dtype == sin.dtype == k_pool.dtype):
        raise TypeError("q, k, v, cos, sin and the pools must share one dtype")
    fn = _lib.load().b200_rope_kv_append_prompt if rotate_k_in_place else _lib.load().b200_rope_kv_append
    rc = fn(
        q.data_ptr(), _row_stride(q, Hq, D), k.data_ptr(), _row_stride(k, Hkv, D), v.data_ptr(), _row_stride(v, Hkv, D),
        cos.data_ptr(), sin.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), slot_mapping.data_ptr(), n, Hq, Hkv, D, bs,
        max_blocks, _dtype_id(k_pool), _stream())
    _lib.check(rc, "b200_rope_kv_append")


def silu_mul(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """``silu(gate) * up`` for ``[..., cols]`` tensors with contiguous last dim; returns a contiguous tensor."""
    _need_cuda(gate, up)
    cols = gate.shape[-1]
    g2, u2 = gate.flatten(0, -2), up.flatten(0, -2)  # views when the leading dims collapse (uniform row stride)
    rows = g2.shape[0]
    if g2.stride(-1) != 1 or u2.stride(-1) != 1:
        raise ValueError("gate / up need a contiguous last dimension")
    out = torch.empty((rows, cols), dtype=gate.dtype, device=gate.device)
    rc = _lib.load().b200_silu_mul(out.data_ptr(), g2.data_ptr(), u2.data_ptr(), rows, cols,
                                  g2.stride(0) if rows > 1 else cols, u2.stride(0) if rows > 1 else cols,
                                  _dtype_id(gate), _stream())
    _lib.check(rc, "b200_silu_mul")
    return out.view(gate.shape)


def argmax_rows(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``torch.argmax(x, dim=-1)`` for a 2-D ``[rows, cols]`` tensor with a contiguous last dim (the greedy token of
    a decode step); ``out``: int64 ``[rows]``."""
    _need_cuda(x)
    if x.dim() != 2 or x.stride(1) != 1:
        raise ValueError("x must be [rows, cols] with a contiguous last dimension")
    rows, cols = x.shape
    if out is None:
        out = torch.empty((rows,), dtype=torch.int64, device=x.device)
    elif out.dtype != torch.int64 or out.shape != (rows,) or not out.is_contiguous() or not out.is_cuda:
        raise ValueError("out must be a contiguous CUDA int64 [rows]")
    rc = _lib.load().b200_argmax_rows(out.data_ptr(), x.data_ptr(), rows, cols, x.stride(0) if rows > 1 else cols,
                                     _dtype_id(x), _stream())
    _lib.check(rc, "b200_argmax_rows")
    return out


def argmax_rows_keys(x: torch.Tensor, col_offset: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Per-row best (value, ``col_offset`` + column) of a column slice of the logits as an int64 key; the element-wise
    MAX of the ranks' key vectors (one all-reduce) is the key of ``torch.argmax`` over the whole row."""
    _need_cuda(x)
    if x.dim() != 2 or x.stride(1) != 1:
        raise ValueError("x must be [rows, cols] with a contiguous last dimension")
    rows, cols = x.shape
    if out is None:
        out = torch.empty((rows,), dtype=torch.int64, device=x.device)
    elif out.dtype != torch.int64 or out.shape != (rows,) or not out.is_contiguous() or not out.is_cuda:
        raise ValueError("out must be a contiguous CUDA int64 [rows]")
    rc = _lib.load().b200_argmax_rows_keys(out.data_ptr(), x.data_ptr(), rows, cols, x.stride(0) if rows > 1 else cols,
                                          int(col_offset), _dtype_id(x), _stream())
    _lib.check(rc, "b200_argmax_rows_keys")
    return out


def argmax_keys_decode(keys: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Token ids of (max-reduced) ``argmax_rows_keys`` keys."""
    _need_cuda(keys)
    if keys.dtype != torch.int64 or keys.dim() != 1 or not keys.is_contiguous():
        raise ValueError("keys must be a contiguous int64 vector")
    if out is None:
        out = torch.empty_like(keys)
    elif out.dtype != torch.int64 or out.shape != keys.shape or not out.is_contiguous() or not out.is_cuda:
        raise ValueError("out must be a contiguous CUDA int64 vector of the keys' length")
    rc = _lib.load().b200_argmax_keys_decode(out.data_ptr(), keys.data_ptr(), keys.shape[0], _stream())
    _lib.check(rc, "b200_argmax_keys_decode")
    return out


def attn_plan(B: int, Hq: int, Hkv: int, D: int, block_size: int, max_context_len: int, dtype: torch.dtype) -> int:
    """Number of context partitions the library would pick for this shape."""
    return _lib.load().b200_paged_decode_attn_plan(B, Hq, Hkv, D, block_size, max_context_len, DTYPE_ID[dtype])


def attn_workspace_bytes(B: int, Hq: int, D: int, num_splits: int) -> int:
    return _lib.load().b200_paged_decode_attn_workspace(B, Hq, D, num_splits)


def _attn_args(q, k_pool, v_pool, block_tables, context_lens, out):
    """Shared checks of the attention entry points; returns ``(dims, q_stride, table_stride, out)``."""
    _need_cuda(q, k_pool, v_pool, block_tables, context_lens)
    max_blocks, Hkv, bs, D = _pool_dims(k_pool)
    if v_pool.shape != k_pool.shape or v_pool.dtype != k_pool.dtype or not v_pool.is_contiguous():
        raise ValueError("k_pool and v_pool must match")
    if q.dtype != k_pool.dtype:
        raise TypeError("q dtype must equal the pool dtype")
    B, Hq = q.shape[0], q.shape[1]
    q_stride = _row_stride(q, Hq, D)
    if block_tables.dtype != torch.int32 or block_tables.dim() != 2 or block_tables.shape[0] != B:
        raise ValueError("block_tables must be int32 [B, n]")
    if block_tables.stride(1) != 1 and block_tables.shape[1] > 1:
        raise ValueError("block_tables rows must be contiguous")
    if context_lens.dtype != torch.int32 or context_lens.shape != (B,) or not context_lens.is_contiguous():
        raise ValueError("context_lens must be a contiguous int32 [B]")
    if out is None:
        out = torch.empty((B, Hq, D), dtype=q.dtype, device=q.device)
    elif tuple(out.shape) != (B, Hq, D) or not out.is_contiguous() or out.dtype != q.dtype or not out.is_cuda:
        raise ValueError("out must be a contiguous CUDA [B, Hq, D] tensor of q's dtype")
    table_stride = block_tables.stride(0) if B > 1 else block_tables.shape[1]
    return (B, Hq, Hkv, D, bs, max_blocks), q_stride, table_stride, out


def _split_workspace(lib, B, Hq, D, n_splits, workspace, device, always=False):
    """``(ptr, bytes, tensor)`` of the split-K scratch: the caller's, or one from torch's caching allocator."""
    if n_splits <= 1 and not always:
        return 0, 0, None
    need = max(16, lib.b200_paged_decode_attn_workspace(B, Hq, D, n_splits))
    if workspace is None:
        workspace = torch.empty(need, dtype=torch.uint8, device=device)
    elif not workspace.is_cuda:
        raise _lib.B200LibraryError("workspace must be a CUDA tensor")
    return workspace.data_ptr(), workspace.numel() * workspace.element_size(), workspace


class _timed:
    """Brackets a launch with CUDA events when ``ATTN_EVENTS`` is a list (bench instrumentation), else does nothing."""

    def __enter__(self):
        self.events = ATTN_EVENTS
        if self.events is not None:
            self.ev0, self.ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            self.ev0.record()
        return self

    def __exit__(self, *exc):
        if self.events is not None:
            self.ev1.record()
            self.events.append((self.ev0, self.ev1))
        return False


def paged_decode_attention(
    q: torch.Tensor,
    k_pool: torch.Tensor,
    v_pool: torch.Tensor,
    block_tables: torch.Tensor,
    context_lens: torch.Tensor,
    max_context_len: int,
    scale: Optional[float] = None,
    out: Optional[torch.Tensor] = None,
    num_splits: int = 0,
    workspace: Optional[torch.Tensor] = None,
    flags: int = 0,
) -> torch.Tensor:
    """One-query-per-sequence GQA attention over paged KV.

    ``q [B, Hq, D]`` (row stride free), pools ``[max_blocks, Hkv, bs, D]``, ``block_tables int32 [B, n]``,
    ``context_lens int32 [B]``.  ``max_context_len`` is the host-side bound used to plan split-K.
    ``workspace`` (any dtype, CUDA) is needed when more than one split is used; when it is ``None`` and
    ``num_splits`` is 0 a workspace is taken from torch's caching allocator.
    """
    (B, Hq, Hkv, D, bs, max_blocks), q_stride, table_stride, out = _attn_args(q, k_pool, v_pool, block_tables,
                                                                               context_lens, out)
    lib = _lib.load()
    dt = _dtype_id(k_pool)
    if num_splits == 0:
        num_splits = lib.b200_paged_decode_attn_plan(B, Hq, Hkv, D, bs, max_context_len, dt)
    ws_ptr, ws_bytes, _ws = _split_workspace(lib, B, Hq, D, num_splits, workspace, q.device)
    with _timed():
        rc = lib.b200_paged_decode_attn(
            out.data_ptr(), q.data_ptr(), q_stride, k_pool.data_ptr(), v_pool.data_ptr(), block_tables.data_ptr(),
            table_stride, context_lens.data_ptr(), B, Hq, Hkv, D, bs, max_blocks,
            float(D) ** -0.5 if scale is None else float(scale), int(max_context_len), int(num_splits), ws_ptr, ws_bytes,
            dt, int(flags), _stream(),
        )
    _lib.check(rc, "b200_paged_decode_attn")
    return out


def fused_decode_supported(k_pool: torch.Tensor, Hq: int) -> bool:
    """Shapes the one-launch RoPE + append + attention entry point covers."""
    _, Hkv, bs, D = k_pool.shape
    return k_pool.dtype in (torch.bfloat16, torch.float16) and bs == 16 and D in (64, 128) and Hq // Hkv <= 8


def rope_append_decode_attention(
    q: torch.Tensor,
    k_new: torch.Tensor,
    v_new: torch.Tensor,
    cos: torch.Tensor,
    sin: torch.Tensor,
    k_pool: torch.Tensor,
    v_pool: torch.Tensor,
    block_tables: torch.Tensor,
    context_lens: torch.Tensor,
    max_context_len: int,
    scale: Optional[float] = None,
    out: Optional[torch.Tensor] = None,
    num_splits: int = 0,
    workspace: Optional[torch.Tensor] = None,
    flags: int = 0,
) -> torch.Tensor:
    """RoPE(q), RoPE(k_new), append of k_new / v_new at each sequence's last slot, and paged attention — one launch.

    ``q [B, Hq, D]`` un-rotated (left untouched), ``k_new / v_new [B, Hkv, D]``, ``cos / sin [B, D]``;
    ``context_lens`` already counts the new token.
    """
    _need_cuda(k_new, v_new, cos, sin)
    (B, Hq, Hkv, D, bs, max_blocks), q_stride, table_stride, out = _attn_args(q, k_pool, v_pool, block_tables,
                                                                               context_lens, out)
    if cos.shape != (B, D) or sin.shape != (B, D) or not cos.is_contiguous() or not sin.is_contiguous():
        raise ValueError("cos / sin must be contiguous [B, D]")
    if not (k_new.dtype == v_new.dtype == cos.dtype == sin.dtype == q.dtype):
        raise TypeError("q, k_new, v_new, cos and sin must share the pool dtype")
    if k_new.shape[0] != B or v_new.shape[0] != B:
        raise ValueError("k_new / v_new must hold one token per sequence")
    lib = _lib.load()
    dt = _dtype_id(k_pool)
    if num_splits == 0:
        num_splits = lib.b200_paged_decode_attn_plan(B, Hq, Hkv, D, bs, max_context_len, dt)
    ws_ptr, ws_bytes, _ws = _split_workspace(lib, B, Hq, D, num_splits, workspace, q.device)
    with _timed():
        rc = lib.b200_rope_append_decode_attn(
            out.data_ptr(), q.data_ptr(), q_stride, k_new.data_ptr(), _row_stride(k_new, Hkv, D),
            v_new.data_ptr(), _row_stride(v_new, Hkv, D), cos.data_ptr(), sin.data_ptr(), k_pool.data_ptr(),
            v_pool.data_ptr(), block_tables.data_ptr(), table_stride, context_lens.data_ptr(), B, Hq, Hkv, D, bs,
            max_blocks, float(D) ** -0.5 if scale is None else float(scale), int(max_context_len), int(num_splits),
            ws_ptr, ws_bytes, dt, int(flags), _stream())
    _lib.check(rc, "b200_rope_append_decode_attn")
    return out


def attn_prefix_plan(B: int, Hq: int, Hkv: int, D: int, block_size: int, prefix_blocks: int, max_context_len: int,
                     dtype: torch.dtype):
    """``(prefix_splits, suffix_splits)`` the library would choose for the shared-prefix variant."""
    import ctypes

    ps, ss = ctypes.c_int(0), ctypes.c_int(0)
    rc = _lib.load().b200_paged_decode_attn_prefix_plan(B, Hq, Hkv, D, block_size, prefix_blocks, max_context_len,
                                                       DTYPE_ID[dtype], ctypes.byref(ps), ctypes.byref(ss))
    _lib.check(rc, "b200_paged_decode_attn_prefix_plan")
    return ps.value, ss.value


def paged_decode_attention_prefix(
    q: torch.Tensor,
    k_pool: torch.Tensor,
    v_pool: torch.Tensor,
    block_tables: torch.Tensor,
    context_lens: torch.Tensor,
    max_context_len: int,
    prefix_blocks: int,
    scale: Optional[float] = None,
    out: Optional[torch.Tensor] = None,
    prefix_splits: int = 0,
    suffix_splits: int = 0,
    workspace: Optional[torch.Tensor] = None,
    flags: int = 0,
) -> torch.Tensor:
    """``paged_decode_attention`` for a batch whose first ``prefix_blocks`` table entries are shared by all
    sequences: the shared blocks are read from HBM once for the whole batch."""
    (B, Hq, Hkv, D, bs, max_blocks), q_stride, table_stride, out = _attn_args(q, k_pool, v_pool, block_tables,
                                                                               context_lens, out)
    dt = _dtype_id(k_pool)
    if prefix_splits == 0 or suffix_splits == 0:
        ps, ss = attn_prefix_plan(B, Hq, Hkv, D, bs, prefix_blocks, max_context_len, q.dtype)
        prefix_splits = prefix_splits or ps
        suffix_splits = suffix_splits or ss
    lib = _lib.load()
    ws_ptr, ws_bytes, _ws = _split_workspace(lib, B, Hq, D, prefix_splits + suffix_splits, workspace, q.device,
                                             always=True)
    with _timed():
        rc = lib.b200_paged_decode_attn_prefix(
            out.data_ptr(), q.data_ptr(), q_stride, k_pool.data_ptr(), v_pool.data_ptr(), block_tables.data_ptr(),
            table_stride, context_lens.data_ptr(), B, Hq, Hkv, D, bs, max_blocks,
            float(D) ** -0.5 if scale is None else float(scale), int(max_context_len), int(prefix_blocks),
            int(prefix_splits), int(suffix_splits), ws_ptr, ws_bytes, dt, int(flags), _stream())
    _lib.check(rc, "b200_paged_decode_attn_prefix")
    return out


def rope_append_decode_attention_prefix(
    q: torch.Tensor,
    k_new: torch.Tensor,
    v_new: torch.Tensor,
    cos: torch.Tensor,
    sin: torch.Tensor,
    k_pool: torch.Tensor,
    v_pool: torch.Tensor,
    block_tables: torch.Tensor,
    context_lens: torch.Tensor,
    max_context_len: int,
    prefix_blocks: int,
    scale: Optional[float] = None,
    out: Optional[torch.Tensor] = None,
    prefix_splits: int = 0,
    suffix_splits: int = 0,
    workspace: Optional[torch.Tensor] = None,
    flags: int = 0,
) -> torch.Tensor:
    """``rope_append_decode_attention`` for a batch that shares its first ``prefix_blocks`` blocks: RoPE, the append
    of the new token and the cascade attention in one launch (+ the combine).  Every sequence must already be past
    the shared prefix (shared blocks are never written)."""
    _need_cuda(k_new, v_new, cos, sin)
    (B, Hq, Hkv, D, bs, max_blocks), q_stride, table_stride, out = _attn_args(q, k_pool, v_pool, block_tables,
                                                                               context_lens, out)
    if cos.shape != (B, D) or sin.shape != (B, D) or not cos.is_contiguous() or not sin.is_contiguous():
        raise ValueError("cos / sin must be contiguous [B, D]")
    if not (k_new.dtype == v_new.dtype == cos.dtype == sin.dtype == q.dtype):
        raise TypeError("q, k_new, v_new, cos and sin must share the pool dtype")
    if k_new.shape[0] != B or v_new.shape[0] != B:
        raise ValueError("k_new / v_new must hold one token per sequence")
    dt = _dtype_id(k_pool)
    if prefix_splits == 0 or suffix_splits == 0:
        ps, ss = attn_prefix_plan(B, Hq, Hkv, D, bs, prefix_blocks, max_context_len, q.dtype)
        prefix_splits = prefix_splits or ps
        suffix_splits = suffix_splits or ss
    lib = _lib.load()
    ws_ptr, ws_bytes, _ws = _split_workspace(lib, B, Hq, D, prefix_splits + suffix_splits, workspace, q.device,
                                             always=True)
    with _timed():
        rc = lib.b200_rope_append_decode_attn_prefix(
            out.data_ptr(), q.data_ptr(), q_stride, k_new.data_ptr(), _row_stride(k_new, Hkv, D), v_new.data_ptr(),
            _row_stride(v_new, Hkv, D), cos.data_ptr(), sin.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(),
            block_tables.data_ptr(), table_stride, context_lens.data_ptr(), B, Hq, Hkv, D, bs, max_blocks,
            float(D) ** -0.5 if scale is None else float(scale), int(max_context_len), int(prefix_blocks),
            int(prefix_splits), int(suffix_splits), ws_ptr, ws_bytes, dt, int(flags), _stream())
    _lib.check(rc, "b200_rope_append_decode_attn_prefix")
    return out


STATIC_SCRATCH_BYTES = 8 * 1024 * 1024


def paged_decode_attention_static(
    q: torch.Tensor,
    k_pool: torch.Tensor,
    v_pool: torch.Tensor,
    block_tables: torch.Tensor,
    context_lens: torch.Tensor,
    max_context_len: int,
    out: torch.Tensor,
    scale: Optional[float] = None,
    num_splits: int = 0,
    flags: int = 0,
) -> torch.Tensor:
    """``paged_decode_attention`` for callers that may not allocate: nothing is allocated here or below.

    Meant for code that only holds pointers it was given — an extension living inside another serving runtime that
    bans allocation on its hot path, or a binding with no tensor library behind it.  ``out`` must be supplied; the
    split-K partials go to the library's built-in 8 MiB scratch array instead of a workspace tensor.  Shapes whose
    partials exceed that raise; the first call per device performs a one-time address look-up, so issue it before
    any stream capture; calls that overlap on the GPU (different streams) race on the shared array.
    """
    if out is None:
        raise ValueError("out must be supplied: this entry point allocates nothing")
    # the same argument checks as every other attention wrapper; with `out` given they allocate nothing
    (B, Hq, Hkv, D, bs, max_blocks), q_stride, table_stride, out = _attn_args(q, k_pool, v_pool, block_tables,
                                                                               context_lens, out)
    rc = _lib.load().b200_paged_decode_attn_static(
        out.data_ptr(), q.data_ptr(), q_stride, k_pool.data_ptr(), v_pool.data_ptr(), block_tables.data_ptr(),
        table_stride, context_lens.data_ptr(), B, Hq, Hkv, D, bs, max_blocks,
        float(D) ** -0.5 if scale is None else float(scale), int(max_context_len), int(num_splits), _dtype_id(k_pool),
        int(flags), _stream(),
    )
    _lib.check(rc, "b200_paged_decode_attn_static")
    return out


def _check_table(block_table: torch.Tensor) -> int:
    if block_table.dtype != torch.int32 or block_table.dim() != 1 or not block_table.is_contiguous():
        raise ValueError("block_table must be a contiguous int32 vector")
    return block_table.shape[0]


def kv_gather_dense(k_pool: torch.Tensor, v_pool: torch.Tensor, block_tables: torch.Tensor, n_tokens: int):
    """The first ``n_tokens`` tokens of every row of ``block_tables`` ``[B, >= ceil(n_tokens / bs)]`` as dense
    ``(k, v)``, each ``[B, Hkv, n_tokens, D]``, in one launch (bit-exact copy out of the pools)."""
    _need_cuda(k_pool, v_pool, block_tables)
    max_blocks, Hkv, bs, D = _pool_dims(k_pool)
    _pool_dims(v_pool)
    if block_tables.dtype != torch.int32 or block_tables.dim() != 2 or (block_tables.shape[1] > 1 and block_tables.stride(1) != 1):
        raise ValueError("block_tables must be int32 [B, max_blocks_per_seq] with contiguous rows")
    B = block_tables.shape[0]
    if n_tokens < 0 or (n_tokens + bs - 1) // bs > block_tables.shape[1]:
        raise ValueError(f"{n_tokens} tokens do not fit rows of {block_tables.shape[1]} blocks")
    k_out = torch.empty((B, Hkv, n_tokens, D), dtype=k_pool.dtype, device=k_pool.device)
    v_out = torch.empty_like(k_out)
    stride = block_tables.stride(0) if B > 1 else block_tables.shape[1]
    rc = _lib.load().b200_kv_gather_dense(k_out.data_ptr(), v_out.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(),
                                         block_tables.data_ptr(), stride, B, n_tokens, Hkv, D, bs, max_blocks,
                                         _dtype_id(k_pool), _stream())
    _lib.check(rc, "b200_kv_gather_dense")
    return k_out, v_out


def kv_pack(k_pool: torch.Tensor, v_pool: torch.Tensor, block_table: torch.Tensor,
            out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Gather one layer's blocks: returns ``[2, n_blocks, Hkv, bs, D]`` (K blocks, then V blocks)."""
    _need_cuda(k_pool, v_pool, block_table)
    max_blocks, Hkv, bs, D = _pool_dims(k_pool)
    _pool_dims(v_pool)
    n = _check_table(block_table)
    shape = (2, n, Hkv, bs, D)
    if out is None:
        out = torch.empty(shape, dtype=k_pool.dtype, device=k_pool.device)
    elif tuple(out.shape) != shape or not out.is_contiguous() or out.dtype != k_pool.dtype:
        raise ValueError(f"out must be contiguous {shape}")
    rc = _lib.load().b200_kv_pack(out.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), block_table.data_ptr(), n, Hkv,
                                 D, bs, max_blocks, _dtype_id(k_pool), _stream())
    _lib.check(rc, "b200_kv_pack")
    return out


def kv_unpack(packed: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor, block_table: torch.Tensor) -> None:
    """Scatter ``packed [2, n_blocks, Hkv, bs, D]`` into the pools at the slots of ``block_table``."""
    _need_cuda(packed, k_pool, v_pool, block_table)
    max_blocks, Hkv, bs, D = _pool_dims(k_pool)
    _pool_dims(v_pool)
    n = _check_table(block_table)
    if tuple(packed.shape) != (2, n, Hkv, bs, D) or not packed.is_contiguous() or packed.dtype != k_pool.dtype:
        raise ValueError("packed must be contiguous [2, n_blocks, Hkv, bs, D] of the pool dtype")
    rc = _lib.load().b200_kv_unpack(k_pool.data_ptr(), v_pool.data_ptr(), packed.data_ptr(), block_table.data_ptr(), n,
                                   Hkv, D, bs, max_blocks, _dtype_id(k_pool), _stream())
    _lib.check(rc, "b200_kv_unpack")


def pool_pointer_array(pools: Sequence[torch.Tensor]) -> torch.Tensor:
    """Device int64 vector of ``data_ptr()``s — the ``k_pools`` / ``v_pools`` argument of the *_layers calls."""
    dev = pools[0].device
    return torch.tensor([p.data_ptr() for p in pools], dtype=torch.int64).to(dev)


def kv_pack_layers(k_ptrs: torch.Tensor, v_ptrs: torch.Tensor, pool_shape, dtype: torch.dtype,
                   block_table: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Gather the blocks of every layer in one launch: ``[L, 2, n_blocks, Hkv, bs, D]``."""
    _need_cuda(k_ptrs, v_ptrs, block_table)
    max_blocks, Hkv, bs, D = pool_shape
    L = k_ptrs.shape[0]
    n = _check_table(block_table)
    shape = (L, 2, n, Hkv, bs, D)
    if out is None:
        out = torch.empty(shape, dtype=dtype, device=block_table.device)
    elif tuple(out.shape) != shape or not out.is_contiguous() or out.dtype != dtype:
        raise ValueError(f"out must be contiguous {shape}")
    rc = _lib.load().b200_kv_pack_layers(out.data_ptr(), k_ptrs.data_ptr(), v_ptrs.data_ptr(), L, block_table.data_ptr(),
                                        n, Hkv, D, bs, max_blocks, DTYPE_ID[dtype], _stream())
    _lib.check(rc, "b200_kv_pack_layers")
    return out


def kv_unpack_layers(packed: torch.Tensor, k_ptrs: torch.Tensor, v_ptrs: torch.Tensor, pool_shape,
                     block_table: torch.Tensor) -> None:
    _need_cuda(packed, k_ptrs, v_ptrs, block_table)
    max_blocks, Hkv, bs, D = pool_shape
    L = k_ptrs.shape[0]
    n = _check_table(block_table)
    if tuple(packed.shape) != (L, 2, n, Hkv, bs, D) or not packed.is_contiguous():
        raise ValueError("packed must be contiguous [L, 2, n_blocks, Hkv, bs, D]")
    rc = _lib.load().b200_kv_unpack_layers(k_ptrs.data_ptr(), v_ptrs.data_ptr(), L, packed.data_ptr(),
                                          block_table.data_ptr(), n, Hkv, D, bs, max_blocks, _dtype_id(packed), _stream())
    _lib.check(rc, "b200_kv_unpack_layers")


def kv_move_blocks_layers(dst_k_ptrs: torch.Tensor, dst_v_ptrs: torch.Tensor, dst_pool_shape,
                          src_k_ptrs: torch.Tensor, src_v_ptrs: torch.Tensor, src_pool_shape, dtype: torch.dtype,
                          src_table: torch.Tensor, dst_table: torch.Tensor) -> None:
    """Copy block ``src_table[j]`` of every source layer to block ``dst_table[j]`` of the matching destination layer.

    The ``*_ptrs`` tensors are device int64 vectors of pool addresses (``pool_pointer_array``); destination
    addresses may point into another GPU's memory (peer-mapped), which turns the call into a push over NVLink.
    """
    _need_cuda(dst_k_ptrs, dst_v_ptrs, src_k_ptrs, src_v_ptrs, src_table, dst_table)
    n = _check_table(src_table)
    if _check_table(dst_table) != n:
        raise ValueError("source and destination tables differ in length")
    smb, Hkv, bs, D = src_pool_shape
    dmb = dst_pool_shape[0]
    if tuple(dst_pool_shape[1:]) != (Hkv, bs, D):
        raise ValueError("source and destination blocks differ in shape")
    L = src_k_ptrs.shape[0]
    if dst_k_ptrs.shape[0] != L:
        raise ValueError("layer counts differ")
    rc = _lib.load().b200_kv_move_blocks_layers(dst_k_ptrs.data_ptr(), dst_v_ptrs.data_ptr(), src_k_ptrs.data_ptr(),
                                               src_v_ptrs.data_ptr(), L, src_table.data_ptr(), dst_table.data_ptr(), n,
                                               Hkv, D, bs, smb, dmb, DTYPE_ID[dtype], _stream())
    _lib.check(rc, "b200_kv_move_blocks_layers")


def kv_pack_layers_wire(k_ptrs: torch.Tensor, v_ptrs: torch.Tensor, pool_shape, block_table: torch.Tensor,
                        out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """bf16 pools → float16 ``[L, 2, n_blocks, Hkv, bs, D]``: the reference serializer's bytes for these blocks."""
    _need_cuda(k_ptrs, v_ptrs, block_table)
    max_blocks, Hkv, bs, D = pool_shape
    L = k_ptrs.shape[0]
    n = _check_table(block_table)
    shape = (L, 2, n, Hkv, bs, D)
    if out is None:
        out = torch.empty(shape, dtype=torch.float16, device=block_table.device)
    elif tuple(out.shape) != shape or out.dtype != torch.float16 or not out.is_contiguous():
        raise ValueError(f"out must be contiguous float16 {shape}")
    rc = _lib.load().b200_kv_pack_layers_wire(out.data_ptr(), k_ptrs.data_ptr(), v_ptrs.data_ptr(), L,
                                             block_table.data_ptr(), n, Hkv, D, bs, max_blocks, _stream())
    _lib.check(rc, "b200_kv_pack_layers_wire")
    return out


def kv_unpack_layers_wire(packed: torch.Tensor, k_ptrs: torch.Tensor, v_ptrs: torch.Tensor, pool_shape,
                          block_table: torch.Tensor) -> None:
    """float16 wire payload → bf16 pools at the slots of ``block_table``."""
    _need_cuda(packed, k_ptrs, v_ptrs, block_table)
    max_blocks, Hkv, bs, D = pool_shape
    L = k_ptrs.shape[0]
    n = _check_table(block_table)
    if tuple(packed.shape) != (L, 2, n, Hkv, bs, D) or packed.dtype != torch.float16 or not packed.is_contiguous():
        raise ValueError("packed must be contiguous float16 [L, 2, n_blocks, Hkv, bs, D]")
    rc = _lib.load().b200_kv_unpack_layers_wire(k_ptrs.data_ptr(), v_ptrs.data_ptr(), L, packed.data_ptr(),
                                               block_table.data_ptr(), n, Hkv, D, bs, max_blocks, _stream())
    _lib.check(rc, "b200_kv_unpack_layers_wire")
