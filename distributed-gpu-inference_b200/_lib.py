"""ctypes loader for libb200_kv.so — the C-ABI library declared in include/b200_kv.h.

There is no CPU or pure-torch fallback: if the library is missing or does not load, every op raises.
Build it with ``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C <package>/csrc``.
"""
from __future__ import annotations

import ctypes
import os
import threading
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_uint, c_void_p

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG_DIR, "libb200_kv.so")

# name -> (restype, argtypes); mirrors include/b200_kv.h one to one
SIGNATURES = {
    "b200_abi_version": (c_int, []),
    "b200_last_error": (c_char_p, []),
    "b200_kv_append": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_int, c_void_p],
    ),
    "b200_paged_decode_attn": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int,
         c_int64, c_float, c_int, c_int, c_void_p, c_size_t, c_int, c_uint, c_void_p],
    ),
    "b200_rope_append_decode_attn": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int64, c_float, c_int, c_int, c_void_p, c_size_t,
         c_int, c_uint, c_void_p],
    ),
    "b200_paged_decode_attn_prefix": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int,
         c_int64, c_float, c_int, c_int, c_int, c_int, c_void_p, c_size_t, c_int, c_uint, c_void_p],
    ),
    "b200_rope_append_decode_attn_prefix": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int64, c_float, c_int, c_int, c_int, c_int,
         c_void_p, c_size_t, c_int, c_uint, c_void_p],
    ),
    "b200_paged_decode_attn_prefix_plan": (
        c_int, [c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "b200_paged_decode_attn_static": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int,
         c_int64, c_float, c_int, c_int, c_int, c_uint, c_void_p],
    ),
    "b200_paged_decode_attn_workspace": (c_size_t, [c_int, c_int, c_int, c_int]),
    "b200_paged_decode_attn_plan": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int, c_int]),
    "b200_kv_pack_layers_wire": (
        c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_void_p]),
    "b200_kv_unpack_layers_wire": (
        c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_void_p]),
    "b200_kv_move_blocks_layers": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_int64,
         c_int, c_void_p],
    ),
    "b200_add_rmsnorm": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_int, c_int, c_int64, c_int64, c_int, c_void_p]),
    "b200_rope_kv_append": (
        c_int,
        [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
         c_int, c_int, c_int, c_int, c_int64, c_int, c_void_p],
    ),
    "b200_rope_kv_append_prompt": (
        c_int,
        [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
         c_int, c_int, c_int, c_int, c_int64, c_int, c_void_p],
    ),
    "b200_argmax_rows": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_void_p]),
    "b200_argmax_rows_keys": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int, c_void_p]),
    "b200_argmax_keys_decode": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "b200_silu_mul": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int64, c_int64, c_int, c_void_p]),
    "b200_kv_gather_dense": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int64, c_int,
         c_void_p],
    ),
    "b200_kv_pack": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_int, c_void_p]),
    "b200_kv_unpack": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_int, c_void_p]),
    "b200_kv_pack_layers": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_int, c_void_p],
    ),
    "b200_kv_unpack_layers": (
        c_int,
        [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_int, c_void_p],
    ),
}

_lock = threading.Lock()
_lib = None


class B200LibraryError(RuntimeError):
    """The CUDA extension is missing, failed to load, or an entry point reported an error."""


def load() -> ctypes.CDLL:
    """Load (once) and return the library with argument types set.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise B200LibraryError(
                f"{LIB_PATH} not found: the sm_100a extension is not built (run __graft_entry__.build()). "
                "There is no CPU fallback for the paged-KV ops."
            )
        try:
            lib = ctypes.CDLL(LIB_PATH)
        except OSError as exc:
            raise B200LibraryError(f"cannot load {LIB_PATH}: {exc}") from exc
        for name, (restype, argtypes) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError here means header and library drifted apart
            fn.restype = restype
            fn.argtypes = argtypes
        if lib.b200_abi_version() != 1:
            raise B200LibraryError(f"ABI version {lib.b200_abi_version()} != 1")
        _lib = lib
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().b200_last_error()
        raise B200LibraryError(f"{what} failed (code {rc}): {msg.decode() if msg else ''}")
