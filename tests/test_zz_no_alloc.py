"""GPU: the allocation-free attention entry point (b200_paged_decode_attn_static) against the regular one.

Some users of this library cannot allocate device memory at all: they sit inside another runtime's serving loop
where cudaMalloc is banned because it stalls the device, or they reach the library through a bare FFI binding and
only ever hold pointers somebody else gave them.  For them the split-K partials live in an array built into the
library.  The array is shared by every call and keeps whatever the previous call left in it, so this test makes
two different calls back to back and expects both to agree with the workspace-based entry point.

The file name sorts last on purpose: if this test fails under -x, everything else has already run.
"""
import ctypes

import pytest
import torch

import dgi_b200  # noqa: F401
from dgi_b200 import _lib, ops
from oracle import paged_ref as R

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(B, Hq, Hkv, D, ctx_lens, seed, dtype=torch.bfloat16):
    bs = 16
    n_blocks = [(c + bs - 1) // bs for c in ctx_lens]
    max_blocks = sum(n_blocks) + 4
    g = torch.Generator(device=DEV).manual_seed(seed)
    k_pool = torch.randn((max_blocks, Hkv, bs, D), generator=g, device=DEV).to(dtype)
    v_pool = torch.randn((max_blocks, Hkv, bs, D), generator=g, device=DEV).to(dtype)
    q = torch.randn((B, Hq, D), generator=g, device=DEV).to(dtype)
    tables, _ = R.build_block_tables(n_blocks, max_blocks, seed)
    return q, k_pool, v_pool, tables.to(DEV), torch.tensor(ctx_lens, dtype=torch.int32, device=DEV)


def _call_static_raw(q, k_pool, v_pool, tables, lens, max_ctx, out, splits, flags=0):
    """Straight through the C ABI with ctypes — the way a caller without torch helpers would do it."""
    lib = _lib.load()
    max_blocks, Hkv, bs, D = k_pool.shape
    B, Hq = q.shape[0], q.shape[1]
    return lib.b200_paged_decode_attn_static(
        ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(q.data_ptr()), q.stride(0), ctypes.c_void_p(k_pool.data_ptr()),
        ctypes.c_void_p(v_pool.data_ptr()), ctypes.c_void_p(tables.data_ptr()), tables.stride(0),
        ctypes.c_void_p(lens.data_ptr()), B, Hq, Hkv, D, bs, max_blocks, float(D) ** -0.5, max_ctx, splits,
        ops.DTYPE_ID[q.dtype], flags, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))


def test_static_scratch_entry_matches_workspace_entry_twice_in_a_row():
    # call 1: long contexts, 5 splits; call 2: other shapes and contents, 3 splits — stale partials of call 1
    # (more splits, more rows) are still sitting in the array when call 2 runs.
    cases = [
        (_case(4, 32, 8, 128, [2048, 1500, 40, 777], seed=101), 2048, 5),
        (_case(2, 8, 2, 64, [300, 1000], seed=202), 1000, 3),
    ]
    for (q, k_pool, v_pool, tables, lens), max_ctx, splits in cases:
        want = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, max_ctx, num_splits=splits)
        exact = R.paged_attention_ref(q, k_pool, v_pool, tables, lens, mode="fp32")
        got = torch.empty_like(want)
        rc = _call_static_raw(q, k_pool, v_pool, tables, lens, max_ctx, got, splits)
        assert rc == 0, _lib.load().b200_last_error()
        torch.cuda.synchronize()
        # same kernels, same partition: identical bits; and within the bf16 tolerance of the other GPU tests
        assert torch.equal(got.view(torch.int16), want.view(torch.int16))
        num = (got.float() - exact).abs().amax(-1)
        den = exact.abs().amax(-1).clamp_min(1e-6)
        assert (num / den).max().item() <= 1e-2
    # the python wrapper goes through the same entry point
    (q, k_pool, v_pool, tables, lens), max_ctx, splits = cases[0]
    out = torch.empty((q.shape[0], q.shape[1], q.shape[2]), dtype=q.dtype, device=DEV)
    ops.paged_decode_attention_static(q, k_pool, v_pool, tables, lens, max_ctx, out, num_splits=0)
    want = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, max_ctx, num_splits=0)
    assert torch.equal(out.view(torch.int16), want.view(torch.int16))


def test_static_scratch_entry_rejects_oversized_requests():
    # 64 sequences x 64 heads x 128 dims x 8 splits of fp32 partials = 17 MB > the 8 MiB array
    q, k_pool, v_pool, tables, lens = _case(64, 64, 8, 128, [256] * 64, seed=303)
    out = torch.empty_like(q)
    rc = _call_static_raw(q, k_pool, v_pool, tables, lens, 256, out, splits=8)
    assert rc == -3
    msg = _lib.load().b200_last_error().decode()
    assert "scratch" in msg and "8388608" in msg
    with pytest.raises(_lib.B200LibraryError, match="scratch"):
        ops.paged_decode_attention_static(q, k_pool, v_pool, tables, lens, 256, out, num_splits=8)
    # one split needs no scratch and is always accepted
    assert _call_static_raw(q, k_pool, v_pool, tables, lens, 256, out, splits=1) == 0
    torch.cuda.synchronize()
    want = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, 256, num_splits=1)
    assert torch.equal(out.view(torch.int16), want.view(torch.int16))


def test_static_scratch_entry_sizes_its_check_like_the_regular_entry_and_validates_arguments():
    # a generous split count on a short context is clamped to the number of blocks (3) before the size check, exactly
    # as the regular entry point does — 64 splits of this shape would not fit the array, 3 do
    q, k_pool, v_pool, tables, lens = _case(64, 64, 8, 128, [48] * 64, seed=404)
    out = torch.empty_like(q)
    assert _call_static_raw(q, k_pool, v_pool, tables, lens, 48, out, splits=64) == 0, _lib.load().b200_last_error()
    torch.cuda.synchronize()
    want = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, 48, num_splits=64)
    assert torch.equal(out.view(torch.int16), want.view(torch.int16))
    # the wrapper refuses mismatched arguments instead of reinterpreting them
    with pytest.raises(TypeError, match="dtype"):
        ops.paged_decode_attention_static(q.to(torch.float16), k_pool, v_pool, tables, lens, 48, out.to(torch.float16))
    with pytest.raises(ValueError, match="block_tables"):
        ops.paged_decode_attention_static(q, k_pool, v_pool, tables.t().contiguous().t(), lens, 48, out)
    with pytest.raises(ValueError, match="out must be supplied"):
        ops.paged_decode_attention_static(q, k_pool, v_pool, tables, lens, 48, None)
