"""Oracle (test infrastructure): torch restatement of the paged-KV device ops on whatever device the inputs are.

The reference has no paged append / attention / pack (last paragraph); it only fixes the pool
layout ``[max_blocks, heads, block_size, head_dim]`` (worker/distributed/kv_cache.py:130-144) and copies whole
blocks in ``_promote_to_gpu`` (:447-462).  These functions define the device ops in terms of torch indexing on
that layout, which is what the CUDA kernels must reproduce:

* append  == ``pool[slot // bs, :, slot % bs, :] = new[token]``                     (bit-exact)
* pack    == ``stack(key_pool.index_select(0, table), value_pool.index_select(0, table))``   (bit-exact)
* unpack  == ``pool.index_copy_(0, table, packed[i])``                               (bit-exact)
* attention == gather the sequence's K/V through its block table into the dense ``[Hkv, ctx, D]`` cache the
  reference's HF layer would hold, then (a) the fp32 formula, (b) the eager HF recipe of oracle/llama_ref.py.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from .llama_ref import eager_attention


def append_ref(k_new: torch.Tensor, v_new: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor,
               slot_mapping: torch.Tensor) -> None:
    """k_new/v_new [n_tok, Hkv, D]; pools [max_blocks, Hkv, bs, D]; slot_mapping int64 [n_tok], negative = skip."""
    bs = k_pool.shape[2]
    keep = slot_mapping >= 0
    slots = slot_mapping[keep]
    blk, off = slots // bs, slots % bs
    k_pool[blk, :, off, :] = k_new[keep]
    v_pool[blk, :, off, :] = v_new[keep]


def pack_ref(k_pool: torch.Tensor, v_pool: torch.Tensor, block_table: torch.Tensor) -> torch.Tensor:
    idx = block_table.to(torch.int64)
    return torch.stack((k_pool.index_select(0, idx), v_pool.index_select(0, idx)), dim=0)


def unpack_ref(packed: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor, block_table: torch.Tensor) -> None:
    idx = block_table.to(torch.int64)
    k_pool.index_copy_(0, idx, packed[0])
    v_pool.index_copy_(0, idx, packed[1])


def gather_dense(pool: torch.Tensor, table_row: torch.Tensor, ctx: int) -> torch.Tensor:
    """Blocks of one sequence -> the dense [Hkv, ctx, D] cache."""
    bs = pool.shape[2]
    n = (ctx + bs - 1) // bs
    blocks = pool.index_select(0, table_row[:n].to(torch.int64))  # [n, Hkv, bs, D]
    Hkv, D = pool.shape[1], pool.shape[3]
    return blocks.permute(1, 0, 2, 3).reshape(Hkv, n * bs, D)[:, :ctx, :]


def paged_attention_ref(q: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor, block_tables: torch.Tensor,
                        context_lens: torch.Tensor, scale: Optional[float] = None, mode: str = "fp32") -> torch.Tensor:
    """q [B, Hq, D] -> [B, Hq, D].

    mode "fp32": softmax(q k^T * scale) v evaluated in fp32 from the stored values, result in fp32.
    mode "eager": the HF eager recipe in the storage dtype (scores rounded to it, fp32 softmax, cast, PV).
    """
    B, Hq, D = q.shape
    Hkv = k_pool.shape[1]
    g = Hq // Hkv
    scale = D ** -0.5 if scale is None else scale
    out_dtype = torch.float32 if mode == "fp32" else q.dtype
    out = torch.zeros((B, Hq, D), dtype=out_dtype, device=q.device)
    lens = context_lens.tolist()
    for b in range(B):
        ctx = lens[b]
        if ctx == 0:
            continue
        k = gather_dense(k_pool, block_tables[b], ctx)
        v = gather_dense(v_pool, block_tables[b], ctx)
        if mode == "fp32":
            qf = q[b].to(torch.float32).view(Hkv, g, D)
            s = torch.einsum("hgd,hsd->hgs", qf, k.to(torch.float32)) * scale
            p = torch.softmax(s, dim=-1)
            out[b] = torch.einsum("hgs,hsd->hgd", p, v.to(torch.float32)).reshape(Hq, D)
        else:
            assert abs(scale - D ** -0.5) < 1e-12
            o = eager_attention(q[b][None, :, None, :], k[None], v[None], g, None)  # [1, 1, Hq*D]
            out[b] = o.view(Hq, D)
    return out


def build_block_tables(n_blocks_per_seq: List[int], max_blocks: int, seed: int, share_prefix_blocks: int = 0
                       ) -> Tuple[torch.Tensor, int]:
    """Scattered block tables: a seeded permutation of pool slots dealt to the sequences in order.

    With ``share_prefix_blocks`` > 0 every sequence's first that-many entries are the same slots (radix-style
    prefix sharing).  Returns (int32 [B, max_n] padded with -1... no: padded with 0, and the slots used).
    """
    g = torch.Generator().manual_seed(seed)
    perm = torch.randperm(max_blocks, generator=g).tolist()
    cursor = 0
    shared = perm[cursor:cursor + share_prefix_blocks]
    cursor += share_prefix_blocks
    width = max(n_blocks_per_seq) if n_blocks_per_seq else 0
    rows = []
    for n in n_blocks_per_seq:
        own = max(0, n - share_prefix_blocks)
        row = (shared[:n] + perm[cursor:cursor + own])
        cursor += own
        if cursor > max_blocks:
            raise ValueError("pool too small for the requested tables")
        rows.append(row + [0] * (width - len(row)))
    return torch.tensor(rows, dtype=torch.int32).reshape(len(rows), width), cursor


def golden_case_inputs(dtype: torch.dtype):
    """The seeded case behind tests/golden/paged_ops.json (CPU RNG, shared by oracle/make_golden.py and the tests):
    ragged contexts incl. an empty sequence, a single token and an exact block multiple; scattered block tables."""
    B, Hq, Hkv, D, bs, max_blocks = 5, 8, 2, 64, 16, 24
    ctx = [37, 0, 1, 48, 90]
    g = torch.Generator().manual_seed(20240607)
    k_pool = torch.randn((max_blocks, Hkv, bs, D), generator=g).to(dtype)
    v_pool = torch.randn((max_blocks, Hkv, bs, D), generator=g).to(dtype)
    q = torch.randn((B, Hq, D), generator=g).to(dtype)
    tables, _ = build_block_tables([(c + bs - 1) // bs for c in ctx], max_blocks, seed=11)
    return q, k_pool, v_pool, tables, torch.tensor(ctx, dtype=torch.int32)
