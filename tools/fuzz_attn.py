#!/usr/bin/env python
"""Randomised cross-check of every attention entry point against the oracle and against each other.

For each random case (batch, head counts, head_dim, ragged contexts, optional shared prefix, split counts, dtype):
  plain  b200_paged_decode_attn                (TMA kernel and generic kernel)  vs  oracle fp32 formula
  fused  b200_rope_append_decode_attn          == rope_kv_append + plain        (bit-identical outputs and pools)
  prefix b200_paged_decode_attn_prefix         vs  oracle
  fused  b200_rope_append_decode_attn_prefix   == rope_kv_append + prefix       (bit-identical)
usage: python tools/fuzz_attn.py [--cases N] [--seed S]
"""
import argparse
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import dgi_b200  # noqa: E402,F401
from dgi_b200 import ops  # noqa: E402
from oracle import paged_ref as R  # noqa: E402

DEV = "cuda:0"


def rel_err(got, want):
    num = (got.float() - want.float()).abs().amax(dim=-1)
    den = want.float().abs().amax(dim=-1).clamp_min(1e-6)
    return (num / den).max().item()


def bits(t):
    return t.contiguous().view(torch.uint8)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=200)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    rng = random.Random(a.seed)
    worst = {"plain": 0.0, "generic": 0.0, "prefix": 0.0}
    n_fused = n_prefix = 0
    for case in range(a.cases):
        dtype = rng.choice([torch.bfloat16, torch.bfloat16, torch.float16])
        D = rng.choice([64, 128])
        Hkv = rng.choice([1, 2, 4, 8])
        G = rng.choice([1, 2, 4, 8])
        Hq = Hkv * G
        B = rng.choice([1, 2, 3, 5, 8, 17, 40])  # 40 x Hkv 8 = 320 suffix items: the one-item-per-warp role
        bs = 16
        use_prefix = B > 1 and rng.random() < 0.5
        prefix_blocks = rng.choice([1, 3, 16, 40]) if use_prefix else 0
        tails = [rng.choice([1, 2, 15, 16, 17, 100, 333, 700]) for _ in range(B)]
        ctx = [prefix_blocks * bs + t for t in tails]
        n_blocks = [(c + bs - 1) // bs for c in ctx]
        max_blocks = prefix_blocks + sum(max(0, n - prefix_blocks) for n in n_blocks) + rng.randint(1, 9)
        g = torch.Generator(device=DEV).manual_seed(rng.randrange(1 << 30))
        k_pool = torch.randn((max_blocks, Hkv, bs, D), generator=g, device=DEV).to(dtype)
        v_pool = torch.randn((max_blocks, Hkv, bs, D), generator=g, device=DEV).to(dtype)
        tables, _ = R.build_block_tables(n_blocks, max_blocks, rng.randrange(1 << 30), prefix_blocks)
        tables = tables.to(DEV)
        lens = torch.tensor(ctx, dtype=torch.int32, device=DEV)
        big = torch.randn((B, (Hq + 2 * Hkv) * D), generator=g, device=DEV).to(dtype)
        q = big[:, :Hq * D].view(B, Hq, D)
        k_new = big[:, Hq * D:(Hq + Hkv) * D].view(B, Hkv, D)
        v_new = big[:, (Hq + Hkv) * D:].view(B, Hkv, D)
        pos = (lens - 1).to(torch.int64)
        inv_freq = 1.0 / (500000.0 ** (torch.arange(0, D, 2, device=DEV).float() / D))
        emb = torch.cat((pos.float()[:, None] * inv_freq[None],) * 2, -1)
        cos, sin = emb.cos().to(dtype), emb.sin().to(dtype)
        slots = tables.long().gather(1, (pos // bs)[:, None]).squeeze(1) * bs + pos % bs
        splits = rng.choice([0, 1, 2, 5])
        tol = 1e-2 if dtype == torch.bfloat16 else 2e-3
        tag = f"case {case}: B={B} Hq={Hq} Hkv={Hkv} D={D} ctx={ctx} prefix={prefix_blocks} splits={splits} {dtype}"

        # reference pipeline: separate RoPE + append, then attention
        k_ref, v_ref, q_ref = k_pool.clone(), v_pool.clone(), q.clone()
        ops.rope_kv_append(q_ref, k_new, v_new, cos, sin, k_ref, v_ref, slots)
        exact = R.paged_attention_ref(q_ref, k_ref, v_ref, tables, lens, mode="fp32")
        plain = ops.paged_decode_attention(q_ref, k_ref, v_ref, tables, lens, max(ctx), num_splits=splits)
        gen = ops.paged_decode_attention(q_ref, k_ref, v_ref, tables, lens, max(ctx), num_splits=splits,
                                         flags=ops.ATTN_FORCE_GENERIC)
        e1, e2 = rel_err(plain, exact), rel_err(gen, exact)
        assert e1 <= tol and e2 <= tol, f"{tag}: plain {e1:.2e} generic {e2:.2e}"
        worst["plain"], worst["generic"] = max(worst["plain"], e1), max(worst["generic"], e2)
        if ops.fused_decode_supported(k_pool, Hq):
            kf, vf = k_pool.clone(), v_pool.clone()
            fused = ops.rope_append_decode_attention(q, k_new, v_new, cos, sin, kf, vf, tables, lens, max(ctx),
                                                     num_splits=splits)
            assert torch.equal(bits(fused), bits(plain)), f"{tag}: fused != separate"
            assert torch.equal(bits(kf), bits(k_ref)) and torch.equal(bits(vf), bits(v_ref)), f"{tag}: fused pools differ"
            n_fused += 1
            if prefix_blocks:
                ps, ss = rng.choice([0, 1, 3]), rng.choice([0, 1, 2])
                pre = ops.paged_decode_attention_prefix(q_ref, k_ref, v_ref, tables, lens, max(ctx), prefix_blocks,
                                                        prefix_splits=ps, suffix_splits=ss)
                e3 = rel_err(pre, exact)
                assert e3 <= tol, f"{tag}: prefix {e3:.2e} (ps={ps} ss={ss})"
                worst["prefix"] = max(worst["prefix"], e3)
                kf, vf = k_pool.clone(), v_pool.clone()
                fpre = ops.rope_append_decode_attention_prefix(q, k_new, v_new, cos, sin, kf, vf, tables, lens, max(ctx),
                                                               prefix_blocks, prefix_splits=ps, suffix_splits=ss)
                assert torch.equal(bits(fpre), bits(pre)), f"{tag}: fused prefix != separate (ps={ps} ss={ss})"
                assert torch.equal(bits(kf), bits(k_ref)) and torch.equal(bits(vf), bits(v_ref)), f"{tag}: fused prefix pools"
                n_prefix += 1
    torch.cuda.synchronize()
    print(f"{a.cases} cases ok (seed {a.seed}): fused checked {n_fused}, shared-prefix checked {n_prefix}; worst error vs "
          f"exact: plain {worst['plain']:.2e}, generic {worst['generic']:.2e}, prefix {worst['prefix']:.2e}")


if __name__ == "__main__":
    main()
