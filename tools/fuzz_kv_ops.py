#!/usr/bin/env python
"""Randomised bit-exact cross-check of the KV data-movement kernels against torch indexing (the oracle's definitions):
append, RoPE + append (decode and prompt forms), pack / unpack (one layer and all layers), gather to dense, block move,
wire pack / unpack.
Odd shapes on purpose (head_dim 16..256 incl. non-multiples of 64, block sizes 4..32, any head count, three dtypes,
skipped slots, strided source rows).   usage: python tools/fuzz_kv_ops.py [--cases N] [--seed S]
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


def bits(t):
    return t.contiguous().view(torch.uint8)


def rot_half(x):
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=300)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    rng = random.Random(a.seed)
    counts = {"append": 0, "rope": 0, "pack": 0, "gather": 0, "layers": 0, "move": 0, "wire": 0}
    for case in range(a.cases):
        dtype = rng.choice([torch.bfloat16, torch.float16, torch.float32])
        D = rng.choice([16, 32, 64, 80, 96, 128, 256])
        Hkv = rng.choice([1, 2, 3, 8])
        Hq = Hkv * rng.choice([1, 2, 4])
        bs = rng.choice([4, 8, 16, 32])
        L_ = rng.choice([1, 2, 5])
        max_blocks = rng.randint(6, 40)
        n_tok = rng.choice([1, 2, 7, 33, 64, 200])
        g = torch.Generator(device=DEV).manual_seed(rng.randrange(1 << 30))
        tag = f"case {case}: {dtype} D={D} Hkv={Hkv} Hq={Hq} bs={bs} L={L_} max_blocks={max_blocks} n_tok={n_tok}"
        backing = torch.randn((L_, 2, max_blocks, Hkv, bs, D), generator=g, device=DEV).to(dtype)
        n_tok = min(n_tok, max_blocks * bs)
        slots = torch.randperm(max_blocks * bs, generator=g, device=DEV)[:n_tok].to(torch.int64)
        if n_tok > 2:
            slots[rng.randrange(n_tok)] = -1  # a padded token
        big = torch.randn((n_tok, (Hq + 2 * Hkv) * D + 8), generator=g, device=DEV).to(dtype)
        q = big[:, :Hq * D].view(n_tok, Hq, D)
        k = big[:, Hq * D:(Hq + Hkv) * D].view(n_tok, Hkv, D)
        v = big[:, (Hq + Hkv) * D:(Hq + 2 * Hkv) * D].view(n_tok, Hkv, D)

        # append
        kp, vp = backing[0, 0].clone(), backing[0, 1].clone()
        kr, vr = kp.clone(), vp.clone()
        R.append_ref(k, v, kr, vr, slots)
        ops.kv_append(k, v, kp, vp, slots)
        assert torch.equal(bits(kp), bits(kr)) and torch.equal(bits(vp), bits(vr)), f"{tag}: append"
        counts["append"] += 1

        # RoPE + append, both forms
        es = backing.element_size()
        if (D // 2) % (16 // es) == 0:
            pos = torch.randint(0, 9000, (n_tok,), generator=g, device=DEV)
            inv_freq = 1.0 / (10000.0 ** (torch.arange(0, D, 2, device=DEV).float() / D))
            emb = torch.cat((pos.float()[:, None] * inv_freq[None],) * 2, -1)
            cos, sin = emb.cos().to(dtype), emb.sin().to(dtype)
            q_want = q * cos[:, None] + rot_half(q) * sin[:, None]
            k_want = k * cos[:, None] + rot_half(k) * sin[:, None]
            kr, vr = backing[0, 0].clone(), backing[0, 1].clone()
            R.append_ref(k_want, v, kr, vr, slots)
            for in_place in (False, True):
                b2 = big.clone()
                q2 = b2[:, :Hq * D].view(n_tok, Hq, D)
                k2 = b2[:, Hq * D:(Hq + Hkv) * D].view(n_tok, Hkv, D)
                v2 = b2[:, (Hq + Hkv) * D:(Hq + 2 * Hkv) * D].view(n_tok, Hkv, D)
                kp, vp = backing[0, 0].clone(), backing[0, 1].clone()
                ops.rope_kv_append(q2, k2, v2, cos, sin, kp, vp, slots, rotate_k_in_place=in_place)
                assert torch.equal(bits(q2), bits(q_want)), f"{tag}: rope q (in_place={in_place})"
                assert torch.equal(bits(k2), bits(k_want if in_place else k)), f"{tag}: rope k (in_place={in_place})"
                assert torch.equal(bits(kp), bits(kr)) and torch.equal(bits(vp), bits(vr)), f"{tag}: rope pools"
            counts["rope"] += 1

        # pack / unpack, one layer
        n = rng.randint(1, max_blocks)
        table = torch.randperm(max_blocks, generator=g, device=DEV)[:n].to(torch.int32)
        packed = ops.kv_pack(backing[0, 0], backing[0, 1], table)
        assert torch.equal(bits(packed), bits(R.pack_ref(backing[0, 0], backing[0, 1], table))), f"{tag}: pack"
        dk, dv = torch.zeros_like(backing[0, 0]), torch.zeros_like(backing[0, 1])
        dst = torch.randperm(max_blocks, generator=g, device=DEV)[:n].to(torch.int32)
        ops.kv_unpack(packed, dk, dv, dst)
        rk, rv = torch.zeros_like(dk), torch.zeros_like(dv)
        R.unpack_ref(packed, rk, rv, dst)
        assert torch.equal(bits(dk), bits(rk)) and torch.equal(bits(dv), bits(rv)), f"{tag}: unpack"
        counts["pack"] += 1

        # gather to dense through a table with several rows (prompt pass continuing cached tokens)
        rows_b = rng.randint(1, 3)
        per_row = max(1, min(max_blocks // rows_b, rng.randint(1, 12)))
        n_dense = rng.randint(1, per_row * bs)
        tab2 = torch.randperm(max_blocks, generator=g, device=DEV)[:rows_b * per_row].to(torch.int32).view(rows_b, per_row)
        kd, vd = ops.kv_gather_dense(backing[0, 0], backing[0, 1], tab2, n_dense)
        for b in range(rows_b):
            assert torch.equal(bits(kd[b]), bits(R.gather_dense(backing[0, 0], tab2[b], n_dense))), f"{tag}: gather k"
            assert torch.equal(bits(vd[b]), bits(R.gather_dense(backing[0, 1], tab2[b], n_dense))), f"{tag}: gather v"
        counts["gather"] += 1

        # all layers in one launch, and the block move
        shape = (max_blocks, Hkv, bs, D)
        kps = ops.pool_pointer_array([backing[l, 0] for l in range(L_)])
        vps = ops.pool_pointer_array([backing[l, 1] for l in range(L_)])
        pl = ops.kv_pack_layers(kps, vps, shape, dtype, table)
        want = torch.stack([R.pack_ref(backing[l, 0], backing[l, 1], table) for l in range(L_)])
        assert torch.equal(bits(pl), bits(want)), f"{tag}: pack_layers"
        other = torch.zeros_like(backing)
        okp = ops.pool_pointer_array([other[l, 0] for l in range(L_)])
        ovp = ops.pool_pointer_array([other[l, 1] for l in range(L_)])
        ops.kv_unpack_layers(pl, okp, ovp, shape, dst)
        assert torch.equal(bits(other[:, :, dst.long()]), bits(backing[:, :, table.long()])), f"{tag}: unpack_layers"
        rest = [i for i in range(max_blocks) if i not in set(dst.tolist())]
        assert not other[:, :, rest].any(), f"{tag}: unpack_layers touched other slots"
        counts["layers"] += 1
        moved = torch.zeros_like(backing)
        mkp = ops.pool_pointer_array([moved[l, 0] for l in range(L_)])
        mvp = ops.pool_pointer_array([moved[l, 1] for l in range(L_)])
        ops.kv_move_blocks_layers(mkp, mvp, shape, kps, vps, shape, dtype, table, dst)
        assert torch.equal(bits(moved), bits(other)), f"{tag}: move"
        counts["move"] += 1

        # wire form (bf16 pools travel as fp16)
        if dtype == torch.bfloat16:
            wire = ops.kv_pack_layers_wire(kps, vps, shape, table)
            assert torch.equal(bits(wire), bits(want.to(torch.float16))), f"{tag}: wire pack"
            back = torch.zeros_like(backing)
            bkp = ops.pool_pointer_array([back[l, 0] for l in range(L_)])
            bvp = ops.pool_pointer_array([back[l, 1] for l in range(L_)])
            ops.kv_unpack_layers_wire(wire, bkp, bvp, shape, dst)
            assert torch.equal(bits(back[:, :, dst.long()]), bits(want.to(torch.float16).to(torch.bfloat16))), f"{tag}: wire unpack"
            counts["wire"] += 1
    torch.cuda.synchronize()
    print(f"{a.cases} cases ok (seed {a.seed}): " + ", ".join(f"{k} {v}" for k, v in counts.items()))


if __name__ == "__main__":
    main()
