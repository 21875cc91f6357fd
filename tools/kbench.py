#!/usr/bin/env python
"""Kernel micro-benchmarks on one B200: paged decode attention, append, pack/unpack as GB/s of algorithmic bytes.

Pools are rotated over `--layers` independent copies so consecutive launches never re-read L2-resident data
(one layer's KV at config 2 is 537 MB, the L2 is 126 MB).  Timing: CUDA events on the launching stream, after
warm-up, mean over the timed launches.  Prints one JSON line per case.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import dgi_b200  # noqa: E402,F401
from dgi_b200 import ops  # noqa: E402

DEV = "cuda:0"
PEAK = 6546.6
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:  # noqa: BLE001
    pass


def timed(fn, n, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(warm + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def attn_case(name, B, Hq, Hkv, D, ctx_lens, layers, splits, flags=0, share_prefix=0, iters=64, dtype=torch.bfloat16,
              cascade=False, fused=False):
    bs = 16
    nblk = [(c + bs - 1) // bs for c in ctx_lens]
    uniq_blocks = share_prefix + sum(max(0, n - share_prefix) for n in nblk)
    max_blocks = uniq_blocks + 16
    g = torch.Generator().manual_seed(7)
    perm = torch.randperm(max_blocks, generator=g)
    rows, cur = [], share_prefix
    for n in nblk:
        own = max(0, n - share_prefix)
        rows.append(torch.cat([perm[:min(n, share_prefix)], perm[cur:cur + own]]))
        cur += own
    width = max(nblk)
    tables = torch.zeros((B, width), dtype=torch.int32)
    for b, r in enumerate(rows):
        tables[b, :len(r)] = r.to(torch.int32)
    tables = tables.to(DEV)
    lens = torch.tensor(ctx_lens, dtype=torch.int32, device=DEV)
    pools = [(torch.randn((max_blocks, Hkv, bs, D), device=DEV, dtype=dtype),
              torch.randn((max_blocks, Hkv, bs, D), device=DEV, dtype=dtype)) for _ in range(layers)]
    q = torch.randn((B, Hq, D), device=DEV, dtype=dtype)
    out = torch.empty_like(q)
    max_ctx = max(ctx_lens)
    if splits == 0:
        splits = ops.attn_plan(B, Hq, Hkv, D, bs, max_ctx, dtype)
    ws = torch.empty(max(16, ops.attn_workspace_bytes(B, Hq, D, splits)), dtype=torch.uint8, device=DEV)

    if cascade:
        ps, ss = ops.attn_prefix_plan(B, Hq, Hkv, D, bs, share_prefix, max_ctx, dtype)
        ps = int(os.environ.get("KBENCH_PSPLITS", ps))
        ss = int(os.environ.get("KBENCH_SSPLITS", ss))
        splits = f"{ps}+{ss}"
        ws = torch.zeros(max(16, ops.attn_workspace_bytes(B, Hq, D, ps + ss)), dtype=torch.uint8, device=DEV)
        if not os.environ.get("KBENCH_NO_WSFLAG"):
            flags |= ops.ATTN_WS_COUNTERS_ZERO  # zero-filled workspace of our own: no per-call clearing of the counters

    if fused:  # the launch the decode step actually uses: RoPE(q, k_new) + append + attention
        kv_new = torch.randn((B, (Hq + 2 * Hkv) * D), device=DEV, dtype=dtype)
        q = kv_new[:, :Hq * D].view(B, Hq, D)  # strided rows, as slices of the fused q|k|v projection
        k_new = kv_new[:, Hq * D:(Hq + Hkv) * D].view(B, Hkv, D)
        v_new = kv_new[:, (Hq + Hkv) * D:].view(B, Hkv, D)
        cs = torch.randn((B, D), device=DEV, dtype=dtype)
        sn = torch.randn((B, D), device=DEV, dtype=dtype)

    def run(i):
        k, v = pools[i % layers]
        if fused and cascade:
            ops.rope_append_decode_attention_prefix(q, k_new, v_new, cs, sn, k, v, tables, lens, max_ctx, share_prefix,
                                                    out=out, workspace=ws, flags=flags, prefix_splits=ps, suffix_splits=ss)
        elif fused:
            ops.rope_append_decode_attention(q, k_new, v_new, cs, sn, k, v, tables, lens, max_ctx, out=out,
                                             num_splits=splits, workspace=ws, flags=flags)
        elif cascade:
            ops.paged_decode_attention_prefix(q, k, v, tables, lens, max_ctx, share_prefix, out=out, workspace=ws, flags=flags,
                                              prefix_splits=ps, suffix_splits=ss)
        else:
            ops.paged_decode_attention(q, k, v, tables, lens, max_ctx, out=out, num_splits=splits, workspace=ws, flags=flags)

    ms = timed(run, iters)
    es = 2
    logical = sum(ctx_lens) * 2 * Hkv * D * es + 2 * B * Hq * D * es
    unique = uniq_blocks * 2 * Hkv * bs * D * es + 2 * B * Hq * D * es
    gbs = (unique if share_prefix else logical) / ms / 1e6
    print(json.dumps({"kernel": "paged_decode_attn", "case": name, "B": B, "Hq": Hq, "Hkv": Hkv, "D": D,
                      "ctx_max": max_ctx, "splits": splits, "flags": flags, "ms": round(ms, 5),
                      "logical_bytes": logical, "unique_bytes": unique, "GBps": round(gbs, 1),
                      "frac_of_measured_peak": round(gbs / PEAK, 4)}), flush=True)
    del pools


def copy_cases():
    # pack / unpack at config 3: one 4k-token sequence of the 8B shape = 256 blocks x 32 layers x 64 KiB = 512 MiB
    L_, Hkv, D, bs, n = 32, 8, 128, 16, 256
    max_blocks = 1024
    backing = torch.randn((L_, 2, max_blocks, Hkv, bs, D), device=DEV, dtype=torch.bfloat16)
    kp = ops.pool_pointer_array([backing[l, 0] for l in range(L_)])
    vp = ops.pool_pointer_array([backing[l, 1] for l in range(L_)])
    tables = [torch.randperm(max_blocks, generator=torch.Generator().manual_seed(s))[:n].to(torch.int32).to(DEV) for s in range(4)]
    packed = torch.empty((L_, 2, n, Hkv, bs, D), dtype=torch.bfloat16, device=DEV)
    shape = (max_blocks, Hkv, bs, D)
    payload = packed.numel() * 2
    ms = timed(lambda i: ops.kv_pack_layers(kp, vp, shape, torch.bfloat16, tables[i % 4], out=packed), 20)
    print(json.dumps({"kernel": "kv_pack_layers", "payload_MiB": payload / 2 ** 20, "ms": round(ms, 4),
                      "GBps_read_plus_write": round(2 * payload / ms / 1e6, 1),
                      "frac_of_measured_peak": round(2 * payload / ms / 1e6 / PEAK, 4)}), flush=True)
    ms = timed(lambda i: ops.kv_unpack_layers(packed, kp, vp, shape, tables[i % 4]), 20)
    print(json.dumps({"kernel": "kv_unpack_layers", "payload_MiB": payload / 2 ** 20, "ms": round(ms, 4),
                      "GBps_read_plus_write": round(2 * payload / ms / 1e6, 1),
                      "frac_of_measured_peak": round(2 * payload / ms / 1e6 / PEAK, 4)}), flush=True)
    a, b_ = torch.empty(payload, dtype=torch.uint8, device=DEV), torch.empty(payload, dtype=torch.uint8, device=DEV)
    ms = timed(lambda i: b_.copy_(a), 20)
    print(json.dumps({"kernel": "torch_copy_same_bytes", "ms": round(ms, 4), "GBps_read_plus_write": round(2 * payload / ms / 1e6, 1)}), flush=True)
    # host tier: the same sequence out to pinned memory and back (pack + D2H, H2D + unpack); bound = the host link
    host = torch.empty(packed.shape, dtype=packed.dtype, pin_memory=True)

    def out_(i):
        ops.kv_pack_layers(kp, vp, shape, torch.bfloat16, tables[i % 4], out=packed)
        host.copy_(packed, non_blocking=True)

    def in_(i):
        packed.copy_(host, non_blocking=True)
        ops.kv_unpack_layers(packed, kp, vp, shape, tables[i % 4])

    for name, fn in (("host_swap_out", out_), ("host_swap_in", in_)):
        ms = timed(fn, 6, warm=2)
        print(json.dumps({"kernel": name, "payload_MiB": payload / 2 ** 20, "ms": round(ms, 3),
                          "GBps_payload": round(payload / ms / 1e6, 1)}), flush=True)
    # append at decode (64 tokens) and at prefill (8 x 2048 tokens)
    for n_tok in (64, 16384):
        k_new = torch.randn((n_tok, Hkv, D), device=DEV, dtype=torch.bfloat16)
        v_new = torch.randn((n_tok, Hkv, D), device=DEV, dtype=torch.bfloat16)
        slots = torch.randperm(max_blocks * bs, device=DEV)[:n_tok].to(torch.int64)
        ms = timed(lambda i: ops.kv_append(k_new, v_new, backing[i % L_, 0], backing[i % L_, 1], slots), 50)
        byt = 2 * (2 * n_tok * Hkv * D * 2) + 8 * n_tok
        print(json.dumps({"kernel": "kv_append", "n_tok": n_tok, "ms": round(ms, 5), "GBps": round(byt / ms / 1e6, 1)}), flush=True)


def append_cases(iters):
    """k1 alone: the paged KV append of a 16k-token prompt chunk (8B shape), and the prompt-pass launch that does RoPE
    on q and k plus that append (b200_rope_kv_append_prompt)."""
    Hq, Hkv, D, bs, max_blocks, n_tok, L_ = 32, 8, 128, 16, 4096, 16384, 4
    pools = [(torch.zeros((max_blocks, Hkv, bs, D), device=DEV, dtype=torch.bfloat16),
              torch.zeros((max_blocks, Hkv, bs, D), device=DEV, dtype=torch.bfloat16)) for _ in range(L_)]
    qkv = torch.randn((n_tok, (Hq + 2 * Hkv) * D), device=DEV, dtype=torch.bfloat16)  # strided rows, as in the layer
    q = qkv[:, :Hq * D].view(n_tok, Hq, D)
    k_new = qkv[:, Hq * D:(Hq + Hkv) * D].view(n_tok, Hkv, D)
    v_new = qkv[:, (Hq + Hkv) * D:].view(n_tok, Hkv, D)
    slots = torch.randperm(max_blocks * bs, device=DEV)[:n_tok].to(torch.int64)
    cos = torch.randn((n_tok, D), device=DEV, dtype=torch.bfloat16)
    sin = torch.randn((n_tok, D), device=DEV, dtype=torch.bfloat16)
    es = 2
    ms = timed(lambda i: ops.kv_append(k_new, v_new, pools[i % L_][0], pools[i % L_][1], slots), iters)
    byt = 2 * (2 * n_tok * Hkv * D * es) + 8 * n_tok  # K and V rows read + written, + the slot mapping
    print(json.dumps({"kernel": "kv_append", "n_tok": n_tok, "ms": round(ms, 5), "algorithmic_bytes": byt,
                      "GBps": round(byt / ms / 1e6, 1), "frac_of_measured_peak": round(byt / ms / 1e6 / PEAK, 4)}), flush=True)
    ms = timed(lambda i: ops.rope_kv_append(q, k_new, v_new, cos, sin, pools[i % L_][0], pools[i % L_][1], slots,
                                            rotate_k_in_place=True), iters)
    # q read + written, k read + written in place + written to the pool, v read + written to the pool, cos / sin, slots
    byt = (2 * n_tok * Hq * D + 3 * n_tok * Hkv * D + 2 * n_tok * Hkv * D + 2 * n_tok * D) * es + 8 * n_tok
    print(json.dumps({"kernel": "rope_kv_append_prompt", "n_tok": n_tok, "ms": round(ms, 5), "algorithmic_bytes": byt,
                      "GBps": round(byt / ms / 1e6, 1), "frac_of_measured_peak": round(byt / ms / 1e6 / PEAK, 4)}), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="cfg2,cfg4,cfg5,copy")
    ap.add_argument("--splits", type=int, default=0)
    ap.add_argument("--iters", type=int, default=64)
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--generic", action="store_true")
    a = ap.parse_args()
    fl = ops.ATTN_FORCE_GENERIC if a.generic else 0
    for c in a.cases.split(","):
        if c == "cfg2":
            attn_case("cfg2 8B B64 ctx2k", 64, 32, 8, 128, [2048] * 64, a.layers, a.splits, fl, iters=a.iters)
        elif c == "cfg2f":
            attn_case("cfg2 8B B64 ctx2k, fused RoPE + append + attention", 64, 32, 8, 128, [2048] * 64, a.layers, a.splits,
                      fl, iters=a.iters, fused=True)
        elif c == "b1_32k":
            attn_case("8B B1 ctx32k (single long sequence)", 1, 32, 8, 128, [32768], a.layers, a.splits, fl, iters=a.iters)
        elif c in ("b1_16k", "b1_64k", "b2_32k"):
            nb, nctx = (2, 32768) if c == "b2_32k" else (1, 16384 if c == "b1_16k" else 65536)
            attn_case(f"8B B{nb} ctx{nctx // 1024}k", nb, 32, 8, 128, [nctx] * nb, a.layers, a.splits, fl, iters=a.iters)
        elif c in ("b1_2k", "b1_4k", "b8_2k"):
            nb, nctx = {"b1_2k": (1, 2048), "b1_4k": (1, 4096), "b8_2k": (8, 2048)}[c]
            attn_case(f"8B B{nb} ctx{nctx // 1024}k", nb, 32, 8, 128, [nctx] * nb, a.layers, a.splits, fl, iters=a.iters)
        elif c == "b1_128k":
            attn_case("8B B1 ctx128k (single long sequence)", 1, 32, 8, 128, [131072], min(a.layers, 4), a.splits, fl,
                      iters=a.iters)
        elif c == "b4_8k":
            attn_case("8B B4 ctx8k", 4, 32, 8, 128, [8192] * 4, a.layers, a.splits, fl, iters=a.iters)
        elif c == "cfg4":
            attn_case("cfg4 70B B128 ctx4k (per layer)", 128, 64, 8, 128, [4096] * 128, min(a.layers, 4), a.splits, fl, iters=a.iters)
        elif c == "cfg4mb":
            attn_case("cfg4 70B micro-batch 16 ctx4k", 16, 64, 8, 128, [4096] * 16, a.layers, a.splits, fl, iters=a.iters)
        elif c == "cfg4mbf":
            attn_case("cfg4 70B micro-batch 16 ctx4k, fused RoPE + append + attention", 16, 64, 8, 128, [4096] * 16, a.layers,
                      a.splits, fl, iters=a.iters, fused=True)
        elif c == "cfg5":
            lens = [24576 + 8192 - 512 * i for i in range(8)]
            attn_case("cfg5 8B B8 ctx32k ragged, shared 24576 prefix", 8, 32, 8, 128, lens, a.layers, a.splits, fl,
                      share_prefix=24576 // 16, iters=a.iters)
            attn_case("cfg5 8B B8 ctx32k ragged, shared 24576 prefix, CASCADE kernel", 8, 32, 8, 128, lens, a.layers,
                      a.splits, fl, share_prefix=24576 // 16, iters=a.iters, cascade=True)
            attn_case("cfg5 8B B8 ctx32k ragged, unshared", 8, 32, 8, 128, lens, min(a.layers, 4), a.splits, fl, iters=a.iters)
        elif c == "cfg5_prefix_only":   # diagnostic: the prefix role alone (every context ends with the shared blocks)
            attn_case("cfg5 cascade, prefix role only", 8, 32, 8, 128, [24576] * 8, a.layers, a.splits, fl,
                      share_prefix=24576 // 16, iters=a.iters, cascade=True)
        elif c == "cfg5_tails_plain":   # diagnostic: the tails as an unshared batch through the plain kernel
            attn_case("cfg5 tails only, plain kernel", 8, 32, 8, 128, [8192 - 512 * i for i in range(8)], a.layers,
                      a.splits, fl, iters=a.iters)
        elif c == "cfg5c":
            lens = [24576 + 8192 - 512 * i for i in range(8)]
            attn_case("cfg5 cascade", 8, 32, 8, 128, lens, a.layers, a.splits, fl, share_prefix=24576 // 16, iters=a.iters,
                      cascade=True)
        elif c == "cfg5cf":
            lens = [24576 + 8192 - 512 * i for i in range(8)]
            attn_case("cfg5 cascade, fused RoPE + append", 8, 32, 8, 128, lens, a.layers, a.splits, fl,
                      share_prefix=24576 // 16, iters=a.iters, cascade=True, fused=True)
        elif c == "prefix64":
            lens64 = [2048 + 64 + 16 * (i % 8) for i in range(64)]
            attn_case("8B B64, shared 2048-token system prompt + short tails, plain", 64, 32, 8, 128, lens64, a.layers,
                      a.splits, fl, share_prefix=128, iters=a.iters)
            attn_case("8B B64, shared 2048-token system prompt + short tails, CASCADE", 64, 32, 8, 128, lens64, a.layers,
                      a.splits, fl, share_prefix=128, iters=a.iters, cascade=True)
        elif c == "copy":
            copy_cases()
        elif c == "append":
            append_cases(a.iters)
        elif c.startswith("grid:"):  # e.g. grid:20x4096 or grid:20x4096x64 = batch x context [x q heads], 8 kv heads
            parts = [int(x) for x in c[5:].split("x")]
            nb, nctx, hq = parts[0], parts[1], (parts[2] if len(parts) > 2 else 32)
            attn_case(f"B{nb} ctx{nctx} Hq{hq}", nb, hq, 8, 128, [nctx] * nb, a.layers, a.splits, fl, iters=a.iters)
