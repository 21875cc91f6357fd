"""Shared-prefix attention at several shapes: run once with B200_PREFIX_SEPARATE_COMBINE=0 (merge by the last-arriving
CTA) and once with =1 (separate, programmatically dependent combine launch) to see which the library should pick."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import dgi_b200  # noqa: E402,F401
from dgi_b200 import ops  # noqa: E402

dev, dt = "cuda:0", torch.bfloat16
Hq, Hkv, D, bs = 32, 8, 128, 16
for B, prefix_tok, tail in ((8, 512, 64), (8, 2048, 128), (16, 4096, 256), (64, 512, 32), (64, 2048, 176), (32, 8192, 512),
                            (8, 24576, 6400), (128, 1024, 64)):
    pb = prefix_tok // bs
    tb = (tail + bs - 1) // bs
    max_blocks = pb + B * tb + 4
    k_pool = torch.randn((max_blocks, Hkv, bs, D), device=dev).to(dt)
    v_pool = torch.randn_like(k_pool)
    tables = torch.zeros((B, pb + tb), dtype=torch.int32, device=dev)
    tables[:, :pb] = torch.arange(pb, dtype=torch.int32, device=dev)
    tables[:, pb:] = (pb + torch.arange(B * tb, dtype=torch.int32, device=dev)).view(B, tb)
    lens = torch.full((B,), prefix_tok + tail, dtype=torch.int32, device=dev)
    q = torch.randn((B, Hq, D), device=dev).to(dt)
    ctx = prefix_tok + tail
    ps, ss = ops.attn_prefix_plan(B, Hq, Hkv, D, bs, pb, ctx, dt)
    # the caller owns the workspace and vouches for zeroed counters, as the engine does (no memset per call)
    ws = torch.zeros(max(ops.attn_workspace_bytes(B, Hq, D, ps + ss), 16), dtype=torch.uint8, device=dev)
    out = torch.empty_like(q)

    def call():
        ops.paged_decode_attention_prefix(q, k_pool, v_pool, tables, lens, ctx, pb, out=out, prefix_splits=ps,
                                          suffix_splits=ss, workspace=ws, flags=ops.ATTN_WS_COUNTERS_ZERO)

    for _ in range(5):
        call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tot = 0.0
    n = 20
    for _ in range(n):
        flush.zero_()
        e0.record()
        call()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    print(json.dumps({"B": B, "prefix": prefix_tok, "tail": tail, "splits": f"{ps}+{ss}", "us": round(tot / n * 1e3, 2),
                      "separate_combine": os.environ.get("B200_PREFIX_SEPARATE_COMBINE", "library")}), flush=True)
