"""Sweep of (prefix_splits, suffix_splits) for the shared-prefix attention at batch-behind-a-system-prompt shapes,
next to the library's own plan (profiles/r02_prefix_split_sweep.jsonl)."""
import json, os, sys
sys.path.insert(0, os.getcwd())
import torch
import dgi_b200
from dgi_b200 import ops
dev, dt = "cuda:0", torch.bfloat16
Hq, Hkv, D, bs = 32, 8, 128, 16
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for B, prefix_tok, tail in ((64, 2048, 176), (64, 512, 32), (128, 1024, 64), (16, 4096, 256), (32, 8192, 512), (8, 2048, 128)):
    pb = prefix_tok // bs; tb = (tail + bs - 1) // bs
    max_blocks = pb + B * tb + 4
    k_pool = torch.randn((max_blocks, Hkv, bs, D), device=dev).to(dt); v_pool = torch.randn_like(k_pool)
    tables = torch.zeros((B, pb + tb), dtype=torch.int32, device=dev)
    tables[:, :pb] = torch.arange(pb, dtype=torch.int32, device=dev)
    tables[:, pb:] = (pb + torch.arange(B * tb, dtype=torch.int32, device=dev)).view(B, tb)
    lens = torch.full((B,), prefix_tok + tail, dtype=torch.int32, device=dev)
    q = torch.randn((B, Hq, D), device=dev).to(dt); ctx = prefix_tok + tail
    plan = ops.attn_prefix_plan(B, Hq, Hkv, D, bs, pb, ctx, dt)
    res = {}
    for ps in (1, 2, 3, 4, 6, 8, 12, 16):
        if pb // ps < 4: continue
        for ss in (1, 2):
            ws = torch.zeros(max(ops.attn_workspace_bytes(B, Hq, D, ps + ss), 16), dtype=torch.uint8, device=dev)
            out = torch.empty_like(q)
            call = lambda: ops.paged_decode_attention_prefix(q, k_pool, v_pool, tables, lens, ctx, pb, out=out, prefix_splits=ps, suffix_splits=ss, workspace=ws)
            for _ in range(3): call()
            torch.cuda.synchronize()
            tot = 0.0
            for _ in range(12):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); call(); e1.record(); torch.cuda.synchronize(); tot += e0.elapsed_time(e1)
            res[f"{ps}+{ss}"] = round(tot / 12 * 1e3, 1)
    print(json.dumps({"B": B, "prefix": prefix_tok, "tail": tail, "plan": f"{plan[0]}+{plan[1]}", "us": res}), flush=True)
