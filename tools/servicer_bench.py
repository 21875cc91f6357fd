#!/usr/bin/env python
"""Decode steps through the servicer (InferenceServicer.Forward, in-process tensors) next to the engine's graph step,
same shard, same batch and context: what the reference's own per-hop call path costs on top of the kernels.
usage: python tools/servicer_bench.py [--shape 8b] [--layers 8] [--batch 64] [--ctx 2048] [--steps 20]"""
import argparse
import asyncio
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import dgi_b200  # noqa: E402,F401
from bench import SHAPES, ShapeConfig  # noqa: E402
from dgi_b200.engine import PagedDecodeEngine  # noqa: E402
from dgi_b200.worker.distributed.kv_cache import DistributedKVCacheManager  # noqa: E402
from dgi_b200.worker.distributed.model_shard import ModelShard  # noqa: E402
from dgi_b200.worker.distributed.servicer import InferenceServicer  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="8b")
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    dev = "cuda:0"
    shape = dict(SHAPES[a.shape])
    shape["num_hidden_layers"] = a.layers
    cfg = ShapeConfig(**shape)
    shard = ModelShard.from_config(cfg, 0, a.layers, device=dev, dtype=torch.bfloat16, seed=0)
    B, ctx, steps = a.batch, a.ctx, a.steps
    n_blk = (ctx + 15) // 16
    Hkv = cfg.num_key_value_heads
    D = cfg.hidden_size // cfg.num_attention_heads
    mgr = DistributedKVCacheManager(num_layers=a.layers, num_heads=Hkv, head_dim=D, gpu_cache_blocks=B * n_blk + 8,
                                    cpu_cache_gb=0.01, block_size=16, device=dev, dtype=torch.bfloat16)
    sv = InferenceServicer("w0", shard, mgr, max_sessions=2)
    start = ctx - steps - 4
    ids = torch.randint(0, cfg.vocab_size, (B, 1), device=dev)

    async def run(graphs):
        sv.decode_graphs = graphs
        sid = (await sv.CreateSession({"model_name": a.shape, "max_length": ctx}))["session_id"]
        sess = sv._sessions[sid]
        sv._session_kv(sess, B, start)  # blocks taken; pretend `start` tokens are cached (synthetic KV)
        mgr.gpu_cache.backing.normal_(0.0, 1.0)
        sess["position"] = start
        pos = start
        for _ in range(4):  # warm-up
            r = await sv.Forward({"session_id": sid, "input": ids, "position": pos})
            assert r["success"], r
            pos += 1
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            r = await sv.Forward({"session_id": sid, "input": ids, "position": pos})
            pos += 1
        e1.record()
        torch.cuda.synchronize()
        await sv.CloseSession({"session_id": sid})
        return e0.elapsed_time(e1) / steps

    ms_eager = asyncio.run(run(False))
    ms_servicer = asyncio.run(run(True))
    eng = PagedDecodeEngine([shard], B, ctx, external_head=True)
    eng.allocate(B, ctx)
    for pool in eng.pools:
        pool.backing.normal_(0.0, 1.0)
    eng.context_lens[:B] = start
    eng._host_lens[:B] = [start] * B
    eng.ids_buf[:B].copy_(ids)
    for _ in range(4):
        eng.stage_step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        eng.stage_step()
    e1.record()
    torch.cuda.synchronize()
    ms_engine = e0.elapsed_time(e1) / steps
    sv.fuse_projections = True  # last: it rewrites the shard's weight storage
    ms_fused = asyncio.run(run(True))
    print(json.dumps({"shape": a.shape, "layers": a.layers, "batch": B, "context": ctx, "steps": steps,
                      "servicer_forward_ms_per_step": round(ms_servicer, 3),
                      "servicer_forward_eager_launches_ms_per_step": round(ms_eager, 3),
                      "servicer_forward_fused_projections_ms_per_step": round(ms_fused, 3),
                      "engine_graph_ms_per_step": round(ms_engine, 3),
                      "note": "same shard (embedding + layers + final norm), bf16, synthetic KV; the servicer step is the "
                              "session's decode graph behind the RPC envelope (decode_graphs=False: eager launches), "
                              "the engine step the bare graph"}))


if __name__ == "__main__":
    main()
