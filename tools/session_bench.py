#!/usr/bin/env python
"""Greedy decode through the reference's own call stack — DistributedInferenceSession.step -> WorkerSession.forward ->
InferenceServicer.Forward -> ModelShard.forward, then get_logits + argmax on the caller's side (reference:
worker/distributed/session.py:271-337, grpc_server.py:142-188) — with the hops in-process (ServicerTransport) and the
shards of one model on ONE GPU.  Tokens/s of that API next to the engine's chained-graph step.
usage: python tools/session_bench.py [--hops 2] [--layers 32] [--batch 64] [--ctx 2048] [--steps 32]"""
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
from dgi_b200 import ops  # noqa: E402
from dgi_b200.common.data_structures import SessionConfig, WorkerInfo, WorkerState  # noqa: E402
from dgi_b200.worker.distributed import session as S  # noqa: E402
from dgi_b200.worker.distributed.kv_cache import DistributedKVCacheManager  # noqa: E402
from dgi_b200.worker.distributed.model_shard import ModelShard, get_layer_range_for_worker  # noqa: E402
from dgi_b200.worker.distributed.servicer import InferenceServicer, ServicerTransport  # noqa: E402


def run(shape_name="8b", layers=32, hops=2, batch=64, ctx=2048, steps=32, fuse=True, host_io=False, dev="cuda:0"):
    """One measurement; ``host_io``: the caller's token ids start in pinned host memory and every step's tokens are
    read back to the host (the end-to-end form of bench.py's ``e2e``)."""
    a = argparse.Namespace(shape=shape_name, layers=layers, hops=hops, batch=batch, ctx=ctx, steps=steps,
                           no_fuse=not fuse)
    shape = dict(SHAPES[a.shape])
    shape["num_hidden_layers"] = a.layers
    cfg = ShapeConfig(**shape)
    B, ctx, steps = a.batch, a.ctx, a.steps
    n_blk = (ctx + 15) // 16
    Hkv = cfg.num_key_value_heads
    D = cfg.hidden_size // cfg.num_attention_heads
    start = ctx - steps - 6
    shards, svs, route = [], [], []
    for r in range(a.hops):
        lo, hi = get_layer_range_for_worker(a.layers, a.hops, r)
        sh = ModelShard.from_config(cfg, lo, hi, device=dev, dtype=torch.bfloat16, seed=0)
        mgr = DistributedKVCacheManager(num_layers=hi - lo, num_heads=Hkv, head_dim=D, gpu_cache_blocks=B * n_blk + 8,
                                        cpu_cache_gb=0.01, block_size=16, device=dev, dtype=torch.bfloat16)
        sv = InferenceServicer(f"sb-worker-{r}", sh, mgr, max_sessions=2)
        sv.fuse_projections = not a.no_fuse
        S.TRANSPORTS[f"sb-worker-{r}"] = ServicerTransport(sv, max_length=ctx)
        route.append(WorkerInfo(worker_id=f"sb-worker-{r}", state=WorkerState.ONLINE))
        shards.append(sh)
        svs.append(sv)
    ids = torch.randint(0, cfg.vocab_size, (B, 1), device=dev)

    async def drive():
        sess = S.DistributedInferenceSession(SessionConfig(model_name=a.shape, max_length=ctx, max_retries=1), route)
        await sess.setup()
        await sess.step(ids)  # the first call creates the hops' sessions and their blocks
        # synthetic history: every hop's session holds `start` tokens of N(0, 1) KV
        for sv in svs:
            for s in sv._sessions.values():
                s["position"] = start
            sv.kv_cache.gpu_cache.backing.normal_(0.0, 1.0)
        sess.position = start  # fans out to the hop sessions
        nxt = ids

        host_ids = torch.empty((B, 1), dtype=torch.int64).pin_memory() if host_io else None
        if host_io:
            host_ids.copy_(ids.cpu())

        async def one():
            nonlocal nxt
            if host_io:
                nxt = host_ids.to(dev, non_blocking=True)
            h = await sess.step(nxt)
            logits = shards[-1].get_logits(h[:, -1:, :])[:, 0, :]
            nxt = ops.argmax_rows(logits)[:, None]
            if host_io:
                host_ids.copy_(nxt)  # device -> pinned host; synchronises the step

        for _ in range(5):
            await one()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            await one()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        pos = sess.position
        await sess.close()
        return ms, pos

    ms, pos = asyncio.run(drive())
    for r in range(a.hops):
        S.TRANSPORTS.pop(f"sb-worker-{r}", None)
    return {"shape": a.shape, "layers": a.layers, "hops": a.hops, "batch": B, "context_at_end": pos,
            "steps": steps, "fused_projections": not a.no_fuse, "host_ids_in_and_out": bool(host_io),
            "ms_per_step": round(ms, 3), "tokens_per_second": round(B / ms * 1e3, 1),
            "path": "DistributedInferenceSession.step -> WorkerSession.forward -> ServicerTransport -> "
                    "InferenceServicer.Forward (session decode graph) per hop; get_logits + argmax by the caller"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="8b")
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--hops", type=int, default=2)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=32)
    ap.add_argument("--no-fuse", action="store_true")
    ap.add_argument("--host-io", action="store_true")
    a = ap.parse_args()
    print(json.dumps(run(a.shape, a.layers, a.hops, a.batch, a.ctx, a.steps, not a.no_fuse, a.host_io)))


if __name__ == "__main__":
    main()
