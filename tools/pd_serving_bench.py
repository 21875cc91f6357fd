#!/usr/bin/env python
"""P/D-separated serving vs hybrid serving on GPU pairs, with the request metrics of the reference's benchmark
(benchmarks/pd_separation.py:342-368): TTFT = prompt pass end - submit, TPOT = decode time / tokens generated,
E2E = decode end - submit, plus the KV migration time per chunk and tokens/s.

Run under torchrun with N a multiple of K + 1 (``--prefill-ranks K``, default 1).  All requests are submitted at t0
(closed burst); a group of K + 1 GPUs serves W "super-waves" of K*B requests.
  pd         : K ranks run prompt passes (waves of B) and push the KV blocks to the group's decode rank, which
               decodes K*B sequences per step (dgi_b200.pd_serving.PDGroupWorker; prompt passes overlap the
               previous wave's decode steps)
  hybrid     : every rank serves whole waves of B (prompt pass, then decode), round-robin over the group
  hybrid_big : (K > 1) every rank serves whole waves of K*B — the decode batch of the pd mode
Times are CUDA-event times relative to a per-rank t0 event recorded right after a barrier.
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import dgi_b200  # noqa: E402,F401
from bench import SHAPES, ShapeConfig  # noqa: E402
from dgi_b200.engine import PagedDecodeEngine  # noqa: E402
from dgi_b200.pd_serving import PDGroupWorker, make_fanin_groups, make_pair_groups  # noqa: E402
from dgi_b200.worker.distributed.model_shard import ModelShard  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="8b")
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (0 = the shape's)")
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--prompt", type=int, default=4096)
    ap.add_argument("--new", type=int, default=128)
    ap.add_argument("--waves", type=int, default=4)
    ap.add_argument("--prefill-chunk", type=int, default=4)
    ap.add_argument("--prefill-ranks", type=int, default=1, help="K prefill GPUs feed one decode GPU")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    dist.init_process_group("nccl", device_id=torch.device(dev))
    K = a.prefill_ranks
    group = make_pair_groups(world) if K == 1 else make_fanin_groups(world, K)
    shape = dict(SHAPES[a.shape])
    if a.layers:
        shape["num_hidden_layers"] = a.layers
    cfg = ShapeConfig(**shape)
    shard = ModelShard.from_config(cfg, 0, cfg.num_hidden_layers, device=dev, dtype=torch.bfloat16, seed=0)
    w = PDGroupWorker(shard, group, a.batch, a.prompt, a.new, prefill_chunk=a.prefill_chunk, fuse_projections=True)
    B, P, n_new, W = a.batch, a.prompt, a.new, a.waves
    gen = torch.Generator().manual_seed(1234)
    # super-wave i: K*B prompts; prefill rank p of a group takes rows p*B .. (p+1)*B
    prompts = [torch.randint(0, cfg.vocab_size, (K * B, P), generator=gen).to(dev) for _ in range(W)]
    engines = {w.engine.max_batch: w.engine}

    def engine_for(batch):
        """Hybrid serving at a batch this rank's own engine was not built for: a second engine over the same pools
        (so its split plan and graphs are the ones a dedicated engine of that batch would have)."""
        if batch not in engines:
            engines[batch] = PagedDecodeEngine([shard], batch, w.total, prefill_chunk=a.prefill_chunk,
                                               fuse_projections=True, kv_pools=[w.endpoint.pool])
        return engines[batch]

    def start():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        t0 = torch.cuda.Event(enable_timing=True)
        t0.record()
        w._side.wait_stream(torch.cuda.current_stream())
        return t0

    def pd_round(record):
        ev = {} if record else None
        t0 = start()
        toks = []
        for i in range(W):
            if w.is_prefill:
                w.prefill_wave(i, prompts[i][w.me * B:(w.me + 1) * B], ev)
            else:
                toks.append(w.decode_wave(i, B, W, ev))
        torch.cuda.synchronize()
        w._side.synchronize()
        out = {}
        if record and w.is_prefill:
            out["ttft_ms"] = [t0.elapsed_time(e) for (_, lo, hi, e) in ev["prefill_end"] for _ in range(hi - lo)]
            out["migration_ms"] = [p0.elapsed_time(p1) for (_, _, _, p0, p1) in ev["push"]]
            out["migration_bytes"] = w.endpoint.payload_bytes(w.n_blk_prompt) * a.prefill_chunk
        if record and not w.is_prefill:
            out["tpot_ms"] = [e0.elapsed_time(e1) / (n_new - 1) for (_, e0, e1) in ev["decode"]]
            out["e2e_ms"] = [t0.elapsed_time(e1) for (_, _, e1) in ev["decode"] for _ in range(K * B)]
            out["end_ms"] = max(t0.elapsed_time(e1) for (_, _, e1) in ev["decode"])
            out["checksum"] = int(sum(int(t.sum()) for t in toks))
        return out

    def hybrid_round(record, per_wave=1):
        """Waves of per_wave * B requests, round-robin over the K + 1 ranks of the group."""
        t0 = start()
        nb = per_wave * B
        eng = engine_for(nb)
        n_waves = W * K // per_wave
        mine = [j for j in range(n_waves) if j % (K + 1) == w.me]
        marks, sums = [], []
        for n_served, j in enumerate(mine):
            i, r0 = (j * nb) // (K * B), (j * nb) % (K * B)  # super-wave and first row of this wave's prompts
            # this rank's blocks are all in its slot sets: serve the wave out of one of them
            s = n_served % w.sets
            eng.adopt_sequences(w._slot_blocks[s][:nb], [0] * nb, table_rows=w._my_tables[s, :nb])
            first = torch.empty((nb,), dtype=torch.int64, device=dev)
            for lo in range(0, nb, a.prefill_chunk):
                hi = min(nb, lo + a.prefill_chunk)
                first[lo:hi] = eng.prefill_chunk(lo, hi, prompts[i][r0 + lo:r0 + hi])
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                marks.append(("p", hi - lo, e))
            eng.ids_buf[:nb, 0].copy_(first)
            d0 = torch.cuda.Event(enable_timing=True)
            d0.record()
            csum = first.sum()
            for _ in range(1, n_new):
                eng.decode_step_device()
                csum = csum + eng.next_tokens[:nb].sum()
            sums.append(csum)
            d1 = torch.cuda.Event(enable_timing=True)
            d1.record()
            marks.append(("d", d0, d1, nb))
        torch.cuda.synchronize()
        out = {}
        if record:
            out["ttft_ms"] = [t0.elapsed_time(m[2]) for m in marks if m[0] == "p" for _ in range(m[1])]
            out["tpot_ms"] = [m[1].elapsed_time(m[2]) / (n_new - 1) for m in marks if m[0] == "d"]
            out["e2e_ms"] = [t0.elapsed_time(m[2]) for m in marks if m[0] == "d" for _ in range(m[3])]
            out["end_ms"] = max(t0.elapsed_time(m[2]) for m in marks if m[0] == "d")
            out["checksum"] = int(sum(int(c) for c in sums))
        return out

    results = {}
    modes = [("pd", pd_round), ("hybrid", hybrid_round)]
    if K > 1:
        modes.append(("hybrid_big", lambda record: hybrid_round(record, per_wave=K)))
    for mode, fn in modes:
        fn(False)  # warm-up round: graph capture, cuBLAS plans, allocator
        mine = fn(True)
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        if rank == 0:
            def cat(key):
                return [x for g in gathered for x in g.get(key, [])]
            end_ms = max(g.get("end_ms", 0.0) for g in gathered)
            n_req = (world // (K + 1)) * W * K * B
            res = {
                "mode": mode, "n_gpus": world, "prefill_ranks_per_decode_rank": K, "requests": n_req, "prompt": P,
                "new_tokens": n_new, "batch": B, "decode_batch": K * B if mode != "hybrid" else B,
                "waves_per_pair": W, "shape": a.shape, "layers": cfg.num_hidden_layers,
                "tokens_per_second": n_req * n_new / (end_ms / 1e3),
                "avg_ttft_ms": statistics.mean(cat("ttft_ms")), "p99_ttft_ms": max(cat("ttft_ms")),
                "avg_tpot_ms": statistics.mean(cat("tpot_ms")), "avg_e2e_ms": statistics.mean(cat("e2e_ms")),
                "total_time_s": end_ms / 1e3,
            }
            mig = cat("migration_ms")
            if mig:
                nbytes = max(g.get("migration_bytes", 0) for g in gathered)
                res["avg_migration_ms"] = statistics.mean(mig)
                res["migration_GBps_per_pair"] = nbytes / (statistics.mean(mig) * 1e-3) / 1e9
            res["checksum"] = sum(g.get("checksum", 0) for g in gathered)  # sum of all generated token ids
            results[mode] = res
            if mode != "pd":
                # same prompts, same kernels, same decode batch (so the same attention split plan): the modes must
                # generate the same tokens.  hybrid with K > 1 decodes at batch B against pd's K*B — another split
                # plan, so near-tie logits may fall the other way and the sums need not agree.
                res["tokens_match_pd_mode"] = res["checksum"] == results["pd"]["checksum"]
            print(json.dumps(res), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
