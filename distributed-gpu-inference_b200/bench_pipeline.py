"""bench.py's multi-GPU leg: layer-sharded decode over N ranks (BASELINE config 4 by default).

Launched by torchrun, one rank per GPU.  Rank r builds layers ``get_layer_range_for_worker(L, N, r)`` of the
named shape (random-init, same seed scheme on every partition), one ``PagedDecodeEngine`` per micro-batch, and
joins a ``LayerPipeline``.  Timed region: K decode steps of all micro-batches, bracketed by barrier +
synchronize, CUDA events on every rank, MAX over ranks.
"""
from __future__ import annotations

import json
import os
import statistics
import time

import torch
import torch.distributed as dist

from . import ops
from .engine import PagedDecodeEngine
from .pipeline import LayerPipeline, make_pipeline_groups
from .worker.distributed.model_shard import ModelShard, get_layer_range_for_worker


def run_pipeline(args, workload, SHAPES, ShapeConfig, ClockSampler, measured_peaks, cpu_reference, cpu_handoff=None,
                 extra_workload=None):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    # stdout must carry exactly one JSON line; NCCL prints its version banner to fd 1, so everything goes to
    # stderr for the duration of the run and the line is written to the saved descriptor
    import sys
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    dist.init_process_group("nccl", device_id=torch.device(dev))
    try:
        line = _run(args, workload, SHAPES, ShapeConfig, ClockSampler, measured_peaks, rank, world, local, dev)
        if extra_workload is not None:
            import gc

            gc.collect()
            torch.cuda.empty_cache()
            more = _run(args, extra_workload, SHAPES, ShapeConfig, ClockSampler, measured_peaks, rank, world, local,
                        dev, with_pd=False)
            if line is not None and more is not None:
                keep = ("value", "unit", "ms_per_step", "scaling", "e2e", "gpu_launches", "roofline", "clocks")
                line["north_star_cfg4"] = {"workload": more["config"]["workload"],
                                           "parallelism": more["config"]["parallelism"],
                                           "bound_tokens_per_s": more["config"].get("bound_tokens_per_s"),
                                           **{k: more[k] for k in keep}}
        if line is not None and cpu_handoff is not None and isinstance(line.get("pd_transfer"), dict):
            line["pd_transfer"]["cpu_reference_handoff"] = cpu_handoff()  # bench.py's host-side baseline leg
        if line is not None:
            os.write(saved_stdout, (json.dumps(line) + "\n").encode())
    finally:
        dist.destroy_process_group()
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)


def _measure_pd_handoff(rank, world, dev):
    """Third part of BASELINE's metric: prefill -> decode KV hand-off, rank i -> rank i + N/2 on all pairs at once,
    one fused peer-push launch per sequence (8B-shape blocks, 8 layers x 4 sequences of 4k tokens = 512 MiB per
    pair), CUDA events between barriers, max over ranks."""
    if world < 2 or world % 2:
        return None
    try:
        from .kv_transfer import PeerPushEndpoint

        half = world // 2
        layers, Hkv, D, bs, n_blocks, seqs = 8, 8, 128, 16, 256, 4
        max_blocks = n_blocks * seqs + 32
        pp = PeerPushEndpoint(layers, Hkv, D, bs, max_blocks, dev, torch.bfloat16)
        sender = rank < half
        peer = rank + half if sender else rank - half
        if sender:
            pp.pool.backing.normal_()
        mk = lambda seed: [torch.randperm(max_blocks, generator=torch.Generator().manual_seed(seed * 100 + i))[:n_blocks]
                           .to(torch.int32).to(dev) for i in range(seqs)]
        src_tables, dst_tables = mk(rank), mk(peer)

        def go():
            if sender:
                for st, dt in zip(src_tables, dst_tables):
                    pp.push(st, dt, peer)
            pp.barrier()

        for _ in range(2):
            go()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        go()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        payload = pp.payload_bytes(n_blocks) * seqs
        gbs = payload / (ms.item() * 1e-3) / 1e9
        del pp
        return {"value": gbs, "unit": "GB/s per (prefill, decode) pair", "aggregate": gbs * half, "pairs": half,
                "payload_MiB_per_pair": payload / 2 ** 20, "ms": ms.item(), "frac_of_nvlink_900": gbs / 900.0,
                "path": "fused peer push (b200_kv_move_blocks_layers) into the decode rank's pools"}
    except Exception as exc:  # noqa: BLE001 - the hand-off figure is an extra; never lose the decode line over it
        return {"unavailable": repr(exc)}


def _run(args, workload, SHAPES, ShapeConfig, ClockSampler, measured_peaks, rank, world, local, dev, with_pd=True):
    cfg = ShapeConfig(**SHAPES[workload["shape"]])
    B, ctx = workload["batch"], workload["context"]
    K, W = args.steps, args.warmup
    M = getattr(args, "micro_batches", 0) or world
    if B % M:
        raise SystemExit(f"batch {B} not divisible into {M} micro-batches")
    mb = B // M
    total_steps = 2 * (K + W) + 4
    P = ctx - total_steps
    start, end = get_layer_range_for_worker(cfg.num_hidden_layers, world, rank)
    t0 = time.time()
    shard = ModelShard.from_config(cfg, start, end, device=dev, dtype=torch.bfloat16, seed=0)
    engines = [PagedDecodeEngine([shard], max_batch=mb, max_seq_len=ctx, block_size=16, prefill_chunk=2,
                                 fuse_projections=not getattr(args, "no_fuse_projections", False)) for _ in range(M)]
    for e in engines:
        e.allocate(mb, ctx)
    groups = make_pipeline_groups(world, warm_up_device=torch.device(dev))
    pipe = LayerPipeline(engines, rank, world, groups)
    synthetic = not getattr(args, "real_prefill", False)
    gen = torch.Generator().manual_seed(1234)
    prompts = [torch.randint(0, cfg.vocab_size, (mb, P), generator=gen) for _ in range(M)]
    if synthetic:
        for e in engines:
            for pool in e.pools:
                pool.backing.normal_(0.0, 1.0)
            e.context_lens[:mb] = P
        firsts = [p[:, -1].to(dev) for p in prompts]
    else:
        firsts = pipe.prefill([p.to(dev) for p in prompts] if rank == 0 else None, P, chunk=2)
    if rank == 0:
        pipe.set_first_tokens(firsts)
    torch.cuda.synchronize()
    setup_s = time.time() - t0

    def timed(steps, host_roundtrip):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        pipe.decode(steps, host_roundtrip=host_roundtrip)
        e1.record()
        torch.cuda.synchronize()
        dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    pipe.decode(W)  # warm-up: graph capture, communicator bring-up
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dev_ms = timed(K, host_roundtrip=False)
    pipe.decode(W, host_roundtrip=True)
    e2e_ms = timed(K, host_roundtrip=True)
    clocks = sampler.stop() if rank == 0 else None

    # attention launch time on this rank's first micro-batch, eagerly launched (no comm; state advances locally)
    eng = engines[0]
    eng.use_cuda_graph = False
    ops.ATTN_EVENTS = []
    for _ in range(2):
        torch.cuda._sleep(60_000_000)  # let the host run ahead so the event pairs bracket device time only
        eng.stage_step()
    torch.cuda.synchronize()
    ev, ops.ATTN_EVENTS = ops.ATTN_EVENTS, None
    attn_ms = statistics.mean(a.elapsed_time(b) for a, b in ev)
    ctx_now = P + 2 * (K + W) + 1
    algo = mb * ctx_now * 2 * eng.Hkv * eng.D * 2 + 2 * mb * eng.Hq * eng.D * 2
    pd = _measure_pd_handoff(rank, world, dev) if with_pd else None
    if rank != 0:
        return None
    peaks, kind = measured_peaks()
    achieved = algo / (attn_ms * 1e-3) / 1e9
    layers_here = end - start
    # HBM bound of layer-range decode: every micro-batch re-streams the stage's weights, the KV is read once per step
    es = 2
    per_layer_w = (cfg.hidden_size * (eng.Hq + 2 * eng.Hkv) * eng.D + cfg.hidden_size * eng.Hq * eng.D
                   + 3 * cfg.hidden_size * cfg.intermediate_size) * es
    w_total = per_layer_w * cfg.num_hidden_layers + cfg.vocab_size * cfg.hidden_size * es
    kv_total = B * ctx_now * 2 * eng.Hkv * eng.D * es * cfg.num_hidden_layers
    bound = B * world * peaks["hbm_gbs"] * 1e9 / (M * w_total + kv_total)
    per_step = M * layers_here * (4 + (1 if eng.num_splits > 1 else 0))  # rank 0's kernels from this repo
    line = {
        "metric": "decode_tokens_per_s",
        "value": B * K / (dev_ms * 1e-3),
        "unit": "tokens/s",
        "n_gpus": world,
        "steps": K,
        "warmup": W,
        "ms_per_step": dev_ms / K,
        "higher_is_better": True,
        "scaling": workload["scaling"],
        "vs_baseline": None,
        "dtype": "bf16",
        "data": "synthetic",
        "config": {
            "workload": workload["name"],
            "parallelism": f"layer ranges over {world} ranks (pp{world}), {M} micro-batches of {mb}",
            "block_size": 16,
            "batch": B,
            "context": f"{P + W + 1}..{P + W + K} during the timed steps",
            "kv_fill": "synthetic normal(0,1)" if synthetic else "real prefill of random prompts",
            "l2": "inputs larger than L2 (each rank streams its weights once per micro-batch plus its KV per step)",
            "attn_splits": eng.num_splits,
            "setup_s": round(setup_s, 1),
            "bound_tokens_per_s": bound,
            "frac_of_bound": (B * K / (dev_ms * 1e-3)) / bound,
            "note": ("bound for layer-range decode = B*N*BW / (M*W + KV) with M micro-batches, BW the measured HBM "
                     "copy peak.  The N = 1 line (config 2) and the weak-scaled N > 1 lines run the same model and "
                     "the same per-GPU batch, so they compare directly; the 70B shape of configs[3] (141 GB of "
                     "weights + 172 GB of KV, does not fit one GPU) is under north_star_cfg4"),
        },
        "clocks": clocks,
        "e2e": {"value": B * K / (e2e_ms * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": B * 8,
                "d2h_bytes_per_step": B * 8, "ms_per_step": e2e_ms / K},
        "gpu_launches": per_step * K,
        "pd_transfer": pd,
        "roofline": {
            "bound": "hbm",
            "kernel": "paged_attn_tma_kernel (b200_paged_decode_attn), one micro-batch on rank 0",
            "achieved": achieved,
            "peak": peaks["hbm_gbs"],
            "peak_source": f"MEASURED_PEAKS.json ({kind})",
            "unit": "GB/s",
            "frac": achieved / peaks["hbm_gbs"],
            "traffic": None,
            "algorithmic_bytes_per_launch": algo,
            "launch_ms": attn_ms,
            "launches_timed": len(ev),
        },
    }
    return line
