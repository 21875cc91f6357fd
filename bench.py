#!/usr/bin/env python
"""Headline benchmark: greedy decode tokens/s of the layer-range path over paged KV (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's torch CPU path (oracle/_ref bytecode, else the oracle port)

N = 1 : BASELINE configs[1] — Llama-3-8B shape, random-init bf16, block 16, batch 64, context 2k, decode; the line
        also carries configs[4] (batch 8, context 32k, shared prefix) as `cfg5` and configs[0] (tiny model, in
        full, GPU and host side by side) as `cpu_baseline_cfg1`, and configs[1] once more through the reference's own call
        stack (DistributedInferenceSession.step -> servicer hop, host ids in / out) as `reference_api`.
N > 1 : BASELINE configs[3] — Llama-3-70B shape layer-sharded over N ranks (torchrun, NCCL send/recv of
        activations, vocabulary-parallel head), batch 128, context 4k, strong scaling; the 8B shape weak-scaled over
        the same ranks rides along as `weak_scaled_8b`, the P/D hand-off as `pd_transfer`, and `parity` holds the
        token checks of the run (dgi_b200/bench_pipeline.py).
A "step" is one decode step of the whole batch.  `value` times K chained CUDA-graph steps with inputs resident
in HBM (CUDA events, max over ranks); `e2e` times the public per-step call with host token ids in and out.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

SHAPES = {
    "tiny": dict(hidden_size=512, num_attention_heads=8, num_key_value_heads=2, intermediate_size=1408,
                 vocab_size=1024, num_hidden_layers=4),
    "8b": dict(hidden_size=4096, num_attention_heads=32, num_key_value_heads=8, intermediate_size=14336,
               vocab_size=128256, num_hidden_layers=32),
    "70b": dict(hidden_size=8192, num_attention_heads=64, num_key_value_heads=8, intermediate_size=28672,
                vocab_size=128256, num_hidden_layers=80),
}


class ShapeConfig:
    """Attribute bag with the HF config names the shards read."""

    def __init__(self, **kw):
        self.rms_norm_eps = 1e-5
        self.rope_theta = 500000.0
        self.initializer_range = 0.02
        self.__dict__.update(kw)
        self.head_dim = self.hidden_size // self.num_attention_heads


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


# ---------------------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:  # noqa: BLE001 - no nvidia-smi: report nulls
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                parts = [p.strip() for p in line.split(",")]
                if len(parts) < 9:
                    continue
                try:
                    sm.append(float(parts[1]))
                    mx.append(float(parts[2]))
                except ValueError:
                    continue
                for name, val in zip(names, parts[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:  # noqa: BLE001
            pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ---------------------------------------------------------------------------------------------------------
# reference CPU path (oracle/_ref bytecode of the reference's ModelShard, else the oracle port) — bounded sample
# ---------------------------------------------------------------------------------------------------------
def _reference_modules():
    """The reference's own ModelShard module from oracle/_ref (bytecode compiled from /root/reference by
    oracle/build_ref.py; it travels with the snapshot), or None when it was not built."""
    from oracle import ref_loader

    return ref_loader.load("model_shard")


def cpu_reference_tokens_per_s(shape: str, batch: int, ctx: int, steps: int, warmup: int, sample_layers: int = 2,
                               sample_batch: int = 0):
    """Time the reference's torch CPU path on the host cores for `sample_layers` layers of the named shape at the full
    batch and context, plus the embedding / final norm / lm_head once per step, and extrapolate to the full depth.

    kind "reference": the reference's own ``ModelShard.forward`` / ``get_logits`` (oracle/_ref bytecode) run the step,
    over the legacy-convention HF layer of oracle/llama_ref.py with dense (k, v) tuples, fp32 (the installed
    transformers no longer speaks the call convention of model_shard.py:209-215).  kind "port": the same
    step spelled out here, when oracle/_ref was not built."""
    from oracle import llama_ref as L

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    full_batch = batch
    if sample_batch and sample_batch < batch:
        batch = sample_batch  # bounded sample: fewer sequences, same shape and context
    cfg = L.RefConfig(**SHAPES[shape])
    total_layers = cfg.num_hidden_layers
    g = torch.Generator().manual_seed(0)
    layers = []
    for _ in range(sample_layers):
        layer = L.RefDecoderLayer(cfg)
        for p in layer.parameters():
            if p.dim() == 2:
                p.data.normal_(0.0, 0.02, generator=g)
        layers.append(layer)
    head = torch.empty(cfg.vocab_size, cfg.hidden_size).normal_(0.0, 0.02, generator=g)
    embed = torch.empty(cfg.vocab_size, cfg.hidden_size).normal_(0.0, 0.02, generator=g)
    norm_w = torch.ones(cfg.hidden_size)
    start_ctx = ctx - (steps + warmup)
    pasts = [(torch.randn(batch, cfg.num_key_value_heads, start_ctx, cfg.head_dim, generator=g),
              torch.randn(batch, cfg.num_key_value_heads, start_ctx, cfg.head_dim, generator=g)) for _ in layers]
    ids = torch.randint(0, cfg.vocab_size, (batch, 1), generator=g)
    ref_ms = _reference_modules()
    full = ends = None
    if ref_ms is not None:
        def assemble(owned):  # one shard that is first and last: embedding, `owned` layers, norm, lm_head
            sh = ref_ms.ModelShard(shape, 0, len(owned), device="cpu", dtype=torch.float32)
            for layer in owned:
                sh.layers.append(layer)
            sh.is_first_shard = sh.is_last_shard = True
            sh.total_layers = len(owned)
            sh.embed_tokens = torch.nn.Embedding.from_pretrained(embed, freeze=True)
            sh.embed_positions = None
            sh.norm = L.RefNorm(norm_w, cfg.rms_norm_eps)
            sh.lm_head = torch.nn.Linear(cfg.hidden_size, cfg.vocab_size, bias=False)
            sh.lm_head.weight.data = head
            return sh.eval()

        full, ends = assemble(layers), assemble([])  # `ends`: the per-step work that does not grow with depth
    t_layers, t_ends, t_extra = [], [], []
    with torch.no_grad():
        for it in range(warmup + steps):
            pos = torch.full((batch, 1), start_ctx + it, dtype=torch.long)
            if full is not None:
                t0 = time.perf_counter()
                h, pasts = full.forward(ids, position_ids=pos, past_key_values=pasts, use_cache=True)
                nxt = full.get_logits(h)[:, 0].argmax(-1, keepdim=True)
                t1 = time.perf_counter()
                h0, _ = ends.forward(ids, position_ids=pos, past_key_values=None, use_cache=True)
                ends.get_logits(h0)[:, 0].argmax(-1, keepdim=True)
                t2 = time.perf_counter()
                ids = nxt
                if it >= warmup:
                    t_ends.append(t2 - t1)
                    t_layers.append(max((t1 - t0) - (t2 - t1), 0.0))
                    t_extra.append(t2 - t1)  # the ends-only pass is work this step really did, beyond the sample
                continue
            t0 = time.perf_counter()
            h = torch.nn.functional.embedding(ids, embed)
            t1 = time.perf_counter()
            for li, layer in enumerate(layers):
                h, pasts[li] = layer(h, position_ids=pos, past_key_value=pasts[li], use_cache=True)
            t2 = time.perf_counter()
            logits = torch.nn.functional.linear(L.rms_norm(h, norm_w, cfg.rms_norm_eps), head)
            ids = logits[:, 0].argmax(-1, keepdim=True)
            t3 = time.perf_counter()
            if it >= warmup:
                t_layers.append(t2 - t1)
                t_ends.append((t1 - t0) + (t3 - t2))
    per_layer = statistics.mean(t_layers) / sample_layers
    step_s = per_layer * total_layers + statistics.mean(t_ends)
    how = ("the reference's ModelShard.forward + get_logits (oracle/_ref bytecode) over the legacy HF layer"
           if full is not None else "oracle port of ModelShard.forward + the legacy HF layer")
    return {
        "value": batch / step_s,
        "unit": "tokens/s",
        "cores": cores,
        "kind": "reference" if full is not None else "port",
        "extrapolated": True,
        "factor": total_layers / sample_layers,
        "ms_per_sampled_step": (sum(t_layers) + sum(t_ends) + sum(t_extra)) / max(1, len(t_layers)) * 1e3,
        "sample": (f"{how}: {sample_layers} of {total_layers} layers of the {shape} shape at batch {batch}"
                   f"{'' if batch == full_batch else f' (of {full_batch}; throughput reported at the sampled batch)'}, context "
                   f"{start_ctx + warmup}..{ctx}, fp32, dense (k,v) cache, {steps} decode steps after {warmup} "
                   f"warm-up; per-layer time x {total_layers} + embed/norm/lm_head per step"
                   f"{' (timed by a second pass through a shard that owns no layer)' if full is not None else ''}"),
        "ms_per_step_extrapolated": step_s * 1e3,
        "sample_seconds": sum(t_layers) + sum(t_ends) + sum(t_extra),
    }


def run_reference(args, workload, rank, world):
    if rank != 0:
        return
    shape, batch, ctx = workload["shape"], workload["batch"], workload["context"]
    big = shape == "70b"
    res = cpu_reference_tokens_per_s(shape, batch, ctx, args.steps, args.warmup, sample_layers=1 if big else 2,
                                     sample_batch=16 if big else min(batch, 64))
    line = {
        "impl": "reference",
        "metric": "decode_tokens_per_s",
        "value": res["value"],
        "unit": "tokens/s",
        "n_gpus": args.gpus,
        "steps": args.steps,
        "warmup": args.warmup,
        # what one timed step of this arm really took (a step here is the bounded sample described in
        # cpu_baseline.sample); the full-depth step it stands for is ms_per_step_extrapolated
        "ms_per_step": res["ms_per_sampled_step"],
        "ms_per_step_extrapolated": res["ms_per_step_extrapolated"],
        "higher_is_better": True,
        "scaling": workload["scaling"],
        "vs_baseline": None,
        "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload["name"], "l2": "inputs larger than L2"},
        "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample", "extrapolated", "factor")},
        "e2e": {"value": res["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
# this repo's path
# ---------------------------------------------------------------------------------------------------------
def workload_for(args):
    if args.workload:
        name = args.workload
    else:
        name = "cfg2" if args.gpus == 1 else "cfg4"
    table = {
        "cfg2": dict(shape="8b", batch=64, context=2048, scaling="weak",
                     name="configs[1]: Llama-3-8B-shape random-init bf16, 1xB200, block 16, batch 64, context 2k, decode"),
        "cfg2pp": dict(shape="8b", batch=64 * args.gpus, context=2048, scaling="weak", real_prefill=True,
                       check_sequences=64,
                       name=f"configs[1] weak-scaled: Llama-3-8B-shape random-init bf16, layer ranges over {args.gpus} "
                            f"B200, block 16, batch 64 per GPU = {64 * args.gpus}, context 2k, decode"),
        # sequences of micro-batch 0 re-run on one GPU for the token check: the whole micro-batch when it fits beside the
        # 141 GB of weights (16 sequences at N = 8: same GEMM row counts in both runs), else the first 8
        "cfg4": dict(shape="70b", batch=128, context=4096, scaling="strong",
                     check_sequences=16 if 128 // max(1, args.gpus) <= 16 else 8,
                     name=f"configs[3]: Llama-3-70B-shape layer-sharded over {args.gpus} B200, batch 128, context 4k, decode"),
        "cfg5": dict(shape="8b", batch=8, context=32768, scaling="weak",
                     name="configs[4]: Llama-3-8B-shape long-context decode, batch 8, context 32k"),
        "tiny": dict(shape="tiny", batch=4, context=192, scaling="weak",
                     name="configs[0]: tiny 4-layer model, batch 4, prompt 128, 64 decode tokens"),
    }
    w = dict(table[name])
    w["key"] = name
    if args.batch:
        w["batch"] = args.batch
    if args.context:
        w["context"] = args.context
    return w


def fill_pools_synthetic(engine, batch, ctx_len):
    """Kernel-level fill: K/V ~ N(0, 1) in every allocated slot (used with --synthetic-kv; default is a real prefill)."""
    for pool in engine.pools:
        pool.backing.normal_(0.0, 1.0)
    engine.context_lens[:batch] = ctx_len
    engine._host_lens[:batch] = [min(ctx_len, engine.max_seq_len - 1)] * batch


def cpu_handoff_reference(layers: int = 2, n_blocks: int = 256):
    """The reference's own hand-off (serialize → bytes → deserialize → copy into a pool slot, per block per layer,
    on the host; oracle/handoff_ref.py) timed on a bounded sample of the 512 MiB sequence: ``layers`` of its 32."""
    from oracle import handoff_ref as H, ref_loader

    ref_ser = ref_loader.load("serialization")  # the reference's own TensorSerializer when oracle/_ref is built
    ser = ref_ser.TensorSerializer if ref_ser is not None else None
    Hkv, D, bs, max_blocks = 8, 128, 16, n_blocks + 16
    gen = torch.Generator().manual_seed(7)
    src_k = [torch.randn(max_blocks, Hkv, bs, D, generator=gen).to(torch.bfloat16) for _ in range(layers)]
    src_v = [torch.randn(max_blocks, Hkv, bs, D, generator=gen).to(torch.bfloat16) for _ in range(layers)]
    dst_k = [torch.zeros_like(t) for t in src_k]
    dst_v = [torch.zeros_like(t) for t in src_v]
    st = torch.randperm(max_blocks, generator=gen)[:n_blocks].tolist()
    dt = torch.randperm(max_blocks, generator=gen)[:n_blocks].tolist()
    H.handoff_sequence(src_k[:1], src_v[:1], st[:16], dst_k[:1], dst_v[:1], dt[:16], serializer=ser)  # warm-up
    t0 = time.perf_counter()
    moved = H.handoff_sequence(src_k, src_v, st, dst_k, dst_v, dt, serializer=ser)
    sec = time.perf_counter() - t0
    # the device ops' own definition on the host: index_select (pack) and index_copy_ (unpack), oracle/paged_ref.py
    from oracle import paged_ref as R

    tbl = torch.tensor(st, dtype=torch.int64)
    t0 = time.perf_counter()
    packed = [R.pack_ref(k, v, tbl) for k, v in zip(src_k, src_v)]
    t1 = time.perf_counter()
    for pk, k, v in zip(packed, dst_k, dst_v):
        R.unpack_ref(pk, k, v, tbl)
    t2 = time.perf_counter()
    nbytes = 2 * sum(p.numel() * p.element_size() for p in packed)  # read + written, as for the GPU figures
    return {"GBps": moved / sec / 1e9, "kind": "reference" if ser is not None else "port", "cores": 1,
            "cpu_pack_GBps": nbytes / (t1 - t0) / 1e9, "cpu_unpack_GBps": nbytes / (t2 - t1) / 1e9,
            "cpu_pack_threads": torch.get_num_threads(),
            "sample": f"{layers} of 32 layers x {n_blocks} blocks x 64 KiB = {moved / 2**20:.0f} MiB on the wire "
                      f"(fp16 bytes), host serializer per block, {sec:.2f} s"}


def measure_pack_unpack(dev):
    """Second hand-off figure available on one GPU: pack / unpack of one 4k-token 8B-shape sequence
    (32 layers x 256 blocks x 64 KiB = 512 MiB), GB/s of bytes read + written."""
    from dgi_b200 import ops

    L_, Hkv, D, bs, n, max_blocks = 32, 8, 128, 16, 256, 320
    backing = torch.empty((L_, 2, max_blocks, Hkv, bs, D), device=dev, dtype=torch.bfloat16).normal_()
    kp = ops.pool_pointer_array([backing[l, 0] for l in range(L_)])
    vp = ops.pool_pointer_array([backing[l, 1] for l in range(L_)])
    table = torch.randperm(max_blocks, generator=torch.Generator().manual_seed(3))[:n].to(torch.int32).to(dev)
    packed = torch.empty((L_, 2, n, Hkv, bs, D), dtype=torch.bfloat16, device=dev)
    out = {}
    for name, fn in (("pack", lambda: ops.kv_pack_layers(kp, vp, (max_blocks, Hkv, bs, D), torch.bfloat16, table, out=packed)),
                     ("unpack", lambda: ops.kv_unpack_layers(packed, kp, vp, (max_blocks, Hkv, bs, D), table))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        out[name + "_GBps"] = 2 * packed.numel() * 2 / ms / 1e6
    out["payload_MiB"] = packed.numel() * 2 / 2 ** 20
    out["unit"] = "GB/s of bytes read + written (HBM roofline: the same copy peak)"
    return out


def _single_leg(args, workload):
    """One timed single-GPU leg; returns the JSON line as a dict."""
    import dgi_b200  # noqa: F401
    from dgi_b200 import ops
    from dgi_b200.engine import PagedDecodeEngine
    from dgi_b200.worker.distributed.model_shard import ModelShard

    dev = "cuda:0"
    torch.cuda.set_device(0)
    cfg = ShapeConfig(**SHAPES[workload["shape"]])
    B, ctx = workload["batch"], workload["context"]
    K, W = args.steps, args.warmup
    total_steps = 2 * (K + W) + K + 2  # device loop + e2e loop + instrumented pass
    P = ctx - total_steps
    if P < 1:
        raise SystemExit("context too short for the requested steps")
    t_setup = time.time()
    shard = ModelShard.from_config(cfg, 0, cfg.num_hidden_layers, device=dev, dtype=torch.bfloat16, seed=0)
    eng = PagedDecodeEngine([shard], max_batch=B, max_seq_len=ctx, block_size=16, prefill_chunk=args.prefill_chunk,
                            fuse_projections=not args.no_fuse_projections)
    ragged = None
    synthetic_kv = args.synthetic_kv
    if workload["key"] == "cfg5":
        # config 5: one shared prefix of 24576 tokens (1536 blocks, reference-counted) + unique tails of
        # 8192 - 512 i tokens: ragged block tables, cascade attention, synthetic KV
        shared = (ctx * 3 // 4) // 16 * 16
        eng.allocate(B, ctx, shared_prefix_len=shared)
        ragged = [shared + (ctx - shared) - 512 * i * (ctx // 32768 or 1) - total_steps for i in range(B)]
        synthetic_kv = True
    else:
        eng.allocate(B, ctx)
    prompts = torch.randint(0, cfg.vocab_size, (B, P), generator=torch.Generator().manual_seed(1234))
    if synthetic_kv:
        fill_pools_synthetic(eng, B, P)
        if ragged is not None:
            eng.context_lens[:B] = torch.tensor(ragged, dtype=torch.int32)
        first = prompts[:, -1].to(dev)
    else:
        first = eng.prefill(prompts)
    torch.cuda.synchronize()
    setup_s = time.time() - t_setup

    sampler = ClockSampler(0)
    # ---- device-resident loop: K chained graph replays --------------------------------------------------
    eng.ids_buf[:B, 0].copy_(first)
    for _ in range(W):
        eng.decode_step_device()
    torch.cuda.synchronize()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        eng.decode_step_device()
    e1.record()
    torch.cuda.synchronize()
    dev_ms = e0.elapsed_time(e1)

    # ---- end to end: host ids in (pinned H2D) and out (D2H) every step through the public call ------------
    nxt = eng.next_tokens[:B].cpu()
    for _ in range(W):
        nxt = eng.decode_step(nxt)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(K):
        nxt = eng.decode_step(nxt)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop()

    # ---- instrumented pass: same steps launched eagerly with CUDA events around every attention launch ----
    eng.use_cuda_graph = False
    ops.ATTN_EVENTS = []
    eng.ids_buf[:B, 0].copy_(nxt.to(dev))
    n_inst = min(K, 4)
    for _ in range(n_inst):
        # park the GPU for ~30 ms first so the host is a whole step ahead: the event pairs then bracket pure
        # device time instead of host launch latency
        torch.cuda._sleep(60_000_000)
        eng.decode_step_device()
    torch.cuda.synchronize()
    ev = ops.ATTN_EVENTS
    ops.ATTN_EVENTS = None
    attn_ms = [a.elapsed_time(b) for a, b in ev]
    attn_ms_mean = statistics.mean(attn_ms) if attn_ms else float("nan")
    ctx_now = P + 1 + 2 * (K + W) + n_inst // 2  # mean context during the instrumented pass
    es = 2
    algo_bytes = B * ctx_now * 2 * eng.Hkv * eng.D * es + 2 * B * eng.Hq * eng.D * es
    if ragged is not None:
        # headline on UNIQUE bytes (shared prefix counted once), logical bytes reported beside it
        steps_done = 2 * (K + W) + n_inst // 2 + 1
        shared_tok = eng.prefix_blocks * 16
        logical = sum(r + steps_done for r in ragged) * 2 * eng.Hkv * eng.D * es + 2 * B * eng.Hq * eng.D * es
        algo_bytes = logical - (B - 1) * shared_tok * 2 * eng.Hkv * eng.D * es
    peaks, peak_kind = measured_peaks()
    achieved = algo_bytes / (attn_ms_mean * 1e-3) / 1e9
    traffic, traffic_source = None, None
    tpath = os.path.join(ROOT, "profiles", "attn_traffic.json")
    if os.path.exists(tpath):
        try:
            tj = json.load(open(tpath))
            traffic = tj.get(workload["key"])
            src = tj.get("source")
            traffic_source = (src.get(workload["key"]) if isinstance(src, dict) else src) if traffic is not None else None
        except Exception:  # noqa: BLE001
            traffic = None
    # this repo's kernels per step: per layer add_rmsnorm x2, the one-launch RoPE + append + attention (+ combine when
    # split), silu_mul; + the final norm and the argmax over the logits.  With a shared prefix the one-launch step goes
    # through the cascade kernel, which always has a combine.
    per_layer = 5 if eng.state.prefix_blocks else 4 + (1 if eng.num_splits > 1 else 0)
    launches_per_step = cfg.num_hidden_layers * per_layer + 2

    line = {
        "metric": "decode_tokens_per_s",
        "value": B * K / (dev_ms * 1e-3),
        "unit": "tokens/s",
        "n_gpus": 1,
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
            "block_size": 16,
            "batch": B,
            "context": (f"{P + 1 + W}..{P + W + K} during the timed steps" if ragged is None else
                        f"shared prefix {eng.prefix_blocks * 16} + ragged tails, longest sequence {ragged[0] + 1 + W}.."
                        f"{ragged[0] + W + K}, shortest {ragged[-1] + 1 + W}..{ragged[-1] + W + K} during the timed steps"),
            "kv_fill": "synthetic normal(0,1)" if synthetic_kv else "real prefill of random prompts",
            "l2": "inputs larger than L2 (all weights and the batch's KV are streamed every step)",
            "attn_splits": (f"cascade {eng.state.prefix_splits}+{eng.state.suffix_splits}, {eng.prefix_blocks} shared blocks"
                            if eng.state.prefix_blocks else eng.num_splits),
            "projections": "separate q/k/v and gate/up GEMMs" if args.no_fuse_projections else "q|k|v and gate|up fused into one GEMM each",
            "setup_s": round(setup_s, 1),
        },
        "clocks": clocks,
        "e2e": {"value": B * K / e2e_s, "unit": "tokens/s", "h2d_bytes_per_step": B * 8, "d2h_bytes_per_step": B * 8,
                "ms_per_step": e2e_s / K * 1e3},
        "gpu_launches": launches_per_step * K,
        "roofline": {
            "bound": "hbm",
            "kernel": ("paged_attn_cascade_kernel + attn_combine_kernel (b200_paged_decode_attn_prefix), unique bytes"
                       if eng.state.prefix_blocks else "paged_attn_tma_kernel (b200_paged_decode_attn)"),
            "achieved": achieved,
            "peak": peaks["hbm_gbs"],
            "peak_source": f"MEASURED_PEAKS.json ({peak_kind})",
            "unit": "GB/s",
            "frac": achieved / peaks["hbm_gbs"],
            "frac_of_nominal_8TBs": achieved / 8000.0,
            "traffic": traffic,
            "traffic_source": traffic_source,
            "algorithmic_bytes_per_launch": algo_bytes,
            "launch_ms": attn_ms_mean,
            "launches_timed": len(attn_ms),
            "how": ("CUDA events around each attention launch of eagerly launched decode steps that continue the "
                    f"timed loop's state (context {ctx_now - n_inst // 2}..{ctx_now - n_inst // 2 + n_inst - 1}, "
                    f"{2 * (K + W)} tokens past the start of the device-timed window)"),
        },
    }
    del eng, shard
    import gc

    gc.collect()
    torch.cuda.empty_cache()
    return line


def config1_side_by_side():
    """BASELINE configs[0] IN FULL, un-extrapolated: tiny 4-layer model in 2 layer-range shards, batch 4, prompt 128,
    64 greedy tokens, fp32 — the reference's torch CPU path (its own ModelShard from oracle/_ref, else the oracle port, + legacy
    HF layer with dense (k, v) tuples) on the host cores, and this repo's paged CUDA path on the GPU, both timed
    over the 63 decode steps after the prompt pass, tokens compared."""
    import dgi_b200  # noqa: F401
    from dgi_b200.engine import PagedDecodeEngine
    from dgi_b200.testing import LlamaShapeConfig, build_product_shards, seeded_state_dict
    from oracle import llama_ref as L

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = L.RefConfig(**L.TINY)
    sd = L.make_weights(cfg, seed=0)
    ranges = [(0, 2), (2, 4)]
    B, P, N = 4, 128, 64
    prompts = torch.randint(0, cfg.vocab_size, (B, P), generator=torch.Generator().manual_seed(1234))
    from oracle import ref_loader

    by_reference = ref_loader.available("model_shard")
    # the reference's own ModelShard objects (oracle/_ref bytecode) when built, else the oracle's restatement of them
    shards = ref_loader.reference_shards(cfg, sd, ranges, L) if by_reference else L.build_shards(cfg, sd, ranges)
    L.greedy_decode(shards, prompts[:, :16], 2)  # warm-up
    t0 = time.perf_counter()
    L.greedy_decode(shards, prompts, 1)  # prompt pass + first token
    t_prompt = time.perf_counter() - t0
    t0 = time.perf_counter()
    want = L.greedy_decode(shards, prompts, N)
    t_all = time.perf_counter() - t0
    cpu_decode_s = max(t_all - t_prompt, 1e-9)
    gcfg = LlamaShapeConfig(**L.TINY)
    gshards = build_product_shards(gcfg, seeded_state_dict(gcfg, 0), ranges, "cuda:0", torch.float32)
    eng = PagedDecodeEngine(gshards, max_batch=B, max_seq_len=P + N, prefill_mode="eager")
    eng.generate(prompts, 4)  # warm-up: graph capture
    eng.allocate(B, P + N)
    nxt = eng.prefill(prompts).cpu()
    got = [nxt]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(N - 1):
        nxt = eng.decode_step(nxt)  # host ids in and out every step
        got.append(nxt)
    torch.cuda.synchronize()
    gpu_s = time.perf_counter() - t0
    got = torch.stack(got, dim=1)
    # the same 63 steps with the tokens staying on the device (chained CUDA-graph steps, as `value` is measured)
    eng.allocate(B, P + N)
    first = eng.prefill(prompts)
    eng.ids_buf[:B, 0].copy_(first)
    dev_toks = [first.clone()]
    eng.decode_step_device()  # captures the chained graph
    dev_toks.append(eng.next_tokens[:B].clone())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(N - 2):
        eng.decode_step_device()
        dev_toks.append(eng.next_tokens[:B].clone())
    e1.record()
    torch.cuda.synchronize()
    gpu_dev_s = e0.elapsed_time(e1) / 1e3 * (N - 1) / (N - 2)  # 62 timed steps scaled to the 63 of the other legs
    got_dev = torch.stack(dev_toks, dim=1).cpu()
    return {
        "workload": "configs[0]: tiny 4-layer model (d=512, 8 heads, 2 KV heads), 2 layer-range shards, batch 4, prompt 128, "
                    "64 greedy tokens, fp32",
        "value": B * (N - 1) / cpu_decode_s, "unit": "tokens/s", "cores": cores,
        "kind": "reference" if by_reference else "port", "extrapolated": False,
        "sample": f"the whole configuration: {N - 1} decode steps of batch {B} after the prompt pass, {cpu_decode_s:.2f} s",
        "gpu_value": B * (N - 1) / gpu_s, "gpu_path": "PagedDecodeEngine.decode_step (host ids in / out per step), 1 GPU",
        "gpu_value_device_resident": B * (N - 1) / gpu_dev_s,
        "gpu_path_device_resident": "PagedDecodeEngine.decode_step_device (chained CUDA-graph steps, CUDA events)",
        "tokens_identical": bool(torch.equal(got, want)) and bool(torch.equal(got_dev, want)),
    }


def reference_api_leg(workload):
    """configs[1] driven the way a user of the reference drives it (tools/session_bench.py): one hop over all layers,
    in-process transport, token ids from pinned host memory in and tokens back to the host every step."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("session_bench", os.path.join(ROOT, "tools", "session_bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    torch.cuda.empty_cache()
    res = mod.run(workload["shape"], SHAPES[workload["shape"]]["num_hidden_layers"], 1, workload["batch"],
                  workload["context"], steps=32, fuse=True, host_io=True)
    torch.cuda.empty_cache()
    return {"value": res["tokens_per_second"], "unit": "tokens/s", "ms_per_step": res["ms_per_step"],
            "h2d_bytes_per_step": 8 * workload["batch"], "d2h_bytes_per_step": 8 * workload["batch"],
            "kv_fill": "synthetic normal(0,1)", "path": res["path"]}


def run_single(args, workload):
    line = _single_leg(args, workload)
    dev = "cuda:0"
    if not args.workload and not args.batch and not args.context and not args.no_riders:
        # configs[4] rides along on the N = 1 line: long-context decode over a shared prefix (cascade kernel)
        args5 = argparse.Namespace(**vars(args))
        args5.workload = "cfg5"
        five = _single_leg(args5, workload_for(args5))
        line["cfg5"] = {"workload": five["config"]["workload"], "config": five["config"],
                        **{k: five[k] for k in ("value", "unit", "ms_per_step", "e2e", "gpu_launches", "roofline")}}
    line["kv_pack"] = measure_pack_unpack(dev)
    if not args.workload and not args.batch and not args.context and not args.no_riders:
        # the same workload through the reference's own call stack (DistributedInferenceSession.step -> WorkerSession ->
        # InferenceServicer.Forward -> ModelShard.forward; get_logits + argmax by the caller), host ids in and out
        line["reference_api"] = reference_api_leg(workload)
    if not args.no_cpu_baseline:
        res = cpu_reference_tokens_per_s(workload["shape"], workload["batch"], workload["context"], steps=10,
                                         warmup=2)  # about 10 s of host work
        line["cpu_baseline"] = {k: res[k] for k in ("value", "unit", "cores", "kind", "sample", "extrapolated", "factor")}
        line["kv_pack"]["cpu_reference_handoff"] = cpu_handoff_reference()
        line["cpu_baseline_cfg1"] = config1_side_by_side()
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="", help="cfg2 | cfg4 | cfg5 | tiny (default: by --gpus)")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--context", type=int, default=0)
    ap.add_argument("--synthetic-kv", action="store_true", help="fill KV pools with N(0,1) instead of a real prefill")
    ap.add_argument("--prefill-chunk", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-fuse-projections", action="store_true",
                    help="keep q/k/v and gate/up as separate GEMMs (the reference's GEMM shapes)")
    ap.add_argument("--micro-batches", type=int, default=0, help="pipeline micro-batches (default: one per rank)")
    ap.add_argument("--real-prefill", action="store_true", help="multi-GPU: run a real prompt pass instead of synthetic KV")
    ap.add_argument("--no-riders", action="store_true",
                    help="default runs: skip the extra legs appended to the line (N = 1: configs[4]; N > 1: the "
                         "weak-scaled 8B pipeline)")
    ap.add_argument("--no-parity", action="store_true", help="multi-GPU: skip the token checks of the run")
    ap.add_argument("--no-vocab-parallel-head", action="store_true",
                    help="multi-GPU: keep lm_head + argmax on the last stage (the round-1 arrangement)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    workload = workload_for(args)
    if args.impl == "reference":
        run_reference(args, workload, rank, world)
        return
    if world == 1 and args.gpus == 1:
        run_single(args, workload)
    else:
        from dgi_b200.bench_pipeline import run_pipeline  # multi-GPU layer-sharded decode

        extra = None
        if not args.workload and not args.no_riders and not args.batch and not args.context:
            args.workload = "cfg2pp"  # the N = 1 model weak-scaled over the same ranks rides along in the line
            extra = workload_for(args)
            args.workload = ""
        run_pipeline(args, workload, SHAPES, ShapeConfig, ClockSampler, measured_peaks, cpu_reference_tokens_per_s,
                     cpu_handoff=cpu_handoff_reference, extra_workload=extra)


if __name__ == "__main__":
    main()
