#!/usr/bin/env python
"""P/D KV hand-off benchmark (BASELINE config 3): 8B-shape, 4k-token prompts = 256 blocks x 32 layers x 64 KiB =
512 MiB per sequence, prefill rank i -> decode rank i + N/2 over NVLink.  Run under torchrun with an even N.

Reports, per pair (max over the pair's two ranks, CUDA-event timed between barriers) and aggregated over pairs:
  nccl path : pack -> dist.isend/irecv -> unpack, for several layer-group counts (pipelining)
  push path : one b200_kv_move_blocks_layers launch writing into the peer's pools (symmetric memory)
GB/s = payload bytes / time, payload counted once.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import dgi_b200  # noqa: E402,F401
from dgi_b200.kv_transfer import KVTransferEndpoint, PeerPushEndpoint  # noqa: E402
from dgi_b200.worker.distributed.kv_cache import KVCachePool  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seqs", type=int, default=8)
    ap.add_argument("--prompt", type=int, default=4096)
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--skip-push", action="store_true")
    ap.add_argument("--quick", action="store_true", help="NCCL path with one layer group only")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    dist.init_process_group("nccl", device_id=torch.device(dev))
    half = world // 2
    sender = rank < half
    peer = rank + half if sender else rank - half
    pair_group = None
    for i in range(half):  # one communicator per (prefill, decode) pair
        g = dist.new_group([i, i + half])
        if rank in (i, i + half):
            pair_group = g
    Hkv, D, bs = 8, 128, 16
    n_blocks = (a.prompt + bs - 1) // bs
    max_blocks = n_blocks * a.seqs + 64
    results = []

    def run(name, fn, payload):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.barrier()
        if rank == 0:
            gbs = payload / (ms.item() * 1e-3) / 1e9
            results.append({"path": name, "ms": round(ms.item(), 3), "payload_MiB_per_pair": payload / 2 ** 20,
                            "GBps_per_pair": round(gbs, 1), "GBps_aggregate": round(gbs * half, 1),
                            "frac_of_900GBps_nvlink": round(gbs / 900.0, 3)})

    # ---- NCCL path ------------------------------------------------------------------------------------------
    pool = KVCachePool(a.layers, Hkv, D, bs, max_blocks, device=dev, dtype=torch.bfloat16)
    if sender:
        pool.backing.normal_()
    gen = torch.Generator().manual_seed(rank)
    tables = [torch.randperm(max_blocks, generator=gen)[:n_blocks].to(torch.int32).to(dev) for _ in range(a.seqs)]
    # bare wire rate: the same bytes as ready-made contiguous buffers, no pack / unpack — what NCCL's send/recv itself
    # sustains between the two GPUs (the ceiling of the packed path)
    ep = KVTransferEndpoint(pool, chunks=1)
    flat = [ep.staging(n_blocks)[0] for _ in range(2)]

    def bare():
        works = []
        for i in range(a.seqs):
            if sender:
                works.append(dist.isend(flat[i % 2], dst=peer, group=pair_group))
            else:
                works.append(dist.irecv(flat[i % 2], src=peer, group=pair_group))
        for w in works:
            w.wait()

    run("bare nccl send/recv of the packed bytes (no pack / unpack)", bare, ep.payload_bytes(n_blocks) * a.seqs)
    del flat
    for chunks in ((1,) if a.quick else (1, 2, 4, 8)):
        ep = KVTransferEndpoint(pool, chunks=chunks)
        bufs = [ep.staging(n_blocks) for _ in range(2)]
        payload = ep.payload_bytes(n_blocks) * a.seqs

        def nccl_path():
            works = []
            for i, t in enumerate(tables):
                if sender:
                    if len(works) >= 2:  # double-buffered staging
                        for w in works.pop(0):
                            w.wait()
                    works.append(ep.send(t, dst=peer, group=pair_group, bufs=bufs[i % 2]))
                else:
                    ep.recv(t, src=peer, group=pair_group, bufs=bufs[i % 2])
            for ws in works:
                for w in ws:
                    w.wait()

        run(f"nccl pack->send/recv->unpack, {chunks} layer group(s)", nccl_path, payload)
    del pool, ep, bufs
    torch.cuda.empty_cache()

    # ---- peer push path -------------------------------------------------------------------------------------
    if not a.skip_push:
        pp = PeerPushEndpoint(a.layers, Hkv, D, bs, max_blocks, dev, torch.bfloat16)
        if sender:
            pp.pool.backing.normal_()
        dst_tables = [torch.randperm(max_blocks, generator=torch.Generator().manual_seed(peer))[:n_blocks].to(torch.int32).to(dev)
                      for _ in range(a.seqs)]  # the peer's tables (same seed scheme as above)
        src_tables = tables
        payload = pp.payload_bytes(n_blocks) * a.seqs

        def push_path():
            if sender:
                for st, dt in zip(src_tables, dst_tables):
                    pp.push(st, dt, peer)
            pp.barrier()

        run("peer push (b200_kv_move_blocks_layers into symmetric-memory pools)", push_path, payload)
    if rank == 0:
        env = {k: v for k, v in os.environ.items() if k.startswith("NCCL_") and k != "NCCL_DEBUG"}
        for r in results:
            if env:
                r["nccl_env"] = env
            print(json.dumps(r), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
