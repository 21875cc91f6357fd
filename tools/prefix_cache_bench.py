#!/usr/bin/env python
"""What the cross-request prefix cache buys: time of the prompt pass (TTFT) for a batch whose prompts start with the same
system prompt, cold (nothing cached), warm (an earlier call left the system prompt's blocks in the pool) and without
any sharing.  8B shape, bf16; CUDA-event timed."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import dgi_b200  # noqa: E402,F401
from bench import SHAPES, ShapeConfig  # noqa: E402
from dgi_b200.engine import PagedDecodeEngine  # noqa: E402
from dgi_b200.worker.distributed.model_shard import ModelShard  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--system", type=int, default=3072)
    ap.add_argument("--tail", type=int, default=1024)
    ap.add_argument("--new", type=int, default=16)
    a = ap.parse_args()
    dev = "cuda:0"
    cfg = ShapeConfig(**SHAPES["8b"])
    shard = ModelShard.from_config(cfg, 0, cfg.num_hidden_layers, device=dev, dtype=torch.bfloat16, seed=0)
    P = a.system + a.tail
    eng = PagedDecodeEngine([shard], a.batch, P + a.new, prefill_chunk=4, fuse_projections=True, prefix_cache=True,
                            spare_blocks=(a.system // 16) * 2)
    gen = torch.Generator().manual_seed(1)
    system = torch.randint(0, cfg.vocab_size, (a.system,), generator=gen)

    def batch_of(seed):
        tails = torch.randint(0, cfg.vocab_size, (a.batch, a.tail), generator=torch.Generator().manual_seed(seed))
        return torch.cat([system[None].expand(a.batch, -1), tails], dim=1)

    def ttft(prompts, shared, use_cache):
        pre = prompts[0, :shared].tolist() if (use_cache and shared) else None
        eng.allocate(a.batch, P + a.new, shared_prefix_len=shared, prefix_tokens=pre)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        first = eng.prefill(prompts)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), eng.cached_blocks, first.tolist()

    def forget():
        eng.release()
        for idx in eng.prefix_index:
            idx.drop_all()

    ttft(batch_of(7), 0, False)  # warm-up of every shape that is timed below (library plan selection is one-off)
    ttft(batch_of(7), a.system, True)
    ttft(batch_of(7), a.system, True)
    forget()
    res = {"batch": a.batch, "system_tokens": a.system, "tail_tokens": a.tail}
    res["no_sharing_ms"], _, t0 = ttft(batch_of(11), 0, False)
    res["shared_cold_ms"], c, t1 = ttft(batch_of(11), a.system, True)
    assert c == 0
    res["shared_warm_ms"], c, t2 = ttft(batch_of(11), a.system, True)
    res["cached_blocks_warm"] = c
    res["other_tails_warm_ms"], _, _ = ttft(batch_of(12), a.system, True)
    res["first_tokens_equal_cold_warm"] = t1 == t2
    res["first_tokens_equal_no_sharing"] = t0 == t1
    print(json.dumps(res))


if __name__ == "__main__":
    main()
