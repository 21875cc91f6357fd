#!/usr/bin/env python
"""Randomised end-to-end check of the decode engine against the oracle (tiny fp32 model, tokens must be identical).

One long-lived engine with a prefix cache and a deliberately small pool serves a random stream of calls: 1-4 prompts of
random lengths, with or without one of three shared system prompts, 1-10 new tokens.  Between calls blocks are shared,
cached, evicted under pressure and reused.  Every returned sequence is compared with the oracle's greedy decode of
that prompt alone (dense (k, v) cache, CPU).  A mismatch is accepted only where the oracle's own top-2 logit margin at
that step is below 1e-4 (an arithmetic tie); the run reports how many such ties it saw.
usage: python tools/fuzz_engine.py [--calls N] [--seed S]
"""
import argparse
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import dgi_b200  # noqa: E402,F401
from dgi_b200.engine import PagedDecodeEngine  # noqa: E402
from dgi_b200.testing import build_product_shards  # noqa: E402
from oracle import llama_ref as L  # noqa: E402

DEV = "cuda:0"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--calls", type=int, default=40)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    rng = random.Random(a.seed)
    g = json.load(open(os.path.join(ROOT, "tests", "golden", "tiny_decode.json")))
    cfg = L.RefConfig(**g["config"])
    sd = L.make_weights(cfg, seed=0)
    ref = L.build_shards(cfg, sd, [(0, 4)])
    shards = build_product_shards(cfg, sd, [(0, 2), (2, 4)], DEV, torch.float32)
    eng = PagedDecodeEngine(shards, 4, 176, prefill_mode="eager", prefix_cache=True, spare_blocks=0)
    systems = [[rng.randrange(cfg.vocab_size) for _ in range(n)] for n in (23, 64, 90)]
    memo = {}
    ties = seqs = hits = 0
    for call in range(a.calls):
        n = rng.randint(1, 4)
        sysp = rng.choice(systems + [[]])
        longest = 60 if sysp else 160  # long unrelated prompts put the pool under pressure: cached blocks must give way
        prompts = [sysp + [rng.randrange(cfg.vocab_size) for _ in range(rng.randint(1, longest))] for _ in range(n)]
        n_new = rng.randint(1, 10)
        share = len(sysp)
        got = eng.generate_ragged(prompts, n_new, shared_prefix_len=share)
        hits += eng.cached_blocks
        for p, toks in zip(prompts, got):
            key = (tuple(p), n_new)
            if key not in memo:
                memo[key] = L.greedy_decode(ref, torch.tensor([p]), n_new, return_logits=True)
            want, logits = memo[key]
            want = want[0].tolist()
            seqs += 1
            if toks != want:
                j = next(i for i, (x, y) in enumerate(zip(toks, want)) if x != y)
                top2 = logits[0, j].topk(2).values
                margin = float(top2[0] - top2[1])
                assert margin < 1e-4, (f"call {call}: tokens differ at step {j} with oracle margin {margin:.3e}; "
                                       f"prompt len {len(p)}, shared {share}, cached blocks {eng.cached_blocks}")
                ties += 1
    ev = sum(c.get_stats()["evictions"] for pool in eng.pools for c in pool._layer_caches)
    print(f"{a.calls} calls, {seqs} sequences identical to the oracle (seed {a.seed}); arithmetic ties {ties}; "
          f"cached blocks reused {hits}; evictions {ev}")


if __name__ == "__main__":
    main()
