#!/usr/bin/env python
"""Where a prompt pass spends its time (8B shape, one chunk of n x P tokens): per-piece CUDA-event timings and a
torch-profiler table of one layer.  Diagnostic only."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import dgi_b200  # noqa: E402,F401
from bench import SHAPES, ShapeConfig  # noqa: E402
from dgi_b200.engine import PagedDecodeEngine  # noqa: E402
from dgi_b200.worker.distributed.model_shard import ModelShard  # noqa: E402


def timeit(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    dev = "cuda:0"
    n, P = int(os.environ.get("N", 4)), int(os.environ.get("P", 4096))
    Hq, Hkv, D = 32, 8, 128
    q = torch.randn(n, Hq, P, D, device=dev, dtype=torch.bfloat16)
    k = torch.randn(n, Hkv, P, D, device=dev, dtype=torch.bfloat16)
    v = torch.randn(n, Hkv, P, D, device=dev, dtype=torch.bfloat16)
    print("sdpa enable_gqa      ms", timeit(lambda: F.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=True)))
    ke, ve = k.repeat_interleave(4, 1), v.repeat_interleave(4, 1)
    print("sdpa expanded kv     ms", timeit(lambda: F.scaled_dot_product_attention(q, ke, ve, is_causal=True)))
    print("  + the expansion    ms", timeit(lambda: (k.repeat_interleave(4, 1), v.repeat_interleave(4, 1))))
    from torch.nn.attention import SDPBackend, sdpa_kernel
    for be in (SDPBackend.FLASH_ATTENTION, SDPBackend.CUDNN_ATTENTION, SDPBackend.EFFICIENT_ATTENTION):
        for gqa in (True, False):
            try:
                with sdpa_kernel(be):
                    kk, vv = (k, v) if gqa else (ke, ve)
                    t = timeit(lambda: F.scaled_dot_product_attention(q, kk, vv, is_causal=True, enable_gqa=gqa))
                print(f"sdpa {be.name:22s} gqa={gqa} ms", t)
            except Exception as exc:  # noqa: BLE001
                print(f"sdpa {be.name:22s} gqa={gqa} unavailable: {str(exc)[:80]}")
    x = torch.randn(n * P, 4096, device=dev, dtype=torch.bfloat16)
    for name, out_f in (("qkv", 6144), ("o", 4096), ("gate_up", 28672)):
        w = torch.randn(out_f, 4096, device=dev, dtype=torch.bfloat16) * 0.02
        wk = w.t().contiguous().t()
        t1, t2 = timeit(lambda: F.linear(x, w)), timeit(lambda: F.linear(x, wk))
        fl = 2 * n * P * 4096 * out_f / 1e9
        print(f"gemm {name:8s} row-major {t1:.3f} ms ({fl / t1:.0f} TF/s)   K-major {t2:.3f} ms ({fl / t2:.0f} TF/s)")
    xd = torch.randn(n * P, 14336, device=dev, dtype=torch.bfloat16)
    w = torch.randn(4096, 14336, device=dev, dtype=torch.bfloat16) * 0.02
    wk = w.t().contiguous().t()
    t1, t2 = timeit(lambda: F.linear(xd, w)), timeit(lambda: F.linear(xd, wk))
    fl = 2 * n * P * 4096 * 14336 / 1e9
    print(f"gemm down     row-major {t1:.3f} ms ({fl / t1:.0f} TF/s)   K-major {t2:.3f} ms ({fl / t2:.0f} TF/s)")
    del q, k, v, ke, ve, x, xd, w, wk

    shape = dict(SHAPES["8b"])
    shape["num_hidden_layers"] = 4
    cfg = ShapeConfig(**shape)
    for fuse in (False, True):
        shard = ModelShard.from_config(cfg, 0, 4, device=dev, dtype=torch.bfloat16, seed=0)
        eng = PagedDecodeEngine([shard], n, P + 16, prefill_chunk=n, fuse_projections=fuse)
        eng.allocate(n, P + 16)
        ids = torch.randint(0, cfg.vocab_size, (n, P), device=dev)
        print(f"prefill_chunk 4 layers fuse={fuse}: {timeit(lambda: eng.prefill_chunk(0, n, ids), 3) / 4:.3f} ms per layer")
        if fuse:
            from torch.profiler import ProfilerActivity, profile
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                eng.prefill_chunk(0, n, ids)
                torch.cuda.synchronize()
            print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
        del eng, shard


if __name__ == "__main__":
    main()
