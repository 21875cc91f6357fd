#!/usr/bin/env python
"""How fast does torch stream the decode-time weight matrices?  F.linear at M = 16 / 64 rows for the 8B- and
70B-shape projections, default cuBLAS vs cuBLASLt, GB/s of weight bytes.  (The projections stay torch GEMMs by
design; this only tells which library path bench.py should select.)"""
import json

import torch
import torch.nn.functional as F

DEV = "cuda:0"


def bench(fn, n=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


shapes = {"8b qkv": (6144, 4096), "8b o": (4096, 4096), "8b gate_up": (28672, 4096), "8b down": (4096, 14336),
          "70b qkv": (10240, 8192), "70b o": (8192, 8192), "70b gate_up": (57344, 8192), "70b down": (8192, 28672)}
for lib in ("cublas", "cublaslt"):
    torch.backends.cuda.preferred_blas_library(lib)
    for name, (n, k) in shapes.items():
        # rotate over several copies so the weights never sit in L2
        copies = max(2, int(400e6 // (n * k * 2)) + 1)
        ws = [torch.randn(n, k, device=DEV, dtype=torch.bfloat16) for _ in range(copies)]
        for m in (16, 64):
            x = torch.randn(m, k, device=DEV, dtype=torch.bfloat16)
            i = [0]

            def run():
                F.linear(x, ws[i[0] % copies])
                i[0] += 1

            ms = bench(run)
            print(json.dumps({"lib": lib, "gemm": name, "M": m, "N": n, "K": k, "us": round(ms * 1e3, 2),
                              "weight_GBps": round(n * k * 2 / ms / 1e6, 1)}), flush=True)
        del ws
