#!/usr/bin/env python
"""Does the weight layout change how fast torch streams a decode-time projection?  y = x @ W^T with W stored
[N, K] (nn.Linear's layout, cuBLAS 'TN') against W stored [K, N] ('NN'), M = 16 / 64 rows, bf16."""
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
          "70b qkv": (10240, 8192), "70b o": (8192, 8192)}
for name, (n, k) in shapes.items():
    copies = max(2, int(400e6 // (n * k * 2)) + 1)
    w_nk = [torch.randn(n, k, device=DEV, dtype=torch.bfloat16) for _ in range(copies)]
    w_kn = [w.t().contiguous() for w in w_nk]
    for m in (16, 64):
        x = torch.randn(m, k, device=DEV, dtype=torch.bfloat16)
        out = torch.empty(m, n, device=DEV, dtype=torch.bfloat16)
        i = [0]

        def tn():
            F.linear(x, w_nk[i[0] % copies])
            i[0] += 1

        def nn_():
            torch.matmul(x, w_kn[i[0] % copies], out=out)
            i[0] += 1

        a, b = bench(tn), bench(nn_)
        same = torch.equal(F.linear(x, w_nk[0]), torch.matmul(x, w_kn[0]))
        print(json.dumps({"gemm": name, "M": m, "TN_us": round(a * 1e3, 2), "NN_us": round(b * 1e3, 2),
                          "TN_GBps": round(n * k * 2 / a / 1e6), "NN_GBps": round(n * k * 2 / b / 1e6),
                          "bit_identical": same}), flush=True)
    del w_nk, w_kn
