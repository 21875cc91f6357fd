"""Prompt pass that continues a cached prefix: how should the T new tokens attend to S = past + T tokens?
(a) dense [T, S] additive mask into torch SDPA (what paged_layer._prefill_attention did),
(b) torch.nn.attention.bias.causal_lower_right(T, S) (dispatches the flash kernel, no mask tensor),
(c) two SDPA calls (prefix part unmasked, new part causal) merged by their log-sum-exps.
Prints time per call and the distance to (a)."""
import json
import sys

import torch
import torch.nn.functional as F
from torch.nn.attention.bias import causal_lower_right

dev = "cuda:0"


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters, out


def two_part(q, k, v, past):
    """LSE merge of (q vs the cached prefix, unmasked) and (q vs the new tokens, causal)."""
    g = q.shape[1] // k.shape[1]
    kk, vv = k.repeat_interleave(g, dim=1), v.repeat_interleave(g, dim=1)
    o1, l1 = torch.ops.aten._scaled_dot_product_flash_attention(q, kk[:, :, :past], vv[:, :, :past], 0.0, False)[:2]
    o2, l2 = torch.ops.aten._scaled_dot_product_flash_attention(q, kk[:, :, past:], vv[:, :, past:], 0.0, True)[:2]
    m = torch.maximum(l1, l2)
    w1, w2 = torch.exp(l1 - m), torch.exp(l2 - m)
    return ((o1.float() * w1[..., None] + o2.float() * w2[..., None]) / (w1 + w2)[..., None]).to(q.dtype)


for B, T, past in ((4, 512, 3584), (4, 2048, 2048), (16, 64, 4032), (1, 4096, 4096), (8, 16, 1024)):
    Hq, Hkv, D = 32, 8, 128
    S = past + T
    g = torch.Generator(device=dev).manual_seed(0)
    q = torch.randn(B, Hq, T, D, device=dev, generator=g).bfloat16()
    k = torch.randn(B, Hkv, S, D, device=dev, generator=g).bfloat16()
    v = torch.randn(B, Hkv, S, D, device=dev, generator=g).bfloat16()
    mask = torch.triu(torch.full((T, S), float("-inf"), dtype=q.dtype, device=dev), diagonal=S - T + 1)[None, None]
    res = {"B": B, "T": T, "past": past}
    ta, oa = timeit(lambda: F.scaled_dot_product_attention(q, k, v, attn_mask=mask, enable_gqa=True))
    res["dense_mask_ms"] = round(ta, 3)
    try:
        tb, ob = timeit(lambda: F.scaled_dot_product_attention(q, k, v, attn_mask=causal_lower_right(T, S), enable_gqa=True))
        res["lower_right_ms"] = round(tb, 3)
        res["lower_right_max_abs_diff"] = float((ob.float() - oa.float()).abs().max())
    except Exception as exc:  # noqa: BLE001
        res["lower_right_error"] = str(exc)[:200]
    try:
        tc, oc = timeit(lambda: two_part(q, k, v, past))
        res["two_part_ms"] = round(tc, 3)
        res["two_part_max_abs_diff"] = float((oc.float() - oa.float()).abs().max())
    except Exception as exc:  # noqa: BLE001
        res["two_part_error"] = str(exc)[:200]
    flops = 4.0 * B * Hq * T * D * (past + T / 2)
    res["TFLOPs_dense_mask"] = round(flops / ta / 1e9, 1)
    if "lower_right_ms" in res:
        res["TFLOPs_lower_right"] = round(flops / res["lower_right_ms"] / 1e9, 1)
    print(json.dumps(res), flush=True)
sys.exit(0)
