#!/usr/bin/env python
"""Turn an `ncu --metrics gpu__time_duration.sum --csv` launch list of bench.py into the per-kernel share of one
decode step.  Usage: python tools/ncu_step_share.py gpurun_out/r01_launches.csv profiles/r01 [step_index]"""
import collections
import csv
import sys


def short(n):
    for k in ("add_rmsnorm", "rope_append", "paged_attn_tma", "paged_attn_generic", "silu_mul", "attn_combine", "kv_append",
              "kv_block"):
        if k in n:
            return "b200::" + k
    if n.startswith("nvjet") or "gemm" in n.lower() or "cutlass" in n.lower():
        return "cuBLAS " + n[:60]
    if "CUDAFunctor_add<c10::BFloat16>" in n:
        return "torch residual add (bf16)"
    return "torch other: " + n[:60]


def main():
    src, out_prefix = sys.argv[1], sys.argv[2]
    step = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    with open(src) as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.reader(lines)
    hdr = next(rd)
    idx = {n: i for i, n in enumerate(hdr)}
    rows = [r for r in rd if len(r) >= len(hdr)]
    names = [r[idx["Kernel Name"]] for r in rows]
    durs = [float(r[idx["Metric Value"]].replace(",", "")) for r in rows]
    pos = [i for i, n in enumerate(names) if "rope_append" in n]
    if len(pos) < 64:  # RoPE + append are fused into the attention kernel: anchor on that one instead
        pos = [i for i, n in enumerate(names) if "paged_attn_tma" in n]
    per_step = 32
    # a step starts a few launches (slot mapping, RoPE tables, first norm) before its first rope_append
    first = pos[per_step * step]
    nxt = pos[per_step * (step + 1)]
    lead = 0
    for j in range(first - 1, max(first - 40, 0), -1):
        if "reduce_kernel" in names[j] or "argmax" in names[j].lower():
            lead = first - j - 1
            break
    s0, s1 = first - lead, nxt - lead
    with open(out_prefix + "_ncu_launches_one_step.csv", "w") as f:
        w = csv.writer(f)
        w.writerow(["ID", "Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum (ns)"])
        for r in rows[s0:s1]:
            w.writerow([r[idx["ID"]], r[idx["Kernel Name"]][:160], r[idx["Block Size"]], r[idx["Grid Size"]], r[idx["Metric Value"]]])
    agg = collections.defaultdict(lambda: [0, 0.0])
    for n, d in zip(names[s0:s1], durs[s0:s1]):
        a = agg[short(n)]
        a[0] += 1
        a[1] += d
    tot = sum(durs[s0:s1])
    print(f"kernels in step: {s1 - s0}; sum of durations: {tot / 1e6:.3f} ms\n")
    print("| share | total us | launches | avg us | kernel |\n|---:|---:|---:|---:|---|")
    for k, (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if d / tot >= 0.0005:
            print(f"| {100 * d / tot:.1f} % | {d / 1e3:.1f} | {c} | {d / c / 1e3:.2f} | {k} |")


if __name__ == "__main__":
    main()
