#!/usr/bin/env python
"""Group a sass_<n>_<kernel>.txt (ncu_digest.py --sass) into runs of equal execution count.
    python scripts/sass_blocks.py gpurun_out/sass_v8/sass_0_texture_forward_kernel.txt [min_total_M]"""
import sys
rows = []
for ln in open(sys.argv[1]):
    f = ln.split(None, 4)
    rows.append((int(f[0]), float(f[1][:-1]), int(f[2]), float(f[3]), f[4].strip()))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
runs, s = [], 0
for i in range(1, len(rows) + 1):
    if i == len(rows) or abs(rows[i][1] - rows[s][1]) > 0.02 * max(rows[s][1], 0.05):
        blk = rows[s:i]
        runs.append((s, i - 1, blk[0][1], sum(r[1] for r in blk), sum(r[2] for r in blk), sum(r[3] for r in blk) / len(blk), blk[0][4][:60]))
        s = i
tot = sum(r[3] for r in runs); samp = sum(r[4] for r in runs)
print(f"total {tot:.1f}M warp-instr, {samp} samples")
for r in runs:
    if r[3] >= thr:
        print(f"{r[0]:5d}-{r[1]:5d} n={r[1]-r[0]+1:4d} exec={r[2]:7.2f}M total={r[3]:7.1f}M ({100*r[3]/tot:4.1f}%) samp={100*r[4]/samp:4.1f}% thr={r[5]:4.1f}  {r[6]}")
