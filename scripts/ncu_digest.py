#!/usr/bin/env python
"""Digest an .ncu-rep: headline metrics per kernel + per-SASS-instruction execution counts.

    python scripts/ncu_digest.py gpurun_out/ncu_v4.ncu-rep [--sass DIR] [--min-m 0.5]

Writes DIR/sass_<n>_<kernel>.txt (index, executed M warp-instr, stall samples, avg threads, SASS)."""
import csv
import io
import os
import re
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_active.avg", "sm__cycles_elapsed.avg",
        "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum", "l1tex__t_set_accesses_pipe_lsu_mem_global_op_red.sum",
        "lts__t_sectors_srcunit_tex_op_write.sum", "lts__t_sectors_srcunit_tex_op_read.sum"]


def run(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    sass_dir = None
    if "--sass" in sys.argv:
        sass_dir = sys.argv[sys.argv.index("--sass") + 1]
        os.makedirs(sass_dir, exist_ok=True)
    rows = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("----")
        print(r[hdr.index("Kernel Name")][:100])
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print(f"{w:78s} {r[i]} {units[i]}")
        for i, h in enumerate(hdr):
            if "smsp__average_warps_issue_stalled" in h and h.endswith("_per_issue_active.ratio"):
                try:
                    v = float(r[i])
                except ValueError:
                    continue
                if v > 0.15:
                    nm = h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]
                    print(f"   stall {nm:30s} {v:.2f} warps/issue")
    if sass_dir is None:
        return
    rows = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "source", "--csv"]))))
    kern, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            kern.append(cur)
        elif r and r[0] == "Address":
            cur["hdr"] = r
        elif cur is not None and r and "hdr" in cur:
            cur["rows"].append(r)
    for n, k in enumerate(kern):
        h = k["hdr"]
        isrc, ie, isamp, it = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples"), h.index("Avg. Threads Executed")
        short = re.sub(r"[^A-Za-z0-9_]", "", k["name"].split("(")[0].split("::")[-1])[:40]
        path = os.path.join(sass_dir, f"sass_{n}_{short}.txt")
        tot = 0
        with open(path, "w") as f:
            for i, r in enumerate(k["rows"]):
                tot += int(r[ie])
                f.write(f"{i:4d} {int(r[ie]) / 1e6:8.2f}M {int(r[isamp]):6d} {float(r[it]):5.1f}  {r[isrc]}\n")
        print(path, "total warp-instr M", tot / 1e6)


if __name__ == "__main__":
    main()
