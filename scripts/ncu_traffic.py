#!/usr/bin/env python
"""Write profiles/traffic.json from an ncu capture of the build being benchmarked.

    python scripts/ncu_traffic.py gpurun_out/raster_<tag>.ncu-rep [more.ncu-rep ...]

Per C-ABI entry point the dominant kernel's DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum),
warp-instructions per launch, duration and the instruction-issue roofline fraction
inst / (SMs x 4 schedulers x SM clock) / duration, stamped with a hash of the CUDA sources (bench.py recomputes the
hash and ignores a capture taken from other sources -- the GPU box has no .git to ask)."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
KERNELS = {"texture_forward_kernel": "gstex_texture_forward", "texture_backward_stream_kernel": "gstex_texture_backward",
           "unpack_grads_kernel": "unpack_grads_kernel", "tile_local_sort_kernel": "gstex_tile_bucket_sort"}


def main():
    from bench import kernel_source_stamp

    out = {"source_stamp": kernel_source_stamp(), "_source": [os.path.basename(p) for p in sys.argv[1:]]}
    for rep in sys.argv[1:]:
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        hdr = rows[0]
        col = {h: i for i, h in enumerate(hdr)}
        acc = {}
        for r in rows[2:]:
            name = r[col["Kernel Name"]]
            key = next((v for k, v in KERNELS.items() if k in name), None)
            if key is None:
                continue

            def f(m):
                return float(r[col[m]].replace(",", "")) if m in col and r[col[m]] not in ("", "n/a") else None

            dur_us = f("gpu__time_duration.sum")
            units = rows[1][col["gpu__time_duration.sum"]]
            if units == "ns":
                dur_us /= 1e3
            elif units == "ms":
                dur_us *= 1e3
            rd, wr = f("dram__bytes_read.sum"), f("dram__bytes_write.sum")
            scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            rd *= scale[rows[1][col["dram__bytes_read.sum"]]]
            wr *= scale[rows[1][col["dram__bytes_write.sum"]]]
            inst = f("smsp__inst_executed.sum")
            cyc = f("sm__cycles_elapsed.avg")
            sms = f("launch__sm_count") or 148
            acc.setdefault(key, []).append(dict(dram_bytes=rd + wr, inst=inst, duration_us=dur_us,
                                                issue_frac=inst / (sms * 4 * cyc) if cyc else None))
        for key, lst in acc.items():
            out[key] = {k: sum(x[k] for x in lst) / len(lst) for k in lst[0]}
            out[key]["launches_averaged"] = len(lst)
    with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as fjs:
        json.dump(out, fjs, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
