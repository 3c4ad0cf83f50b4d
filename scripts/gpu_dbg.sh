#!/bin/bash
# debugging: bench per-entry times for several GSTEX_DBG values
for d in "$@"; do
  GSTEX_DBG=$d python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-ref-cuda > gpurun_out/dbg_$d.json 2>gpurun_out/dbg_$d.err
  python - <<PY
import json
d=json.load(open("gpurun_out/dbg_$d.json"))
e=d["roofline"]["per_entry_ms_per_view"]
print("dbg=$d ms/view",round(d["config"]["ms_per_view"],3),"fwd",round(e.get("gstex_texture_forward_act", e.get("gstex_texture_forward")),3),"bwd",round(e.get("gstex_texture_backward_act", e.get("gstex_texture_backward")),3))
PY
done
