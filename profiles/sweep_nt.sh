#!/bin/bash
# tuning helper: vertical-path kernel threads per CTA
for nt in 1024 512 256; do
  CVR_PATHV_NT=$nt timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --batch ${1:-36} > gpurun_out/bb_$nt.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/bb_$nt.json")); s=d["stage_ms_per_step"]
print("NT", $nt, "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "path_v", round(s["path_v_down"],2))
PY
done
