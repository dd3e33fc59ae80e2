#!/bin/bash
# A/B builds on the GPU box: tools/ab.sh name "-DFLAG=1 ..."  -> gpurun_out/ab_<name>.json (bench line without the CPU arm)
name=$1; defs=$2
B2_NVCC_DEFS="$defs" python __graft_entry__.py > gpurun_out/ab_${name}.build.log 2>&1 || { echo "build failed: $name"; exit 1; }
python bench.py --no-cpu > gpurun_out/ab_${name}.json 2> gpurun_out/ab_${name}.err
python - <<PY
import json
d = json.loads(open("gpurun_out/ab_${name}.json").read().strip().splitlines()[-1])
print("${name}", round(d["value"]), "Msamples/s", round(d["ms_per_step"], 3), "ms", {k: round(v, 2) for k, v in d["stages_ms"].items()}, "e2e", round(d["e2e"]["value"]))
PY
