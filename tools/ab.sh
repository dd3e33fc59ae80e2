#!/bin/bash
# A/B builds on the GPU box: tools/ab.sh name "-DFLAG=1 ..." [c3]  -> gpurun_out/ab_<name>.json (bench line without the
# CPU arm); with a third argument also the reduced C3 render (1 M triangles, 32 spp) -> gpurun_out/ab_<name>_c3.json
name=$1; defs=$2
B2_NVCC_DEFS="$defs" python __graft_entry__.py > gpurun_out/ab_${name}.build.log 2>&1 || { echo "build failed: $name"; exit 1; }
python bench.py --no-cpu > gpurun_out/ab_${name}.json 2> gpurun_out/ab_${name}.err
[ -n "$3" ] && timeout 900 python tools/c3_render.py --spp 32 > gpurun_out/ab_${name}_c3.json 2> gpurun_out/ab_${name}_c3.err
python - <<PY
import json, os
d = json.loads(open("gpurun_out/ab_${name}.json").read().strip().splitlines()[-1])
print("${name}", round(d["value"]), "Msamples/s", round(d["ms_per_step"], 3), "ms", {k: round(v, 2) for k, v in d["stages_ms"].items()}, "e2e", round(d["e2e"]["value"]))
p = "gpurun_out/ab_${name}_c3.json"
if "$3" and os.path.exists(p):
    c = json.loads(open(p).read().strip().splitlines()[-1])
    print("   C3", {k: (round(c[k]["eye_msamples_per_s"], 2), round(c[k]["mrays_per_s"])) for k in ("spl1", "spl8")})
PY
