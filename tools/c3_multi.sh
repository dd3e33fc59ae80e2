#!/bin/bash
# C3 on N GPUs of one node through pt_cli (in-process multi-GPU entry, NCCL film reduce): tools/c3_multi.sh N [spp]
N=${1:-8}; SPP=${2:-256}
D=$(mktemp -d)
python - <<PY
import sys; sys.path.insert(0, ".")
from path_tracing_b200 import gen
print(gen.write_obj_scene("$D", n_tris=1 << 20))
PY
for spl in 1 8; do
  timeout 900 path_tracing_b200/pt_cli --mode bdpt --input $D/scene.txt --output gpurun_out/c3_n${N}_spl${spl}.ppm --gpus $N --spp $SPP --spl $spl --eye-depth 16 --light-depth 16 --width 1920 --height 1080 | grep -E "Finished|Triangle|Error|GPU"
done
rm -f gpurun_out/c3_n${N}_spl*.ppm
rm -rf $D
