#!/bin/bash
# A/B on the GPU box: tools/r02_ab.sh <tag> name=-DDEFS ...   (name "base" with empty defs = the committed build)
# each variant is built into its own object directory, then timed on C3 (8 spp, spl 8), C2 and C1 (twice each: the first run pages the box in)
tag=$1; shift
O=gpurun_out
for v in "$@"; do
  name=${v%%=*}; defs=${v#*=}
  B2_NVCC_DEFS="$defs" python __graft_entry__.py > $O/${tag}_${name}.build.log 2>&1 || { echo "build failed: $name"; tail -5 $O/${tag}_${name}.build.log; continue; }
  echo "== $name ($defs)"
  for k in 1 2; do python tools/c3_once.py 1920 1080 8 8 16 2>&1 | tail -1; done | tee $O/${tag}_${name}_c3.json
  python tools/render_once.py 64 2>&1 | tail -1 | tee $O/${tag}_${name}_c2.json
  python tools/c1_once.py 2>&1 | tail -1 | tee $O/${tag}_${name}_c1.json
done
