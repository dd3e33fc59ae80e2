#!/bin/bash
# One measurement round on the GPU box: tools/r02_iter.sh <tag> [variant "-DDEFS" ...]
# default build: parity tests, C3 (8 spp, spl 8 and 1), C2 bench line, C5; then each variant (name=-DFLAGS) is built
# into its own object directory and timed on C3 / C5 only.  Results: gpurun_out/<tag>_*.
tag=$1; shift
O=gpurun_out
run_perf() {   # $1 = label
  python tools/c3_once.py 1920 1080 8 8 16 > $O/${tag}_$1_c3_spl8.json 2> $O/${tag}_$1_c3_spl8.err
  python tools/c3_once.py 1920 1080 8 1 16 > $O/${tag}_$1_c3_spl1.json 2>> $O/${tag}_$1_c3_spl8.err
  python tools/c5_once.py 10000000 20000000 > $O/${tag}_$1_c5.log 2>&1
  echo "== $1"; cat $O/${tag}_$1_c3_spl8.json $O/${tag}_$1_c3_spl1.json $O/${tag}_$1_c5.log
}
timeout 900 python -m pytest tests -m gpu -x -q > $O/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/${tag}_pytest.log
run_perf base
python tools/render_once.py 64 > $O/${tag}_c2.json 2>&1; cat $O/${tag}_c2.json | tail -2
python tools/c1_once.py > $O/${tag}_c1.json 2>&1; tail -1 $O/${tag}_c1.json
for v in "$@"; do
  name=${v%%=*}; defs=${v#*=}
  B2_NVCC_DEFS="$defs" python __graft_entry__.py > $O/${tag}_${name}.build.log 2>&1 || { echo "build failed: $name"; tail -5 $O/${tag}_${name}.build.log; continue; }
  run_perf $name
  python tools/render_once.py 64 > $O/${tag}_${name}_c2.json 2>&1; tail -1 $O/${tag}_${name}_c2.json
done
