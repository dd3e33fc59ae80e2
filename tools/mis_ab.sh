#!/bin/bash
# A/B builds of the B2_STRATEGIES_MIS kernel on the GPU box: tools/mis_ab.sh name "-DB2_MIS_THREADS=64 ..."
# Only b2_integrators.cu depends on these flags: the other objects are copied from the default build (build/obj).
name=$1; defs=$2
od=build/$(python -c "import sys; print('obj_' + '_'.join(e.strip('-').replace('=', '') for e in sys.argv[1].split()))" "$defs")
mkdir -p $od; for f in build/obj/*.o; do case $f in *b2_integrators*) ;; *) cp -p $f $od/;; esac; done
B2_NVCC_DEFS="$defs" python __graft_entry__.py > gpurun_out/misab_${name}.build.log 2>&1 || { echo "build failed: $name"; tail -5 gpurun_out/misab_${name}.build.log; exit 1; }
echo "$name: cornell 256x256x16 / mis_variant 512x512x64 (ms, closest, shadow)"
timeout 200 python tools/mis_profile_once.py cornell 256 256 16 8
timeout 200 python tools/mis_profile_once.py mis_variant 512 512 64 8
