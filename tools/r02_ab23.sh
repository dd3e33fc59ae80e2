#!/bin/bash
# A/B of prebuilt library variants (variants/libb2pt_<name>.so, built in the container with B2_NVCC_DEFS): C3 at 8 spp with the
# eye-vertex sort forced on (steady state), then C1
O=gpurun_out
cp path_tracing_b200/libb2pt.so /tmp/libb2pt_keep.so
for name in "$@"; do
  cp variants/libb2pt_$name.so path_tracing_b200/libb2pt.so
  echo "== $name"
  B2PT_CONN_SORT=1 python tools/c3_once.py 1920 1080 8 8 16 2>&1 | tail -1
  python tools/c1_once.py 2>&1 | tail -1
done
cp /tmp/libb2pt_keep.so path_tracing_b200/libb2pt.so
