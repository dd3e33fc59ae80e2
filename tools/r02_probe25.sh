#!/bin/bash
# (1) product path: a ~1930-triangle OBJ scene (non-pooled connection kernel, sorted eye vertices) against the host compilation
# (2) the non-pooled kernel forced onto a 60 k-triangle scene under compute-sanitizer (it failed on C3 in r02_ab24.log)
python tools/conn_variant_probe.py 2350 256 256 2 hostsim 2>&1 | tail -1
cp path_tracing_b200/libb2pt.so /tmp/libb2pt_keep.so
cp variants/libb2pt_nopool.so path_tracing_b200/libb2pt.so
timeout 150 compute-sanitizer --tool memcheck --print-limit 3 python tools/conn_variant_probe.py 60000 96 64 1 2>&1 | grep -v "^$" | tail -30
cp /tmp/libb2pt_keep.so path_tracing_b200/libb2pt.so
