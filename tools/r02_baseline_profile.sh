#!/bin/bash
# Round-2 starting point on the GPU box: stage times + ncu captures of the C3 render and the C5 traversal kernels as
# round 1 left them.  Everything lands in gpurun_out/r02_base_*.
set -x
O=gpurun_out
python tools/c3_once.py 1920 1080 8 8 16 > $O/r02_base_c3_spl8.json 2> $O/r02_base_c3_spl8.err
python tools/c3_once.py 1920 1080 8 1 16 > $O/r02_base_c3_spl1.json 2> $O/r02_base_c3_spl1.err
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none --csv \
    --log-file $O/r02_base_c3_launches.csv python tools/c3_once.py 1920 1080 2 8 16 > $O/r02_base_c3_ncu_list.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_connect --launch-skip 1 --launch-count 1 -f -o $O/r02_base_c3_connect \
    python tools/c3_once.py 1920 1080 2 8 16 > $O/r02_base_c3_ncu_connect.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_extend --launch-skip 1 --launch-count 1 -f -o $O/r02_base_c3_extend \
    python tools/c3_once.py 1920 1080 2 8 16 > $O/r02_base_c3_ncu_extend.log 2>&1
python tools/c5_once.py 10000000 20000000 > $O/r02_base_c5.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_closest_pooled --launch-skip 1 --launch-count 1 -f -o $O/r02_base_c5_closest \
    python tools/c5_once.py 10000000 4000000 > $O/r02_base_c5_ncu_closest.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_intersect --launch-skip 1 --launch-count 1 -f -o $O/r02_base_c5_any \
    python tools/c5_once.py 10000000 4000000 > $O/r02_base_c5_ncu_any.log 2>&1
ls -la $O | tail -20
