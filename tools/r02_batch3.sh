#!/bin/bash
# parity numbers behind the test thresholds, the share of the CUB sort in the C5 build, full captures of the two C2 kernels
O=gpurun_out
python tools/r02_parity_explore.py c2 fast tie c3 c1 c5 > $O/r02_parity_explore.log 2>&1; tail -3 $O/r02_parity_explore.log
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r02_c5_build_launches.csv python tools/c5_once.py 10000000 1000 > $O/r02_c5_build.log 2>&1
python tools/render_once.py 64 > $O/r02_c2_plain.json 2>&1 &&
ncu --set full --import-source on --clock-control none -k regex:'k_extend|k_connect' --launch-count 3 -f -o $O/r02_c2_full python tools/render_once.py 64 > $O/r02_c2_full.log 2>&1
python tools/c1_once.py > $O/r02_c1_plain.json 2>&1 &&
ncu --set full --import-source on --clock-control none -k regex:k_connect --launch-skip 1 --launch-count 1 -f -o $O/r02_c1_connect python tools/c1_once.py > $O/r02_c1_full.log 2>&1
ls -la $O | grep "r02_c[125]_\|parity"
