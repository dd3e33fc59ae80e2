#!/bin/bash
# last check of the committed build on the GPU box: the GPU suite, smoke(), and the per-launch counters of C1's connection kernel
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > $O/r02v_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r02v_smoke.log 2>&1
N=gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,sm__warps_active.avg.pct_of_peak_sustained_active
ncu --metrics $N --clock-control none -k regex:k_connect --csv --log-file $O/r02p_c1_counters.csv python tools/c1_once.py > $O/r02p_c1_counters.log 2>&1
cat $O/r02v_pytest.log; tail -2 $O/r02v_smoke.log; tail -1 $O/r02p_c1_counters.log | cut -c1-200
