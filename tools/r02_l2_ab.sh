#!/bin/bash
# A/B of the persisting-L2 window over the top of the node array (B2_L2_PERSIST_MB: 0 = off, unset = default = half of L2):
# C5 traversal and the C3 render, plus the ncu counters of the C5 kernels with and without it.  -> gpurun_out/r02_l2_*
O=gpurun_out
for mb in 0 16 32 default; do
  if [ $mb = default ]; then unset B2_L2_PERSIST_MB; else export B2_L2_PERSIST_MB=$mb; fi
  echo "== window $mb MB"
  python tools/c5_once.py 10000000 20000000 2>&1 | tail -2
  python tools/c3_once.py 1920 1080 8 8 16 2>&1 | tail -1
done
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_srcunit_tex_op_read.sum,l1tex__t_sector_hit_rate.pct,smsp__inst_executed.sum,smsp__inst_executed_pipe_fma.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_xu.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio
export B2_L2_PERSIST_MB=0
python tools/c5_once.py 10000000 4000000 > /dev/null 2>&1 &&
ncu --metrics $M --clock-control none -k regex:'k_closest_pooled|k_intersect' --csv --log-file $O/r02_l2_off_c5.csv python tools/c5_once.py 10000000 4000000 > $O/r02_l2_off_c5.log 2>&1
unset B2_L2_PERSIST_MB
ncu --metrics $M --clock-control none -k regex:'k_closest_pooled|k_intersect' --csv --log-file $O/r02_l2_on_c5.csv python tools/c5_once.py 10000000 4000000 > $O/r02_l2_on_c5.log 2>&1
ncu --metrics $M --clock-control none -k regex:'k_connect|k_extend' --csv --log-file $O/r02_c3_counters.csv python tools/c3_once.py 1920 1080 2 8 16 > $O/r02_c3_counters.log 2>&1
ls -la $O | grep r02_l2
