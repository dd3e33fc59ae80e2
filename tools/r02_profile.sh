#!/bin/bash
# Round-2 ncu evidence for the committed kernels (run on the GPU box; everything lands in gpurun_out/r02p_*):
#   launch lists (time + DRAM bytes + instructions) of one C3 render (2 spp = 1/16 of the bench step), one C2 step, one C1 render
#   one --set full capture of the dominant kernel of C3 (k_connect<0,1>, second launch) and of C5's two traversal kernels
#   the counters BASELINE.json's north_star names for C3 k_connect / k_extend and the C5 kernels (L2 / DRAM bytes, hit rates, FMA / ALU pipes)
O=gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum
python tools/c3_once.py 1920 1080 2 8 16 > $O/r02p_c3_plain.json 2>&1 &&
ncu --metrics $M --clock-control none --csv --log-file $O/r02p_c3_launches.csv python tools/c3_once.py 1920 1080 2 8 16 > $O/r02p_c3_list.log 2>&1
python tools/render_once.py 64 > $O/r02p_c2_plain.json 2>&1 &&
ncu --metrics $M --clock-control none --csv --log-file $O/r02p_c2_launches.csv python tools/render_once.py 64 > $O/r02p_c2_list.log 2>&1
python tools/c1_once.py > $O/r02p_c1_plain.json 2>&1 &&
ncu --metrics $M --clock-control none --csv --log-file $O/r02p_c1_launches.csv python tools/c1_once.py > $O/r02p_c1_list.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_connect --launch-skip 1 --launch-count 1 -f -o $O/r02p_c3_connect python tools/c3_once.py 1920 1080 2 8 16 > $O/r02p_c3_full.log 2>&1
python tools/c5_once.py 10000000 4000000 > $O/r02p_c5_plain.log 2>&1 &&
ncu --set full --import-source on --clock-control none -k regex:'k_closest_pooled|k_any_pooled' --launch-skip 1 --launch-count 2 -f -o $O/r02p_c5 python tools/c5_once.py 10000000 4000000 > $O/r02p_c5_full.log 2>&1
N=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,l1tex__t_bytes.sum,smsp__inst_executed.sum,smsp__inst_executed_pipe_fma.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_xu.sum,smsp__inst_executed_pipe_lsu.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum
ncu --metrics $N --clock-control none -k regex:'k_connect|k_extend' --csv --log-file $O/r02p_c3_counters.csv python tools/c3_once.py 1920 1080 2 8 16 > $O/r02p_c3_counters.log 2>&1
ncu --metrics $N --clock-control none -k regex:'k_closest_pooled|k_any_pooled' --csv --log-file $O/r02p_c5_counters.csv python tools/c5_once.py 10000000 4000000 > $O/r02p_c5_counters.log 2>&1
ls -la $O | grep r02p
