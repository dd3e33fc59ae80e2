#!/bin/bash
# Round-2 closing run on the GPU box for the committed build: the GPU suite, the launch lists + traffic summaries the bench's roofline
# reads (stamped with b2_version()), both bench arms at N = 1, then the --set full capture and the counters of the connection stage.
# C3 under ncu runs with B2PT_CONN_SORT=1: a single render in a fresh process would spend its first launch on the decision.
O=gpurun_out
V=$(python -c "from path_tracing_b200 import _abi; print(_abi.lib().b2_version().decode())")
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > $O/r02f_pytest.log
B2PT_CONN_SORT=1 ncu --metrics $M --clock-control none --csv --log-file $O/r02p_c3_launches.csv python tools/c3_once.py 1920 1080 2 8 16 > $O/r02p_c3_list.log 2>&1
ncu --metrics $M --clock-control none --csv --log-file $O/r02p_c2_launches.csv python tools/render_once.py 64 > $O/r02p_c2_list.log 2>&1
ncu --metrics $M --clock-control none --csv --log-file $O/r02p_c1_launches.csv python tools/c1_once.py > $O/r02p_c1_list.log 2>&1
for w in 1 2 3; do
  cp $O/r02p_c${w}_launches.csv profiles/r02_c${w}_launches.csv
  s=1; [ $w = 3 ] && s=16
  python tools/ncu_traffic.py profiles/r02_c${w}_launches.csv profiles/r02_c${w}_traffic.json --from-last k_light_paths --version "$V" --workload C$w --scale $s > $O/r02f_traffic_c$w.txt 2>&1
  cp profiles/r02_c${w}_traffic.json $O/r02_c${w}_traffic.json
done
python bench.py > $O/r02_bench_n1_v9.json 2> $O/r02_bench_n1_v9.err
python bench.py --impl reference > $O/r02_bench_ref_v9.json 2> $O/r02_bench_ref_v9.err
B2PT_CONN_SORT=1 ncu --set full --import-source on --clock-control none -k regex:k_connect --launch-skip 1 --launch-count 1 -f -o $O/r02p_c3_connect python tools/c3_once.py 1920 1080 2 8 16 > $O/r02p_c3_full.log 2>&1
N=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,l1tex__t_bytes.sum,smsp__inst_executed.sum,smsp__inst_executed_pipe_fma.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_xu.sum,smsp__inst_executed_pipe_lsu.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum
B2PT_CONN_SORT=1 ncu --metrics $N --clock-control none -k regex:'k_connect|k_extend' --csv --log-file $O/r02p_c3_counters.csv python tools/c3_once.py 1920 1080 2 8 16 > $O/r02p_c3_counters.log 2>&1
cat $O/r02f_pytest.log; head -c 600 $O/r02_bench_n1_v9.json; echo; head -c 400 $O/r02_bench_ref_v9.json; echo; tail -3 $O/r02_bench_n1_v9.err
