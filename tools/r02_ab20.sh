B2_NVCC_DEFS="-DB2_CONN_SORT_SMALL=1" python __graft_entry__.py > gpurun_out/ab20.build.log 2>&1 || tail -5 gpurun_out/ab20.build.log
for k in 1 2; do python tools/c1_once.py 2>&1 | tail -1; done
python tools/render_once.py 64 2>&1 | tail -1
python tools/render_once.py 64 mis_variant 2>&1 | tail -1
python tools/render_once.py 64 cornell 2>&1 | tail -1
python __graft_entry__.py > /dev/null 2>&1
python tools/render_once.py 64 mis_variant 2>&1 | tail -1
python tools/render_once.py 64 cornell 2>&1 | tail -1
