set -x
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_tc.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"cbf_forward_tc|ctrl_edge_tc" -s 3 -c 2 -o gpurun_out/prof_tc python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_tc.log 2>&1
tail -2 gpurun_out/ncu_tc.log
