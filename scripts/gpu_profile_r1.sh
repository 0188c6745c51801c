# Round-1 profile capture: launch list of the default bench + ncu --set full of the dominant kernels.
set -x
timeout 120 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_ll.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_ll.log 2>&1
timeout 120 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_tc.log 2>&1 && timeout 500 ncu --set full --clock-control none --import-source on -k regex:"cbf_forward_tc|ctrl_edge_tc|ctrl_node_tc|knn_grid_tpq" -s 4 -c 5 -o gpurun_out/prof_r1 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_tc.log 2>&1
tail -2 gpurun_out/ncu_tc.log
