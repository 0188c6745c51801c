set -x
timeout 120 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_tc.log 2>&1 && timeout 500 ncu --set full --clock-control none --import-source on -k regex:"cbf_forward_tc|ctrl_edge_tc|ctrl_node_tc|knn_grid_tpq|loss_partial|grid_build" -s 12 -c 7 -o gpurun_out/prof_fwd_r1 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_tc.log 2>&1
tail -2 gpurun_out/ncu_tc.log
