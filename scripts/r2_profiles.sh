# round-2 profile capture: launch list (forward + train blocks of the default bench) and ncu --set full of the main kernels
set -x
mkdir -p gpurun_out
bash scripts/r2_launchlist.sh launches_r2
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-train-block > gpurun_out/plain_full_fwd.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:"cbf_forward_tc|ctrl_edge_tc|ctrl_node_tc|knn_grid_tpq|grid_build|loss_partial" -s 14 -c 9 -o gpurun_out/prof_r2_forward python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-train-block > gpurun_out/ncu_full_fwd.log 2>&1
tail -2 gpurun_out/ncu_full_fwd.log
timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --mode train > gpurun_out/plain_full_train.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:"dgrad|wgrad|dump|bwd|cbf_forward_tc_kernel<\(int\)[12]|ctrl_node_tc_kernel<\(bool\)1|ctrl_edge_tc_kernel<\(bool\)1" -s 40 -c 14 -o gpurun_out/prof_r2_train python bench.py --steps 2 --warmup 3 --no-cpu-baseline --mode train > gpurun_out/ncu_full_train.log 2>&1
tail -2 gpurun_out/ncu_full_train.log
