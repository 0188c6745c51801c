# ncu launch list of one bench run (cold-cache, serialised: compare shares, not absolutes). usage: r2_launchlist.sh <out-name> [bench args]
set -x
mkdir -p gpurun_out
O=$1; shift
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/plain_$O.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -c 600 --csv --log-file gpurun_out/$O.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/ncu_$O.log 2>&1
tail -2 gpurun_out/ncu_$O.log
