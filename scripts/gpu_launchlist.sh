# ncu launch list of one bench run (cold-cache, serialised: compare shares, not absolutes)
set -x
timeout 120 python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/plain_ll.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/ncu_ll.log 2>&1
