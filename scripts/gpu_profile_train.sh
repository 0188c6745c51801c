# usage: bash scripts/gpu_profile_train.sh <kernel-regex> <out-name> [count]   (ncu --set full on train-mode kernels)
set -x
timeout 120 python bench.py --mode train --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain_train.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k regex:"$1" -s ${4:-8} -c ${3:-4} -o gpurun_out/$2 python bench.py --mode train --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_train.log 2>&1
tail -1 gpurun_out/ncu_train.log
