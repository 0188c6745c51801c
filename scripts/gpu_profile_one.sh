# usage: bash scripts/gpu_profile_one.sh <kernel-regex> <out-name>
set -x
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_one.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"$1" -s 2 -c 1 -o gpurun_out/$2 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_one.log 2>&1
tail -1 gpurun_out/ncu_one.log
