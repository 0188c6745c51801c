# usage: bash scripts/r2_ncu_one.sh <kernel-regex> <out-name> [skip] [count] [extra bench args]
set -x
mkdir -p gpurun_out
R=$1; O=$2; S=${3:-1}; C=${4:-1}; shift 4
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/plain_$O.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$R" -s $S -c $C -o gpurun_out/$O python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/ncu_$O.log 2>&1
tail -2 gpurun_out/ncu_$O.log
