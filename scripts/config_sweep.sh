# forward / train throughput on the other BASELINE configs (1 GPU): C3, the per-GPU share of C4, the C5 density sweep
run() { timeout 250 python bench.py "$@" --steps 8 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', '->', round(d['ms_per_step'],3), 'ms', round(d['value']/1e6,2), 'M agent-steps/s, e2e', round(d['e2e']['value']/1e6,2))"; }
run --mode forward --agents 4096 --envs 1 --obstacles 384
run --mode train --agents 4096 --envs 1 --obstacles 384
run --mode forward --agents 16384 --envs 32
run --mode train --agents 16384 --envs 32
for k in 8 12 32 64 128; do run --mode forward --agents 65536 --envs 1 --top-k $k; done
run --mode train --agents 65536 --envs 1 --top-k 12
run --mode train --agents 65536 --envs 1 --top-k 64
