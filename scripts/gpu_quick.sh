# quick GPU check: tests (optional) + bench + short per-kernel event timing
set -x
[ "$1" = "test" ] && timeout 400 python -m pytest tests -m gpu -q -x 2>&1 | tail -4
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err || tail -5 gpurun_out/bench_quick.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
print('ms_per_step',round(d['ms_per_step'],4),'value',round(d['value']/1e6,2),'M/s e2e',round(d['e2e']['value']/1e6,2),'cbf_ms',round(d['roofline']['ms_per_launch'],4),'TF',round(d['roofline']['achieved'],1),'clk',d['clocks'])
PY
timeout 200 python scripts/kernel_times.py
