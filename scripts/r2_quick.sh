# quick GPU check: selected tests + default bench with per-kernel table.  usage: r2_quick.sh "<pytest -k expr>" [bench args]
set -x
mkdir -p gpurun_out
K="$1"; shift
timeout 900 python -m pytest tests -m gpu -q -x -k "$K" 2>&1 | tee gpurun_out/pytest_quick.log | tail -6
timeout 400 python bench.py --no-cpu-baseline "$@" > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err || tail -20 gpurun_out/bench_quick.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
print('fwd ms',round(d['ms_per_step'],4),'value',round(d['value']/1e6,2),'M/s e2e',round(d['e2e']['value']/1e6,2))
print('roofline',{k:d['roofline'][k] for k in ('achieved','frac','ms_per_launch','frac_of_3xtf32_ceiling')})
for r in d['kernel_breakdown']['kernels']: print('  F',r['kernel'],r['launches_per_step'],round(r['ms_per_launch']*1e3,1),'us',round(r['share'],3),round(r.get('tflops',0),1))
t=d.get('train')
if t:
    print('train ms',round(t['ms_per_step'],4),'value',round(t['value']/1e6,2),'e2e',round(t['e2e']['value']/1e6,2), 'ar_exposed', t.get('allreduce_exposed_ms'), 'parity', t.get('grad_parity_max_rel'))
    for r in t['kernel_breakdown']['kernels']: print('  T',r['kernel'],r['launches_per_step'],round(r['ms_per_launch']*1e3,1),'us',round(r['share'],3),round(r.get('tflops',0),1))
PY
