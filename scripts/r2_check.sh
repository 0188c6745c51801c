# round-2 GPU check: peak probe, GPU tests, default bench (forward headline + train block)
set -x
mkdir -p gpurun_out
timeout 120 macbf-gnn_b200/tools/peak_probe > gpurun_out/peak_probe.json 2> gpurun_out/peak_probe.err; cat gpurun_out/peak_probe.json
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 2>&1 | tee gpurun_out/pytest_gpu.log | tail -30
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err || tail -20 gpurun_out/bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
print('fwd ms',round(d['ms_per_step'],4),'value',round(d['value']/1e6,2),'M/s e2e',round(d['e2e']['value']/1e6,2),'clk',d['clocks'])
print('roofline',{k:d['roofline'][k] for k in ('achieved','frac','ms_per_launch','frac_of_3xtf32_ceiling','share_of_step_live')})
for r in d['kernel_breakdown']['kernels']: print('  F',r['kernel'],r['launches_per_step'],round(r['ms_per_launch']*1e3,1),'us',round(r['share'],3),round(r.get('tflops',0),1))
t=d.get('train')
if t:
    print('train ms',round(t['ms_per_step'],4),'value',round(t['value']/1e6,2),'e2e',round(t['e2e']['value']/1e6,2))
    for r in t['kernel_breakdown']['kernels']: print('  T',r['kernel'],r['launches_per_step'],round(r['ms_per_launch']*1e3,1),'us',round(r['share'],3),round(r.get('tflops',0),1))
    print(t.get('roofline'))
print('cpu',d.get('cpu_baseline'))
PY
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
