# A/B of the controller edge kernel: v1 (MACBF_CTRL_EDGE_V1=1) vs the transposed-layer-2 kernel; parity tests with the new one
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "controller or step_forward or fused or literal or batched or c2_shape or backward or index_reuse or rollout or c5 or c3" 2>&1 | tee gpurun_out/pytest_edge.log | tail -15
for v in 1 0; do
MACBF_CTRL_EDGE_V1=$v timeout 300 python bench.py --no-cpu-baseline --steps 20 > gpurun_out/bench_edge_v$v.json 2> gpurun_out/bench_edge_v$v.err || tail -5 gpurun_out/bench_edge_v$v.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_edge_v$v.json').read().strip().splitlines()[-1])
print('V1=$v fwd ms',round(d['ms_per_step'],4),'train ms',round(d['train']['ms_per_step'],4))
for r in d['kernel_breakdown']['kernels']: print('  F',r['kernel'],r['launches_per_step'],round(r['ms_per_launch']*1e3,1),'us',round(r['share'],3),round(r.get('tflops',0),1))
for r in d['train']['kernel_breakdown']['kernels']:
    if 'ctrl_edge_tc' in r['kernel']: print('  T',r['kernel'],r['launches_per_step'],round(r['ms_per_launch']*1e3,1),'us')
PY
done
