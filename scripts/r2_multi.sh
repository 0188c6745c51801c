# multi-GPU bench check: usage r2_multi.sh <N>
set -x
N=$1
mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err || tail -30 gpurun_out/bench_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n$N.json').read().strip().splitlines()[-1])
print('N',d['n_gpus'],'fwd ms',round(d['ms_per_step'],4),'value',round(d['value']/1e6,2),'e2e',round(d['e2e']['value']/1e6,2),'per-rank',[round(x,4) for x in d['per_rank_ms_per_step']])
t=d['train']
print('train ms',round(t['ms_per_step'],4),'value',round(t['value']/1e6,2),'e2e',round(t['e2e']['value']/1e6,2),'ar_exposed_ms',t.get('allreduce_exposed_ms'),'parity',t.get('grad_parity_max_rel'),'per-rank',[round(x,4) for x in t['per_rank_ms_per_step']])
PY
