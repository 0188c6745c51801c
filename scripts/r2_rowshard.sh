# row-sharded C5 step at N GPUs: usage r2_rowshard.sh <N> [K...]
set -x
N=$1; shift
mkdir -p gpurun_out
for K in "$@"; do
  if [ "$N" = "1" ]; then
    timeout 200 python bench.py --row-shard --agents 65536 --envs 1 --top-k $K --steps 10 --warmup 3 > gpurun_out/rowshard_n${N}_k$K.json 2> gpurun_out/rowshard_n${N}_k$K.err || tail -20 gpurun_out/rowshard_n${N}_k$K.err
  else
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --row-shard --agents 65536 --envs 1 --top-k $K --steps 10 --warmup 3 > gpurun_out/rowshard_n${N}_k$K.json 2> gpurun_out/rowshard_n${N}_k$K.err || tail -20 gpurun_out/rowshard_n${N}_k$K.err
  fi
  python - <<PY
import json
d=json.loads(open('gpurun_out/rowshard_n${N}_k$K.json').read().strip().splitlines()[-1])
print('ROWSHARD N',d['n_gpus'],'K',$K,'ms',round(d['ms_per_step'],4),'value',round(d['value']/1e6,2),'M/s parity',d['rows_equal_unsharded_step'],'per-rank',[round(x,3) for x in d['per_rank_ms_per_step']])
PY
done
