# A/B of builds of the library on the default bench.  usage: r2_ab_lib.sh <libA.so> [<libB.so> ...]
for L in "$@"; do
  echo "== $L"
  MACBF_B200_LIB=$PWD/$L timeout 300 python bench.py --no-cpu-baseline > gpurun_out/ab.json 2> gpurun_out/ab.err || tail -5 gpurun_out/ab.err
  python - <<'PY'
import json
d=json.loads(open('gpurun_out/ab.json').read().strip().splitlines()[-1])
print('fwd ms',round(d['ms_per_step'],4))
t=d.get('train')
if t:
    print('train ms',round(t['ms_per_step'],4))
    for r in t['kernel_breakdown']['kernels']:
        if 'cbf' in r['kernel'] or 'wgrad' in r['kernel'] or 'bwd' in r['kernel'] or 'dgrad' in r['kernel']: print('  T',r['kernel'],r['launches_per_step'],round(r['ms_per_launch']*1e3,1),'us')
PY
done
