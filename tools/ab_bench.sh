python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for c in cfg2 cfg5 cfg3; do
  n=""; [ $c = cfg5 ] && n="--n 200000"; [ $c = cfg3 ] && n="--n 50000"
  python bench.py --workload $c $n --steps 5 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read()); print('$c', 'ms/step %.2f'%j['ms_per_step'], {k:round(v['mean_ms'],2) for k,v in j['kernels'].items()})"
done
