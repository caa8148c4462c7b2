# usage: bash tools/sweep_T.sh "T values" "workloads"   (tuning aid: E-step threads per datapoint)
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for c in ${2:-cfg2 cfg5 cfg3 cfg4}; do
  for T in ${1:-0}; do
    n=""; [ $c = cfg5 ] && n="--n 200000"; [ $c = cfg3 ] && n="--n 50000"; [ $c = cfg4 ] && n="--n 50000"
    EVO_ESTEP_T=$T python bench.py --workload $c $n --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read()); print('$c T=$T', 'ms/step %.2f'%j['ms_per_step'], {k:round(v['mean_ms'],2) for k,v in j['kernels'].items()})"
  done
done
