for c in cfg2 cfg5; do
  for T in 0 32 64 128 256; do
    n=""; [ $c = cfg5 ] && n="--n 200000"
    EVO_ESTEP_T=$T python bench.py --workload $c $n --steps 5 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read()); print('$c T=$T', 'ms/step %.2f'%j['ms_per_step'], {k:round(v['mean_ms'],2) for k,v in j['kernels'].items()})"
  done
done
