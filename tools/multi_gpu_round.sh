# usage (multi-GPU box): bash tools/multi_gpu_round.sh <tag> [max_gpus=8]  -- sharding parity at 2, 4 and 8 ranks, then the 1/2/4/8-GPU bench lines
tag=${1:-r01i}; M=${2:-8}; NS=""; for n in 2 4 8; do [ $n -le $M ] && NS="$NS $n"; done
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in $NS; do
  $TR --nproc-per-node $n --master-port $((29600 + n)) tests/run_multigpu_parity.py > gpurun_out/${tag}_parity_${n}gpu.log 2>&1; echo parity$n=$?
done
for n in $NS; do
  $TR --nproc-per-node $n --master-port $((29700 + n)) bench.py --gpus $n --no-cpu > gpurun_out/${tag}_bench_cfg2_${n}gpu.json 2> gpurun_out/${tag}_bench_cfg2_${n}gpu.err; echo cfg2_$n=$?
done
$TR --nproc-per-node $M --master-port 29808 bench.py --gpus $M --workload cfg5 --steps 5 --warmup 3 --no-cpu > gpurun_out/${tag}_bench_cfg5_${M}gpu.json 2> gpurun_out/${tag}_bench_cfg5_${M}gpu.err; echo cfg5_8=$?
for f in gpurun_out/${tag}_bench_cfg2_?gpu.json gpurun_out/${tag}_bench_cfg5_${M}gpu.json; do python - $f <<'P'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith("{"):
        d = json.loads(l); print(sys.argv[1], d["n_gpus"], "ms %.2f" % d["ms_per_step"], "%.3g lpj/s" % d["value"], "e2e ms %.2f" % d["e2e"]["ms_per_step"])
P
done
tail -n 3 gpurun_out/${tag}_parity_?gpu.log
