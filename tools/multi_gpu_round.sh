# usage (multi-GPU box): bash tools/multi_gpu_round.sh <tag> [max_gpus=8]
# sharding parity (pytest, 2/4/8 ranks as visible), then the strong-scaling bench lines of the default workload
tag=${1:-r02}; M=${2:-8}
python -m pytest -m gpu -q tests/test_multigpu_gpu.py 2>&1 | tail -3
for n in 2 4 8; do [ -f gpurun_out/multigpu_parity_${n}gpu.log ] && cp gpurun_out/multigpu_parity_${n}gpu.log gpurun_out/${tag}_multigpu_parity_${n}gpu.log; done
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
python bench.py --steps 5 --warmup 3 --no-cpu --no-also 2>/dev/null | grep "^{" > gpurun_out/${tag}_bench_cfg5_1gpu.json
for n in 2 4 8; do
  [ $n -le $M ] || continue
  $TR --nproc-per-node $n --master-port $((29700 + n)) bench.py --gpus $n --steps 5 --warmup 3 --no-cpu 2> gpurun_out/${tag}_bench_cfg5_${n}gpu.err | grep "^{" > gpurun_out/${tag}_bench_cfg5_${n}gpu.json; echo cfg5_$n=$?
done
python tools/show_bench.py gpurun_out/${tag}_bench_cfg5_?gpu.json | grep -E "value|comm|at_density"
