tag=r05d
python -m pytest -m gpu -q tests/test_multigpu_gpu.py 2>&1 | tail -3
for n in 2 4 8; do [ -f gpurun_out/multigpu_parity_${n}gpu.log ] && cp gpurun_out/multigpu_parity_${n}gpu.log gpurun_out/${tag}_multigpu_parity_${n}gpu.log; done
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29700 + n)) bench.py --gpus $n --steps 10 --warmup 3 --no-cpu 2> gpurun_out/${tag}_bench_cfg5_${n}gpu.err | grep "^{" > gpurun_out/${tag}_bench_cfg5_${n}gpu.json; echo cfg5_$n=$?
done
