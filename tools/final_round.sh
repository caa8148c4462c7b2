# usage (GPU box): bash tools/final_round.sh <tag>   -- the round-end evidence set under gpurun_out/
tag=${1:-r01h}
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo pytest=$?
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo smoke=$?
python bench.py > gpurun_out/${tag}_bench_cfg2.json 2> gpurun_out/${tag}_bench_cfg2.err; echo bench=$?
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_ref_cfg2.json 2>/dev/null; echo ref=$?
for c in cfg5 cfg3 cfg4 cfg1; do
  python bench.py --workload $c > gpurun_out/${tag}_bench_$c.json 2> gpurun_out/${tag}_bench_$c.err; echo $c=$?
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_cfg2.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/${tag}_launches_ncu.log 2>&1; echo launches=$?
bash tools/ncu_estep.sh ${tag}_cfg2_full_estep cfg2 100000; echo ncu=$?
rm -f gpurun_out/${tag}_launches_ncu.log
