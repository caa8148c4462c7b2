# usage (on the GPU box): bash tools/ncu_estep.sh <tag> [workload] [n]   -- one `ncu --set full` capture of
# the E-step launches of EM iteration 5 (list kernel + packed-bits pass), summarised with source lines
tag=${1:-estep}; wl=${2:-cfg2}; n=${3:-20000}
python bench.py --workload $wl --n $n --steps 2 --warmup 3 --no-cpu > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:estep_kernel --launch-skip 8 -c 2 -f -o gpurun_out/${tag} \
  python bench.py --workload $wl --n $n --steps 2 --warmup 3 --no-cpu > gpurun_out/${tag}_ncu.log 2>&1
python tools/ncu_summary.py gpurun_out/${tag}.ncu-rep estep_kernel > gpurun_out/${tag}_ncu_summary.txt 2>&1
rm -f gpurun_out/${tag}.ncu-rep
