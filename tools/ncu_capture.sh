# usage (on the GPU box): bash tools/ncu_capture.sh <tag> <kernel-regex> <launch-skip> <launch-count> <bench args...>
# One `ncu --set full` capture of the chosen launches of a bench command that has first exited 0 WITHOUT ncu;
# the report is summarised (headline metrics, stall breakdown, top source lines) and removed.
tag=$1; rx=$2; skip=$3; cnt=$4; shift 4
python bench.py "$@" > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${tag}_plain.err; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$rx" --launch-skip $skip -c $cnt -f -o gpurun_out/${tag} \
  python bench.py "$@" > gpurun_out/${tag}_ncu.log 2>&1
python tools/ncu_summary.py gpurun_out/${tag}.ncu-rep "$rx" > gpurun_out/${tag}_ncu_summary.txt 2>&1
rm -f gpurun_out/${tag}.ncu-rep
