# ncu launch list of one bench run (every launch a kernel of its own: no CUDA graph) + one tree's timeline; run under gpurun,
# then: cp gpurun_out/r02_launches.csv profiles/ && python tools/summarize_ncu.py profiles/r02_launches.csv profiles/r02_full_raw.csv > profiles/r02_summary.txt
set -x
export YTK_GBDT_NO_GRAPH=1
ARGS="--steps 2 --warmup 1 --no-cpu-baseline --no-loss-wise --no-parity-check --no-other-configs --no-strong"
python bench.py $ARGS > gpurun_out/plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches.csv python bench.py $ARGS > gpurun_out/ncu1.log 2>&1
tail -2 gpurun_out/ncu1.log | cut -c1-200
unset YTK_GBDT_NO_GRAPH
YTK_GBDT_TRACE_STAMPS=1 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-loss-wise --no-parity-check --no-other-configs --no-strong 2> gpurun_out/timeline_all.txt | tail -1 | cut -c1-100
