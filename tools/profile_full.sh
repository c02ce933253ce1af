# ncu --set full of the three main kernels (root + three levels); run under gpurun, then export with
# ncu -i gpurun_out/r02_full.ncu-rep --page raw --csv > profiles/r02_full_raw.csv
set -x
export YTK_GBDT_NO_GRAPH=1
ARGS="--steps 2 --warmup 1 --no-cpu-baseline --no-loss-wise --no-parity-check --no-other-configs --no-strong"
python bench.py $ARGS > gpurun_out/plain.log 2>&1 || exit 1
ncu --set full --import-source on --clock-control none -k regex:'hist_span_k|partition_fused_k|split_scan_k' --launch-skip 24 --launch-count 12 -o gpurun_out/r02_full -f python bench.py $ARGS > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log | cut -c1-200
