set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
bash tools/ab.sh "X=1" 2>&1 | tail -3
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-loss-wise --no-parity-check --no-other-configs --no-strong > gpurun_out/plain2.log 2>&1 || exit 1
ncu --set full --import-source on --clock-control none -k regex:split_scan_k --launch-skip 24 --launch-count 8 -o gpurun_out/scan_new -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-loss-wise --no-parity-check --no-other-configs --no-strong > gpurun_out/ncu_scan.log 2>&1
tail -3 gpurun_out/ncu_scan.log
