#!/bin/bash
# DRAM bytes and time of the histogram / partition launches of one tree under different L2 fetch granularities
# (cudaLimitMaxL2FetchGranularity via YTK_GBDT_L2_FETCH): tools/l2fetch_probe.sh > gpurun_out/l2fetch.txt
for g in 32 64 128; do
  YTK_GBDT_L2_FETCH=$g YTK_GBDT_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum --clock-control none \
     -k regex:"hist_span_k|partition_fused_k" -s 16 -c 16 --csv --log-file gpurun_out/l2f_$g.csv \
     python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-loss-wise --no-parity-check --no-other-configs 2>&1 > /dev/null | grep -m1 cudaLimit
  python - $g <<'PY'
import csv, collections, sys
g = sys.argv[1]
rows = [r for r in csv.reader(l for l in open("gpurun_out/l2f_%s.csv" % g) if not l.startswith("=="))]
hdr = rows[0]; ix = {n: i for i, n in enumerate(hdr)}
t = collections.OrderedDict()
for r in rows[1:]:
    k = r[ix["Kernel Name"]].split("(")[0]
    v = float(r[ix["Metric Value"]].replace(",", "")); u = r[ix["Metric Unit"]]
    a = t.setdefault(k, [0.0, 0.0])
    if r[ix["Metric Name"]] == "gpu__time_duration.sum":
        a[0] += v / 1000 if u == "ns" else v
    else:
        a[1] += v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}[u]
for k, v in t.items():
    print("L2 fetch granularity %3s B  %-20s %8.1f us  %8.0f MB DRAM read (one tree: 8 launches)" % (g, k, v[0], v[1]))
PY
done
