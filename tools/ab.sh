#!/bin/bash
# A/B runs of bench.py under different env settings: tools/ab.sh "NAME=VALUE ..." "NAME=VALUE ..." ...
# prints ms per tree, histogram-kernel ms per tree, roofline fraction for each.
for cfg in "$@"; do
  out=$(env $cfg python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-loss-wise --no-parity-check 2>/dev/null | tail -1)
  python - "$cfg" "$out" <<'PY'
import json,sys
d=json.loads(sys.argv[2])
print("%-50s ms/tree %.3f  e2e %.3f  hist ms %.3f  frac %.3f  dev_round %.3f" % (sys.argv[1], d["ms_per_step"], d["e2e"]["seconds_per_tree"]*1e3, d["roofline"]["kernel_ms_per_step"], d["roofline"]["frac"], d["device_gradient_round"]["ms_per_round"]))
print("      ", {k: (round(v, 3) if isinstance(v, float) else v) for k, v in d["time_stats_ms_per_tree"].items() if k != "ms_of_each_step"})
PY
done
