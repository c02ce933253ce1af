#!/usr/bin/env python
"""hist_build_k on one node of 2.75 M rows reached through different row-id lists (run under
`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum -k regex:hist_build_k`): how much of the
gathered launches' time is the indirection, how much the scatter of the rows over HBM."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _pkg
import workloads as wl

yb = _pkg.load()
n, F, nb = 11_000_000, 28, 255
x8, y = wl.c2_higgs(n, F, nb, 1234)
start = (np.arange(F + 1) * nb).astype(np.int32)
gp = wl.sigmoid_grads(np.zeros(n, np.float32), y, np.empty(2 * n, np.float32))
tm = yb.GpuTreeMaker(yb.make_params(max_depth=8, min_child_hessian_sum=100.0, l2=0.0), start, n)
for r0 in range(0, n, 1 << 20):
    tm.upload_rows(x8[r0:r0 + (1 << 20)].astype(np.int32), r0)
tm.set_grads(gp)
rng = np.random.default_rng(0)
m = n // 4
lists = {
    "identity (no list), 11M rows": None,
    "list = first quarter, contiguous": np.arange(m, dtype=np.int32),
    "list = every 4th row": np.arange(0, n, 4, dtype=np.int32)[:m],
    "list = random quarter, ascending": np.sort(rng.choice(n, m, replace=False)).astype(np.int32),
    "list = random quarter, shuffled": rng.choice(n, m, replace=False).astype(np.int32),
    "list = random 1/16, ascending (x4 launches of rows)": np.sort(rng.choice(n, n // 16, replace=False)).astype(np.int32),
    "list = split on a feature (x < 64), ascending": np.nonzero(x8[:, 0] < 64)[0].astype(np.int32),
}
for name, rows in lists.items():
    for _ in range(2):
        tm.build_hist(0, rows=rows)
    print(name, 0 if rows is None else len(rows), flush=True)
