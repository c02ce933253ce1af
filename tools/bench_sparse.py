#!/usr/bin/env python
"""BASELINE configs[3] shape (10 000 features at ~1 % density) on one GPU through the sparse-row path:
ms per tree with the gradients computed on the device.  For information; bench.py keeps the headline
workload.  python tools/bench_sparse.py [--rows N] [--features F] [--density D] [--depth 6] [--steps K]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def synth_csr(n, F, density, seed):
    """k = F*density stored entries per row, one per stratum of F/k features (sorted, distinct);
    slots 1..b_f-1 with b_f in [2,255], default slot 0; labels depend on a few features."""
    rng = np.random.default_rng(seed)
    k = max(1, int(round(F * density)))
    width = F // k
    nbs = rng.integers(2, 256, size=F)
    fids = (np.arange(k, dtype=np.int32)[None, :] * width + rng.integers(0, width, size=(n, k), dtype=np.int32))
    slots = 1 + (rng.random((n, k), dtype=np.float32) * (nbs[fids] - 1)).astype(np.int32)
    w = (rng.normal(size=F) * (rng.random(F) < 0.05)).astype(np.float32)
    z = (w[fids] * (slots / nbs[fids])).sum(axis=1)
    y = (rng.random(n) < 1.0 / (1.0 + np.exp(-3 * z))).astype(np.float32)
    row_ptr = (np.arange(n + 1, dtype=np.int64) * k)
    start = np.concatenate([[0], np.cumsum(nbs)]).astype(np.int32)
    return start, row_ptr, fids.reshape(-1), slots.reshape(-1), y


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=2_000_000)
    ap.add_argument("--features", type=int, default=10_000)
    ap.add_argument("--density", type=float, default=0.01)
    ap.add_argument("--depth", type=int, default=6)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    a = ap.parse_args()
    import _pkg
    yb = _pkg.load()
    start, row_ptr, fids, slots, y = synth_csr(a.rows, a.features, a.density, 7)
    p = yb.make_params(max_depth=a.depth, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=100.0, l2=0.0,
                       learning_rate=0.1)
    tm = yb.GpuTreeMaker(p, start, a.rows, sparse_default=np.zeros(a.features, np.int32), nnz=len(fids))
    step = 1 << 18
    for r0 in range(0, a.rows, step):
        r1 = min(a.rows, r0 + step)
        tm.upload_csr(row_ptr[r0:r1 + 1], fids, slots, row0=r0)
    tm.set_targets("sigmoid", y)
    tm.compute_grads()
    t = []
    for it in range(a.warmup + a.steps):
        t0 = time.perf_counter()
        tree = tm.make(0)
        t1 = time.perf_counter()
        tm.update_scores(0)
        tm.compute_grads()
        if it >= a.warmup:
            t.append(t1 - t0)
    ts = tm.last_stats
    ms = 1e3 * float(np.mean(t))
    print(json.dumps({"workload": "sparse rows: %d x %d features, density %g, level-wise depth %d" % (a.rows, a.features, a.density, a.depth),
                      "nnz": int(len(fids)), "bins_total": int(start[-1]), "ms_per_tree": ms, "nodes_per_tree": int(tree.n_nodes),
                      "stored_entries_per_sec_root_equivalent": len(fids) / (ms * 1e-3),
                      "hist_ms_per_tree": ts.gpu_hist_ms, "device_allocs_last_tree": ts.device_allocs}))


if __name__ == "__main__":
    main()
