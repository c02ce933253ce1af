"""BASELINE.json shapes.  At full size the oracle is too slow to be the checker, so the CUDA path is
checked through properties that do not depend on size: histogram(parent) =
histogram(left) + histogram(right) exactly, every feature's bins add up to the same totals, the
partition is a permutation consistent with the tree walk, child counts add up.  The other configs'
shapes run against the oracle at reduced row counts."""
import numpy as np
import pytest

import oracle_binding as ob
from gpu_common import assert_same_tree, maker, synth, yb

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("gpu_required")]


def _walk(tree, x):
    """node of every row by walking the tree on slots with the partition's rule (slot < cond)."""
    node = np.zeros(x.shape[0], dtype=np.int32)
    for _ in range(64):
        inner = tree.left[node] != -1
        if not inner.any():
            break
        idx = np.where(inner)[0]
        n = node[idx]
        go_left = x[idx, tree.fid[n]].astype(np.float32) < tree.node_info[n]
        node[idx] = np.where(go_left, tree.left[n], tree.right[n])
    return node


def test_config2_higgs_shape_full_size_properties():
    """11 M x 28, 255 bins, depth 8 (configs[1]) -- the bench workload."""
    import bench
    n, F, nb = 11_000_000, 28, 255
    x8, y = bench.synth_bins(n, F, nb, 1234)
    start = (np.arange(F + 1) * nb).astype(np.int32)
    gp = np.empty(2 * n, np.float32)
    bench.sigmoid_grads(np.zeros(n, np.float32), y, gp)
    kw = dict(max_depth=8, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=100.0, l2=0.0,
              learning_rate=0.1)
    tm = yb.GpuTreeMaker(yb.make_params(**kw), start, n)
    step = 1 << 20
    for r0 in range(0, n, step):
        tm.upload_rows(x8[r0:r0 + step].astype(np.int32), r0)
    tm.set_grads(gp)
    tree = tm.make(0)
    pos = tm.get_positions()
    # tree bookkeeping
    inner = tree.left != -1
    assert tree.sample_cnt[0] == n
    assert np.array_equal(tree.sample_cnt[inner], tree.sample_cnt[tree.left[inner]] + tree.sample_cnt[tree.right[inner]])
    # the partition is exactly what walking the tree gives, and every row is in exactly one leaf
    sub = np.random.default_rng(0).choice(n, 2_000_000, replace=False)
    assert np.array_equal(pos[sub], _walk(tree, x8[sub]))
    leaves, counts = np.unique(pos, return_counts=True)
    assert np.array_equal(counts, tree.sample_cnt[leaves]) and not inner[leaves].any()
    # histogram(parent) = histogram(left) + histogram(right), bit for bit (exact integer accumulation)
    root = tm.build_hist(0)
    l_rows = np.where(x8[:, tree.fid[0]].astype(np.float32) < tree.node_info[0])[0].astype(np.int32)
    r_rows = np.setdiff1d(np.arange(n, dtype=np.int32), l_rows, assume_unique=True)
    assert len(l_rows) == tree.sample_cnt[tree.left[0]]
    hl, hr = tm.build_hist(0, rows=l_rows), tm.build_hist(0, rows=r_rows)
    assert np.array_equal(tm.subtract_hist(root, hl), hr)
    # checksum of checksums: every feature's bins sum to the node totals
    tot = np.array([root[2 * start[f]: 2 * start[f + 1]].reshape(-1, 2).sum(0) for f in range(F)])
    np.testing.assert_allclose(tot, np.tile(tot[0], (F, 1)), rtol=1e-12)
    np.testing.assert_allclose(tot[0], gp.reshape(-1, 2).astype(np.float64).sum(0), rtol=1e-7)
    np.testing.assert_allclose(tree.hess_sum[0], tot[0][1], rtol=1e-6)


def test_config3_shape_128_features_leafwise_255_reduced_rows():
    """100 M x 128 regression, leaf-wise 255 leaves (configs[2]) at 40 k rows: 4 feature groups, l2 loss."""
    n, F = 40000, 128
    x, start, _ = synth(n, F, 255, 3)
    rng = np.random.default_rng(3)
    yv = (x[:, :F] @ rng.normal(size=F)) / 2000.0 + rng.normal(size=n)
    gp = np.stack([(0.0 - yv), np.ones(n)], 1).astype(np.float32).reshape(-1)      # L2: g = score - y, h = 1
    kw = dict(max_depth=-1, max_leaf_cnt=255, grow_policy="loss", min_child_hessian_sum=1.0, l2=1.0, learning_rate=0.1)
    tm = maker(kw, x, start, gp)
    t = tm.make(0)
    o, opos, _ = ob.make_tree(ob.make_params(**kw), x, gp, 0, start)
    assert_same_tree(t, o)
    assert np.array_equal(tm.get_positions(), opos)
    assert t.n_nodes == 509


def test_config5_shape_multiclass_63_bins_depth10_reduced_rows():
    """20 M x 256, 10 classes, 63 bins, depth 10 (configs[4]) at 30 k rows."""
    n, F, K = 30000, 256, 10
    x, start, gp = synth(n, F, 63, 5, K=K)
    kw = dict(max_depth=10, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0, l2=1.0, learning_rate=0.1)
    tm = maker(kw, x, start, gp, K=K)
    p = ob.make_params(class_num=K, **kw)
    for gid in (0, 4, 9):
        t = tm.make(gid)
        o, opos, _ = ob.make_tree(p, x, gp, gid, start)
        assert_same_tree(t, o)
        assert np.array_equal(tm.get_positions(), opos)


def test_config4_sparse_shape_reduced_rows_through_the_dense_path():
    """configs[3]: 10 000 features at ~1 % density (absent -> the default bin 0, present -> 1..b_f-1 with
    b_f in [2,255]), binary labels, level-wise depth 6.  There is no sparse (CSR) layout in the
    reference's GBDTCoreData and none here yet; the shape runs through the dense path -- 313 feature
    groups, 1.3 M bins per node -- at a row count the oracle can check."""
    rng = np.random.default_rng(7)
    n, F = 12000, 10000
    nbs = rng.integers(2, 256, size=F)
    start = np.concatenate([[0], np.cumsum(nbs)]).astype(np.int32)
    x = np.zeros((n, F), dtype=np.int32)
    nnz = rng.random((n, F)) < 0.01
    r, f = np.nonzero(nnz)
    x[r, f] = 1 + (rng.random(len(r)) * (nbs[f] - 1)).astype(np.int64)
    w = rng.normal(size=F) * (rng.random(F) < 0.02)
    y = (rng.random(n) < 1.0 / (1.0 + np.exp(-(nnz * w).sum(axis=1) * 3))).astype(np.float64)
    gp = np.empty((n, 2), np.float32)
    gp[:, 0] = (0.5 - y).astype(np.float32)
    gp[:, 1] = 0.25
    gp = gp.reshape(-1)
    kw = dict(max_depth=6, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0, l2=1.0, learning_rate=0.1)
    tm = maker(kw, x, start, gp)
    t = tm.make(0)
    o, opos, _ = ob.make_tree(ob.make_params(**kw), x, gp, 0, start)
    assert_same_tree(t, o)
    assert np.array_equal(tm.get_positions(), opos)
    assert t.n_nodes > 15
