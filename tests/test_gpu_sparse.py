"""Sparse rows (CSR, BASELINE configs[3] shape): the reference's GBDTCoreData is dense, so the contract
is "what the dense path gives on the densified rows" -- checked against the dense CUDA path and the
oracle, stage by stage and for whole trees."""
import numpy as np
import pytest

import oracle_binding as ob
from gpu_common import assert_hist_close, assert_same_tree, maker, yb

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("gpu_required")]

KW = dict(max_depth=6, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0,
          min_split_loss=0.0, l1=0.0, l2=1.0, learning_rate=0.1)


def sparse_data(n, F, density, seed, K=1, max_bins=255):
    """dense int32 slots + the same rows as CSR; absent entries sit in a per-feature default slot
    that is not always 0, and some stored entries equal the default slot."""
    rng = np.random.default_rng(seed)
    nbs = rng.integers(2, max_bins + 1, size=F)
    start = np.concatenate([[0], np.cumsum(nbs)]).astype(np.int32)
    default = (rng.random(F) * nbs * (rng.random(F) < 0.5)).astype(np.int32)        # half the features: slot 0
    x = np.tile(default, (n, 1)).astype(np.int32)
    stored = rng.random((n, F)) < density
    stored[rng.integers(0, n, 5)] = False                                            # a few empty rows
    r, f = np.nonzero(stored)
    x[r, f] = (rng.random(len(r)) * nbs[f]).astype(np.int32)
    row_ptr = np.concatenate([[0], np.cumsum(stored.sum(axis=1))]).astype(np.int64)
    fids = f.astype(np.int32)
    slots = x[r, f].astype(np.int32)
    w = rng.normal(size=F) * (rng.random(F) < 0.2)
    z = ((x != default) * w).sum(axis=1)
    gp = np.zeros((n, K, 2), np.float32)
    for k in range(K):
        y = (rng.random(n) < 1.0 / (1.0 + np.exp(-(z * (k + 1) + rng.normal(size=n))))).astype(np.float64)
        p0 = 1.0 / (1.0 + np.exp(-rng.normal(size=n) * 0.5))
        gp[:, k, 0] = (p0 - y).astype(np.float32)
        gp[:, k, 1] = (p0 * (1 - p0)).astype(np.float32)
    return x, start, default, row_ptr, fids, slots, gp.reshape(-1)


def sparse_maker(kw, start, default, row_ptr, fids, slots, gp, K=1, chunks=3, **extra):
    n = len(row_ptr) - 1
    tm = yb.GpuTreeMaker(yb.make_params(class_num=K, **kw), start, n, sparse_default=default, nnz=len(fids), **extra)
    edges = np.linspace(0, n, chunks + 1).astype(np.int64)
    for a, b in zip(edges[:-1], edges[1:]):
        tm.upload_csr(row_ptr[a:b + 1], fids, slots, row0=a)          # row_ptr keeps its global offsets
    tm.set_grads(gp)
    return tm


def test_csr_upload_roundtrip_and_checks():
    x, start, default, row_ptr, fids, slots, gp = sparse_data(3000, 40, 0.05, 1)
    tm = sparse_maker(KW, start, default, row_ptr, fids, slots, gp)
    assert np.array_equal(tm.download_rows(), x)
    with pytest.raises(yb.YtkError):                      # dense upload into a sparse handle
        tm.upload_rows(x)
    bad = yb.GpuTreeMaker(yb.make_params(**KW), start, 3000, sparse_default=default, nnz=len(fids))
    f2 = fids.copy()
    a, b = row_ptr[7], row_ptr[8]
    if b - a >= 2:
        f2[a:b] = f2[a:b][::-1]                           # a row that is not sorted by feature
        with pytest.raises(yb.YtkError):
            bad.upload_csr(row_ptr, f2, slots)
    bad2 = yb.GpuTreeMaker(yb.make_params(**KW), start, 3000, sparse_default=default, nnz=len(fids))
    s2 = slots.copy()
    s2[0] = 100000
    with pytest.raises(yb.YtkError):
        bad2.upload_csr(row_ptr, fids, s2)
    with pytest.raises(yb.YtkError):                      # default slot outside the feature's bins
        yb.GpuTreeMaker(yb.make_params(**KW), start, 10, sparse_default=np.full(40, 999, np.int32), nnz=1)


def test_sparse_histogram_partition_equal_dense():
    n, F = 30000, 300
    x, start, default, row_ptr, fids, slots, gp = sparse_data(n, F, 0.03, 2)
    sp = sparse_maker(KW, start, default, row_ptr, fids, slots, gp)
    rows = np.sort(np.random.default_rng(5).permutation(n)[:17000]).astype(np.int32)
    fea = np.sort(np.random.default_rng(6).permutation(F)[:120]).astype(np.int32)
    for r, fi in ((None, None), (rows, None), (rows, fea)):
        got = sp.build_hist(0, rows=r, fea_index=fi)
        want = ob.build_hist(x, gp, 1, 0, np.arange(n) if r is None else r, np.arange(F) if fi is None else fi, start)
        assert_hist_close(got, want, gp)
    dn = maker(KW, x, start, gp)
    assert np.array_equal(sp.build_hist(0, rows=rows), dn.build_hist(0, rows=rows))       # bit for bit
    for fid, cond in [(3, 0.5), (17, 10.5), (250, 1.5), (0, 300.0)]:
        a, la = sp.partition_rows(rows, fid, cond)
        b, lb = dn.partition_rows(rows, fid, cond)
        assert la == lb and np.array_equal(a, b)


CASES = [("level6", dict(KW), 1), ("loss40", dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=40), 1),
         ("multiclass", dict(KW, max_depth=4), 3), ("small_pool", dict(KW, max_depth=5, histogram_pool_capacity=0.5), 1)]


@pytest.mark.parametrize("name,kw,K", CASES, ids=[c[0] for c in CASES])
def test_sparse_whole_tree_equals_dense_and_oracle(name, kw, K):
    n, F = 25000, 400
    x, start, default, row_ptr, fids, slots, gp = sparse_data(n, F, 0.02, 3, K=K)
    sp = sparse_maker(kw, start, default, row_ptr, fids, slots, gp, K=K)
    dn = maker(kw, x, start, gp, K=K)
    p = ob.make_params(class_num=K, **kw)
    for gid in range(K):
        ts, td = sp.make(gid), dn.make(gid)
        for nm, _ in yb.Tree.FIELDS:
            assert np.array_equal(getattr(ts, nm), getattr(td, nm), equal_nan=True), nm
        assert np.array_equal(sp.get_positions(), dn.get_positions())
        o, opos, _ = ob.make_tree(p, x, gp, gid, start)
        assert_same_tree(ts, o)
        assert np.array_equal(sp.get_positions(), opos)
    # feature and row sampling
    rng = np.random.default_rng(8)
    fea = np.sort(rng.permutation(F)[:150]).astype(np.int32)
    mask = (rng.random(n) < 0.6).astype(np.uint8)
    ts = sp.make(0, fea_index=fea, row_mask=mask)
    o, opos, _ = ob.make_tree(p, x, gp, 0, start, fea_index=fea, row_mask=mask)
    assert_same_tree(ts, o)
    assert np.array_equal(sp.get_positions(), opos)


def test_sparse_device_boosting_rounds():
    """score update (incl. rows left out by the row mask, which walk the tree on sparse rows) + gradients on the device"""
    n, F = 20000, 200
    x, start, default, row_ptr, fids, slots, gp = sparse_data(n, F, 0.03, 4)
    rng = np.random.default_rng(1)
    y = (rng.random(n) < 0.4).astype(np.float32)
    sp = sparse_maker(KW, start, default, row_ptr, fids, slots, gp)
    dn = maker(KW, x, start, gp)
    mask = (rng.random(n) < 0.7).astype(np.uint8)
    for tm in (sp, dn):
        tm.set_targets("sigmoid", y)
        tm.compute_grads()
    for _ in range(3):
        for tm in (sp, dn):
            tm.make(0, row_mask=mask)
            tm.update_scores(0)
        assert sp.compute_grads() == dn.compute_grads()
        assert np.array_equal(sp.get_scores(), dn.get_scores())


def test_sparse_values_binned_on_the_device():
    """ytk_gbdt_quantize_csr: stored original values -> slots with the A1 rule, as the dense quantize_rows does"""
    rng = np.random.default_rng(12)
    n, F = 4000, 60
    nbs = rng.integers(1, 300, size=F)
    start = np.concatenate([[0], np.cumsum(nbs)]).astype(np.int32)
    cuts = np.concatenate([np.sort(rng.normal(size=nb)).astype(np.float32) for nb in nbs])
    stored = rng.random((n, F)) < 0.1
    vals = rng.normal(size=(n, F)).astype(np.float32).round(2)
    fill = np.zeros(F, np.float32)                                    # missing entries are filled with 0.0
    dense_vals = np.where(stored, vals, fill[None, :]).astype(np.float32)
    want = ob.find_bins(dense_vals, cuts, start)
    default = ob.find_bins(fill[None, :], cuts, start)[0].astype(np.int32)
    r, f = np.nonzero(stored)
    row_ptr = np.concatenate([[0], np.cumsum(stored.sum(axis=1))]).astype(np.int64)
    tm = yb.GpuTreeMaker(yb.make_params(**KW), start, n, sparse_default=default, nnz=len(r))
    tm.quantize_csr(row_ptr[:2001], f.astype(np.int32), vals[r, f], cuts, row0=0)
    tm.quantize_csr(row_ptr[2000:], f.astype(np.int32), vals[r, f], cuts, row0=2000)
    assert np.array_equal(tm.download_rows(), want)
