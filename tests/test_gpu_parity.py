"""Parity of the CUDA path (through the C ABI) against the CPU oracle, stage by stage
 and end to end.  Bit-exact for bins, partitions and tree
structure; 1e-5 relative for histogram sums and gains."""
import numpy as np
import pytest

import kat
import oracle_binding as ob
from gpu_common import REL, assert_hist_close, assert_same_tree, maker, synth, yb

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("gpu_required")]

KW = dict(max_depth=6, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0,
          min_split_loss=0.0, l1=0.0, l2=1.0, learning_rate=0.1)


# ---------------------------------------------------------------- A1 value -> bin
def test_a1_bins_golden_vectors():
    import os
    import ytk_data as yd
    d = np.load(os.path.join(yd.GOLDEN, "bins_golden.npz"))
    cuts, vals, want = d["cuts"], d["vals"], d["bins"]
    start = np.array([0, len(cuts)], np.int32)
    tm = yb.GpuTreeMaker(yb.make_params(**KW), start, len(vals))
    tm.quantize_rows(vals.reshape(-1, 1), cuts)
    assert np.array_equal(tm.download_rows().reshape(-1), want)


@pytest.mark.parametrize("n,F,S", [(5000, 7, 7), (3001, 33, 40), (1, 1, 1)])
def test_a1_bins_bit_exact(n, F, S):
    rng = np.random.default_rng(n + F)
    nbs = rng.integers(1, 200, size=F)
    cuts, start = [], [0]
    for f in range(F):
        c = np.unique(rng.normal(size=nbs[f]).astype(np.float32).round(2))
        cuts.append(c)
        start.append(start[-1] + len(c))
    cuts = np.concatenate(cuts)
    start = np.array(start, np.int32)
    vals = rng.normal(size=(n, S)).astype(np.float32).round(2)
    vals[rng.random((n, S)) < 0.05] *= 10
    want = ob.find_bins(vals[:, :F], cuts, start)
    tm = yb.GpuTreeMaker(yb.make_params(**KW), start, n)
    tm.quantize_rows(vals, cuts)
    assert np.array_equal(tm.download_rows(), want)


def test_upload_download_roundtrip_and_range_check():
    x, start, gp = synth(999, 37, 200, 1, S=41)
    tm = yb.GpuTreeMaker(yb.make_params(**KW), start, 999)
    tm.upload_rows(x[:500], 0)
    tm.upload_rows(x[500:], 500)
    assert np.array_equal(tm.download_rows(), x[:, :37])
    bad = x.copy()
    bad[3, 2] = 256
    with pytest.raises(yb.YtkError):
        tm.upload_rows(bad)


def test_more_than_65536_bins_is_refused():
    with pytest.raises(yb.YtkError) as ei:
        yb.GpuTreeMaker(yb.make_params(**KW), np.array([0, 70000], np.int32), 10)
    assert ei.value.code == -5


# ---------------------------------------------------------------- features with more than 256 bins
# ("no_sample" on a continuous column, or sample_by_cnt over many threads): uint16 slots, their
# histograms built by hist_wide_k; everything else is the same path
WIDE_NBS = [300, 5, 1000, 256, 257, 4000, 2, 255, 40, 700]


def test_wide_features_upload_quantize_roundtrip():
    rng = np.random.default_rng(11)
    n, F = 20000, len(WIDE_NBS)
    x, start, gp = synth(n, F, WIDE_NBS, 5, S=F + 2)
    tm = yb.GpuTreeMaker(yb.make_params(**KW), start, n)
    tm.upload_rows(x[:7000], 0)
    tm.upload_rows(x[7000:], 7000)
    assert np.array_equal(tm.download_rows(), x[:, :F])
    bad = x.copy()
    bad[5, 0] = 300                      # feature 0 has slots 0..299
    with pytest.raises(yb.YtkError):
        tm.upload_rows(bad)
    cuts = np.concatenate([np.sort(rng.normal(size=nb)).astype(np.float32) for nb in WIDE_NBS])
    vals = rng.normal(size=(n, F)).astype(np.float32)
    tm2 = yb.GpuTreeMaker(yb.make_params(**KW), start, n)
    tm2.quantize_rows(vals, cuts)
    assert np.array_equal(tm2.download_rows(), ob.find_bins(vals, cuts, start))


def test_wide_features_histogram_partition_split():
    n, F = 40000, len(WIDE_NBS)
    x, start, gp = synth(n, F, WIDE_NBS, 9, skew=True)
    tm = maker(KW, x, start, gp)
    rows = np.sort(np.random.default_rng(2).permutation(n)[:25000]).astype(np.int32)
    got = tm.build_hist(0, rows=rows)
    want = ob.build_hist(x, gp, 1, 0, rows, np.arange(F), start)
    assert_hist_close(got, want, gp)
    for fid, cond in [(0, 120.5), (2, 999.5), (5, 0.5), (4, 256.5), (3, 255.5)]:
        out, lc = tm.partition_rows(rows, fid, cond)
        left_mask = x[rows, fid].astype(np.float32) < np.float32(cond)
        assert lc == int(left_mask.sum())
        assert np.array_equal(out[:lc], rows[left_mask]) and np.array_equal(out[lc:], rows[~left_mask])
    got = tm.find_split(want, len(rows))
    best, G, H, rg = _oracle_split(ob.make_params(**KW), want, start, len(rows), np.arange(F, dtype=np.int32))
    assert got.fid == best[1] and got.lo == best[2] and got.hi == best[3], (got.as_tuple(), best)
    assert abs(got.loss_chg - best[0]) <= REL * max(1.0, abs(best[0]))


@pytest.mark.parametrize("kw", [dict(KW, max_depth=6), dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=40)],
                         ids=["level6", "loss40"])
def test_wide_features_whole_tree(kw):
    n, F = 50000, len(WIDE_NBS)
    x, start, gp = synth(n, F, WIDE_NBS, 13, skew=True)
    tm = maker(kw, x, start, gp)
    t = tm.make(0)
    o, opos, _ = ob.make_tree(ob.make_params(**kw), x, gp, 0, start)
    assert_same_tree(t, o)
    assert np.array_equal(tm.get_positions(), opos)
    inner = t.left != -1
    assert (np.asarray(WIDE_NBS)[t.fid[inner]] > 256).any()       # the tree does split on wide features
    # feature and row sampling on top
    rng = np.random.default_rng(4)
    fea = np.sort(rng.permutation(F)[:6]).astype(np.int32)
    mask = (rng.random(n) < 0.5).astype(np.uint8)
    t = tm.make(0, fea_index=fea, row_mask=mask)
    o, opos, _ = ob.make_tree(ob.make_params(**kw), x, gp, 0, start, fea_index=fea, row_mask=mask)
    assert_same_tree(t, o)
    assert np.array_equal(tm.get_positions(), opos)


# ---------------------------------------------------------------- A2 histogram
@pytest.mark.parametrize("n,F,nb,K,skew", [
    (40000, 28, 255, 1, False), (30011, 128, 255, 1, True), (20000, 5, 256, 1, False),
    (7000, 65, [3, 17, 255, 1, 64] * 13, 3, True), (33, 28, 255, 1, False), (1, 2, 2, 1, False)])
def test_a2_hist_all_rows(n, F, nb, K, skew):
    x, start, gp = synth(n, F, nb, 11, K=K, skew=skew)
    tm = maker(KW, x, start, gp, K=K)
    for gid in range(K):
        got = tm.build_hist(gid)
        want = ob.build_hist(x, gp, K, gid, np.arange(n), np.arange(F), start)
        assert_hist_close(got, want, gp.reshape(-1, K, 2)[:, gid], "gid %d" % gid)


def test_a2_hist_row_list_and_feature_subset():
    n, F = 50000, 40
    x, start, gp = synth(n, F, 100, 5)
    tm = maker(KW, x, start, gp)
    rng = np.random.default_rng(0)
    rows = rng.permutation(n)[:17777].astype(np.int32)
    fea = np.sort(rng.permutation(F)[:11]).astype(np.int32)
    got = tm.build_hist(0, rows=rows, fea_index=fea)
    want = ob.build_hist(x, gp, 1, 0, rows, fea, start)
    assert_hist_close(got, want, gp)
    off = np.setdiff1d(np.arange(F), fea)
    for f in off:
        assert not got[2 * start[f]: 2 * start[f + 1]].any()
    assert not tm.build_hist(0, rows=np.zeros(0, np.int32)).any()      # empty node


def test_a2_hist_is_exact_and_order_independent():
    """Fixed-point accumulation: any row order gives the same bits, and every feature's bins
    add up to the same totals (checksum of checksums)."""
    n, F = 60000, 28
    x, start, gp = synth(n, F, 255, 9, skew=True)
    tm = maker(KW, x, start, gp)
    rows = np.arange(n, dtype=np.int32)
    a = tm.build_hist(0, rows=rows)
    b = tm.build_hist(0, rows=rows[::-1].copy())
    c = tm.build_hist(0)
    assert np.array_equal(a, b) and np.array_equal(a, c)
    tot = np.array([a[2 * start[f]: 2 * start[f + 1]].reshape(-1, 2).sum(0) for f in range(F)])
    assert np.allclose(tot, tot[0], rtol=1e-12)
    g64 = gp.reshape(-1, 2).astype(np.float64).sum(0)
    np.testing.assert_allclose(tot[0], g64, rtol=1e-6)


def test_a2_extreme_gradients():
    n, F = 4096, 3
    x, start, _ = synth(n, F, 16, 2)
    rng = np.random.default_rng(1)
    g = rng.normal(size=n).astype(np.float32)
    g[::7] *= 1e-30
    g[5] = 3e6
    g[9] = 0.0
    h = np.abs(rng.normal(size=n)).astype(np.float32) * np.float32(1e-12)
    h[11] = 0.0
    gp = np.stack([g, h], 1).reshape(-1)
    tm = maker(KW, x, start, gp)
    got = tm.build_hist(0)
    want = ob.build_hist(x, gp, 1, 0, np.arange(n), np.arange(F), start)
    assert_hist_close(got, want, gp)
    assert np.array_equal(got != 0, want != 0)      # "bin has samples" is preserved exactly


# ---------------------------------------------------------------- A3 subtraction
def test_a3_subtract():
    n, F = 30000, 28
    x, start, gp = synth(n, F, 255, 4)
    tm = maker(KW, x, start, gp)
    rows = np.random.default_rng(2).permutation(n).astype(np.int32)
    parent = tm.build_hist(0, rows=rows)
    small = tm.build_hist(0, rows=rows[:9000])
    large = tm.build_hist(0, rows=rows[9000:])
    got = tm.subtract_hist(parent, small)
    np.testing.assert_allclose(got, parent - small, rtol=0, atol=1e-12 * np.abs(parent).max())
    assert_hist_close(got, large, gp)


# ---------------------------------------------------------------- A4-A6 split search
def _oracle_split(p, hist, start, sample_cnt, fea):
    G = hist[2 * start[fea[0]]: 2 * start[fea[0] + 1]: 2].sum() if False else 0.0
    g = h = 0.0
    for i in range(start[fea[0]], start[fea[0] + 1]):     # sequential like the reference
        g += hist[2 * i]
        h += hist[2 * i + 1]
    rg = np.float32(ob.calc_gain(p, g, h))
    best = (-np.inf, -1, -1, -1)
    if h >= 2 * p["min_child_hessian_sum"] and sample_cnt >= p["min_split_samples"]:
        for f in fea:
            best = ob.enumerate_split(p, hist, start, int(f), g, h, float(rg), best)
    return best, g, h, rg


@pytest.mark.parametrize("seed", range(6))
def test_a456_find_split_random_hist(seed):
    rng = np.random.default_rng(seed)
    F = int(rng.integers(1, 40))
    nbs = rng.integers(1, 256, size=F)
    start = np.concatenate([[0], np.cumsum(nbs)]).astype(np.int32)
    B = int(start[-1])
    kw = dict(KW, min_child_hessian_sum=float(rng.choice([0.0, 1.0, 30.0])), l1=float(rng.choice([0.0, 0.5])),
              max_abs_leaf_val=float(rng.choice([-1.0, 0.3])))
    # a consistent histogram: every feature's bins partition the same rows
    n = 5000
    g = rng.normal(size=n)
    h = rng.random(n) * 0.25
    hist = np.zeros(2 * B)
    for f in range(F):
        b = np.minimum((rng.random(n) ** 3 * nbs[f]).astype(np.int64), nbs[f] - 1)
        if f % 3 == 0:
            b = (b // 5) * 5                              # empty bins in between
        np.add.at(hist, 2 * (start[f] + b), g)
        np.add.at(hist, 2 * (start[f] + b) + 1, h)
    tm = yb.GpuTreeMaker(yb.make_params(**kw), start, 8)
    p = ob.make_params(**kw)
    fea = np.arange(F, dtype=np.int32) if seed % 2 == 0 else np.sort(rng.permutation(F)[: max(1, F // 2)]).astype(np.int32)
    got = tm.find_split(hist, n, fea_index=None if seed % 2 == 0 else fea)
    want, G, H, rg = _oracle_split(p, hist, start, n, fea)
    assert got.fid == want[1] and got.lo == want[2] and got.hi == want[3], (got.as_tuple(), want)
    if np.isfinite(want[0]):
        assert abs(got.loss_chg - want[0]) <= REL * max(1.0, abs(want[0]))
    np.testing.assert_allclose([got.sum_grad, got.sum_hess], [G, H], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(got.root_gain, rg, rtol=REL)


@pytest.mark.parametrize("lead", [1, 3, 40])
def test_a5_bins_with_gradient_but_zero_hessian_before_the_first_real_bin(lead):
    """enumerateSplit tests "left is initialised" with GradStats.empty() = (H == 0)
    (J/optimizer/gbdt/DataParallelTreeMaker.java:598-637, J/data/gbdt/GradStats.java:73-75): while the
    accumulated left hessian is still exactly 0 -- leading bins with g != 0 and h == 0 -- bins keep being
    added and `last` advances without any candidate being evaluated; the first candidate comes at the
    second bin whose prefix has H != 0.  Features: (0) `lead` degenerate bins first, (1) degenerate bins
    in the middle (prefix H already > 0: ordinary candidates), (2) only degenerate bins, (3) a degenerate
    bin whose g cancels a later one."""
    rng = np.random.default_rng(lead)
    nb = 64
    F = 4
    start = (np.arange(F + 1) * nb).astype(np.int32)
    hist = np.zeros((F, nb, 2))
    # feature 0: degenerate lead, then real bins (some empty in between)
    hist[0, :lead, 0] = rng.normal(size=lead) + 3.0
    real = np.arange(lead, nb)[rng.random(nb - lead) < 0.6]
    hist[0, real, 0] = rng.normal(size=len(real))
    hist[0, real, 1] = rng.random(len(real)) + 0.1
    # feature 1: real bins with a few (g != 0, h == 0) bins in the middle
    hist[1, :, 0] = rng.normal(size=nb)
    hist[1, :, 1] = rng.random(nb) + 0.1
    hist[1, 10:14, 1] = 0.0
    hist[1, 30, :] = 0.0
    # feature 2: nothing but degenerate bins
    hist[2, ::3, 0] = rng.normal(size=len(range(0, nb, 3)))
    # feature 3: +g, h == 0 first; -g later inside a real bin
    hist[3, 0] = (5.0, 0.0)
    hist[3, 5] = (-5.0, 1.0)
    hist[3, 9] = (1.0, 2.0)
    hist[3, 20] = (-2.0, 1.5)
    hist = hist.reshape(-1)
    for kw in (dict(KW, min_child_hessian_sum=0.0, l2=0.0), dict(KW, min_child_hessian_sum=0.5, l2=1.0)):
        tm = yb.GpuTreeMaker(yb.make_params(**kw), start, 8)
        p = ob.make_params(**kw)
        for fea in ([0], [1], [2], [3], [0, 1, 2, 3], [2, 3]):
            fea = np.array(fea, np.int32)
            got = tm.find_split(hist, 1000, fea_index=fea)
            want, G, H, rg = _oracle_split(p, hist, start, 1000, fea)
            assert (got.fid, got.lo, got.hi) == (want[1], want[2], want[3]), (list(fea), got.as_tuple(), want)
            if np.isfinite(want[0]):
                assert abs(got.loss_chg - want[0]) <= REL * max(1.0, abs(want[0]))
            else:
                assert got.loss_chg == want[0]
        tm.close()


def test_a6_tie_rule_smallest_feature_then_lowest_bin():
    # three identical features: every candidate ties across features -> fid 0, first best bin
    nb = 6
    start = np.array([0, nb, 2 * nb, 3 * nb], np.int32)
    one = np.array([[-4, 2], [0, 0], [3, 2], [-4, 2], [3, 2], [2, 2]], dtype=np.float64)
    hist = np.concatenate([one, one, one]).reshape(-1)
    kw = dict(KW, min_child_hessian_sum=0.0, l2=0.0)
    tm = yb.GpuTreeMaker(yb.make_params(**kw), start, 8)
    got = tm.find_split(hist, 100)
    want, *_ = _oracle_split(ob.make_params(**kw), hist, start, 100, [0, 1, 2])
    assert got.as_tuple()[1:] == want[1:]
    assert got.fid == 0
    got2 = tm.find_split(hist, 100, fea_index=np.array([1, 2], np.int32))
    assert got2.fid == 1 and (got2.lo, got2.hi) == (got.lo, got.hi)


def test_a4_cannot_split_leaves_record_cleared():
    start = np.array([0, 4], np.int32)
    hist = np.array([1, .2, -1, .2, 2, .2, 0, 0], dtype=np.float64)
    kw = dict(KW, min_child_hessian_sum=1.0)
    tm = yb.GpuTreeMaker(yb.make_params(**kw), start, 8)
    got = tm.find_split(hist, 10)           # H = 0.6 < 2*1
    assert got.fid == -1 and got.loss_chg == -np.inf
    kw2 = dict(KW, min_child_hessian_sum=0.1, min_split_samples=50)
    tm2 = yb.GpuTreeMaker(yb.make_params(**kw2), start, 8)
    assert tm2.find_split(hist, 10).fid == -1        # sampleCnt < min_split_samples
    assert tm2.find_split(hist, 50).fid == 0


# ---------------------------------------------------------------- A7 partition
@pytest.mark.parametrize("n,m", [(100000, 100000), (70001, 12345), (5000, 1), (5000, 0)])
def test_a7_partition_sets_bit_exact(n, m):
    F = 28
    x, start, gp = synth(n, F, 255, 8)
    tm = maker(KW, x, start, gp)
    rows = np.random.default_rng(3).permutation(n)[:m].astype(np.int32)
    for fid, cond in [(3, 100.5), (27, 0.5), (0, 254.5), (5, 300.0), (9, 17.0)]:
        out, lc = tm.partition_rows(rows, fid, cond)
        pos = np.zeros(2 * m, np.int32)
        pos[1::2] = rows
        olc = ob.reset_position(pos, 0, m, x, fid, cond, 1, 2)
        assert lc == olc
        assert np.array_equal(np.sort(out[:lc]), np.sort(pos[1::2][:olc]))
        assert np.array_equal(np.sort(out[lc:]), np.sort(pos[1::2][olc:]))
        # stable: both sides keep the input order
        left_mask = x[rows, fid].astype(np.float32) < np.float32(cond)
        assert np.array_equal(out[:lc], rows[left_mask]) and np.array_equal(out[lc:], rows[~left_mask])


# ---------------------------------------------------------------- make(): whole trees
CASES = [
    ("level_d6", dict(KW), 30000, 28, 255, 1, False),
    ("loss_31", dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=31), 30000, 28, 255, 1, True),
    ("loss_depth_leafcap", dict(KW, grow_policy="loss", max_depth=5, max_leaf_cnt=16), 20000, 117, 2, 1, False),
    ("level_leafcap", dict(KW, max_depth=6, max_leaf_cnt=20), 20000, 40, 63, 1, False),
    ("wide_128", dict(KW, max_depth=5), 15000, 128, 255, 1, True),
    ("min_split_samples", dict(KW, max_depth=8, min_split_samples=400, min_child_hessian_sum=0.0), 20000, 10, 32, 1, False),
    ("l1_capped", dict(KW, l1=0.7, max_abs_leaf_val=0.4, max_depth=5), 20000, 12, 100, 1, False),
    ("min_split_loss", dict(KW, min_split_loss=2.5, max_depth=7), 20000, 12, 100, 1, False),
    ("multiclass", dict(KW, max_depth=4), 12000, 33, 63, 4, False),
    ("stump_depth0", dict(KW, max_depth=0, max_leaf_cnt=4), 3000, 5, 10, 1, False),
    ("tiny", dict(KW, max_depth=3, min_child_hessian_sum=0.0), 9, 3, 4, 1, False),
    ("loss_small_pool", dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=24, histogram_pool_capacity=0.3), 20000, 12, 100, 1, False),
    ("loss_multiclass", dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=20), 15000, 20, 63, 3, True),
    ("loss_l1_capped", dict(KW, grow_policy="loss", max_depth=6, max_leaf_cnt=30, l1=0.4, max_abs_leaf_val=0.5), 20000, 12, 100, 1, False),
    ("loss_min_split_samples", dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=30, min_split_samples=700, min_child_hessian_sum=0.0), 20000, 10, 32, 1, False),
]


@pytest.mark.parametrize("name,kw,n,F,nb,K,skew", CASES, ids=[c[0] for c in CASES])
def test_make_tree_matches_oracle(name, kw, n, F, nb, K, skew):
    x, start, gp = synth(n, F, nb, 21, K=K, skew=skew)
    tm = maker(kw, x, start, gp, K=K)
    p = ob.make_params(class_num=K, **kw)
    for gid in range(K):
        t = tm.make(gid)
        o, opos, info = ob.make_tree(p, x, gp, gid, start)
        assert_same_tree(t, o)
        assert np.array_equal(tm.get_positions(), opos)
        assert tm.last_stats.hist_miss == info["hist_miss"]
        assert tm.last_stats.kernel_launches > 0


def test_make_tree_feature_and_row_sampling():
    n, F = 30000, 28
    x, start, gp = synth(n, F, 255, 31)
    rng = np.random.default_rng(4)
    fea = np.sort(rng.permutation(F)[:9]).astype(np.int32)
    mask = (rng.random(n) < 0.6).astype(np.uint8)
    tm = maker(KW, x, start, gp)
    p = ob.make_params(**KW)
    t = tm.make(0, fea_index=fea, row_mask=mask)
    o, opos, _ = ob.make_tree(p, x, gp, 0, start, fea_index=fea, row_mask=mask)
    assert_same_tree(t, o)
    assert np.array_equal(tm.get_positions(), opos)
    assert set(np.unique(t.fid[t.left != -1])) <= set(fea.tolist())


def test_make_tree_small_pool_rebuilds_from_rows():
    """histogram_pool_capacity forces evictions: misses rebuild the larger child from rows and
    must give the very same tree as the cached (subtracted) path."""
    n, F = 30000, 28
    x, start, gp = synth(n, F, 255, 41)
    kw = dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=24)
    big = maker(kw, x, start, gp)
    t_big = big.make(0)
    cap_mb = 3 * 64 * int(start[-1]) / 1048576.0          # -> capacity 3
    kw_small = dict(kw, histogram_pool_capacity=cap_mb)
    small = maker(kw_small, x, start, gp)
    t_small = small.make(0)
    o, opos, info = ob.make_tree(ob.make_params(**kw_small), x, gp, 0, start)
    assert info["hist_miss"] > 0 and small.last_stats.hist_miss == info["hist_miss"]
    assert_same_tree(t_small, o)
    assert np.array_equal(small.get_positions(), opos)


def test_repeated_make_is_deterministic():
    n, F = 40000, 28
    x, start, gp = synth(n, F, 255, 51, skew=True)
    kw = dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=63)
    tm = maker(kw, x, start, gp)
    a = tm.make(0)
    pa = tm.get_positions()
    b = tm.make(0)
    for name, _ in yb.Tree.FIELDS:
        assert np.array_equal(getattr(a, name), getattr(b, name), equal_nan=True), name
    assert np.array_equal(pa, tm.get_positions())


# ---------------------------------------------------------------- the reference's own known answers
def _gpu_maker(d, params):
    xb = d["xb"].astype(np.int32)
    p = yb.make_params(**params)
    tm = yb.GpuTreeMaker(p, d["start"], xb.shape[0])
    tm.upload_rows(xb)

    def make_tree(gp, gid):
        tm.set_grads(gp)
        return tm.make(gid)
    return make_tree


def test_kat1_agaricus_losses_on_gpu():
    d = kat.load_kat1()
    out, trees = kat.run_kat1(_gpu_maker(d, kat.KAT1_PARAMS), d)
    for it, (tr, te) in kat.KAT1.items():
        assert abs(out[it][0] - tr) < 1e-9, (it, out[it][0], tr)
        assert abs(out[it][1] - te) < 1e-9, (it, out[it][1], te)
    assert [t.n_nodes for t in trees] == [21, 23, 23]


def test_kat2_dermatology_prediction_on_gpu():
    d = kat.load_kat2()
    pred, trees = kat.run_kat2(_gpu_maker(d, kat.KAT2_PARAMS), d)
    err = np.abs(pred - np.array(kat.KAT2_PRED)[None, :]).max(axis=1)
    # 30-bit fixed-point gradients; north_star bar is 1e-5 relative.  All six rows the README line belongs
    # to, and the tree sizes of all five rounds.
    assert list(np.where(err < 1e-9)[0]) == kat.KAT2_ROWS, (np.where(err < 1e-9)[0], err.min())
    assert [[t.n_nodes for t in g] for g in trees] == kat.KAT2_TREE_SIZES


# ---------------------------------------------------------------- batching and fixed-point width
@pytest.mark.parametrize("kw", [dict(KW, max_depth=7), dict(KW, max_depth=6, max_leaf_cnt=20)],
                         ids=["depth7", "leafcap20"])
def test_level_batching_equals_node_at_a_time(kw, monkeypatch):
    """Level-wise growth launches one kernel set per level; it must give bit-identical trees to the
    reference's node-at-a-time order (YTK_GBDT_SEQUENTIAL=1 forces that order)."""
    n, F = 50000, 28
    x, start, gp = synth(n, F, 255, 61, skew=True)
    a = maker(kw, x, start, gp)
    ta = a.make(0)
    pa = a.get_positions()
    monkeypatch.setenv("YTK_GBDT_SEQUENTIAL", "1")
    b = maker(kw, x, start, gp)
    tb = b.make(0)
    for name, _ in yb.Tree.FIELDS:
        assert np.array_equal(getattr(ta, name), getattr(tb, name), equal_nan=True), name
    assert np.array_equal(pa, b.get_positions())
    assert a.last_stats.kernel_launches < b.last_stats.kernel_launches


@pytest.mark.parametrize("kw", [dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=100),
                                dict(KW, grow_policy="loss", max_depth=6, max_leaf_cnt=40),
                                dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=64, min_split_loss=1.0)],
                         ids=["leaves100", "depth6_leaves40", "min_split_loss"])
@pytest.mark.parametrize("ahead", [3, 64])
def test_loss_wise_lookahead_equals_node_at_a_time(kw, ahead, monkeypatch):
    """Loss-wise growth computes the expansions of the best few queue entries per batch of launches and
    applies them in the reference's order: bit-identical trees to one-node-at-a-time
    (YTK_GBDT_LOOKAHEAD=1), and to the oracle."""
    n, F = 60000, 28
    x, start, gp = synth(n, F, 255, 67, skew=True)
    monkeypatch.setenv("YTK_GBDT_LOOKAHEAD", str(ahead))
    a = maker(kw, x, start, gp)
    ta = a.make(0)
    pa = a.get_positions()
    monkeypatch.setenv("YTK_GBDT_LOOKAHEAD", "1")
    b = maker(kw, x, start, gp)
    tb = b.make(0)
    for name, _ in yb.Tree.FIELDS:
        assert np.array_equal(getattr(ta, name), getattr(tb, name), equal_nan=True), name
    assert np.array_equal(pa, b.get_positions())
    assert a.last_stats.kernel_launches < b.last_stats.kernel_launches
    o, opos, _ = ob.make_tree(ob.make_params(**kw), x, gp, 0, start)
    assert_same_tree(ta, o)
    assert np.array_equal(pa, opos)
    # a second tree on the same handle (expansions left over from the first must not leak into it)
    ta2 = a.make(0)
    for name, _ in yb.Tree.FIELDS:
        assert np.array_equal(getattr(ta, name), getattr(ta2, name), equal_nan=True), name


@pytest.mark.parametrize("kw", [dict(KW, max_depth=8), dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=80)],
                         ids=["level8", "loss80"])
def test_single_pass_partition_equals_stable_partition(kw, monkeypatch):
    """make() partitions rows with one chained-scan kernel that writes right rows in reverse order; the
    trees and the per-row leaves are those of the stable three-kernel version (YTK_GBDT_STABLE_PARTITION=1)."""
    n, F = 300000, 28            # 147 tiles at the root: long look-back chains
    x, start, gp = synth(n, F, 255, 73, skew=True)
    a = maker(kw, x, start, gp)
    ta = a.make(0)
    pa = a.get_positions()
    monkeypatch.setenv("YTK_GBDT_STABLE_PARTITION", "1")
    b = maker(kw, x, start, gp)
    tb = b.make(0)
    for name, _ in yb.Tree.FIELDS:
        assert np.array_equal(getattr(ta, name), getattr(tb, name), equal_nan=True), name
    assert np.array_equal(pa, b.get_positions())
    assert a.last_stats.kernel_launches < b.last_stats.kernel_launches


@pytest.mark.parametrize("kw,nb", [(dict(KW, max_depth=8), 255),
                                   (dict(KW, max_depth=9, min_split_loss=3.0), 255),
                                   (dict(KW, max_depth=7, max_leaf_cnt=37), 63),
                                   (dict(KW, max_depth=6), [300, 5, 1000, 256, 257, 4000, 2, 255, 40, 700])],
                         ids=["depth8", "min_split_loss", "leafcap37", "wide"])
def test_device_planned_partition_equals_host_planned(kw, nb, monkeypatch):
    """Level-wise growth on one GPU: the device plans and starts the next level's partition from the
    split records (and the host cross-checks its decisions).  Same trees and per-row leaves as with
    host-built partition tables (YTK_GBDT_NO_PLAN_AHEAD=1) and as the oracle."""
    n = 120000
    F = 28 if np.isscalar(nb) else len(nb)
    x, start, gp = synth(n, F, nb, 79, skew=True)
    a = maker(kw, x, start, gp)
    ta = a.make(0)
    pa = a.get_positions()
    ta2 = a.make(0)                                      # and again on the same handle
    monkeypatch.setenv("YTK_GBDT_NO_PLAN_AHEAD", "1")
    b = maker(kw, x, start, gp)
    tb = b.make(0)
    for name, _ in yb.Tree.FIELDS:
        assert np.array_equal(getattr(ta, name), getattr(tb, name), equal_nan=True), name
        assert np.array_equal(getattr(ta, name), getattr(ta2, name), equal_nan=True), name
    assert np.array_equal(pa, b.get_positions())
    assert a.last_stats.kernel_launches != b.last_stats.kernel_launches      # (the planned path really ran)
    o, opos, _ = ob.make_tree(ob.make_params(**kw), x, gp, 0, start)
    assert_same_tree(ta, o)
    assert np.array_equal(pa, opos)


@pytest.mark.parametrize("bits", [16, 30])
def test_fixed_point_width(bits, monkeypatch):
    monkeypatch.setenv("YTK_GBDT_FIXED_BITS", str(bits))
    n, F = 30000, 28
    x, start, gp = synth(n, F, 255, 71)
    tm = maker(KW, x, start, gp)
    got = tm.build_hist(0)
    want = ob.build_hist(x, gp, 1, 0, np.arange(n), np.arange(F), start)
    if bits == 30:
        np.testing.assert_allclose(got, want, rtol=0, atol=1e-7 * np.abs(want).max())
        t = tm.make(0)
        o, opos, _ = ob.make_tree(ob.make_params(**KW), x, gp, 0, start)
        assert_same_tree(t, o)
    else:
        np.testing.assert_allclose(got, want, rtol=0, atol=1e-3 * np.abs(want).max())


# ---------------------------------------------------------------- degenerate sizes
@pytest.mark.parametrize("n", [1, 2, 33])
@pytest.mark.parametrize("policy", ["level", "loss"])
def test_handful_of_rows(n, policy):
    kw = dict(KW, grow_policy=policy, max_depth=4 if policy == "level" else -1, max_leaf_cnt=-1 if policy == "level" else 8,
              min_child_hessian_sum=0.0)
    x, start, gp = synth(n, 5, 16, 3)
    tm = maker(kw, x, start, gp)
    t = tm.make(0)
    o, opos, _ = ob.make_tree(ob.make_params(**kw), x, gp, 0, start)
    assert_same_tree(t, o)
    assert np.array_equal(tm.get_positions(), opos)


def test_empty_shard_gives_a_single_leaf():
    """A worker without rows (possible with several ranks) must not crash: one node, no samples."""
    start = np.array([0, 4, 9], np.int32)
    tm = yb.GpuTreeMaker(yb.make_params(**KW), start, 0)
    tm.set_grads(np.zeros(0, np.float32))
    t = tm.make(0)
    assert t.n_nodes == 1 and t.left[0] == -1 and t.sample_cnt[0] == 0
    assert tm.get_positions().size == 0


def test_constant_features_and_extreme_gradient_magnitudes():
    """No feature can separate anything -> a single leaf; gradients of 1e-20 and 1e+20 -> the power-of-two
    scale adapts and the tree equals the oracle's."""
    n, F = 5000, 6
    x, start, gp = synth(n, F, 16, 5)
    xc = np.tile(x[:1], (n, 1))
    tm = maker(KW, xc, start, gp)
    t = tm.make(0)
    o, opos, _ = ob.make_tree(ob.make_params(**KW), xc, gp, 0, start)
    assert t.n_nodes == 1
    assert_same_tree(t, o)
    for scale in (1e-20, 1e20):
        g2 = (gp.astype(np.float64) * scale).astype(np.float32)
        kw = dict(KW, min_child_hessian_sum=0.0, l2=0.0)
        tm = maker(kw, x, start, g2)
        t = tm.make(0)
        o, opos, _ = ob.make_tree(ob.make_params(**kw), x, g2, 0, start)
        assert_same_tree(t, o)
        assert np.array_equal(tm.get_positions(), opos)


@pytest.mark.parametrize("policy", ["level", "loss"])
def test_consecutive_trees_with_changing_sampling_and_gradients(policy):
    """One handle, several make() calls as in boosting rounds: different feature subsets, row masks and
    gradients each time -- nothing may leak from one tree into the next (pool slots, planned partitions,
    expansions computed ahead, leaf ranges)."""
    kw = dict(KW, grow_policy=policy, max_depth=6 if policy == "level" else -1, max_leaf_cnt=-1 if policy == "level" else 40)
    n, F = 40000, 30
    x, start, gp = synth(n, F, 200, 41, skew=True)
    rng = np.random.default_rng(6)
    tm = maker(kw, x, start, gp)
    p = ob.make_params(**kw)
    for it in range(5):
        fea = None if it % 2 == 0 else np.sort(rng.permutation(F)[: 5 + 4 * it]).astype(np.int32)
        mask = None if it in (0, 3) else (rng.random(n) < 0.4 + 0.1 * it).astype(np.uint8)
        g = (gp.reshape(-1, 2) * np.float32(1.0 + 0.7 * it)).astype(np.float32)
        g[:, 0] += rng.normal(size=n).astype(np.float32) * np.float32(0.05 * it)
        g = g.reshape(-1)
        tm.set_grads(g)
        t = tm.make(0, fea_index=fea, row_mask=mask)
        o, opos, _ = ob.make_tree(p, x, g, 0, start, fea_index=fea, row_mask=mask)
        assert_same_tree(t, o)
        assert np.array_equal(tm.get_positions(), opos), it


def test_feature_sampling_that_leaves_whole_feature_groups_out():
    """100 features = 4 groups of 32; the sampled ones sit in groups 0 and 3 only."""
    n, F = 30000, 100
    x, start, gp = synth(n, F, 64, 43)
    fea = np.array([1, 5, 30, 96, 97, 99], np.int32)
    for kw in (dict(KW, max_depth=5), dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=20)):
        tm = maker(kw, x, start, gp)
        t = tm.make(0, fea_index=fea)
        o, opos, _ = ob.make_tree(ob.make_params(**kw), x, gp, 0, start, fea_index=fea)
        assert_same_tree(t, o)
        assert np.array_equal(tm.get_positions(), opos)
        got = tm.build_hist(0, fea_index=fea)
        want = ob.build_hist(x, gp, 1, 0, np.arange(n), fea, start)
        assert_hist_close(got, want, gp)
