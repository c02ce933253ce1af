"""Pins the CPU oracle to the only known answers the reference publishes for this path:
the loss lines of demo/gbdt/binary_classification/README.md:43-49 and the prediction
line of demo/gbdt/multiclass_classification/README.md:53."""
import numpy as np
import pytest

import kat
import oracle_binding as ob


def _oracle_maker(d, params, W=1):
    p = ob.make_params(**params)
    xb = d["xb"].astype(np.int32)

    def make_tree(gp, gid):
        t, _, _ = ob.make_tree(p, xb, gp, gid, d["start"], W=W)
        return t
    return make_tree


@pytest.mark.parametrize("W", [1, 4])
def test_kat1_agaricus_losses(W):
    d = kat.load_kat1()
    out, trees = kat.run_kat1(_oracle_maker(d, kat.KAT1_PARAMS, W), d)
    for it, (tr, te) in kat.KAT1.items():
        assert abs(out[it][0] - tr) < 1e-12, (it, out[it][0], tr)
        assert abs(out[it][1] - te) < 1e-12, (it, out[it][1], te)
    assert [t.n_nodes for t in trees] == [21, 23, 23]


def test_kat2_dermatology_prediction():
    d = kat.load_kat2()
    pred, trees = kat.run_kat2(_oracle_maker(d, kat.KAT2_PARAMS), d)
    want = np.array(kat.KAT2_PRED)
    err = np.abs(pred - want[None, :]).max(axis=1)
    hits = np.where(err < 1e-12)[0]
    assert list(hits) == kat.KAT2_ROWS, (list(hits), err.min())
    # every matching row carries the README's label (fourth class)
    assert all(int(d["yt"][r]) == 3 for r in hits)
    assert [[t.n_nodes for t in g] for g in trees] == kat.KAT2_TREE_SIZES


def test_value_walk_with_missing_values_default_child():
    """orc_score_value_bits (Tree.getLeafIndex/getNext, J/data/gbdt/Tree.java:142-168) against a direct
    numpy restatement, and against the plain value walk where nothing is missing."""
    import oracle_binding as ob
    rng = np.random.default_rng(1)

    class T:
        pass
    t = T()
    t.n_nodes = 7
    t.parent = np.array([-1, 0, 0, 1, 1, 2, 2], np.int32)
    t.left = np.array([1, 3, 5, -1, -1, -1, -1], np.int32)
    t.right = np.array([2, 4, 6, -1, -1, -1, -1], np.int32)
    t.fid = np.array([2, 0, 1, 0, 0, 0, 0], np.int32)
    t.lo = np.zeros(7, np.int32); t.hi = np.zeros(7, np.int32)
    t.node_info = np.array([0, 0, 0, 1.5, -2.25, 0.125, 7.0], np.float32)
    t.loss_chg = np.zeros(7, np.float32); t.hess_sum = np.zeros(7, np.float32); t.sample_cnt = np.zeros(7, np.int64)
    thr = np.array([0.1, -0.3, 0.7, 0, 0, 0, 0], np.float32)
    dc = np.array([2, 3, 6, 0, 0, 0, 0], np.int32)
    v = rng.normal(size=(400, 3)).astype(np.float32)
    bits = v.view(np.int32).copy()
    bits[rng.random(bits.shape) < 0.2] = 0x7FC00000
    vv = bits.view(np.float32)
    want = np.zeros(400, np.float32)
    for r in range(400):
        pid = 0
        while t.left[pid] != -1:
            b = bits[r, t.fid[pid]]
            if b == 0x7FC00000:
                pid = dc[pid]
            else:
                pid = t.left[pid] if vv[r, t.fid[pid]] <= thr[pid] else t.right[pid]
        want[r] = t.node_info[pid]
    got = np.zeros(400, np.float32)
    ob.score_value_bits(t, thr, dc, bits, got)
    assert np.array_equal(got, want)
    a = np.zeros(400, np.float32); b2 = np.zeros(400, np.float32)
    ob.score_value_bits(t, thr, None, v.view(np.int32), a)
    ob.score_values(t, thr, v, b2)
    assert np.array_equal(a, b2)


def test_oracle_weighted_median_of_tree_refiner():
    """orc_refine_leaves against a direct numpy restatement of PreciseQuantile.getWeightedQuantileLocal
    (J/utils/PreciseQuantile.java:156-208) and a hand-worked case."""
    import oracle_binding as ob

    class T:
        pass
    t = T()
    t.n_nodes = 3
    for name, dt in ob.Tree.FIELDS:
        setattr(t, name, np.zeros(3, dt))
    t.left[:] = [1, -1, -1]
    t.right[:] = [2, -1, -1]
    # leaf 1: values 1,2,3,4 with weights 1,1,1,1 -> total 4, target 2 -> first acc >= 2 is value 2
    # leaf 2: values 5,7 with weights 1,3 -> target 2 -> value 7; one unsampled row
    pos = np.array([1, 1, 1, 1, 2, 2, -1], np.int32)
    res = np.array([4, 1, 3, 2, 7, 5, 100], np.float32)
    w = np.array([1, 1, 1, 1, 3, 1, 9], np.float32)
    out = ob.refine_leaves(t, pos, res, w, 0.5)
    assert out.node_info[1] == np.float32(2 * 0.5) and out.node_info[2] == np.float32(7 * 0.5)
    rng = np.random.default_rng(0)
    n = 5000
    pos = rng.integers(1, 3, size=n).astype(np.int32)
    res = rng.normal(size=n).astype(np.float32)
    w = rng.random(n).astype(np.float32)
    out = ob.refine_leaves(t, pos, res, w, 1.0)
    for leaf in (1, 2):
        idx = np.where(pos == leaf)[0]
        order = idx[np.argsort(res[idx], kind="stable")]
        acc = np.cumsum(w[order].astype(np.float64))
        k = int(np.argmax(acc >= acc[-1] * 0.5))
        assert out.node_info[leaf] == res[order][k]
