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
