"""score update, loss and gradient pairs on the device
(GBDTOptimizer.predictAndCalcLossGrad) against the oracle's restatement, and the reference's demo
known answers recomputed with gradients that never leave the GPU."""
import numpy as np
import pytest

import kat
import oracle_binding as ob
import ytk_data as yd
from gpu_common import maker, synth, yb

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("gpu_required")]

KW = dict(max_depth=5, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0,
          min_split_loss=0.0, l1=0.0, l2=1.0, learning_rate=0.3)


def _close(a, b, what):
    # exp/log on the device and in libm may differ in the last bit of a double; after the reference's
    # float casts that shows up as one float ulp on a handful of rows at most
    np.testing.assert_allclose(a, b, rtol=3e-7, atol=1e-12, err_msg=what)


@pytest.mark.parametrize("loss", ["sigmoid", "l2"])
def test_scalar_losses_three_rounds(loss):
    n, F = 30000, 20
    x, start, _ = synth(n, F, 64, 13)
    rng = np.random.default_rng(1)
    y = (rng.random(n) < 0.4).astype(np.float32) if loss == "sigmoid" else rng.normal(size=n).astype(np.float32)
    w = (0.5 + rng.random(n)).astype(np.float32)
    init = np.full(n, 0.25, np.float32)
    tm = yb.GpuTreeMaker(yb.make_params(**KW), start, n)
    tm.upload_rows(x)
    tm.set_targets(loss, y, w, init)
    score = np.zeros(n, np.float32)
    want_loss, gp = ob.loss_grad_scalar(loss, score, init, y, w)
    got_loss = tm.compute_grads()
    assert abs(got_loss - want_loss) <= 1e-9 * abs(want_loss)
    _close(tm.get_grads(), gp, "initial gradients")
    p = ob.make_params(**KW)
    for _ in range(3):
        t = tm.make(0)
        tm.update_scores(0)
        o, opos, _ = ob.make_tree(p, x, gp, 0, start)      # same gradients -> same tree
        assert t.n_nodes == o.n_nodes and np.array_equal(t.fid, o.fid)
        ob.score_slots(o, x, score)
        _close(tm.get_scores(), score, "scores")
        want_loss, gp = ob.loss_grad_scalar(loss, score, init, y, w)
        got_loss = tm.compute_grads()
        assert abs(got_loss - want_loss) <= 1e-9 * abs(want_loss)
        got = tm.get_grads()
        _close(got, gp, "gradients")
        gp = got            # keep both sides on identical inputs for the next tree


def test_softmax_round():
    n, F, K = 8000, 12, 5
    x, start, _ = synth(n, F, 32, 17)
    rng = np.random.default_rng(2)
    cls = rng.integers(0, K, size=n)
    yoh = np.zeros((n, K), np.float32)
    yoh[np.arange(n), cls] = 1
    w = np.ones(n, np.float32)
    tm = yb.GpuTreeMaker(yb.make_params(class_num=K, **KW), start, n)
    tm.upload_rows(x)
    tm.set_targets("softmax", yoh, None, None)
    score = np.zeros(n * K, np.float32)
    init = np.zeros(n * K, np.float32)
    want_loss, gp, _ = ob.loss_grad_softmax(score, init, yoh.reshape(-1), w, K)
    assert abs(tm.compute_grads() - want_loss) <= 1e-9 * abs(want_loss)
    _close(tm.get_grads(), gp, "initial gradients")
    p = ob.make_params(class_num=K, **KW)
    for gid in range(K):
        tm.make(gid)
        tm.update_scores(gid)
        o, _, _ = ob.make_tree(p, x, gp, gid, start)
        ob.score_slots(o, x, score, K, gid)
    _close(tm.get_scores(), score, "scores")
    want_loss, gp2, _ = ob.loss_grad_softmax(score, init, yoh.reshape(-1), w, K)
    assert abs(tm.compute_grads() - want_loss) <= 1e-9 * abs(want_loss)
    _close(tm.get_grads(), gp2, "gradients")


def test_rows_outside_the_sample_still_get_scored():
    n, F = 20000, 10
    x, start, gp = synth(n, F, 50, 19)
    mask = (np.random.default_rng(3).random(n) < 0.5).astype(np.uint8)
    tm = maker(KW, x, start, gp)
    tm.set_targets("l2", np.zeros(n, np.float32))
    t = tm.make(0, row_mask=mask)
    tm.update_scores(0)
    score = np.zeros(n, np.float32)
    ob.score_slots(t, x, score)
    assert np.array_equal(tm.get_scores(), score)


def test_kat1_agaricus_with_device_gradients():
    """demo/gbdt/binary_classification/README.md:43-49, train losses, gradients never leaving the GPU."""
    d = kat.load_kat1()
    xb = d["xb"].astype(np.int32)
    n = xb.shape[0]
    tm = yb.GpuTreeMaker(yb.make_params(**kat.KAT1_PARAMS), d["start"], n)
    tm.upload_rows(xb)
    base = np.float32(-np.log(1.0 / 0.5 - 1.0))
    tm.set_targets("sigmoid", d["y"], d["w"], np.full(n, base, np.float32))
    wsum = float(np.sum(d["w"].astype(np.float64)))
    tm.compute_grads()
    losses = {}
    for it in range(1, 4):
        tm.make(0)
        tm.update_scores(0)
        losses[it] = tm.compute_grads() / wsum
    for it, (tr, _) in kat.KAT1.items():
        assert abs(losses[it] - tr) < 1e-9, (it, losses[it], tr)


# ---------------------------------------------------------------- §8(f) row 3: test-set scoring
def test_score_values_matches_oracle_with_missing_values():
    """GBDTOptimizer.score on original values (isOriginTree=true): value <= threshold goes left, the
    canonical NaN bit pattern (Constants.INT_MISSING_VALUE) takes the default child."""
    rng = np.random.default_rng(3)
    n, F, nb, K = 30000, 12, 64, 3
    x, start, gp = synth(n, F, nb, 17, K=K)
    tm = maker(dict(KW, max_depth=6), x, start, gp, K=K)
    cuts = np.sort(rng.normal(size=(F, nb)).astype(np.float32), axis=1)
    nt, S = 70001, F + 3                                     # not the training shard; padded rows
    vals = rng.normal(size=(nt, S)).astype(np.float32)
    bits = vals.view(np.int32)
    bits[rng.random((nt, S)) < 0.08] = 0x7FC00000            # missing
    got = rng.normal(size=nt * K).astype(np.float32)
    want = got.copy()
    for gid in range(K):
        t = tm.make(gid)
        inner = t.left != -1
        thr = np.zeros(t.n_nodes, np.float32)                # FeatureSplitType MEAN: 0.5f * (c[lo] + c[hi])
        thr[inner] = np.float32(0.5) * (cuts[t.fid[inner], t.lo[inner]] + cuts[t.fid[inner], t.hi[inner]])
        dc = np.where(rng.random(t.n_nodes) < 0.5, t.left, t.right).astype(np.int32)
        tm.score_values(t, thr, vals, got, default_child=dc, gid=gid)
        ob.score_value_bits(t, thr, dc, bits, want, K=K, gid=gid)
    assert np.array_equal(got, want)
    # without a default direction missing values go left; and the plain-value oracle agrees where nothing is missing
    full = rng.normal(size=(5000, S)).astype(np.float32)
    g2 = np.zeros(5000 * K, np.float32); w2 = np.zeros(5000 * K, np.float32)
    tm.score_values(t, thr, full, g2, gid=K - 1)
    ob.score_values(t, thr, full, w2, K=K, gid=K - 1)
    assert np.array_equal(g2, w2)
    with pytest.raises(yb.YtkError):                          # a default child that is not a child
        tm.score_values(t, thr, full, g2, default_child=np.full(t.n_nodes, t.n_nodes + 5, np.int32))


def test_loss_wise_rounds_score_update_with_splits_computed_ahead():
    """Loss-wise trees: leaves whose split was computed ahead but never applied keep their rows in several
    ranges; the score update (and rows left out by the row mask, which walk the tree) must still be exact."""
    kw = dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=25)
    n, F = 40000, 16
    x, start, _ = synth(n, F, 64, 19, skew=True)
    rng = np.random.default_rng(4)
    y = (rng.random(n) < 0.45).astype(np.float32)
    w = np.ones(n, np.float32)
    init = np.zeros(n, np.float32)
    tm = yb.GpuTreeMaker(yb.make_params(**kw), start, n)
    tm.upload_rows(x)
    tm.set_targets("sigmoid", y, w, init)
    score = np.zeros(n, np.float32)
    _, gp = ob.loss_grad_scalar("sigmoid", score, init, y, w)
    tm.compute_grads()
    p = ob.make_params(**kw)
    for it in range(3):
        mask = None if it == 0 else (rng.random(n) < 0.6).astype(np.uint8)
        t = tm.make(0, row_mask=mask)
        tm.update_scores(0)
        o, opos, _ = ob.make_tree(p, x, gp, 0, start, row_mask=mask)
        assert t.n_nodes == o.n_nodes and np.array_equal(t.fid, o.fid)
        ob.score_slots(o, x, score)
        _close(tm.get_scores(), score, "scores")
        tm.compute_grads()
        gp = tm.get_grads()
