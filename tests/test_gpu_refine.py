"""TreeRefiner for loss_function = l1 (J/optimizer/gbdt/TreeRefiner.java:209-255, exact
variant): leaf values = learning_rate x weighted median of (label - score) over the leaf's rows, bit for
bit against the oracle's restatement of PreciseQuantile (J/utils/PreciseQuantile.java:156-208)."""
import numpy as np
import pytest

import oracle_binding as ob
from gpu_common import assert_same_tree, maker, synth, yb

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("gpu_required")]

KW = dict(max_depth=6, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0,
          min_split_loss=0.0, l1=0.0, l2=1.0, learning_rate=0.3)


@pytest.mark.parametrize("kw", [KW, dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=40)], ids=["level", "loss"])
@pytest.mark.parametrize("weights", ["none", "random", "ties"])
def test_refine_equals_oracle(kw, weights):
    n, F = 40000, 12
    x, start, gp = synth(n, F, 64, 3)
    rng = np.random.default_rng(5)
    res = rng.normal(size=n).astype(np.float32)
    w = None
    if weights == "random":
        w = (rng.random(n) * 3).astype(np.float32)
        w[rng.random(n) < 0.1] = 0.0                      # rows of weight zero
    if weights == "ties":
        res = np.round(res * 2).astype(np.float32) / 2    # a handful of distinct values: long runs of equal keys
        w = rng.integers(1, 4, size=n).astype(np.float32)
    mask = (rng.random(n) < 0.8).astype(np.uint8)
    tm = maker(kw, x, start, gp)
    for row_mask in (None, mask):
        t = tm.make(0, row_mask=row_mask)
        pos = tm.get_positions()
        want = ob.refine_leaves(t, pos, res, w, kw["learning_rate"])
        before = t.node_info.copy()
        tm.refine_leaves(t, res, w)
        leaf = t.left == -1
        # (equal as floats: -0.0 and +0.0 compare equal in the reference's sort too, so which of the two a
        # median of zeros is depends on its unstable quicksort)
        assert np.array_equal(t.node_info[leaf], want.node_info[leaf])
        assert np.array_equal(t.node_info[~leaf], before[~leaf])
        assert (t.node_info[leaf] != before[leaf]).any()


def test_l1_boosting_rounds_on_the_device():
    """loss_function = l1 end to end on the device: L1Function gradients (J/loss/L1Function.java:31-54),
    make, TreeRefiner on label - score, score update with the refined leaves -- against the oracle's loop."""
    n, F = 30000, 10
    x, start, _ = synth(n, F, 32, 9)
    rng = np.random.default_rng(2)
    y = (x[:, 0] * 0.1 + rng.normal(size=n) * 2).astype(np.float32)
    w = (0.5 + rng.random(n)).astype(np.float32)
    init = np.zeros(n, np.float32)
    tm = yb.GpuTreeMaker(yb.make_params(**KW), start, n)
    tm.upload_rows(x)
    tm.set_targets("l1", y, w, init)
    score = np.zeros(n, np.float32)
    p = ob.make_params(**KW)
    want_loss, gp = ob.loss_grad_scalar("l1", score, init, y, w)
    got_loss = tm.compute_grads()
    assert abs(got_loss - want_loss) <= 1e-9 * abs(want_loss)
    assert np.array_equal(tm.get_grads(), gp)
    for _ in range(3):
        t = tm.make(0)
        o, opos, _ = ob.make_tree(p, x, gp, 0, start)
        assert_same_tree(t, o)
        pos = tm.get_positions()
        assert np.array_equal(pos, opos)
        want = ob.refine_leaves(t, pos, y - score, w, KW["learning_rate"])
        tm.refine_leaves(t)                               # residuals and weights from the device state
        leaf = t.left == -1
        assert np.array_equal(t.node_info[leaf], want.node_info[leaf])
        tm.update_scores(0)
        score = (score + t.node_info[pos]).astype(np.float32)
        assert np.array_equal(tm.get_scores(), score)
        want_loss, gp = ob.loss_grad_scalar("l1", score, init, y, w)
        got_loss = tm.compute_grads()
        assert abs(got_loss - want_loss) <= 1e-9 * abs(want_loss)
        assert np.array_equal(tm.get_grads(), gp)


def test_refine_checks_its_arguments():
    n = 2000
    x, start, gp = synth(n, 5, 16, 1)
    tm = maker(KW, x, start, gp)
    other = yb.Tree(3)
    other.n_nodes = 3
    with pytest.raises(yb.YtkError):
        tm.refine_leaves(other, np.zeros(n, np.float32))  # no tree yet
    t = tm.make(0)
    with pytest.raises(yb.YtkError):
        tm.refine_leaves(other, np.zeros(n, np.float32))  # not the last tree
    with pytest.raises(yb.YtkError):
        tm.refine_leaves(t)                               # no residuals and no device state
