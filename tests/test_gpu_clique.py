"""The multi-rank path on ONE GPU (SURVEY.md §4 iv): P row shards as a clique of handles on one device
(ytk_gbdt_create_clique), run through the same kernels, the same [owner][node][segment] staging layout,
feat_owner / feat_inner tables, per-rank first owned feature, child-count exchange and record merge as one
process per GPU -- only the flags between GPUs are replaced by stream order.  Integer accumulation makes
the tree independent of the number of ranks: every rank's tree must equal the one-GPU tree bit for bit,
and both must match the oracle run with W = P emulated workers
(J/data/gbdt/HistogramBuilder.java:95, J/optimizer/gbdt/DataParallelTreeMaker.java:543-573,640-653)."""
import numpy as np
import pytest

import oracle_binding as ob
from gpu_common import assert_same_tree, maker, synth, yb

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("gpu_required")]

KW = dict(max_depth=6, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0,
          min_split_loss=0.0, l1=0.0, l2=1.0, learning_rate=0.1)


def _data(case):
    if case == "few_features":                      # fewer features than ranks: some ranks own none
        return synth(20000, 3, [17, 255, 4], 5), dict(KW, max_depth=4)
    if case == "uneven_gradients":                  # the ranks' largest |g|, |h| differ by orders of magnitude
        x, start, gp = synth(30000, 12, 64, 37)
        g2 = gp.reshape(-1, 2).copy()
        g2[15000:] *= np.float32(0.013)
        return (x, start, g2.reshape(-1)), dict(KW, max_depth=5, min_child_hessian_sum=0.01)
    if case == "tiny":                              # 5 rows: most ranks hold nothing
        return synth(5, 4, 8, 31), dict(KW, max_depth=3, min_child_hessian_sum=0.0)
    if case == "sorted_shards":                     # rows sorted by a feature: whole nodes live on one rank only
        x, start, gp = synth(40000, 12, 64, 23)
        order = np.lexsort((x[:, 3], x[:, 0]))
        return (np.ascontiguousarray(x[order]), start, np.ascontiguousarray(gp.reshape(-1, 2)[order]).reshape(-1)), dict(KW, max_depth=7)
    if case == "multiclass":
        return synth(30000, 20, 40, 7, K=3), dict(KW, max_depth=5)
    if case == "depth1":
        return synth(20000, 10, 32, 3), dict(KW, max_depth=1)
    return synth(60000, 28, 255, 91, skew=True), KW


def _clique_trees(kw, x, start, gp, P, K=1, gid=0, fea=None, mask=None):
    n = x.shape[0]
    edges = [n * r // P for r in range(P + 1)]
    members = yb.GpuTreeMaker.clique(yb.make_params(class_num=K, **kw), start, [edges[r + 1] - edges[r] for r in range(P)])
    gp2 = gp.reshape(n, -1)
    for r, m in enumerate(members):
        m.upload_rows(x[edges[r]:edges[r + 1]])
        m.set_grads(gp2[edges[r]:edges[r + 1]].reshape(-1))
    masks = None if mask is None else [mask[edges[r]:edges[r + 1]] for r in range(P)]
    trees = yb.GpuTreeMaker.make_clique(members, gid, fea_index=fea, row_masks=masks)
    pos = np.concatenate([m.get_positions() for m in members])
    stats = [m.last_stats for m in members]
    for m in members:
        m.close()
    return trees, pos, stats


@pytest.mark.parametrize("P", [2, 3, 8])
@pytest.mark.parametrize("case", ["level", "few_features", "uneven_gradients", "tiny", "sorted_shards", "depth1"])
def test_clique_of_p_ranks_equals_one_gpu_and_oracle(case, P):
    (x, start, gp), kw = _data(case)
    one = maker(kw, x, start, gp)
    t1 = one.make(0)
    p1 = one.get_positions()
    trees, pos, _ = _clique_trees(kw, x, start, gp, P)
    for r, t in enumerate(trees):
        assert t.n_nodes == t1.n_nodes, (r, t.n_nodes, t1.n_nodes)
        for name, _ in yb.Tree.FIELDS:
            assert np.array_equal(getattr(t, name), getattr(t1, name), equal_nan=True), (r, name)
    assert np.array_equal(pos, p1)
    o, opos, _ = ob.make_tree(ob.make_params(**kw), x, gp, 0, start, W=P)
    assert_same_tree(t1, o)
    assert np.array_equal(p1, opos)


@pytest.mark.parametrize("P", [2, 5])
def test_clique_with_feature_and_row_sampling(P):
    """The sampled features all belong to rank 0's range (the other ranks have no feature to take node totals
    from, J/optimizer/gbdt/DataParallelTreeMaker.java:451-467) and rows are sampled too."""
    (x, start, gp), kw = _data("level")
    n = x.shape[0]
    mask = (np.random.default_rng(9).random(n) < 0.7).astype(np.uint8)
    fea = np.array([0, 2, 5], np.int32)
    one = maker(kw, x, start, gp)
    t1 = one.make(0, fea_index=fea, row_mask=mask)
    p1 = one.get_positions()
    trees, pos, _ = _clique_trees(kw, x, start, gp, P, fea=fea, mask=mask)
    for t in trees:
        for name, _ in yb.Tree.FIELDS:
            assert np.array_equal(getattr(t, name), getattr(t1, name), equal_nan=True), name
    assert np.array_equal(pos, p1)
    o, opos, _ = ob.make_tree(ob.make_params(**kw), x, gp, 0, start, W=P, fea_index=fea, row_mask=mask)
    assert_same_tree(t1, o)
    assert np.array_equal(p1, opos)


def test_clique_multiclass_consecutive_trees():
    (x, start, gp), kw = _data("multiclass")
    K, P = 3, 4
    n = x.shape[0]
    edges = [n * r // P for r in range(P + 1)]
    members = yb.GpuTreeMaker.clique(yb.make_params(class_num=K, **kw), start, [edges[r + 1] - edges[r] for r in range(P)])
    gp2 = gp.reshape(n, -1)
    for r, m in enumerate(members):
        m.upload_rows(x[edges[r]:edges[r + 1]])
        m.set_grads(gp2[edges[r]:edges[r + 1]].reshape(-1))
    p = ob.make_params(class_num=K, **kw)
    for gid in (0, 2, 1, 2):
        trees = yb.GpuTreeMaker.make_clique(members, gid)
        pos = np.concatenate([m.get_positions() for m in members])
        o, opos, _ = ob.make_tree(p, x, gp, gid, start, W=P)
        for t in trees:
            assert_same_tree(t, o)
        assert np.array_equal(pos, opos)
    for m in members:
        m.close()


def test_clique_refuses_what_the_level_loop_does_not_take():
    (x, start, gp), kw = _data("depth1")
    members = yb.GpuTreeMaker.clique(yb.make_params(**dict(kw, grow_policy="loss", max_depth=-1, max_leaf_cnt=8)), start, [10000, 10000])
    for r, m in enumerate(members):
        m.upload_rows(x[r * 10000:(r + 1) * 10000])
        m.set_grads(gp.reshape(-1, 2)[r * 10000:(r + 1) * 10000].reshape(-1))
    with pytest.raises(yb.YtkError):
        yb.GpuTreeMaker.make_clique(members, 0)
    with pytest.raises(yb.YtkError):
        members[0].make(0)                          # a member cannot build a tree on its own
    for m in members:
        m.close()
