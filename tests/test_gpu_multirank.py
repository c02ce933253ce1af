"""N-GPU run of the real NCCL path (needs >= 2 GPUs; `gpurun --gpus 2`): rows sharded over the ranks,
histogram reduce-scatter by feature range + split allgather.  Integer accumulation makes the tree
independent of the number of ranks, so the 2-GPU tree must equal the 1-GPU tree bit for bit, and both
must match the oracle run with W = 2 emulated workers."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("gpu_required")]

KW = dict(max_depth=6, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0,
          min_split_loss=0.0, l1=0.0, l2=1.0, learning_rate=0.1)
CASES = {"level": KW, "loss": dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=24),
         "few_features": dict(KW, max_depth=4), "wide_features": dict(KW, max_depth=5),
         "sparse_rows": dict(KW, max_depth=5),
         "empty_rank": dict(KW, max_depth=3, min_child_hessian_sum=0.0),
         "uneven_gradients": dict(KW, max_depth=5, min_child_hessian_sum=0.01),
         "zero_rows_rank": dict(KW, max_depth=3, min_child_hessian_sum=0.0),
         "sampling_unowned": dict(KW, max_depth=5),
         "sorted_shards": dict(KW, max_depth=7), "sorted_shards_loss": dict(KW, grow_policy="loss", max_depth=-1, max_leaf_cnt=40),
         "refine_l1": dict(KW, max_depth=5, learning_rate=0.3)}


def _residuals(n):
    """(label - score, weight) of the l1 TreeRefiner case, the same on every rank"""
    rng = np.random.default_rng(77)
    return np.round(rng.normal(size=n) * 4).astype(np.float32) / 4, rng.integers(1, 4, size=n).astype(np.float32)


def _sampling(case, n, F):
    """(fea_index, row_mask) of the make() call: none, except for the case where the sampled features all
    belong to rank 0's range (rank 1 has no feature to take node totals from) and rows are sampled too."""
    if case != "sampling_unowned":
        return None, None
    mask = (np.random.default_rng(9).random(n) < 0.7).astype(np.uint8)
    return np.array([0, 2, 5], np.int32), mask


def _data(case):
    from gpu_common import synth
    if case == "few_features":                      # fewer features than some world sizes would own evenly
        return synth(20000, 3, [17, 255, 4], 5)
    if case == "wide_features":                     # features with more than 256 bins (uint16 slots)
        return synth(40000, 7, [300, 5, 1000, 256, 257, 4000, 40], 13, skew=True)
    if case == "uneven_gradients":                  # the ranks' largest |g|, |h| differ by orders of magnitude:
        x, start, gp = synth(30000, 12, 64, 37)     # the fixed-point scales must still be common to all ranks
        g2 = gp.reshape(-1, 2).copy()
        g2[15000:] *= np.float32(0.013)
        return x, start, g2.reshape(-1)
    if case == "zero_rows_rank":                    # one row in all: rank 0 holds nothing
        return synth(1, 4, 8, 33)
    if case == "empty_rank":                        # 3 rows over 2 ranks, and the first split leaves rank 0 one side only
        return synth(3, 4, 8, 31)
    if case == "sparse_rows":                       # mostly slot 0
        x, start, gp = synth(30000, 60, 40, 29)
        x = x * (np.random.default_rng(3).random(x.shape) < 0.05)
        return np.ascontiguousarray(x.astype(np.int32)), start, gp
    if case.startswith("sorted_shards"):            # rows sorted by a feature: whole nodes live on one GPU only
        x, start, gp = synth(40000, 12, 64, 23)
        order = np.lexsort((x[:, 3], x[:, 0]))
        return np.ascontiguousarray(x[order]), start, np.ascontiguousarray(gp.reshape(-1, 2)[order]).reshape(-1)
    return synth(60000, 28, 255, 91, skew=True)


def _rank(rank, world, nccl_id, case, ret):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    import _pkg
    yb = _pkg.load()
    x, start, gp = _data(case)
    n = x.shape[0]
    lo, hi = n * rank // world, n * (rank + 1) // world
    p = yb.make_params(**CASES[case])
    if case == "sparse_rows":                       # the shard as CSR: everything that is not slot 0 is stored
        xs = x[lo:hi]
        r, f = np.nonzero(xs)
        row_ptr = np.concatenate([[0], np.cumsum((xs != 0).sum(axis=1))]).astype(np.int64)
        tm = yb.GpuTreeMaker(p, start, hi - lo, device=rank, rank=rank, world=world, nccl_id=nccl_id,
                             sparse_default=np.zeros(x.shape[1], np.int32), nnz=len(r))
        tm.upload_csr(row_ptr, f.astype(np.int32), xs[r, f].astype(np.int32))
    else:
        tm = yb.GpuTreeMaker(p, start, hi - lo, device=rank, rank=rank, world=world, nccl_id=nccl_id)
        tm.upload_rows(x[lo:hi])
    tm.set_grads(gp.reshape(n, 2)[lo:hi].reshape(-1))
    fea, mask = _sampling(case, n, x.shape[1])
    tree = tm.make(0, fea_index=fea, row_mask=None if mask is None else mask[lo:hi])
    pos = tm.get_positions()
    if case == "refine_l1":                         # TreeRefiner: medians over the rows of ALL ranks
        res, w = _residuals(n)
        tm.refine_leaves(tree, res[lo:hi], w[lo:hi])
    hist = tm.build_hist(0)                         # reduced over ranks: the global root histogram
    ret[rank] = ({name: getattr(tree, name) for name, _ in yb.Tree.FIELDS}, pos, hist)
    tm.close()


@pytest.mark.parametrize("case", list(CASES))
def test_two_gpus_equal_one_gpu_and_oracle(case):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    import oracle_binding as ob
    from gpu_common import assert_same_tree, maker, yb
    world = 2
    x, start, gp = _data(case)
    n = x.shape[0]
    one = maker(CASES[case], x, start, gp)
    fea, mask = _sampling(case, n, x.shape[1])
    t1 = one.make(0, fea_index=fea, row_mask=mask)
    p1 = one.get_positions()
    if case == "refine_l1":
        res, w = _residuals(n)
        want = ob.refine_leaves(t1, p1, res, w, CASES[case]["learning_rate"])
        one.refine_leaves(t1, res, w)
        assert np.array_equal(t1.node_info, want.node_info)
    h1 = one.build_hist(0)
    ctx = mp.get_context("spawn")
    mgr = ctx.Manager()
    ret = mgr.dict()
    nid = yb.nccl_unique_id()
    procs = [ctx.Process(target=_rank, args=(r, world, nid, case, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
    for p in procs:
        if p.is_alive():
            p.kill()
    assert [p.exitcode for p in procs] == [0] * world
    pos = np.concatenate([ret[r][1] for r in range(world)])
    for r in range(world):
        for name, _ in yb.Tree.FIELDS:
            assert np.array_equal(ret[r][0][name], getattr(t1, name), equal_nan=True), (r, name)
        assert np.array_equal(ret[r][2], h1)
    assert np.array_equal(pos, p1)
    o, opos, _ = ob.make_tree(ob.make_params(**CASES[case]), x, gp, 0, start, W=world, fea_index=fea, row_mask=mask)
    if case == "refine_l1":
        res, w = _residuals(n)
        o = ob.refine_leaves(o, opos, res, w, CASES[case]["learning_rate"])
    assert_same_tree(t1, o)
    assert np.array_equal(p1, opos)
