"""World-size-2 test of the N > 1 host logic on CPU (gloo): feature-range ownership
(ytk_gbdt_avg_assign), the reduce-scatter layout, and the split-record merge
(ytk_gbdt_merge_splits) reproduce the single-process result.  The per-rank compute is the CPU oracle
(as the checker -- there is no GPU here); the collectives are torch.distributed/gloo standing in for
NCCL; the ownership and merge code is the library's own."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    import torch
    import torch.distributed as dist
    import _pkg
    import oracle_binding as ob
    from gpu_common import synth
    yb = _pkg.load()
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        n, F = 6000, 13
        x, start, gp = synth(n, F, [7, 31, 255, 2, 100, 64, 9, 200, 33, 5, 128, 17, 60], 77, skew=True)
        kw = dict(max_depth=4, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0,
                  min_split_loss=0.0, l1=0.0, l2=1.0, learning_rate=0.1)
        p = ob.make_params(**kw)
        lo_r, hi_r = n * rank // world, n * (rank + 1) // world
        rows = np.arange(lo_r, hi_r)
        fea = np.arange(F, dtype=np.int32)
        # local histogram of this rank's shard, then the feature-range reduce-scatter
        local = ob.build_hist(x, gp, 1, 0, rows, fea, start)
        assign = yb.avg_assign(F, world)                       # library code under test
        t = torch.from_numpy(local.copy())
        dist.all_reduce(t, op=dist.ReduceOp.SUM)               # gloo has no reduce_scatter: reduce, keep own slice
        f0, f1 = assign[rank]
        owned = np.zeros_like(local)
        owned[2 * start[f0]: 2 * start[f1]] = t.numpy()[2 * start[f0]: 2 * start[f1]]
        # node totals from the first owned feature, MAX over ranks (DataParallelTreeMaker.java:549-561)
        if f0 < f1:
            seg = owned[2 * start[f0]: 2 * start[f0 + 1]].reshape(-1, 2)
            G = H = 0.0
            for g, h in seg:
                G += g
                H += h
        else:
            G = H = -np.inf
        tot = torch.tensor([G, H], dtype=torch.float64)
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        G, H = tot.tolist()
        rg = np.float32(ob.calc_gain(p, G, H))
        best = (-np.inf, -1, -1, -1)
        for f in range(f0, f1):
            best = ob.enumerate_split(p, owned, start, f, G, H, float(rg), best)
        rec = torch.tensor([best[0], best[1], best[2], best[3], G if f0 < f1 else -np.inf,
                            H if f0 < f1 else -np.inf], dtype=torch.float64)
        allrec = [torch.zeros(6, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(allrec, rec)                            # the "split allgather"
        recs = [yb.Split(float(r[0]), int(r[1]), int(r[2]), int(r[3]), float(r[4]), float(r[5]), 0.0, 0.0)
                for r in allrec]
        merged = yb.merge_splits(recs)                          # library code under test
        # single-process answer
        tree, _, _ = ob.make_tree(ob.make_params(**dict(kw, max_depth=1)), x, gp, 0, start)
        assert merged.fid == tree.fid[0] and merged.lo == tree.lo[0] and merged.hi == tree.hi[0], \
            (merged.as_tuple(), tree.fid[0], tree.lo[0], tree.hi[0])
        assert abs(merged.loss_chg - tree.loss_chg[0]) <= 1e-5 * abs(tree.loss_chg[0])
        assert abs(merged.sum_hess - tree.hess_sum[0]) <= 1e-5 * abs(tree.hess_sum[0])
        # global child counts = allreduce(SUM) of the local partition counts
        cond = np.float32(0.5) * np.float32(merged.lo + merged.hi)
        left = int((x[rows, merged.fid].astype(np.float32) < cond).sum())
        c = torch.tensor([left, len(rows) - left], dtype=torch.int64)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        assert c.tolist() == [int(tree.sample_cnt[1]), int(tree.sample_cnt[2])]
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_two_rank_ownership_and_split_merge_match_single_process():
    import torch.multiprocessing as mp
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    mgr = ctx.Manager()
    ret = mgr.dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(150)
    for p in procs:
        if p.is_alive():
            p.kill()
    assert [p.exitcode for p in procs] == [0] * world
    assert dict(ret) == {0: "ok", 1: "ok"}
