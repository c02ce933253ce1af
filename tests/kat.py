"""Known-answer drivers: the boosting loop around the tree maker, restated from
GBDTOptimizer.train / predictAndCalcLossGrad (J/optimizer/GBDTOptimizer.java:406-480,
513-609) so that the loss values / predictions the reference prints in its demo
READMEs can be recomputed with any `make_tree(gp, gid) -> Tree` (the CPU oracle, or
the CUDA tree maker).  Gradients, scoring and loss come from the oracle in both cases;
only the tree maker is swapped.
"""
import os

import numpy as np

import oracle_binding as ob
import ytk_data as yd

# demo/gbdt/binary_classification/README.md:43-49 (reference's own printed output)
KAT1 = {
    2: (0.17249267375913763, 0.17219528385526198),
    3: (0.09960066518065772, 0.09943817574378232),
}
# demo/gbdt/multiclass_classification/README.md:53
KAT2_LABEL = [0, 0, 0, 1, 0, 0]
KAT2_PRED = [0.12211822085981826, 0.12527414781210444, 0.12204564173421276,
             0.38662393640671316, 0.12199698511704811, 0.12194106807010317]

KAT1_PARAMS = dict(max_depth=5, max_leaf_cnt=16, min_split_samples=-1, grow_policy="loss",
                   min_child_hessian_sum=1.0, min_split_loss=0.0, max_abs_leaf_val=-1.0,
                   l1=0.0, l2=1.0, histogram_pool_capacity=-1.0, learning_rate=0.5, class_num=1)
# bundled conf says min_child_hessian_sum 1E-8; the README line was produced with 1
KAT2_PARAMS = dict(max_depth=6, max_leaf_cnt=128, min_split_samples=-1, grow_policy="level",
                   min_child_hessian_sum=1.0, min_split_loss=0.0, max_abs_leaf_val=-1.0,
                   l1=0.0, l2=1.0, histogram_pool_capacity=-1.0, learning_rate=0.1, class_num=6)


KAT2_TREE_SIZES = [[7, 17, 5, 21, 3, 3], [5, 17, 5, 15, 3, 3], [9, 19, 5, 15, 5, 3], [7, 17, 5, 17, 5, 3], [5, 21, 5, 17, 5, 3]]
KAT2_ROWS = [0, 14, 23, 50, 52, 69]      # 0-based rows of dermatology.test.ytklearn that yield KAT2_PRED


def load_kat1():
    """Agaricus demo inputs: from the committed fixture (made by tests/golden/make_golden.py),
    or straight from the reference's demo data when it is mounted."""
    npz = os.path.join(yd.GOLDEN, "agaricus_kat.npz")
    if os.path.exists(npz):
        d = np.load(npz)
        return {k: d[k] for k in d.files}
    return make_kat1_inputs()


def make_kat1_inputs():
    w, y, feats = yd.parse_ytklearn(os.path.join(yd.REF, "demo/data/ytklearn/agaricus.train.ytklearn"))
    wt, yt, featst = yd.parse_ytklearn(os.path.join(yd.REF, "demo/data/ytklearn/agaricus.test.ytklearn"))
    fidx = yd.feature_index(feats)
    xv = yd.dense(feats, fidx)
    xvt = yd.dense(featst, fidx)
    cuts, start = yd.no_sample_cuts(xv)
    xb = ob.find_bins(xv, cuts, start)
    return dict(w=w, y=y, xb=xb.astype(np.uint8), wt=wt, yt=yt, xvt=xvt.astype(np.uint8),
                cuts=cuts, start=start)


def load_kat2():
    npz = os.path.join(yd.GOLDEN, "dermatology_kat.npz")
    if os.path.exists(npz):
        d = np.load(npz)
        return {k: d[k] for k in d.files}
    return make_kat2_inputs()


def make_kat2_inputs():
    w, y, feats = yd.parse_ytklearn(os.path.join(yd.REF, "demo/data/ytklearn/dermatology.train.ytklearn"))
    wt, yt, featst = yd.parse_ytklearn(os.path.join(yd.REF, "demo/data/ytklearn/dermatology.test.ytklearn"))
    names = [str(i) for i in range(33)]          # demo/gbdt/multiclass_classification/feat_dict
    fidx = yd.feature_index(feats, names)
    xv = yd.dense(feats, fidx)
    xvt = yd.dense(featst, fidx)
    cuts, start = yd.no_sample_cuts(xv)
    xb = ob.find_bins(xv, cuts, start)
    return dict(w=w, y=y, xb=xb.astype(np.uint8), wt=wt, yt=yt, xvt=xvt, cuts=cuts, start=start)


def run_kat1(make_tree, d, rounds=3):
    """Returns {iter: (train_loss, test_loss)} and the trees. make_tree(gp, gid) -> Tree."""
    xb = d["xb"].astype(np.int32)
    xvt = d["xvt"].astype(np.float32)
    y, w, yt, wt = (d[k].astype(np.float32) for k in ("y", "w", "yt", "wt"))
    n, nt = len(y), len(yt)
    base = np.float32(-np.log(1.0 / 0.5 - 1.0))              # (float) pred2Score(0.5)
    init, initt = np.full(n, base, np.float32), np.full(nt, base, np.float32)
    score, scoret = np.zeros(n, np.float32), np.zeros(nt, np.float32)
    wsum, wsumt = float(np.sum(w.astype(np.float64))), float(np.sum(wt.astype(np.float64)))
    _, gp = ob.loss_grad_scalar("sigmoid", score, init, y, w)    # initPred (:329-334)
    out, trees = {}, []
    for it in range(1, rounds + 1):
        tree = make_tree(gp, 0)
        trees.append(tree)
        ob.score_slots(tree, xb, score)                       # train: slot data, `<=`
        loss, gp = ob.loss_grad_scalar("sigmoid", score, init, y, w)
        thr = yd.real_thresholds(tree, d["cuts"], d["start"])  # convertModel (:663-677)
        ob.score_values(tree, thr, xvt, scoret)               # test: real values
        losst, _ = ob.loss_grad_scalar("sigmoid", scoret, initt, yt, wt, want_grad=False)
        out[it] = (loss / wsum, losst / wsumt)
    return out, trees


def run_kat2(make_tree, d, rounds=5, K=6):
    """Returns softmax predictions [nt, K] of the online predictor and the trees."""
    xb = d["xb"].astype(np.int32)
    xvt = d["xvt"].astype(np.float32)
    w = d["w"].astype(np.float32)
    n, nt = len(w), len(d["yt"])
    yoh = np.zeros((n, K), np.float32)
    yoh[np.arange(n), d["y"].astype(np.int64)] = 1.0
    yoh = yoh.reshape(-1)
    init = np.zeros(n * K, np.float32)                        # pred2Score(0.0) = 0
    score = np.zeros(n * K, np.float32)
    _, gp, _ = ob.loss_grad_softmax(score, init, yoh, w, K)
    sums = np.zeros((nt, K), np.float64)
    trees = []
    for _ in range(rounds):
        group = []
        for gid in range(K):                                  # doBoost (:482-488)
            tree = make_tree(gp, gid)
            group.append(tree)
        for gid, tree in enumerate(group):
            ob.score_slots(tree, xb, score, K, gid)
        _, gp, _ = ob.loss_grad_softmax(score, init, yoh, w, K)
        for gid, tree in enumerate(group):
            # GBDTOnlinePredictor.scores (J/predictor/GBDTOnlinePredictor.java:186-218):
            # leaf values summed in double, `value <= threshold` goes left
            thr = yd.real_thresholds(tree, d["cuts"], d["start"])
            one = np.zeros(nt, np.float32)
            for r in range(nt):
                pid = 0
                while tree.left[pid] != -1:
                    pid = tree.left[pid] if xvt[r, tree.fid[pid]] <= thr[pid] else tree.right[pid]
                one[r] = tree.node_info[pid]
            sums[:, gid] += one.astype(np.float64)
        trees.append(group)
    mx = sums.max(axis=1, keepdims=True)
    e = np.exp(sums - mx)
    return e * (1.0 / e.sum(axis=1, keepdims=True)), trees
