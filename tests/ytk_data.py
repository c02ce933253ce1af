"""Test helpers restating the parts of the reference that stay in Java but produce the
hot path's inputs: the ytklearn text format, feature indexing, `no_sample` bin
boundaries and slot -> threshold conversion.  (J/ = src/main/java/com/fenbi/ytklearn/.)
"""
import os

import numpy as np

REF = "/root/reference"
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def parse_ytklearn(path):
    """`weight###label###name:val,name:val` (docs/data_format.md:1-40)."""
    w, y, feats = [], [], []
    with open(path) as fh:
        for line in fh:
            line = line.strip()
            if not line:
                continue
            cols = line.split("###")
            w.append(float(cols[0]))
            y.append(float(cols[1]))
            d = {}
            for kv in cols[2].split(","):
                k, v = kv.split(":")
                d[k] = np.float32(v)
            feats.append(d)
    return np.array(w, dtype=np.float32), np.array(y, dtype=np.float32), feats


def feature_index(train_feats, dict_names=None):
    """No dict: rank of the name in lexicographic String order over the train set
    (DataFlow.reduceFeature, J/dataflow/DataFlow.java:299-329).  With a dict: file order."""
    if dict_names is not None:
        return {n: i for i, n in enumerate(dict_names)}
    names = sorted({k for d in train_feats for k in d})
    return {n: i for i, n in enumerate(names)}


def dense(feats, fidx, fill=0.0):
    """Dense float matrix; absent -> fill (FillMissingValue, `missing_value: "value"` -> 0)."""
    x = np.full((len(feats), len(fidx)), fill, dtype=np.float32)
    for r, d in enumerate(feats):
        for k, v in d.items():
            j = fidx.get(k)
            if j is not None:
                x[r, j] = v
    return x


def no_sample_cuts(xtrain):
    """`type: no_sample`: every distinct train value is a bin centre
    (J/feature/gbdt/approximate/sampler/NoSample.java:43-50), sorted
    (FeatureApprData.genApprFeaSorted, J/data/gbdt/FeatureApprData.java:114-144)."""
    cuts, start = [], [0]
    for f in range(xtrain.shape[1]):
        c = np.unique(xtrain[:, f])
        cuts.append(c.astype(np.float32))
        start.append(start[-1] + len(c))
    return np.concatenate(cuts), np.array(start, dtype=np.int32)


def real_thresholds(tree, cuts, start):
    """FeatureSplitType.MEAN (J/feature/gbdt/FeatureSplitType.java:34-38) applied by
    Tree.convertFeatureSplitValueInModel (J/data/gbdt/Tree.java:293-309)."""
    thr = np.zeros(tree.n_nodes, dtype=np.float32)
    for nid in range(tree.n_nodes):
        if tree.left[nid] == -1:
            continue
        c = cuts[start[tree.fid[nid]]: start[tree.fid[nid] + 1]]
        thr[nid] = np.float32(0.5) * np.float32(c[tree.lo[nid]] + c[tree.hi[nid]])
    return thr
