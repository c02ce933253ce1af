"""Synthetic inputs of the BASELINE.json shapes, shared by bench.py and the full-size
parity tests.  Slots are generated directly (bin boundaries are an input of the path)."""
import numpy as np


def sigmoid_grads(score, y, out):
    """SigmoidFunction (J/loss/SigmoidFunction.java:63-80) with the reference's float casts."""
    p = (1.0 / (1.0 + np.exp(-score.astype(np.float64)))).astype(np.float32).astype(np.float64)
    out[0::2] = (p - y).astype(np.float32)
    out[1::2] = (p * (1.0 - p)).astype(np.float32)
    return out


def c2_higgs(n, F=28, nb=255, seed=1234, uniform=False):
    """configs[1]: half the features uniform over the bins, half skewed (geometric); uint8 slots [n,F] and
    labels that depend on a few features so that trees have real structure."""
    rng = np.random.default_rng(seed)
    x = np.empty((n, F), dtype=np.uint8)
    for f in range(F):
        if f % 2 == 0 or uniform:
            x[:, f] = rng.integers(0, nb, size=n, dtype=np.uint8)
        else:
            x[:, f] = np.minimum(rng.geometric(0.04, size=n) - 1, nb - 1).astype(np.uint8)
    z = np.zeros(n, dtype=np.float32)
    for f in range(0, min(F, 12)):
        z += (x[:, f].astype(np.float32) / nb - 0.5) * np.float32(((-1) ** f) * (1.0 + 0.3 * f))
    z += (x[:, 1] > 20) * np.float32(0.8) - (x[:, 3] < 3) * np.float32(0.5)
    p = 1.0 / (1.0 + np.exp(-z))
    y = (rng.random(n, dtype=np.float32) < p).astype(np.float32)
    return x, y


def c3_regression(n, F=128, nb=255, seed=42):
    """configs[2]: uniform slots, y = sum_j a_j * slot_j + N(0,1); returns slots uint8 [n,F], start, and the
    l2 gradient pairs at score 0 (g = -y, h = 1; J/loss/L2Function.java:41-49)."""
    rng = np.random.default_rng(seed)
    x = rng.integers(0, nb, size=(n, F), dtype=np.uint8)
    a = (rng.normal(size=F) * (rng.random(F) < 0.25)).astype(np.float32) / np.float32(nb)
    y = np.zeros(n, np.float32)
    for f in np.nonzero(a)[0]:
        y += x[:, f].astype(np.float32) * a[f]
    y += rng.normal(size=n).astype(np.float32)
    gp = np.empty(2 * n, np.float32)
    gp[0::2] = -y
    gp[1::2] = 1.0
    start = (np.arange(F + 1) * nb).astype(np.int32)
    return x, start, gp


def c4_sparse(n, F=10_000, density=0.01, seed=7):
    """configs[3]: k = F*density stored entries per row, one per stratum of F/k features (sorted, distinct),
    slots 1..b_f-1 with b_f in [2,255], default slot 0; binary labels from a few features.
    Returns start, row_ptr, fids, slots (int32), y."""
    rng = np.random.default_rng(seed)
    k = max(1, int(round(F * density)))
    width = F // k
    nbs = rng.integers(2, 256, size=F)
    fids = (np.arange(k, dtype=np.int32)[None, :] * width + rng.integers(0, width, size=(n, k), dtype=np.int32))
    slots = 1 + (rng.random((n, k), dtype=np.float32) * (nbs[fids] - 1)).astype(np.int32)
    w = (rng.normal(size=F) * (rng.random(F) < 0.05)).astype(np.float32)
    z = (w[fids] * (slots / nbs[fids])).sum(axis=1)
    y = (rng.random(n) < 1.0 / (1.0 + np.exp(-3 * z))).astype(np.float32)
    row_ptr = (np.arange(n + 1, dtype=np.int64) * k)
    start = np.concatenate([[0], np.cumsum(nbs)]).astype(np.int32)
    return start, row_ptr, fids.reshape(-1), slots.reshape(-1), y


def c5_multiclass(n, F=256, nb=63, K=10, seed=99):
    """configs[4]: uniform slots, K classes; returns slots uint8 [n,F], start, labels int [n] and the softmax
    gradient pairs at score 0 (p = 1/K: g = p - y, h = 2p(1-p); J/loss/SoftmaxFunction.java:108-113)."""
    rng = np.random.default_rng(seed)
    x = rng.integers(0, nb, size=(n, F), dtype=np.uint8)
    w = rng.normal(size=(8, K)).astype(np.float32)
    z = np.zeros((n, K), np.float32)
    for j in range(8):
        z += (x[:, j].astype(np.float32) / nb - 0.5)[:, None] * w[j][None, :] * 3
    z += rng.gumbel(size=(n, K)).astype(np.float32)
    label = z.argmax(axis=1)
    p = np.float32(1.0 / K)
    gp = np.empty((n, K, 2), np.float32)
    gp[:, :, 0] = p
    gp[np.arange(n), label, 0] = p - np.float32(1.0)
    gp[:, :, 1] = np.float32(2.0) * p * (np.float32(1.0) - p)
    start = (np.arange(F + 1) * nb).astype(np.int32)
    return x, start, label, gp.reshape(-1)
