"""Shared helpers of the GPU parity tests: synthetic inputs and tree comparison."""
import numpy as np

import _pkg
import oracle_binding as ob

yb = _pkg.load()

REL = 1e-5   # north_star: histogram sums and split gains within 1e-5 relative


def synth(n, F, nb, seed, K=1, skew=False, S=None):
    """Random slot matrix (int32 [n,S]), bin layout and logistic-like gradients."""
    rng = np.random.default_rng(seed)
    S = F if S is None else S
    nbs = np.full(F, nb) if np.isscalar(nb) else np.asarray(nb)
    start = np.concatenate([[0], np.cumsum(nbs)]).astype(np.int32)
    x = np.zeros((n, S), dtype=np.int32)
    for f in range(F):
        if skew and f % 2:
            v = np.minimum((rng.pareto(1.2, size=n)).astype(np.int64), nbs[f] - 1)
        else:
            v = rng.integers(0, nbs[f], size=n)
        x[:, f] = v
    if S > F:
        x[:, F:] = -7  # padding columns the path must never read
    w = (x[:, :F] @ rng.normal(size=F)) / (F * max(1, nbs.max()) ** 0.5)
    gp = np.zeros((n, K, 2), dtype=np.float32)
    for k in range(K):
        pr = 1.0 / (1.0 + np.exp(-(w * (k + 1) + rng.normal(size=n))))
        y = (rng.random(n) < pr).astype(np.float64)
        p0 = 1.0 / (1.0 + np.exp(-rng.normal(size=n) * 0.5))
        gp[:, k, 0] = (p0 - y).astype(np.float32)
        gp[:, k, 1] = (p0 * (1 - p0)).astype(np.float32)
    return x, start, gp.reshape(-1)


def assert_hist_close(got, want, gp, msg=""):
    """|got-want| <= REL * max(|want|, tiny share of the gradient mass): fp64-accumulated sums."""
    g = np.abs(gp.reshape(-1, 2)).astype(np.float64)
    floor = np.array([g[:, 0].max(), g[:, 1].max()]) if len(g) else np.zeros(2)
    tol = REL * np.maximum(np.abs(want).reshape(-1, 2), floor[None, :])
    err = np.abs(got - want).reshape(-1, 2)
    assert (err <= tol).all(), "%s max err %g" % (msg, (err / np.maximum(tol, 1e-300)).max())


def assert_same_tree(t, o, check_stats=True):
    assert t.n_nodes == o.n_nodes, (t.n_nodes, o.n_nodes)
    for name in ("parent", "left", "right", "sample_cnt"):
        assert np.array_equal(getattr(t, name), getattr(o, name)), name
    inner = t.left != -1
    for name in ("fid", "lo", "hi"):
        assert np.array_equal(getattr(t, name)[inner], getattr(o, name)[inner]), name
    assert np.array_equal(t.node_info[inner], o.node_info[inner])
    if check_stats:
        leaf = ~inner
        np.testing.assert_allclose(t.node_info[leaf], o.node_info[leaf], rtol=REL, atol=1e-9)
        np.testing.assert_allclose(t.hess_sum, o.hess_sum, rtol=REL)
        fin = np.isfinite(o.loss_chg)
        assert np.array_equal(np.isfinite(t.loss_chg), fin)
        scale = max(1.0, float(np.abs(o.loss_chg[fin]).max())) if fin.any() else 1.0
        np.testing.assert_allclose(t.loss_chg[fin], o.loss_chg[fin], rtol=REL, atol=REL * scale)


def maker(kw, x, start, gp, K=1, **extra):
    p = yb.make_params(class_num=K, **kw)
    tm = yb.GpuTreeMaker(p, start, x.shape[0], **extra)
    tm.upload_rows(x)
    tm.set_grads(gp)
    return tm
