"""The oracle's virtual dense matrix (oracle/ytk_oracle.h: orc_set_virtual_csr) -- the dense loops of the
reference over CSR storage, used for the 10 000-feature shape whose int32 matrix does not fit in
memory -- must give exactly what the same loops give on the materialised matrix."""
import numpy as np
import pytest

import oracle_binding as ob


@pytest.mark.parametrize("W", [1, 3])
@pytest.mark.parametrize("sampled", [False, True])
def test_virtual_csr_equals_dense(W, sampled):
    rng = np.random.default_rng(10 * W + sampled)
    n, F = 3000, 50
    nbs = rng.integers(2, 40, size=F)
    start = np.concatenate([[0], np.cumsum(nbs)]).astype(np.int32)
    default = (rng.random(F) * nbs * (rng.random(F) < 0.5)).astype(np.int32)
    x = np.tile(default, (n, 1)).astype(np.int32)
    nz = rng.random((n, F)) < 0.1
    r, f = np.nonzero(nz)
    x[r, f] = (rng.random(len(r)) * nbs[f]).astype(np.int32)
    gp = rng.normal(size=(n, 2)).astype(np.float32)
    gp[:, 1] = np.abs(gp[:, 1])
    gp = gp.reshape(-1)
    p = ob.make_params(max_depth=5, min_child_hessian_sum=1.0, l2=1.0)
    fea = np.sort(rng.choice(F, 30, replace=False)).astype(np.int32) if sampled else None
    mask = (rng.random(n) < 0.8).astype(np.uint8) if sampled else None
    t, pos, _ = ob.make_tree(p, x, gp, 0, start, W=W, fea_index=fea, row_mask=mask)
    row_ptr = np.concatenate([[0], np.cumsum(nz.sum(1))]).astype(np.int64)
    t2, pos2, _ = ob.make_tree(p, None, gp, 0, start, W=W, fea_index=fea, row_mask=mask,
                               csr=(row_ptr, f.astype(np.int32), x[r, f], default))
    assert t.n_nodes == t2.n_nodes > 3
    for name, _ in ob.Tree.FIELDS:
        assert np.array_equal(getattr(t, name), getattr(t2, name), equal_nan=True), name
    assert np.array_equal(pos, pos2)
