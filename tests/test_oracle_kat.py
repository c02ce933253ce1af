"""Pins the CPU oracle to the only known answers the reference publishes for this path:
the loss lines of demo/gbdt/binary_classification/README.md:43-49 and the prediction
line of demo/gbdt/multiclass_classification/README.md:53."""
import numpy as np
import pytest

import kat
import oracle_binding as ob


def _oracle_maker(d, params, W=1):
    p = ob.make_params(**params)
    xb = d["xb"].astype(np.int32)

    def make_tree(gp, gid):
        t, _, _ = ob.make_tree(p, xb, gp, gid, d["start"], W=W)
        return t
    return make_tree


@pytest.mark.parametrize("W", [1, 4])
def test_kat1_agaricus_losses(W):
    d = kat.load_kat1()
    out, trees = kat.run_kat1(_oracle_maker(d, kat.KAT1_PARAMS, W), d)
    for it, (tr, te) in kat.KAT1.items():
        assert abs(out[it][0] - tr) < 1e-12, (it, out[it][0], tr)
        assert abs(out[it][1] - te) < 1e-12, (it, out[it][1], te)
    assert [t.n_nodes for t in trees] == [21, 23, 23]


def test_kat2_dermatology_prediction():
    d = kat.load_kat2()
    pred, trees = kat.run_kat2(_oracle_maker(d, kat.KAT2_PARAMS), d)
    want = np.array(kat.KAT2_PRED)
    err = np.abs(pred - want[None, :]).max(axis=1)
    hits = np.where(err < 1e-12)[0]
    assert len(hits) >= 1, err.min()
    # every matching row carries the README's label (fourth class)
    assert all(int(d["yt"][r]) == 3 for r in hits)
    assert [[t.n_nodes for t in g] for g in trees][0] == [7, 17, 5, 21, 3, 3]
