"""CPU-side checks of the C ABI: the library loads, exports every symbol the header
declares, its host-only helpers agree with the oracle, and compute entry points refuse to
run without a GPU (no fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import _pkg
import oracle_binding as ob
from conftest import ROOT, has_gpu

yb = _pkg.load()


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ytk_gbdt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ytk_gbdt_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    names = declared_symbols()
    assert len(names) >= 18
    raw = C.CDLL(yb.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), "libytkgbdt.so does not export %s" % n
    yb.lib()  # and the binding's signatures resolve


def test_header_cites_reference_for_every_entry_point():
    src = open(os.path.join(ROOT, "include", "ytk_gbdt.h")).read()
    assert src.count("J/") >= 15


@pytest.mark.parametrize("amount,workers", [(28, 8), (117, 4), (5, 8), (128, 3), (0, 2), (33, 1)])
def test_avg_assign_matches_oracle(amount, workers):
    assert np.array_equal(yb.avg_assign(amount, workers), ob.avg_assign(amount, workers))


@pytest.mark.parametrize("d,l", [(8, -1), (5, 16), (3, 100), (-1, 255), (0, -1), (10, 1024)])
def test_normalize_max_leaf_cnt(d, l):
    assert yb.normalize_max_leaf_cnt(d, l) == ob.lib().orc_normalize_max_leaf_cnt(d, l)


def test_merge_splits_is_need_replace_fold():
    rng = np.random.default_rng(3)
    for _ in range(200):
        n = int(rng.integers(1, 9))
        recs, cur = [], None
        fids = rng.permutation(40)[:n]
        vals = rng.choice([0.5, 1.0, 1.0, 2.0, -np.inf], size=n)
        for i in range(n):
            lc = np.float32(vals[i])
            fid = int(fids[i]) if np.isfinite(lc) else -1
            s = yb.Split(lc, fid, 1, 2, float(i), float(-i), 0.0, 0.0)
            recs.append(s)
            t = (float(lc), fid, 1, 2)
            if cur is None or ob.split_need_replace(cur, t[0], t[1]):
                cur = t
        out = yb.merge_splits(recs)
        assert (out.loss_chg, out.fid) == (np.float32(cur[0]), cur[1])
        assert out.sum_grad == max(r.sum_grad for r in recs)
        assert out.sum_hess == max(r.sum_hess for r in recs)


def test_argument_errors_are_reported():
    p = yb.make_params()
    with pytest.raises(yb.YtkError):
        yb.GpuTreeMaker(p, [0, 3], 10, rank=2, world=2)
    with pytest.raises(yb.YtkError):
        yb.make_params(grow_policy="depth")


@pytest.mark.skipif(has_gpu(), reason="only meaningful on a machine without a GPU")
def test_no_cpu_fallback():
    with pytest.raises(yb.YtkError) as ei:
        yb.GpuTreeMaker(yb.make_params(), [0, 3, 5], 10)
    assert ei.value.code == -2
    assert "no CPU path" in str(ei.value)
