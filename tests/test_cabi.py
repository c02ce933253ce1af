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


def _header_struct(name):
    """(field, C type) pairs of `typedef struct { ... } name;` in include/ytk_gbdt.h"""
    src = open(os.path.join(ROOT, "include", "ytk_gbdt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    body = re.search(r"typedef struct\s*\{([^}]*)\}\s*%s\s*;" % name, src).group(1)
    out = []
    for decl in body.split(";"):
        decl = " ".join(decl.split())
        if not decl:
            continue
        ctype, names = decl.rsplit(" ", 1) if "," not in decl else (decl.split(" ", 1)[0], decl.split(" ", 1)[1])
        for nm in names.split(","):
            nm = nm.strip()
            ptr = nm.startswith("*") or ctype.endswith("*")
            out.append((nm.lstrip("*"), ctype.rstrip("* ").strip() + ("*" if ptr else "")))
    return out


@pytest.mark.parametrize("cname,pyname", [("ytk_gbdt_params", "Params"), ("ytk_tree_out", "TreeOut"),
                                          ("ytk_time_stats", "TimeStats"), ("ytk_split", "Split")])
def test_ctypes_structs_mirror_the_header(cname, pyname):
    """The ctypes mirror used by the tests and bench.py has the header's fields, in order, with the same types."""
    ctypes_of = {"int32_t": C.c_int32, "int64_t": C.c_int64, "double": C.c_double, "float": C.c_float}
    want = _header_struct(cname)
    got = getattr(yb.binding if hasattr(yb, "binding") else yb, pyname)._fields_
    assert [n for n, _ in want] == [n for n, _ in got]
    for (n, ct), (_, pt) in zip(want, got):
        assert pt is (C.c_void_p if ct.endswith("*") else ctypes_of[ct]), (n, ct, pt)


def test_tree_buffer_is_reused_but_returned_trees_are_copies():
    """GpuTreeMaker.make keeps one full-capacity node buffer per maker and hands out trimmed COPIES
    (binding.Tree._trimmed_copy): a tree returned earlier must not change when the buffer is written again."""
    yb = _pkg.load()
    buf = yb.binding.Tree(7) if hasattr(yb, "binding") else None
    if buf is None:
        import importlib
        buf = importlib.import_module("ytk-learn_b200.binding").Tree(7)
    buf.fid[:] = np.arange(7)
    buf.node_info[:] = 1.5
    t1 = buf._trimmed_copy(3)
    assert t1.n_nodes == 3 and t1.capacity == 7 and list(t1.fid) == [0, 1, 2]
    buf.fid[:] = -1                      # the next make() overwrites the buffer
    buf.node_info[:] = 9.0
    assert list(t1.fid) == [0, 1, 2] and float(t1.node_info[0]) == 1.5
    t2 = buf._trimmed_copy(5)
    assert t2.n_nodes == 5 and list(t2.fid) == [-1] * 5
    assert buf.fid.shape[0] == 7         # the buffer itself keeps its capacity
