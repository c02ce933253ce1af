"""the model text format (GBDTModel.dumpModel / loadModel) and the ytklearn row
format, through the C ABI.  Host-only entry points: no GPU needed."""
import os
import re

import numpy as np
import pytest

import _pkg
import ytk_data as yd

yb = _pkg.load()
from ytk_learn_b200 import formats as fm  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# the reference's own node patterns (J/data/gbdt/Tree.java:47-48)
INNER = re.compile(r"(\S+):\[f_(\S+)<=(\S+)] yes=(\S+),no=(\S+),missing=(\S+),gain=(\S+),hess_sum=(\S+),sample_cnt=(\S+)")
LEAF = re.compile(r"(\S+):leaf=(\S+),hess_sum=(\S+),sample_cnt=(\S+)")


@pytest.mark.parametrize("v,s", [
    (1.0, "1.0"), (0.5, "0.5"), (100.0, "100.0"), (-2.5, "-2.5"), (0.0, "0.0"), (-0.0, "-0.0"),
    (1.0e7, "1.0E7"), (9999999.0, "9999999.0"), (1.6777216e7, "1.6777216E7"), (0.001, "0.001"), (9.999e-4, "9.999E-4"),
    (0.1, "0.1"), (1.0 / 3.0, "0.33333334"), (3.4028235e38, "3.4028235E38"),
    (1.4e-45, "1.4E-45"), (123456.79, "123456.79"), (1.0e-5, "1.0E-5"), (12345678.0, "1.2345678E7"),
    (float("inf"), "Infinity"), (float("-inf"), "-Infinity"), (float("nan"), "NaN"), (0.17249267, "0.17249267")])
def test_java_float_to_string(v, s):
    """java.lang.Float.toString (what the model dump prints): documented outputs."""
    assert fm.float_to_string(np.float32(v)) == s


def test_float_to_string_round_trips():
    rng = np.random.default_rng(0)
    bits = rng.integers(0, 2 ** 32, size=20000, dtype=np.uint64).astype(np.uint32)
    vals = bits.view(np.float32)
    for v in vals[np.isfinite(vals)][:5000]:
        s = fm.float_to_string(v)
        assert np.float32(float(s.replace("E", "e"))) == v, (v, s)
        assert "." in s and not s.endswith(".")


def _tree():
    t = yb.Tree(7)
    t.n_nodes = 7
    t.parent[:] = [-1, 0, 0, 1, 1, 2, 2]
    t.left[:] = [1, 3, -1, -1, -1, -1, -1]
    t.right[:] = [2, 4, -1, -1, -1, -1, -1]
    t.fid[:] = [2, 0, 0, 0, 0, 0, 0]
    t.node_info[:] = [0, 0, 0.25, -0.125, 1.5e-4, 0, 0]
    t.loss_chg[:] = [12.5, 3.0, 0, 0, 0, 0, 0]
    t.hess_sum[:] = [100.0, 60.5, 39.5, 20.25, 40.25, 0, 0]
    t.sample_cnt[:] = [400, 242, 158, 81, 161, 0, 0]
    t = t._trim(5)
    thr = np.array([0.5, 17.25, 0, 0, 0], np.float32)
    dc = np.array([2, 3, -1, -1, -1], np.int32)
    return t, thr, dc


def test_dump_model_is_the_reference_text():
    t, thr, dc = _tree()
    text = fm.dump_model([t], [thr], base_prediction=0.5, class_num=1, loss_function="sigmoid", default_children=[dc],
                         feature_names=["age", "b", "height"])
    want = ("uniform_base_prediction=0.5\nclass_num=1\nloss_function=sigmoid\ntree_num=1\n"
            "booster[1] depth=2,node_num=5,leaf_cnt=3\n"
            "0:[f_height<=0.5] yes=1,no=2,missing=2,gain=12.5,hess_sum=100.0,sample_cnt=400\n"
            "\t1:[f_age<=17.25] yes=3,no=4,missing=3,gain=3.0,hess_sum=60.5,sample_cnt=242\n"
            "\t\t3:leaf=-0.125,hess_sum=20.25,sample_cnt=81\n"
            "\t\t4:leaf=1.5E-4,hess_sum=40.25,sample_cnt=161\n"
            "\t2:leaf=0.25,hess_sum=39.5,sample_cnt=158\n")
    assert text == want
    for line in text.splitlines()[5:]:
        assert (LEAF if "leaf" in line else INNER).search(line), line
    # without statistics (dumpModel(iter, false))
    plain = fm.dump_model([t], [thr], with_stats=False)
    assert "0:[f_2<=0.5] yes=1,no=2,missing=1\n" in plain and "gain" not in plain


def test_load_model_reads_back_what_dump_writes():
    t, thr, dc = _tree()
    text = fm.dump_model([t, t], [thr, thr], base_prediction=-1.25, class_num=2, loss_function="softmax", default_children=[dc, dc],
                         feature_names=["age", "b", "height"])
    base, k, loss, trees = fm.load_model(text)
    assert (base, k, loss, len(trees)) == (-1.25, 2, "softmax", 2)
    for lt, lthr, ldc, names in trees:
        for name in ("parent", "left", "right", "sample_cnt", "hess_sum"):
            assert np.array_equal(getattr(lt, name), getattr(t, name)), name
        inner = t.left != -1
        assert np.array_equal(lthr[inner], thr[inner]) and np.array_equal(ldc[inner], dc[inner])
        assert np.array_equal(lt.node_info[~inner], t.node_info[~inner])
        assert np.array_equal(lt.loss_chg[inner], t.loss_chg[inner])
        assert [names[i] for i in np.where(inner)[0]] == ["height", "age"]
    # numeric feature names come back as indices; dumping the loaded model gives the same bytes
    text2 = fm.dump_model([t], [thr], default_children=[dc])
    _, _, _, (one,) = fm.load_model(text2)
    assert list(one[0].fid[t.left != -1]) == [2, 0]
    assert fm.dump_model([one[0]], [one[1]], default_children=[one[2]]) == text2
    with pytest.raises(yb.YtkError):
        fm.load_model("uniform_base_prediction=0.0\nclass_num=1\nloss_function=l2\ntree_num=0\n")


@pytest.mark.parametrize("name,K", [("agaricus.train.ytklearn.head", 1), ("dermatology.train.ytklearn.head", 6)])
def test_parse_rows_of_the_reference_demo_data(name, K):
    """The first rows of the reference's demo files (tests/golden/*.head, copied by make_golden.py) against the
    harness's own pure-Python reader; feature index = rank of the name in lexicographic String order."""
    path = os.path.join(GOLDEN, name)
    text = open(path).read()
    w, y, feats = yd.parse_ytklearn(path)
    names = sorted({k for d in feats for k in d})
    gw, gy, _, row_ptr, fids, vals, used = fm.parse_rows(text, class_num=K, dict_sorted=names)
    assert used == len(text.encode()) and len(gw) == len(w)
    assert np.array_equal(gw, w)
    if K == 1:
        assert np.array_equal(gy[:, 0], y)
    else:
        assert np.array_equal(gy.argmax(1), y.astype(np.int64)) and np.array_equal(gy.sum(1), np.ones(len(y), np.float32))
    for r, d in enumerate(feats):
        got = {names[f]: v for f, v in zip(fids[row_ptr[r]:row_ptr[r + 1]], vals[row_ptr[r]:row_ptr[r + 1]])}
        assert got == d
        assert list(fids[row_ptr[r]:row_ptr[r + 1]]) == sorted(fids[row_ptr[r]:row_ptr[r + 1]])


def test_parse_rows_details():
    text = (" 2.5###1### b : 3 , a:1.5,zz:9,a:7 ### 0.25 \n"          # blanks, unknown name, repeated name, init prediction
            "\n"
            "1###0###a:-1e-3\n"
            "3###1###b:2")                       # no newline at the end: left for the next chunk
    w, y, init, row_ptr, fids, vals, used = fm.parse_rows(text, dict_sorted=["a", "b"], want_init=True)
    assert list(w) == [2.5, 1.0] and list(y[:, 0]) == [1.0, 0.0] and list(init[:, 0]) == [0.25, 0.0]
    assert list(row_ptr) == [0, 2, 3] and list(fids) == [0, 1, 0] and list(vals) == [7.0, 3.0, np.float32(-1e-3)]
    assert text.encode()[used:] == b"3###1###b:2"
    # no dictionary: the name is the index; capacity stops the parse in front of a row
    w, _, _, row_ptr, fids, vals, used = fm.parse_rows("1;0;3=1|7=2\n1;1;5=4\n", x_delim=";", features_delim="|",
                                                       name_val_delim="=", max_entries=2)
    assert len(w) == 1 and list(fids) == [3, 7] and used == len("1;0;3=1|7=2\n")
    for bad in ("1###0\n", "x###0###a:1\n", "1###0###a:inf\n", "1###0###a\n"):
        with pytest.raises(yb.YtkError):
            fm.parse_rows(bad, dict_sorted=["a"])
    with pytest.raises(yb.YtkError):
        fm.parse_rows("1###7###a:1\n", class_num=3, dict_sorted=["a"])
