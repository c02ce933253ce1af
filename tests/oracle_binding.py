"""ctypes binding of the CPU oracle (oracle/ytk_oracle.{h,c}).

Test infrastructure only: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py, never by the product package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ODIR = os.path.join(_ROOT, "oracle")
_SO = os.path.join(_ODIR, "libytk_oracle.so")


def build_oracle(force=False):
    src = [os.path.join(_ODIR, f) for f in ("ytk_oracle.c", "ytk_oracle.h")]
    if (not force and os.path.exists(_SO)
            and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in src)):
        return _SO
    base = ["gcc", "-O3", "-march=x86-64-v2", "-fPIC", "-ffp-contract=off", "-fno-fast-math",
            "-std=c11", "-D_POSIX_C_SOURCE=200809L", "-shared", "-o", _SO, src[0], "-lm"]
    try:
        subprocess.run(base[:1] + ["-fopenmp"] + base[1:], check=True, capture_output=True)
    except subprocess.CalledProcessError:
        subprocess.run(base, check=True)  # single-threaded oracle still checks parity
    return _SO


class OrcParams(C.Structure):
    _fields_ = [("max_depth", C.c_int32), ("max_leaf_cnt", C.c_int32),
                ("min_split_samples", C.c_int32), ("grow_policy", C.c_int32),
                ("min_child_hessian_sum", C.c_double), ("min_split_loss", C.c_double),
                ("max_abs_leaf_val", C.c_double), ("l1", C.c_double), ("l2", C.c_double),
                ("histogram_pool_capacity", C.c_double), ("learning_rate", C.c_float),
                ("class_num", C.c_int32)]


class OrcTree(C.Structure):
    _fields_ = [("capacity", C.c_int32), ("n_nodes", C.c_int32),
                ("parent", C.c_void_p), ("left", C.c_void_p), ("right", C.c_void_p),
                ("fid", C.c_void_p), ("lo", C.c_void_p), ("hi", C.c_void_p),
                ("node_info", C.c_void_p), ("loss_chg", C.c_void_p), ("hess_sum", C.c_void_p),
                ("sample_cnt", C.c_void_p)]


class OrcSplit(C.Structure):
    _fields_ = [("loss_chg", C.c_float), ("fid", C.c_int32), ("lo", C.c_int32), ("hi", C.c_int32)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build_oracle()
        _lib = C.CDLL(_SO)
        _lib.orc_find_bin.restype = C.c_int32
        _lib.orc_find_bin.argtypes = [C.c_void_p, C.c_int32, C.c_float]
        _lib.orc_calc_gain.restype = C.c_double
        _lib.orc_calc_gain.argtypes = [C.c_void_p, C.c_double, C.c_double]
        _lib.orc_calc_node_value.restype = C.c_double
        _lib.orc_calc_node_value.argtypes = [C.c_void_p, C.c_double, C.c_double]
        _lib.orc_reset_position.restype = C.c_int64
        _lib.orc_loss_grad_scalar.restype = C.c_double
        _lib.orc_loss_grad_softmax.restype = C.c_double
        _lib.orc_normalize_max_leaf_cnt.restype = C.c_int32
        _lib.orc_split_need_replace.restype = C.c_int
        _lib.orc_split_need_replace.argtypes = [C.c_void_p, C.c_float, C.c_int32]
        _lib.orc_set_virtual_csr.restype = None
        _lib.orc_set_virtual_csr.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.orc_enumerate_split.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                             C.c_double, C.c_double, C.c_float, C.c_void_p]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_params(max_depth=6, max_leaf_cnt=-1, min_split_samples=-1, grow_policy="level",
                min_child_hessian_sum=1.0, min_split_loss=0.0, max_abs_leaf_val=-1.0,
                l1=0.0, l2=1.0, histogram_pool_capacity=-1.0, learning_rate=0.1, class_num=1,
                normalize=True):
    if normalize:
        max_leaf_cnt = lib().orc_normalize_max_leaf_cnt(int(max_depth), int(max_leaf_cnt))
    return dict(max_depth=int(max_depth), max_leaf_cnt=int(max_leaf_cnt),
                min_split_samples=int(min_split_samples),
                grow_policy=1 if grow_policy in ("loss", 1) else 0,
                min_child_hessian_sum=float(min_child_hessian_sum),
                min_split_loss=float(min_split_loss), max_abs_leaf_val=float(max_abs_leaf_val),
                l1=float(l1), l2=float(l2),
                histogram_pool_capacity=float(histogram_pool_capacity),
                learning_rate=float(learning_rate), class_num=int(class_num))


def _cparams(p):
    return OrcParams(**p)


class Tree:
    """SoA tree as the tree maker hands it back (see include/ytk_gbdt.h: ytk_tree_out)."""
    FIELDS = (("parent", np.int32), ("left", np.int32), ("right", np.int32), ("fid", np.int32),
              ("lo", np.int32), ("hi", np.int32), ("node_info", np.float32),
              ("loss_chg", np.float32), ("hess_sum", np.float32), ("sample_cnt", np.int64))

    def __init__(self, capacity):
        self.capacity = int(capacity)
        self.n_nodes = 0
        for name, dt in self.FIELDS:
            setattr(self, name, np.zeros(self.capacity, dtype=dt))

    def trim(self):
        for name, _ in self.FIELDS:
            setattr(self, name, getattr(self, name)[: self.n_nodes].copy())
        return self

    def as_orc(self):
        t = OrcTree()
        t.capacity = self.capacity
        t.n_nodes = self.n_nodes
        for name, _ in self.FIELDS:
            setattr(t, name, getattr(self, name).ctypes.data)
        return t


def tree_capacity(p):
    if p["max_leaf_cnt"] > 0:
        return 2 * p["max_leaf_cnt"] - 1
    return 2 ** (p["max_depth"] + 1) - 1


def find_bins(values, cuts, start):
    """values float32 [N,F] -> int32 bins [N,F] (A1), feature-outer like the reference."""
    values = np.ascontiguousarray(values, dtype=np.float32)
    n, f = values.shape
    x = values.view(np.int32).copy()
    cuts = np.ascontiguousarray(cuts, dtype=np.float32)
    start = np.ascontiguousarray(start, dtype=np.int32)
    lib().orc_convert_to_bins(_p(x), C.c_int64(n), C.c_int32(f), C.c_int32(f), _p(cuts), _p(start))
    return x


def avg_assign(amount, bins):
    out = np.zeros(2 * bins, dtype=np.int32)
    lib().orc_avg_assign(C.c_int32(amount), C.c_int32(bins), _p(out))
    return out.reshape(bins, 2)


def build_hist(x, gp, K, gid, rows, fea_index, start):
    """Local histogram (A2) of the listed rows. x int32 [N,S]; returns float64 [2B]."""
    x = np.ascontiguousarray(x, dtype=np.int32)
    gp = np.ascontiguousarray(gp, dtype=np.float32)
    rows = np.asarray(rows, dtype=np.int32)
    pos_idx = np.zeros(2 * len(rows), dtype=np.int32)
    pos_idx[1::2] = rows
    fea_index = np.ascontiguousarray(fea_index, dtype=np.int32)
    start = np.ascontiguousarray(start, dtype=np.int32)
    B = int(start[-1])
    hist = np.zeros(2 * B, dtype=np.float64)
    lib().orc_build_hist(_p(x), C.c_int32(x.shape[1]), _p(gp), C.c_int32(K), C.c_int32(gid),
                         _p(pos_idx), C.c_int64(0), C.c_int64(len(rows)), _p(fea_index),
                         C.c_int32(len(fea_index)), _p(start), C.c_int32(B), _p(hist))
    return hist


def enumerate_split(p, hist, start, fid, sumG, sumH, root_gain, best=None):
    cp = _cparams(p)
    b = OrcSplit()
    if best is None:
        lib().orc_split_clear(C.byref(b))
    else:
        b.loss_chg, b.fid, b.lo, b.hi = best
    hist = np.ascontiguousarray(hist, dtype=np.float64)
    start = np.ascontiguousarray(start, dtype=np.int32)
    lib().orc_enumerate_split(C.byref(cp), _p(hist), _p(start), C.c_int32(fid),
                              C.c_double(sumG), C.c_double(sumH), C.c_float(root_gain), C.byref(b))
    return (b.loss_chg, b.fid, b.lo, b.hi)


def split_need_replace(cur, new_loss_chg, new_fid):
    b = OrcSplit()
    b.loss_chg, b.fid, b.lo, b.hi = cur
    return bool(lib().orc_split_need_replace(C.byref(b), C.c_float(new_loss_chg), C.c_int32(new_fid)))


def calc_gain(p, G, H):
    cp = _cparams(p)
    return lib().orc_calc_gain(C.byref(cp), C.c_double(G), C.c_double(H))


def calc_node_value(p, G, H):
    cp = _cparams(p)
    return lib().orc_calc_node_value(C.byref(cp), C.c_double(G), C.c_double(H))


def reset_position(pos_idx, v1, v2, x, fid, cond, left, right):
    x = np.ascontiguousarray(x, dtype=np.int32)
    return lib().orc_reset_position(_p(pos_idx), C.c_int64(v1), C.c_int64(v2), _p(x),
                                    C.c_int32(x.shape[1]), C.c_int32(fid), C.c_float(cond),
                                    C.c_int32(left), C.c_int32(right))


def make_tree(p, x, gp, gid, start, fea_index=None, W=1, row_off=None, row_mask=None,
              threads=1, want_positions=True, csr=None):
    """DataParallelTreeMaker.make with W emulated workers. Returns (Tree, node_of_row, info).
    csr = (row_ptr int64, fids int32, slots int32, default_slot int32) with x = None: the rows are
    looked up in CSR storage by the same dense loops (oracle/ytk_oracle.h: orc_set_virtual_csr)."""
    if csr is not None:
        assert x is None
        rp, cf, cs, dflt = (np.ascontiguousarray(csr[0], np.int64), np.ascontiguousarray(csr[1], np.int32),
                            np.ascontiguousarray(csr[2], np.int32), np.ascontiguousarray(csr[3], np.int32))
        lib().orc_set_virtual_csr(_p(rp), _p(cf), _p(cs), _p(dflt))
        n, S = len(rp) - 1, len(dflt)
    else:
        x = np.ascontiguousarray(x, dtype=np.int32)
        n, S = x.shape
    gp = np.ascontiguousarray(gp, dtype=np.float32)
    start = np.ascontiguousarray(start, dtype=np.int32)
    F = len(start) - 1
    if fea_index is None:
        fea_index = np.arange(F, dtype=np.int32)
    fea_index = np.ascontiguousarray(fea_index, dtype=np.int32)
    if row_off is None:
        row_off = np.array([(n * w) // W for w in range(W + 1)], dtype=np.int64)
    row_off = np.ascontiguousarray(row_off, dtype=np.int64)
    if row_mask is not None:
        row_mask = np.ascontiguousarray(row_mask, dtype=np.uint8)
    cp = _cparams(p)
    t = Tree(tree_capacity(p))
    ot = t.as_orc()
    node_of_row = np.full(n, -2, dtype=np.int32) if want_positions else None
    miss = C.c_int32(0)
    ts = np.zeros(4, dtype=np.float64)
    cells = C.c_double(0)
    rc = lib().orc_make_tree(C.byref(cp), C.c_int32(W), _p(row_off), _p(x) if csr is None else None, C.c_int32(S),
                             C.c_int32(F), _p(gp), C.c_int32(gid), _p(start), _p(fea_index),
                             C.c_int32(len(fea_index)), _p(row_mask), C.byref(ot),
                             _p(node_of_row), C.byref(miss), C.c_int32(threads), _p(ts),
                             C.byref(cells))
    if csr is not None:
        lib().orc_set_virtual_csr(None, None, None, None)
    if rc != 0:
        raise RuntimeError("orc_make_tree failed rc=%d" % rc)
    t.n_nodes = ot.n_nodes
    t.trim()
    info = dict(hist_miss=miss.value, t_hist=ts[0], t_split=ts[1], t_partition=ts[2],
                t_total=ts[3], hist_cells=cells.value)
    return t, node_of_row, info


def _tree_to_orc(tree):
    t = Tree(tree.n_nodes)
    t.n_nodes = tree.n_nodes
    for name, dt in Tree.FIELDS:
        getattr(t, name)[:] = getattr(tree, name)[: tree.n_nodes]
    return t


def score_slots(tree, x, score, K=1, gid=0):
    x = np.ascontiguousarray(x, dtype=np.int32)
    t = _tree_to_orc(tree)
    ot = t.as_orc()
    lib().orc_score_slots(C.byref(ot), _p(x), C.c_int64(x.shape[0]), C.c_int32(x.shape[1]),
                          _p(score), C.c_int32(K), C.c_int32(gid))


def score_values(tree, thr, xv, score, K=1, gid=0):
    xv = np.ascontiguousarray(xv, dtype=np.float32)
    thr = np.ascontiguousarray(thr, dtype=np.float32)
    t = _tree_to_orc(tree)
    ot = t.as_orc()
    lib().orc_score_values(C.byref(ot), _p(thr), _p(xv), C.c_int64(xv.shape[0]),
                           C.c_int32(xv.shape[1]), _p(score), C.c_int32(K), C.c_int32(gid))


def score_value_bits(tree, thr, def_child, xbits, score, K=1, gid=0):
    xbits = np.ascontiguousarray(xbits, dtype=np.int32)
    thr = np.ascontiguousarray(thr, dtype=np.float32)
    dc = None if def_child is None else np.ascontiguousarray(def_child, dtype=np.int32)
    t = _tree_to_orc(tree)
    ot = t.as_orc()
    f = lib().orc_score_value_bits
    f.restype = None
    f(C.byref(ot), _p(thr), _p(dc) if dc is not None else None, _p(xbits), C.c_int64(xbits.shape[0]),
      C.c_int32(xbits.shape[1]), _p(score), C.c_int32(K), C.c_int32(gid))


def loss_grad_scalar(kind, score, init_score, y, w, want_grad=True):
    n = len(y)
    gp = np.zeros(2 * n, dtype=np.float32) if want_grad else None
    loss = lib().orc_loss_grad_scalar(C.c_int32({"sigmoid": 0, "l2": 1, "l1": 3}[kind]), _p(score),
                                      _p(init_score), _p(y), _p(w), C.c_int64(n), _p(gp))
    return loss, gp


def refine_leaves(tree, node_of_row, residual, weight, learning_rate):
    """TreeRefiner.refine (exact): returns a copy of the tree with refined leaf values."""
    t = _tree_to_orc(tree)
    ot = t.as_orc()
    residual = np.ascontiguousarray(residual, dtype=np.float32)
    weight = None if weight is None else np.ascontiguousarray(weight, dtype=np.float32)
    pos = np.ascontiguousarray(node_of_row, dtype=np.int32)
    f = lib().orc_refine_leaves
    f.restype = None
    f(C.byref(ot), _p(pos), _p(residual), _p(weight) if weight is not None else None, C.c_int64(len(pos)),
      C.c_float(learning_rate))
    return t


def loss_grad_softmax(score, init_score, y_onehot, w, K, want_grad=True):
    n = len(w)
    gp = np.zeros(2 * n * K, dtype=np.float32) if want_grad else None
    pred = np.zeros(n * K, dtype=np.float32)
    loss = lib().orc_loss_grad_softmax(_p(score), _p(init_score), _p(y_onehot), _p(w),
                                       C.c_int64(n), C.c_int32(K), _p(gp), _p(pred))
    return loss, gp, pred
