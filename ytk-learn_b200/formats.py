"""Host-side mirror of the reference's text formats over the C ABI (include/ytk_gbdt.h):
GBDTModel.dumpModel / loadModel (J/data/gbdt/GBDTModel.java:63-97, J/data/gbdt/Tree.java:192-291) and the ytklearn
training-row format (docs/data_format.md).  J/ = src/main/java/com/fenbi/ytklearn/."""
import ctypes as C

import numpy as np

from . import binding as B


def float_to_string(v):
    """java.lang.Float.toString"""
    buf = C.create_string_buffer(48)
    B.check(B.lib().ytk_gbdt_float_to_string(C.c_float(v), buf, 48))
    return buf.value.decode()


def _ctree(tree):
    t = B.Tree(tree.n_nodes)
    for name, _ in B.Tree.FIELDS:
        getattr(t, name)[:] = getattr(tree, name)[: tree.n_nodes]
    ct = t._as_c()
    ct.n_nodes = tree.n_nodes
    return t, ct


def dump_model(trees, thresholds, base_prediction=0.0, class_num=1, loss_function="sigmoid", default_children=None,
               feature_names=None, with_stats=True):
    """The list GBDTModel.dumpModel returns, joined: the header, then one block per tree."""
    names = None
    if feature_names is not None:
        enc = [n.encode() for n in feature_names]
        names = (C.c_char_p * len(enc))(*enc)
    n = B.lib().ytk_gbdt_dump_model_header(C.c_float(base_prediction), int(class_num), loss_function.encode(), len(trees), None, 0)
    buf = C.create_string_buffer(n + 1)
    B.lib().ytk_gbdt_dump_model_header(C.c_float(base_prediction), int(class_num), loss_function.encode(), len(trees), buf, n + 1)
    out = [buf.value.decode()]
    for i, tree in enumerate(trees):
        keep, ct = _ctree(tree)
        thr = np.ascontiguousarray(thresholds[i], dtype=np.float32)
        dc = None if default_children is None else np.ascontiguousarray(default_children[i], dtype=np.int32)
        args = (C.addressof(ct), B.ptr(thr), B.ptr(dc), None if names is None else C.addressof(names), i, 1 if with_stats else 0)
        n = B.lib().ytk_gbdt_dump_tree(*args, None, 0)
        if n < 0:
            B.check(-1)
        buf = C.create_string_buffer(n + 1)
        B.lib().ytk_gbdt_dump_tree(*args, buf, n + 1)
        out.append(buf.value.decode())
    return "".join(out)


def load_model(text, max_nodes=1 << 16, name_stride=64):
    """GBDTModel.loadModel: returns (base_prediction, class_num, loss_function, [(Tree, threshold, default_child, names)])."""
    raw = text.encode()
    base, k, nt, used = C.c_float(), C.c_int32(), C.c_int32(), C.c_int64()
    loss = C.create_string_buffer(64)
    B.check(B.lib().ytk_gbdt_load_model_header(raw, len(raw), C.addressof(base), C.addressof(k), loss, 64, C.addressof(nt),
                                               C.addressof(used)))
    pos = used.value
    trees = []
    for _ in range(nt.value):
        t = B.Tree(max_nodes)
        ct = t._as_c()
        thr = np.zeros(max_nodes, np.float32)
        dc = np.zeros(max_nodes, np.int32)
        names = C.create_string_buffer(max_nodes * name_stride)
        B.check(B.lib().ytk_gbdt_load_tree(raw[pos:], len(raw) - pos, C.addressof(ct), B.ptr(thr), B.ptr(dc), names, name_stride,
                                           C.addressof(used)))
        pos += used.value
        n = ct.n_nodes
        nm = [names.raw[i * name_stride:(i + 1) * name_stride].split(b"\0")[0].decode() for i in range(n)]
        trees.append((t._trim(n), thr[:n].copy(), dc[:n].copy(), nm))
    return base.value, k.value, loss.value.decode(), trees


def parse_rows(text, class_num=1, dict_sorted=None, x_delim="###", y_delim=",", features_delim=",", name_val_delim=":",
               max_rows=None, max_entries=None, want_init=False):
    """Rows of the ytklearn text format -> (weight, y[rows, K], init or None, row_ptr, fids, values float32, bytes consumed)."""
    raw = text.encode() if isinstance(text, str) else text
    max_rows = raw.count(b"\n") + 1 if max_rows is None else max_rows
    max_entries = max(1, raw.count(name_val_delim.encode())) if max_entries is None else max_entries
    w = np.zeros(max_rows, np.float32)
    y = np.zeros(max_rows * class_num, np.float32)
    init = np.zeros(max_rows * class_num, np.float32) if want_init else None
    row_ptr = np.zeros(max_rows + 1, np.int64)
    fids = np.zeros(max_entries, np.int32)
    bits = np.zeros(max_entries, np.int32)
    d = None
    if dict_sorted is not None:
        enc = [s.encode() for s in dict_sorted]
        d = (C.c_char_p * len(enc))(*enc)
    nr, ne, used = C.c_int64(), C.c_int64(), C.c_int64()
    B.check(B.lib().ytk_gbdt_parse_rows(raw, len(raw), x_delim.encode(), y_delim.encode(), features_delim.encode(),
                                        name_val_delim.encode(), int(class_num), None if d is None else C.addressof(d),
                                        0 if d is None else len(d), max_rows, max_entries, B.ptr(w), B.ptr(y), B.ptr(init),
                                        B.ptr(row_ptr), B.ptr(fids), B.ptr(bits), C.addressof(nr), C.addressof(ne),
                                        C.addressof(used)))
    n, e = nr.value, ne.value
    return (w[:n], y[:n * class_num].reshape(n, class_num), None if init is None else init[:n * class_num].reshape(n, class_num),
            row_ptr[:n + 1], fids[:e], bits[:e].view(np.float32), used.value)
