"""ctypes binding of libytkgbdt.so (include/ytk_gbdt.h).  No torch types cross this
boundary: plain pointers and sizes, exactly what a JNI shim would pass."""
import ctypes as C
import os

import numpy as np

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libytkgbdt.so")


class YtkError(RuntimeError):
    """Raised for every non-zero ytk_status (the reference throws YtkLearnException /
    Mp4jException on the same conditions, J/utils/CheckUtils.java)."""

    def __init__(self, code, msg):
        super().__init__("ytk_gbdt error %d: %s" % (code, msg))
        self.code = code


class Params(C.Structure):
    _fields_ = [("max_depth", C.c_int32), ("max_leaf_cnt", C.c_int32),
                ("min_split_samples", C.c_int32), ("grow_policy", C.c_int32),
                ("min_child_hessian_sum", C.c_double), ("min_split_loss", C.c_double),
                ("max_abs_leaf_val", C.c_double), ("l1", C.c_double), ("l2", C.c_double),
                ("histogram_pool_capacity", C.c_double), ("learning_rate", C.c_float),
                ("class_num", C.c_int32)]


class TreeOut(C.Structure):
    _fields_ = [("capacity", C.c_int32), ("n_nodes", C.c_int32),
                ("parent", C.c_void_p), ("left", C.c_void_p), ("right", C.c_void_p),
                ("fid", C.c_void_p), ("lo", C.c_void_p), ("hi", C.c_void_p),
                ("node_info", C.c_void_p), ("loss_chg", C.c_void_p), ("hess_sum", C.c_void_p),
                ("sample_cnt", C.c_void_p)]


class TimeStats(C.Structure):
    _fields_ = [("build_hist", C.c_double), ("compute_hist", C.c_double),
                ("communicate_hist", C.c_double), ("init_stats", C.c_double),
                ("find_best_split", C.c_double), ("sync_best_split", C.c_double),
                ("partition", C.c_double), ("total", C.c_double),
                ("hist_miss", C.c_int32), ("hist_built", C.c_int32),
                ("hist_cells", C.c_double), ("hist_rows", C.c_double), ("hist_rows_indexed", C.c_double),
                ("gpu_hist_ms", C.c_double),
                ("kernel_launches", C.c_int32), ("hist_launches", C.c_int32), ("peer_collectives", C.c_int32), ("batches", C.c_int32), ("device_allocs", C.c_int32), ("max_batch_ms", C.c_double),
                ("enqueue_ms", C.c_double), ("host_wait_ms", C.c_double), ("stage_times_from_gpu", C.c_int32),
                ("graph_replayed", C.c_int32), ("device_span_ms", C.c_double), ("hist_events_ms", C.c_double), ("rendezvous", C.c_int32), ("reserved", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class Split(C.Structure):
    _fields_ = [("loss_chg", C.c_float), ("fid", C.c_int32), ("lo", C.c_int32), ("hi", C.c_int32),
                ("sum_grad", C.c_double), ("sum_hess", C.c_double),
                ("root_gain", C.c_float), ("node_value", C.c_float)]

    def as_tuple(self):
        return (self.loss_chg, self.fid, self.lo, self.hi)


class Tree:
    """Tree / TreeNode / TreeNodeStat as SoA numpy arrays (ytk_tree_out)."""
    FIELDS = (("parent", np.int32), ("left", np.int32), ("right", np.int32), ("fid", np.int32),
              ("lo", np.int32), ("hi", np.int32), ("node_info", np.float32),
              ("loss_chg", np.float32), ("hess_sum", np.float32), ("sample_cnt", np.int64))

    def __init__(self, capacity):
        self.capacity = int(capacity)
        self.n_nodes = 0
        for name, dt in self.FIELDS:
            setattr(self, name, np.zeros(self.capacity, dtype=dt))

    def _as_c(self):
        t = TreeOut()
        t.capacity = self.capacity
        t.n_nodes = 0
        for name, _ in self.FIELDS:
            setattr(t, name, getattr(self, name).ctypes.data)
        return t

    def _trimmed_copy(self, n):
        """A new Tree holding copies of the first n nodes (self stays a reusable full-capacity buffer)."""
        t = Tree.__new__(Tree)
        t.capacity = self.capacity
        t.n_nodes = n
        for name, _ in self.FIELDS:
            setattr(t, name, getattr(self, name)[:n].copy())
        return t

    def _trim(self, n):
        self.n_nodes = n
        for name, _ in self.FIELDS:
            setattr(self, name, getattr(self, name)[:n].copy())
        return self


_lib = None

_SIGS = {
    "ytk_gbdt_last_error": (C.c_char_p, []),
    "ytk_gbdt_normalize_max_leaf_cnt": (C.c_int32, [C.c_int32, C.c_int32]),
    "ytk_gbdt_avg_assign": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p]),
    "ytk_gbdt_merge_splits": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p]),
    "ytk_gbdt_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "ytk_gbdt_create": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "ytk_gbdt_destroy": (C.c_int, [C.c_void_p]),
    "ytk_gbdt_create_clique": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "ytk_gbdt_make_tree_clique": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                            C.c_void_p]),
    "ytk_gbdt_set_bins": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64]),
    "ytk_gbdt_upload_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32]),
    "ytk_gbdt_upload_rows_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32]),
    "ytk_gbdt_quantize_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p]),
    "ytk_gbdt_download_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]),
    "ytk_gbdt_set_grads": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]),
    "ytk_gbdt_make_tree": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ytk_gbdt_get_positions": (C.c_int, [C.c_void_p, C.c_void_p]),
    "ytk_gbdt_build_hist": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p]),
    "ytk_gbdt_subtract_hist": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ytk_gbdt_find_split": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p]),
    "ytk_gbdt_partition_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_float, C.c_void_p, C.c_void_p]),
    "ytk_gbdt_float_to_string": (C.c_int, [C.c_float, C.c_char_p, C.c_int32]),
    "ytk_gbdt_dump_model_header": (C.c_int64, [C.c_float, C.c_int32, C.c_char_p, C.c_int32, C.c_char_p, C.c_int64]),
    "ytk_gbdt_dump_tree": (C.c_int64, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_char_p, C.c_int64]),
    "ytk_gbdt_load_model_header": (C.c_int, [C.c_char_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int32, C.c_void_p,
                                             C.c_void_p]),
    "ytk_gbdt_load_tree": (C.c_int, [C.c_char_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int32, C.c_void_p]),
    "ytk_gbdt_parse_rows": (C.c_int, [C.c_char_p, C.c_int64, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_void_p,
                                      C.c_int32, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ytk_gbdt_refine_leaves": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ytk_gbdt_set_targets": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ytk_gbdt_update_scores": (C.c_int, [C.c_void_p, C.c_int32]),
    "ytk_gbdt_compute_grads": (C.c_int, [C.c_void_p, C.c_void_p]),
    "ytk_gbdt_get_scores": (C.c_int, [C.c_void_p, C.c_void_p]),
    "ytk_gbdt_get_grads": (C.c_int, [C.c_void_p, C.c_void_p]),
    "ytk_gbdt_set_bins_sparse": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int64]),
    "ytk_gbdt_upload_csr": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]),
    "ytk_gbdt_quantize_csr": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]),
    "ytk_gbdt_score_values": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_int32, C.c_int32, C.c_void_p]),
    "ytk_gbdt_partition_rows_static": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_int32, C.c_float,
                                                 C.c_void_p, C.c_void_p, C.c_void_p]),
}


def lib():
    """Loads libytkgbdt.so; fails loudly when it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)" % LIB_PATH)
        _lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(_lib, name)
            fn.restype = res
            fn.argtypes = args
    return _lib


def check(rc):
    if rc != 0:
        raise YtkError(rc, lib().ytk_gbdt_last_error().decode("utf-8", "replace"))


def ptr(a):
    return None if a is None else a.ctypes.data


def normalize_max_leaf_cnt(max_depth, max_leaf_cnt):
    return lib().ytk_gbdt_normalize_max_leaf_cnt(int(max_depth), int(max_leaf_cnt))


def avg_assign(amount, workers):
    out = np.zeros(2 * workers, dtype=np.int32)
    check(lib().ytk_gbdt_avg_assign(int(amount), int(workers), ptr(out)))
    return out.reshape(workers, 2)


def merge_splits(records):
    arr = (Split * len(records))(*records)
    out = Split()
    check(lib().ytk_gbdt_merge_splits(C.addressof(arr), len(records), C.addressof(out)))
    return out


def nccl_unique_id():
    buf = np.zeros(128, dtype=np.uint8)
    check(lib().ytk_gbdt_nccl_unique_id(ptr(buf)))
    return buf
