"""Host-side mirror of the reference's tree-maker interface over the C ABI.

`GpuTreeMaker` plays the role of DataParallelTreeMaker
(J/optimizer/gbdt/DataParallelTreeMaker.java:49-664) behind ITreeMaker
(J/optimizer/gbdt/ITreeMaker.java:34-39): constructed from the optimisation params, the
binned train matrix and the bin layout; `make(group_id)` returns the Tree; the gradient
pairs are pushed before each round as GBDTOptimizer does
(J/optimizer/GBDTOptimizer.java:482-488,550-552).  J/ = src/main/java/com/fenbi/ytklearn/.
"""
import ctypes as C

import numpy as np

from . import binding as B


def make_params(max_depth=6, max_leaf_cnt=-1, min_split_samples=-1, grow_policy="level",
                min_child_hessian_sum=1.0, min_split_loss=0.0, max_abs_leaf_val=-1.0,
                l1=0.0, l2=1.0, histogram_pool_capacity=-1.0, learning_rate=0.1, class_num=1,
                normalize=True):
    """GBDTOptimizationParams fields the tree maker reads
    (J/param/gbdt/GBDTOptimizationParams.java:100-205); `normalize` applies the
    data-parallel max_leaf_cnt rule (:148-154)."""
    if normalize:
        max_leaf_cnt = B.normalize_max_leaf_cnt(max_depth, max_leaf_cnt)
    if isinstance(grow_policy, str):
        if grow_policy not in ("level", "loss"):
            raise B.YtkError(-1, "[GBDT] tree_grow_policy (%s) invalid, loss or level" % grow_policy)
        grow_policy = 1 if grow_policy == "loss" else 0
    return B.Params(int(max_depth), int(max_leaf_cnt), int(min_split_samples), int(grow_policy),
                    float(min_child_hessian_sum), float(min_split_loss), float(max_abs_leaf_val),
                    float(l1), float(l2), float(histogram_pool_capacity), float(learning_rate),
                    int(class_num))


class GpuTreeMaker:
    """One instance per GPU / row shard (the reference: one per worker thread)."""

    def __init__(self, params, start, n_rows, device=0, rank=0, world=1, nccl_id=None, sparse_default=None, nnz=0):
        """sparse_default (int32 [F]) switches the handle to sparse rows (upload_csr): the slot of a
        feature a row does not store; nnz = number of stored entries of this shard."""
        self._h = C.c_void_p()
        self.params = params
        self.start = np.ascontiguousarray(start, dtype=np.int32)
        self.F = len(self.start) - 1
        self.B = int(self.start[-1])
        self.N = int(n_rows)
        self.K = params.class_num
        self.last_stats = None
        idp = None if nccl_id is None else B.ptr(np.ascontiguousarray(nccl_id, dtype=np.uint8))
        self._nccl_id = nccl_id
        B.check(B.lib().ytk_gbdt_create(C.addressof(params), int(device), int(rank), int(world), idp,
                                        C.addressof(self._h)))
        if sparse_default is None:
            B.check(B.lib().ytk_gbdt_set_bins(self._h, B.ptr(self.start), self.F, self.N))
        else:
            d = np.ascontiguousarray(sparse_default, dtype=np.int32)
            B.check(B.lib().ytk_gbdt_set_bins_sparse(self._h, B.ptr(self.start), self.F, self.N, B.ptr(d), int(nnz)))

    @classmethod
    def clique(cls, params, start, n_rows_per_rank, device=0):
        """`len(n_rows_per_rank)` tree makers on ONE device that behave as the ranks of a multi-GPU run
        (ytk_gbdt_create_clique); build their trees together with `make_clique`."""
        world = len(n_rows_per_rank)
        hs = (C.c_void_p * world)()
        B.check(B.lib().ytk_gbdt_create_clique(C.addressof(params), int(device), world, C.addressof(hs)))
        members = []
        for r in range(world):
            m = cls.__new__(cls)
            m._h = C.c_void_p(hs[r])
            m.params = params
            m.start = np.ascontiguousarray(start, dtype=np.int32)
            m.F = len(m.start) - 1
            m.B = int(m.start[-1])
            m.N = int(n_rows_per_rank[r])
            m.K = params.class_num
            m.last_stats = None
            m._nccl_id = None
            B.check(B.lib().ytk_gbdt_set_bins(m._h, B.ptr(m.start), m.F, m.N))
            members.append(m)
        return members

    @staticmethod
    def make_clique(members, group_id=0, fea_index=None, row_masks=None):
        """ITreeMaker.make of every rank of a clique at once; returns the ranks' trees."""
        world = len(members)
        p = members[0].params
        cap = 2 * p.max_leaf_cnt - 1 if p.max_leaf_cnt > 0 else 2 ** (p.max_depth + 1) - 1
        trees = [B.Tree(cap) for _ in range(world)]
        outs = (B.TreeOut * world)(*[t._as_c() for t in trees])
        ts = (B.TimeStats * world)()
        hs = (C.c_void_p * world)(*[m._h for m in members])
        fi = None if fea_index is None else np.ascontiguousarray(fea_index, dtype=np.int32)
        masks, mp = None, None
        if row_masks is not None:
            masks = [None if m is None else np.ascontiguousarray(m, dtype=np.uint8) for m in row_masks]
            mp = (C.c_void_p * world)(*[None if m is None else m.ctypes.data for m in masks])
        B.check(B.lib().ytk_gbdt_make_tree_clique(C.addressof(hs), world, int(group_id), B.ptr(fi), 0 if fi is None else len(fi),
                                                  None if mp is None else C.addressof(mp), C.addressof(outs), C.addressof(ts)))
        for r, m in enumerate(members):
            m.last_stats = B.TimeStats.from_buffer_copy(ts[r])
        return [t._trim(outs[r].n_nodes) for r, t in enumerate(trees)]

    def close(self):
        if self._h:
            B.lib().ytk_gbdt_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- data in -------------------------------------------------------------------
    def upload_csr(self, row_ptr, fids, slots, row0=0):
        """One CSR chunk (rows in order, entries of a row sorted by feature)."""
        rp = np.ascontiguousarray(row_ptr, dtype=np.int64)
        f = np.ascontiguousarray(fids, dtype=np.int32)
        sl = np.ascontiguousarray(slots, dtype=np.int32)
        B.check(B.lib().ytk_gbdt_upload_csr(self._h, B.ptr(rp), B.ptr(f), B.ptr(sl), int(row0), len(rp) - 1))

    def quantize_csr(self, row_ptr, fids, values, cuts, row0=0):
        """CSR chunk of original float values, binned on the device (A1) against `cuts`."""
        rp = np.ascontiguousarray(row_ptr, dtype=np.int64)
        f = np.ascontiguousarray(fids, dtype=np.int32)
        v = np.ascontiguousarray(values, dtype=np.float32)
        c = np.ascontiguousarray(cuts, dtype=np.float32)
        B.check(B.lib().ytk_gbdt_quantize_csr(self._h, B.ptr(rp), B.ptr(f), B.ptr(v.view(np.int32)), int(row0), len(rp) - 1,
                                              B.ptr(c)))

    def upload_rows(self, x, row0=0):
        """x: int32 slots [n, stride] (a GBDTCoreData.x chunk after binning)."""
        x = np.ascontiguousarray(x, dtype=np.int32)
        B.check(B.lib().ytk_gbdt_upload_rows(self._h, B.ptr(x), int(row0), x.shape[0], x.shape[1]))

    def upload_rows_u8(self, x, row0=0):
        """x: uint8 slots [n, stride] (every feature at most 256 bins)."""
        x = np.ascontiguousarray(x, dtype=np.uint8)
        B.check(B.lib().ytk_gbdt_upload_rows_u8(self._h, B.ptr(x), int(row0), x.shape[0], x.shape[1]))

    def quantize_rows(self, values, cuts, row0=0):
        """values: float32 [n, stride] raw feature values; cuts: sorted bin centres."""
        bits = np.ascontiguousarray(values, dtype=np.float32).view(np.int32)
        cuts = np.ascontiguousarray(cuts, dtype=np.float32)
        B.check(B.lib().ytk_gbdt_quantize_rows(self._h, B.ptr(bits), int(row0), bits.shape[0],
                                               bits.shape[1], B.ptr(cuts)))

    def download_rows(self, row0=0, nrows=None):
        nrows = self.N - row0 if nrows is None else nrows
        out = np.zeros((nrows, self.F), dtype=np.int32)
        B.check(B.lib().ytk_gbdt_download_rows(self._h, B.ptr(out), int(row0), int(nrows)))
        return out

    def set_grads(self, gp, row0=0):
        """gp: float32 (g,h) pairs laid out [(row*K + gid)*2 + {0,1}]."""
        gp = np.ascontiguousarray(gp, dtype=np.float32).reshape(-1)
        n = gp.size // (2 * self.K)
        B.check(B.lib().ytk_gbdt_set_grads(self._h, B.ptr(gp), int(row0), n))

    def set_grads_ptr(self, host_ptr, row0, nrows):
        """Same, from a raw host address (e.g. pinned memory)."""
        B.check(B.lib().ytk_gbdt_set_grads(self._h, C.c_void_p(host_ptr), int(row0), int(nrows)))

    # ---- ITreeMaker ------------------------------------------------------------------
    def make(self, group_id=0, fea_index=None, row_mask=None):
        cap = 2 * self.params.max_leaf_cnt - 1 if self.params.max_leaf_cnt > 0 \
            else 2 ** (self.params.max_depth + 1) - 1
        buf = getattr(self, "_tree_buf", None)           # node arrays at full capacity, reused by every call
        if buf is None or buf[0].capacity != cap:
            tree = B.Tree(cap)
            buf = self._tree_buf = (tree, tree._as_c())
        tree, tc = buf
        ts = B.TimeStats()
        fi = None if fea_index is None else np.ascontiguousarray(fea_index, dtype=np.int32)
        rm = None if row_mask is None else np.ascontiguousarray(row_mask, dtype=np.uint8)
        if rm is not None and rm.size != self.N:
            raise B.YtkError(-1, "row_mask must have one byte per local row")
        B.check(B.lib().ytk_gbdt_make_tree(self._h, int(group_id), B.ptr(fi), 0 if fi is None else len(fi),
                                           B.ptr(rm), C.addressof(tc), C.addressof(ts)))
        self.last_stats = ts
        return tree._trimmed_copy(tc.n_nodes)

    def get_total_time_stats(self):
        return self.last_stats

    def get_positions(self):
        out = np.zeros(self.N, dtype=np.int32)
        B.check(B.lib().ytk_gbdt_get_positions(self._h, B.ptr(out)))
        return out

    # ---- the step either side of make(), on the device ------------------
    LOSSES = {"sigmoid": 0, "l2": 1, "softmax": 2, "l1": 3}

    def set_targets(self, loss, y, weight=None, init_score=None):
        """GBDTCoreData.y / weight / initScore; y is [N*K] (one-hot rows when K > 1)."""
        y = np.ascontiguousarray(y, dtype=np.float32).reshape(-1)
        if y.size != self.N * self.K:
            raise B.YtkError(-1, "y must have N*K entries")
        w = None if weight is None else np.ascontiguousarray(weight, dtype=np.float32)
        i0 = None if init_score is None else np.ascontiguousarray(init_score, dtype=np.float32).reshape(-1)
        B.check(B.lib().ytk_gbdt_set_targets(self._h, self.LOSSES[loss], B.ptr(y), B.ptr(w), B.ptr(i0)))

    def refine_leaves(self, tree, residual=None, weight=None):
        """TreeRefiner.refine (l1): leaf values of `tree` (the last make()'s) become learning_rate x the weighted
        median of the residuals of their rows; returns the tree (modified in place)."""
        r = None if residual is None else np.ascontiguousarray(residual, dtype=np.float32)
        w = None if weight is None else np.ascontiguousarray(weight, dtype=np.float32)
        ct = tree._as_c()
        ct.n_nodes = tree.n_nodes
        B.check(B.lib().ytk_gbdt_refine_leaves(self._h, B.ptr(r), B.ptr(w), C.addressof(ct)))
        return tree

    def update_scores(self, group_id=0):
        B.check(B.lib().ytk_gbdt_update_scores(self._h, int(group_id)))

    def compute_grads(self):
        """predictAndCalcLossGrad's loss/gradient loop; returns sum_i w_i * loss_i."""
        loss = C.c_double(0.0)
        B.check(B.lib().ytk_gbdt_compute_grads(self._h, C.addressof(loss)))
        return loss.value

    def get_scores(self):
        out = np.zeros(self.N * self.K, dtype=np.float32)
        B.check(B.lib().ytk_gbdt_get_scores(self._h, B.ptr(out)))
        return out

    def get_grads(self):
        out = np.zeros(self.N * self.K * 2, dtype=np.float32)
        B.check(B.lib().ytk_gbdt_get_grads(self._h, B.ptr(out)))
        return out

    def score_values(self, tree, threshold, values, score, default_child=None, gid=0):
        """Test-set scoring (GBDTOptimizer.score, isOriginTree=true): `values` float32 [n, S] (NaN with the
        canonical bit pattern = missing), `threshold` per node, `score` float32 [n*K] updated in place."""
        v = np.ascontiguousarray(values, dtype=np.float32)
        thr = np.ascontiguousarray(threshold, dtype=np.float32)
        dc = None if default_child is None else np.ascontiguousarray(default_child, dtype=np.int32)
        t = B.Tree(tree.n_nodes)
        for name, _ in B.Tree.FIELDS:
            getattr(t, name)[:] = getattr(tree, name)[: tree.n_nodes]
        ct = t._as_c()
        ct.n_nodes = tree.n_nodes
        assert score.dtype == np.float32 and score.flags.c_contiguous
        B.check(B.lib().ytk_gbdt_score_values(self._h, B.ptr(v.view(np.int32)), v.shape[0], v.shape[1], C.addressof(ct),
                                              B.ptr(thr), B.ptr(dc), self.K, int(gid), B.ptr(score)))
        return score

    # ---- per-stage hooks ---------------------------------------------------------------
    def build_hist(self, group_id=0, rows=None, fea_index=None):
        out = np.zeros(2 * self.B, dtype=np.float64)
        r = None if rows is None else np.ascontiguousarray(rows, dtype=np.int32)
        fi = None if fea_index is None else np.ascontiguousarray(fea_index, dtype=np.int32)
        B.check(B.lib().ytk_gbdt_build_hist(self._h, int(group_id), B.ptr(r), 0 if r is None else len(r),
                                            B.ptr(fi), 0 if fi is None else len(fi), B.ptr(out)))
        return out

    def subtract_hist(self, parent, small):
        parent = np.ascontiguousarray(parent, dtype=np.float64)
        small = np.ascontiguousarray(small, dtype=np.float64)
        out = np.zeros(2 * self.B, dtype=np.float64)
        B.check(B.lib().ytk_gbdt_subtract_hist(self._h, B.ptr(parent), B.ptr(small), B.ptr(out)))
        return out

    def find_split(self, hist, sample_cnt, fea_index=None):
        hist = np.ascontiguousarray(hist, dtype=np.float64)
        fi = None if fea_index is None else np.ascontiguousarray(fea_index, dtype=np.int32)
        out = B.Split()
        B.check(B.lib().ytk_gbdt_find_split(self._h, B.ptr(hist), int(sample_cnt), B.ptr(fi),
                                            0 if fi is None else len(fi), C.addressof(out)))
        return out

    def partition_rows(self, rows, fid, cond):
        rows = np.ascontiguousarray(rows, dtype=np.int32)
        out = np.zeros(len(rows), dtype=np.int32)
        left = C.c_int64(0)
        B.check(B.lib().ytk_gbdt_partition_rows(self._h, B.ptr(rows), len(rows), int(fid), C.c_float(cond),
                                                B.ptr(out), C.addressof(left)))
        return out, left.value
