// Planner kernels of the device-driven level loop (level.cuh, engine.cu): one CTA each, queued between the
// data kernels of a level with programmatic dependent launch.  They restate, on the device, the decisions
// DataParallelTreeMaker.make takes between two kernels of a level
// (J/optimizer/gbdt/DataParallelTreeMaker.java:244-281): which polled nodes are split (:249-252), where
// the children's rows are (SamplePositionData.java:128-132), which child is built from rows and which by
// subtraction (:489-508), whose leaf statistics are summed (:341-385).
#include "level_plan.cuh"

namespace ytk {

__global__ void lv_root_k(LvTables t, long long begin, long long end, int buf, const PartitionCounts* mask_counts) {
    pdl_sync();
    if (threadIdx.x != 0) return;
    if (mask_counts) end = begin + mask_counts[0].left;          // rows kept by the row mask come first
    LvNode nd = {begin, end, 0, 0, 0, 0, end - begin, kLvActive, 0};
    t.nodes[0] = nd;
    const long long n = end - begin;
    t.build[0] = {begin, end, n + t.flush_cost, t.flush_cost, buf, 0};
    *t.n_build = n > 0 ? 1 : 0;
    t.seg_out[0] = t.world == 1 ? t.pool : t.staging;
    if (t.world > 1) {
        t.seg_final[0] = t.pool;
        t.counts[0] = {n, 0};                                    // summed over the ranks beside the root histogram
    }
    t.scans[0] = {t.pool, n, -1, 0, nullptr, nullptr, t.tot};
}

__global__ void __launch_bounds__(kPlanThreads)
lv_plan_split_k(LvTables t, int d, int buf, double min_split_loss, const int32_t* __restrict__ wide_index, int tile_rows) {
    pdl_sync();
    lv_plan_split_dev(t, d, buf, min_split_loss, wide_index, tile_rows);
}

// After the partition of level d.  The children of every split node get their row intervals; the smaller
// child (fewer global rows, ties to the right: :344-350, :493-499 -- with several GPUs: the smaller
// hessian sum of the parent's best split, which every GPU can evaluate without an exchange; integer
// histograms make parent - child exact whichever child is built) goes on the level's build list.
// leaf = false: the smaller child takes a fresh histogram slot, the larger one its parent's (sibling written
// over its parent), and the scan nodes of both children are laid down (the larger child's asks the search for parent - sibling).
__global__ void __launch_bounds__(kPlanThreads)
lv_plan_build_k(LvTables t, int d, int child_buf, int leaf) {
    pdl_sync();
    __shared__ long long s_warp_ll[kPlanThreads / 32];
    __shared__ int s_warp_i[kPlanThreads / 32];
    __shared__ long long s_cost;
    __shared__ int s_cnt;
    const int m = 1 << d, L0 = m - 1;
    if (threadIdx.x == 0) { s_cost = 0; s_cnt = 0; }
    __syncthreads();
    for (int s0 = 0; s0 < m; s0 += kPlanThreads) {
        const int s = s0 + threadIdx.x;
        bool add = false, add2 = false;      // add2: several GPUs, last level: the larger child's sums are taken too
        BuildNode bn = {0, 0, 0, 0, 0, 0}, bn2 = bn;
        if (s < m) {
            const int p = L0 + s;
            LvNode nd = t.nodes[p];
            ScanNode sc_l = {nullptr, 0, -1, 0, nullptr, nullptr, nullptr}, sc_r = sc_l;
            if (nd.flags & kLvSplit) {
                const PartitionCounts c = t.counts[s];
                const PartitionCounts gc = t.world == 1 ? c : t.gcounts[s];
                bool small_left;
                if (t.world == 1) small_left = gc.left < gc.right;
                else {
                    const long long lh = t.left_h[2 * p], th = t.left_h[2 * p + 1];
                    small_left = lh < th - lh;
                }
                nd.cnt_l = c.left; nd.cnt_r = c.right;
                nd.gcnt_l = gc.left; nd.gcnt_r = gc.right;
                if (small_left) nd.flags |= kLvSmallLeft;
                const int slot_small = leaf ? -1 : m + s, slot_large = leaf ? -1 : nd.slot;
                LvNode cl = {nd.begin, nd.begin + c.left, 0, 0, 0, 0, gc.left, kLvActive, small_left ? slot_small : slot_large};
                LvNode cr = {nd.begin + c.left, nd.end, 0, 0, 0, 0, gc.right, kLvActive, small_left ? slot_large : slot_small};
                t.nodes[2 * p + 1] = cl;
                t.nodes[2 * p + 2] = cr;
                t.nodes[p] = nd;
                bn.begin = small_left ? cl.begin : cr.begin;
                bn.end = small_left ? cl.end : cr.end;
                bn.buf = child_buf;
                bn.seg = leaf ? 2 * s + (small_left ? 0 : 1) : s;     // leaf sums are kept per child, histograms per parent slot
                add = bn.end > bn.begin;
                if (leaf && t.world > 1) {
                    // which child is the smaller one follows from the GLOBAL counts, which are only summed beside
                    // the leaf sums themselves: both children are summed and the host takes its pick (:344-350)
                    bn2.begin = small_left ? cr.begin : cl.begin;
                    bn2.end = small_left ? cr.end : cl.end;
                    bn2.buf = child_buf;
                    bn2.seg = 2 * s + (small_left ? 1 : 0);
                    add2 = bn2.end > bn2.begin;
                }
                if (!leaf) {
                    long long* parent_h = t.pool + (long long)nd.slot * t.seg_max;
                    long long* small_h = t.pool + (long long)slot_small * t.seg_max;
                    sc_l.hist = small_left ? small_h : parent_h;
                    sc_r.hist = small_left ? parent_h : small_h;
                    sc_l.sample_cnt = gc.left; sc_r.sample_cnt = gc.right;
                    sc_l.tot_out = t.tot + 2 * (2 * p + 1); sc_r.tot_out = t.tot + 2 * (2 * p + 2);
                    ScanNode& large = small_left ? sc_r : sc_l;            // substract (:101-107): done by the search, in place
                    large.sub = small_h; large.tot_parent = t.tot + 2 * p;
                    t.seg_out[s] = t.world == 1 ? small_h : t.staging + (long long)s * t.seg_max;
                    if (t.world > 1) t.seg_final[s] = small_h;
                }
            } else if (!leaf && t.world > 1) {
                t.seg_final[s] = nullptr;
            }
            if (!leaf) { t.scans[2 * s] = sc_l; t.scans[2 * s + 1] = sc_r; }
        }
        int n_add;
        int j = s_cnt + block_excl_scan<int>((add ? 1 : 0) + (add2 ? 1 : 0), s_warp_i, &n_add);
        long long cost_add;
        const long long c1 = add ? (bn.end - bn.begin) + t.flush_cost : 0, c2 = add2 ? (bn2.end - bn2.begin) + t.flush_cost : 0;
        bn.cost = bn2.cost = t.flush_cost;
        const long long before = s_cost + block_excl_scan<long long>(c1 + c2, s_warp_ll, &cost_add);
        if (add) { bn.cost_end = before + c1; t.build[j++] = bn; }
        if (add2) { bn2.cost_end = before + c1 + c2; t.build[j] = bn2; }
        __syncthreads();
        if (threadIdx.x == 0) { s_cnt += n_add; s_cost += cost_add; }
        __syncthreads();
    }
    if (threadIdx.x == 0) *t.n_build = s_cnt;
}

// computeLeafStats (:352-372) for the last level: exact fixed-point (g,h) sums over the rows of the level's
// build nodes (the smaller child of every node split there), walked in spans of equal cost like
// hist_span_k; build node `seg` = its position in level d + 1 adds into leaf[2 * (2^(d+1) - 1 + seg)].
__global__ void __launch_bounds__(256)
lv_leaf_sums_k(LvTables t, int d, const int2* __restrict__ gq, const int32_t* __restrict__ rowidx0,
               const int32_t* __restrict__ rowidx1) {
    pdl_sync();
    __shared__ long long s_g[8], s_h[8];
    const int nb = *t.n_build;
    if (nb <= 0) return;
    const int L0 = (2 << d) - 1;             // first heap node of the children's level
    const long long total = t.build[nb - 1].cost_end;
    const long long c0 = total * blockIdx.x / gridDim.x, c1 = total * (blockIdx.x + 1) / gridDim.x;
    if (c0 >= c1) return;
    int lo = 0, hi = nb - 1;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (t.build[mid].cost_end > c0) hi = mid; else lo = mid + 1;
    }
    for (int j = lo; j < nb; j++) {
        const BuildNode nd = t.build[j];
        const long long n = nd.end - nd.begin;
        const long long cost_begin = nd.cost_end - n - nd.cost;
        if (cost_begin >= c1) break;
        long long o0 = c0 - cost_begin, o1 = c1 - cost_begin;
        o0 = o0 < 0 ? 0 : o0;
        o1 = o1 > n ? n : o1;
        if (o0 >= o1) continue;
        const int32_t* __restrict__ ridx = nd.buf < 0 ? nullptr : (nd.buf == 0 ? rowidx0 : rowidx1);
        long long g = 0, h = 0;
        for (long long i = nd.begin + o0 + threadIdx.x; i < nd.begin + o1; i += blockDim.x) {
            const int2 q = ld_gather_int2(gq + (ridx ? ridx[i] : (int32_t)i));
            g += q.x; h += q.y;
        }
        for (int k = 16; k > 0; k >>= 1) {
            g += __shfl_xor_sync(0xffffffffu, g, k);
            h += __shfl_xor_sync(0xffffffffu, h, k);
        }
        __syncthreads();
        if ((threadIdx.x & 31) == 0) { s_g[threadIdx.x >> 5] = g; s_h[threadIdx.x >> 5] = h; }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int w = 1; w < 8; w++) { g += s_g[w]; h += s_h[w]; }
            long long* out = t.leaf + 2 * (long long)(L0 + nd.seg);
            if (g) atomicAdd(reinterpret_cast<unsigned long long*>(out), (unsigned long long)g);
            if (h) atomicAdd(reinterpret_cast<unsigned long long*>(out + 1), (unsigned long long)h);
        }
    }
}

cudaError_t launch_lv_root(LvTables t, long long begin, long long end, int buf, const PartitionCounts* mask_counts, cudaStream_t st) {
    return launch_chain(lv_root_k, dim3(1), dim3(32), 0, st, t, begin, end, buf, mask_counts);
}
cudaError_t launch_lv_plan_split(LvTables t, int d, int buf, double min_split_loss, const int32_t* wide_index, int tile_rows,
                                 cudaStream_t st) {
    return launch_chain(lv_plan_split_k, dim3(1), dim3(kPlanThreads), 0, st, t, d, buf, min_split_loss, wide_index, tile_rows);
}
cudaError_t launch_lv_plan_build(LvTables t, int d, int child_buf, bool leaf, cudaStream_t st) {
    return launch_chain(lv_plan_build_k, dim3(1), dim3(kPlanThreads), 0, st, t, d, child_buf, leaf ? 1 : 0);
}
cudaError_t launch_lv_leaf_sums(LvTables t, int d, const int2* gq, const int32_t* rowidx0, const int32_t* rowidx1, int grid,
                                cudaStream_t st) {
    return launch_chain(lv_leaf_sums_k, dim3(grid), dim3(256), 0, st, t, d, gq, rowidx0, rowidx1);
}

}  // namespace ytk
