// Best-split search over the histograms.
// Part of the CUDA side of libytkgbdt.so (sm_100a).  Nothing here is a dense contraction: the work is
// HBM-/shared-memory-bound byte and integer traffic, so there are no tensor cores; what matters is
// sector-aligned row reads, conflict-free shared-memory atomics and grids sized to the 148 SMs.
// Reference semantics restated by each kernel are cited as J/... =
// /root/reference/src/main/java/com/fenbi/ytklearn/...
#include "common.cuh"
#include "device_util.cuh"

#include <math_constants.h>

#include <algorithm>
#include <cstdlib>

namespace ytk {

// ------------------------------------------------------------------------------------
// K4/K5: node totals, split scan and arg-max.
//   updateTreeMakerNodeStats (J/optimizer/gbdt/DataParallelTreeMaker.java:543-573),
//   enumerateSplit (:598-637), SplitInfo.needReplace (J/data/gbdt/SplitInfo.java:99-104),
//   UpdateStrategy.calcGain / calcNodeValue (J/optimizer/gbdt/UpdateStrategy.java:55-100).
// One CTA per node, one warp per feature: a segmented inclusive scan over the feature's
// bins in exact 64-bit integers (so the prefix does not depend on scan order), gains in
// IEEE double with explicit _rn intrinsics (never fused), loss change rounded to float
// as the reference does.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ double threshold_l1(double w, double lambda) {
    if (w > lambda) return __dsub_rn(w, lambda);
    if (w < -lambda) return __dadd_rn(w, lambda);
    return 0.0;
}

__device__ __forceinline__ double calc_node_value(const GainParams& p, double G, double H) {
    if (H < p.min_child_weight) return 0.0;
    double val;
    if (p.l1 == 0.0) val = __ddiv_rn(-G, __dadd_rn(H, p.l2));
    else val = __ddiv_rn(-threshold_l1(G, p.l1), __dadd_rn(H, p.l2));
    if (p.max_abs_leaf_val > 0) {
        if (val > p.max_abs_leaf_val) val = p.max_abs_leaf_val;
        else if (val < -p.max_abs_leaf_val) val = -p.max_abs_leaf_val;
    }
    return val;
}

__device__ __forceinline__ double calc_gain(const GainParams& p, double G, double H) {
    if (H < p.min_child_weight) return 0.0;
    if (p.max_abs_leaf_val <= 0) {
        double t = (p.l1 == 0.0) ? G : threshold_l1(G, p.l1);
        return __ddiv_rn(__dmul_rn(t, t), __dadd_rn(H, p.l2));
    }
    double v = calc_node_value(p, G, H);
    // -2 * (G*v + 0.5*(H+l2)*v^2 + l1*|v|), evaluated left to right as Java does
    double a = __dmul_rn(G, v);
    double b = __dmul_rn(__dmul_rn(0.5, __dadd_rn(H, p.l2)), __dmul_rn(v, v));
    double c = __dmul_rn(p.l1, fabs(v));
    return __dmul_rn(-2.0, __dadd_rn(__dadd_rn(a, b), c));
}

constexpr int kScanWarps = 4;     // features per CTA; a node's features are spread over several CTAs (slices)

// One warp per feature.  Lane L holds bins [8L, 8L + 8) of a block of 256 bins (one block for a feature with uint8
// slots): a lane-local prefix, ONE warp scan per block and eight independent gain evaluations per lane, instead of
// eight dependent scan + gain rounds.  The feature's bins are requested before the node totals are known, so the
// kernel's critical path is one table load, one histogram load and the arithmetic.
// A node with `sub` set has no histogram yet: it is hist - sub (parent - smaller sibling, HistogramBuilder.substract,
// J/data/gbdt/HistogramBuilder.java:101-107), formed here in registers and written over the parent's copy by the
// feature's own warp -- no separate subtraction pass.  Nobody else reads those words during the launch: the node
// totals of such a node come from tot_parent (the parent's exact integer totals) minus the sibling's.
__global__ void __launch_bounds__(kScanWarps * 32)
split_scan_k(const ScanNode* __restrict__ nodes, const int32_t* __restrict__ start,
             const int64_t* __restrict__ feat_off, const uint8_t* __restrict__ fea_selected,
             int32_t f_from, int32_t f_to, int32_t first_fea, const GradScale* __restrict__ scale,
             GainParams gp, const PartitionCounts* __restrict__ cnt_global, SplitPart* __restrict__ parts,
             ytk_split* __restrict__ outs, ytk_split* __restrict__ mirror, unsigned int* __restrict__ done,
             long long* __restrict__ lh_out, unsigned long long* __restrict__ stamp) {
    pdl_sync();
    __shared__ double s_tot[2];
    __shared__ long long s_toth, s_lh[kScanWarps];
    __shared__ float s_lc[kScanWarps];
    __shared__ int s_fid[kScanWarps], s_lo[kScanWarps], s_hi[kScanWarps], s_slice[kScanWarps];
    __shared__ int s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int node = blockIdx.x, slice = blockIdx.y;
    const ScanNode nd = nodes[node];
    if (!nd.hist) return;                   // an unused entry of a device-planned table (the whole CTA leaves)
    stamp_begin(stamp);
    long long* const hist = const_cast<long long*>(nd.hist);       // written only where nd.sub asks for it
    const long long* __restrict__ sub = nd.sub;
    constexpr int kPer = kMaxBins / 32;     // bins per lane and block

    // bins [blk + 8 * lane, +8) of the feature, zero past its end; with `sub`: the difference
    auto load_block = [&](longlong2 (&v)[kPer], const long long* hp, const long long* sp, int blk, int nb) {
        longlong2 b[kPer];
#pragma unroll
        for (int k = 0; k < kPer; k++) {
            const int j = blk + lane * kPer + k;
            v[k] = make_longlong2(0, 0); b[k] = v[k];
            if (j < nb) v[k] = *reinterpret_cast<const longlong2*>(hp + 2 * j);
            if (sp && j < nb) b[k] = *reinterpret_cast<const longlong2*>(sp + 2 * j);
        }
#pragma unroll
        for (int k = 0; k < kPer; k++) { v[k].x -= b[k].x; v[k].y -= b[k].y; }
    };
    const int f = f_from + slice * kScanWarps + warp;       // one feature per warp
    long long* hp = nullptr;
    const long long* sp = nullptr;
    int nb = 0;
    longlong2 cur[kPer];
    bool mine = false;
    if (f < f_to) {
        // (the three table reads are independent: one round trip, not a chain, before the bins can be asked for)
        const uint8_t sel = fea_selected[f];
        const long long off = feat_off[f];
        const int nbf = start[f + 1] - start[f];
        mine = sel != 0;
        if (mine) {
            hp = hist + off;
            sp = sub ? sub + off : nullptr;
            nb = nbf;
            load_block(cur, hp, sp, 0, nb);
        }
    }
    const double ig = scale->inv_g, ih = scale->inv_h;
    long long sample_cnt = nd.sample_cnt;
    if (nd.cseg >= 0) {
        const PartitionCounts g = cnt_global[nd.cseg];
        const long long small = g.left < g.right ? g.left : g.right;
        sample_cnt = nd.which == 0 ? small : (g.left + g.right - small);
    }
    if (warp == 0) {
        double G, H;
        if (first_fea >= 0) {
            // every feature's bins add up to the node's totals: those of the first selected feature are taken
            long long g = 0, h = 0;
            const long long* tp = (sub ? sub : hist) + feat_off[first_fea];
            const int nb0 = start[first_fea + 1] - start[first_fea];
            for (int j = lane; j < nb0; j += 32) {
                longlong2 gh = *reinterpret_cast<const longlong2*>(tp + 2 * j);
                g += gh.x; h += gh.y;
            }
            for (int d = 16; d > 0; d >>= 1) {
                g += __shfl_xor_sync(0xffffffffu, g, d);
                h += __shfl_xor_sync(0xffffffffu, h, d);
            }
            if (sub) { g = nd.tot_parent[0] - g; h = nd.tot_parent[1] - h; }
            if (nd.tot_out && slice == 0 && lane == 0) { nd.tot_out[0] = g; nd.tot_out[1] = h; }
            G = (double)g * ig; H = (double)h * ih;
            if (lane == 0) s_toth = h;
        } else {
            G = -CUDART_INF; H = -CUDART_INF;   // no feature here: neutral for the MAX allreduce (:549)
            if (lane == 0) s_toth = 0;
        }
        if (lane == 0) { s_tot[0] = G; s_tot[1] = H; }
    }
    __syncthreads();
    const double totG = s_tot[0], totH = s_tot[1];
    const float root_gain = (float)calc_gain(gp, totG, totH);
    // UpdateStrategy.canSplit (:50-53)
    const bool can_split = (totH >= __dmul_rn(gp.min_child_weight, 2.0)) && (sample_cnt >= gp.min_split_samples);

    float w_lc = -CUDART_INF_F;
    int w_fid = -1, w_lo = -1, w_hi = -1;
    long long w_lh = 0;                     // fixed-point hessian left of the best split (exact)
    if (can_split && mine) {
        long long carryG = 0, carryH = 0;
        int carry_last = -1;
        float b_lc = -CUDART_INF_F;
        int b_lo = -1, b_hi = 0x7fffffff;
        long long b_lh = 0;
        for (int blk = 0; blk < nb; blk += kMaxBins) {
            if (blk > 0) load_block(cur, hp, sp, blk, nb);
            const int j0 = blk + lane * kPer;
            if (sp) {
#pragma unroll
                for (int k = 0; k < kPer; k++)
                    if (j0 + k < nb) *reinterpret_cast<longlong2*>(hp + 2 * (j0 + k)) = cur[k];
            }
            long long lg = 0, lh = 0;                       // this lane's bins together
            unsigned ne_mask = 0;                           // hasSample, bit k
#pragma unroll
            for (int k = 0; k < kPer; k++) {
                lg += cur[k].x; lh += cur[k].y;
                ne_mask |= (unsigned)((cur[k].x | cur[k].y) != 0) << k;
            }
            long long sg = lg, sh = lh;                     // inclusive over the lanes
            int pm = ne_mask ? j0 + 31 - __clz(ne_mask) : -1;   // last bin with samples up to and including this lane
            for (int d = 1; d < 32; d <<= 1) {
                const long long tg = __shfl_up_sync(0xffffffffu, sg, d);
                const long long th = __shfl_up_sync(0xffffffffu, sh, d);
                const int tm = __shfl_up_sync(0xffffffffu, pm, d);
                if (lane >= d) { sg += tg; sh += th; pm = max(pm, tm); }
            }
            long long exG = carryG + sg - lg, exH = carryH + sh - lh;     // left branch before bin j0 (+ k below)
            int prev = __shfl_up_sync(0xffffffffu, pm, 1);                // lastFeaValue
            prev = lane == 0 ? carry_last : max(prev, carry_last);
            // (a loop, not eight copies: two double-precision gains per bin are ~150 instructions, and a kernel this
            // short pays for every instruction-cache line it touches; the bins move down a register queue)
#pragma unroll 1
            for (int k = 0; k < kPer && (ne_mask >> k) != 0; k++) {
                const long long g = cur[0].x, h = cur[0].y;
                const bool ne = (ne_mask >> k) & 1u;
                if (ne && exH != 0) {                                      // !leftBranchGradStats.empty()
                    double lG = (double)exG * ig, lH = (double)exH * ih;
                    if (lH >= gp.min_child_weight) {
                        double rG = __dsub_rn(totG, lG), rH = __dsub_rn(totH, lH);
                        if (rH >= gp.min_child_weight) {
                            float lc = (float)__dsub_rn(__dadd_rn(calc_gain(gp, lG, lH), calc_gain(gp, rG, rH)),
                                                        (double)root_gain);
                            if (lc > b_lc) { b_lc = lc; b_lo = prev; b_hi = j0 + k; b_lh = exH; }
                        }
                    }
                }
                if (ne) prev = j0 + k;
                exG += g; exH += h;
#pragma unroll
                for (int i = 0; i + 1 < kPer; i++) cur[i] = cur[i + 1];
            }
            carryG += __shfl_sync(0xffffffffu, sg, 31);
            carryH += __shfl_sync(0xffffffffu, sh, 31);
            carry_last = max(carry_last, __shfl_sync(0xffffffffu, pm, 31));
        }
        // first (lowest bin) maximum inside the feature
        for (int d = 16; d > 0; d >>= 1) {
            float o_lc = __shfl_xor_sync(0xffffffffu, b_lc, d);
            int o_lo = __shfl_xor_sync(0xffffffffu, b_lo, d);
            int o_hi = __shfl_xor_sync(0xffffffffu, b_hi, d);
            long long o_lh = __shfl_xor_sync(0xffffffffu, b_lh, d);
            if (o_lc > b_lc || (o_lc == b_lc && o_hi < b_hi)) { b_lc = o_lc; b_lo = o_lo; b_hi = o_hi; b_lh = o_lh; }
        }
        if (b_lc > w_lc) { w_lc = b_lc; w_fid = f; w_lo = b_lo; w_hi = b_hi; w_lh = b_lh; }
    }
    if (lane == 0) { s_lc[warp] = w_lc; s_fid[warp] = w_fid; s_lo[warp] = w_lo; s_hi[warp] = w_hi; s_lh[warp] = w_lh; }
    __syncthreads();
    if (threadIdx.x == 0) {
        // warps hold increasing feature ids: strict > keeps the smaller fid on ties
        SplitPart best = {-CUDART_INF_F, -1, -1, -1, 0, 0};
        for (int w = 0; w < kScanWarps; w++)
            if (s_fid[w] >= 0 && s_lc[w] > best.lc) best = {s_lc[w], s_fid[w], s_lo[w], s_hi[w], s_lh[w], 0};
        parts[(size_t)node * gridDim.y + slice] = best;
        if (slice == 0 && lh_out) lh_out[2 * node + 1] = s_toth;
        if (slice == 0) {
            ytk_split* out = outs + node;
            out->sum_grad = totG; out->sum_hess = totH;
            out->root_gain = root_gain;
            out->node_value = (float)calc_node_value(gp, totG, totH);
        }
        // the last CTA of a node to get here folds the node's slices (no separate merge launch)
        __threadfence();
        s_last = atomicAdd(done + node, 1u) == gridDim.y - 1;
    }
    __syncthreads();
    if (!s_last) { stamp_end(stamp); return; }
    __threadfence();
    // slices hold increasing feature ids: the maximum, ties to the lowest slice = smallest fid
    SplitPart bp = {-CUDART_INF_F, -1, -1, -1, 0, 0};
    int bs = 0x7fffffff;
    for (int s = threadIdx.x; s < (int)gridDim.y; s += blockDim.x) {
        const SplitPart* pp = parts + (size_t)node * gridDim.y + s;
        const int4 v = __ldcg(reinterpret_cast<const int4*>(pp));
        const float lc = __int_as_float(v.x);
        if (v.y >= 0 && lc > bp.lc) { bp = {lc, v.y, v.z, v.w, __ldcg(&pp->lh), 0}; bs = s; }
    }
    for (int d = 16; d > 0; d >>= 1) {
        const float olc = __shfl_xor_sync(0xffffffffu, bp.lc, d);
        const int ofid = __shfl_xor_sync(0xffffffffu, bp.fid, d), olo = __shfl_xor_sync(0xffffffffu, bp.lo, d);
        const int ohi = __shfl_xor_sync(0xffffffffu, bp.hi, d), os = __shfl_xor_sync(0xffffffffu, bs, d);
        const long long olh = __shfl_xor_sync(0xffffffffu, bp.lh, d);
        if (ofid >= 0 && (olc > bp.lc || (olc == bp.lc && os < bs))) { bp = {olc, ofid, olo, ohi, olh, 0}; bs = os; }
    }
    if (lane == 0) { s_lc[warp] = bp.lc; s_fid[warp] = bp.fid; s_lo[warp] = bp.lo; s_hi[warp] = bp.hi; s_slice[warp] = bs; s_lh[warp] = bp.lh; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kScanWarps; w++)
            if (s_fid[w] >= 0 && (s_lc[w] > bp.lc || (s_lc[w] == bp.lc && s_slice[w] < bs))) {
                bp = {s_lc[w], s_fid[w], s_lo[w], s_hi[w], s_lh[w], 0}; bs = s_slice[w];
            }
        if (lh_out) lh_out[2 * node] = bp.lh;
        ytk_split r;                       // totals, root gain and node value: written by the node's slice 0
        r.sum_grad = __ldcg(&outs[node].sum_grad); r.sum_hess = __ldcg(&outs[node].sum_hess);
        r.root_gain = __ldcg(&outs[node].root_gain); r.node_value = __ldcg(&outs[node].node_value);
        r.loss_chg = bp.lc; r.fid = bp.fid; r.lo = bp.lo; r.hi = bp.hi;
        outs[node] = r;
        if (mirror) mirror[node] = r;      // pinned host memory on one GPU: the host reads it without a copy
        done[node] = 0;                    // ready for the next launch
    }
    stamp_end(stamp);
}

int split_scan_slices(int32_t f_from, int32_t f_to) {
    int n = f_to - f_from;
    return n <= 0 ? 1 : (n + kScanWarps - 1) / kScanWarps;
}

cudaError_t launch_split_scan(const ScanNode* nodes, int n_nodes, const int32_t* start, const int64_t* feat_off,
                              const uint8_t* fea_selected, int32_t f_from, int32_t f_to, int32_t first_fea,
                              const GradScale* scale, GainParams gp, const PartitionCounts* cnt_global,
                              SplitPart* parts, ytk_split* out, ytk_split* mirror, unsigned int* done, cudaStream_t st,
                              long long* lh_out, unsigned long long* stamp) {
    if (n_nodes <= 0) return cudaSuccess;
    const int slices = split_scan_slices(f_from, f_to);
    dim3 grid((unsigned)n_nodes, (unsigned)slices);
    return launch_chain(split_scan_k, grid, dim3(kScanWarps * 32), 0, st, nodes, start, feat_off, fea_selected, f_from,
                        f_to, first_fea, scale, gp, cnt_global, parts, out, mirror, done, lh_out, stamp);
}

}  // namespace ytk
