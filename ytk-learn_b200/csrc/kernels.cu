// CUDA kernels of the GBDT hot path (sm_100a).  Nothing here is a dense contraction:
// the work is HBM-/shared-memory-bound byte and integer traffic, so there are no tensor
// cores; what matters is sector-aligned row reads, conflict-free shared-memory atomics
// and grids sized to the 148 SMs.
//
// Reference semantics restated by each kernel are cited as J/... =
// /root/reference/src/main/java/com/fenbi/ytklearn/...
#include "common.cuh"

#include <math_constants.h>

namespace ytk {

// ------------------------------------------------------------------------------------
// K0: int32 slot matrix (GBDTCoreData.x chunk, J/dataflow/GBDTCoreData.java:164-168)
//     -> uint8 row-major, rows padded to a multiple of 32 features (one 32 B sector per
//     feature group and row).
// ------------------------------------------------------------------------------------
__global__ void repack_rows_k(const int32_t* __restrict__ x, int64_t nrows, int32_t stride, int32_t F,
                              uint8_t* __restrict__ bins, int32_t S8, int* bad_flag) {
    int64_t total = nrows * S8;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i / S8;
        int32_t f = (int32_t)(i - r * S8);
        uint8_t b = 0;
        if (f < F) {
            int32_t v = x[r * stride + f];
            if (v < 0 || v >= kMaxBins) { *bad_flag = 1; v = 0; }
            b = (uint8_t)v;
        }
        bins[i] = b;
    }
}

// ------------------------------------------------------------------------------------
// K1: value -> bin.  FeatureApprData.convertFeaVal2ApprFeaIndex
//     (J/data/gbdt/FeatureApprData.java:179-205) with BinarySearch.findLastEqualOrUpper
//     (J/utils/BinarySearch.java:33-56): nearest bin centre, ties to the upper one; the
//     midpoint is a float add widened to double times 0.5.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ int32_t find_bin(const float* __restrict__ c, int32_t m, float v) {
    if (m == 1) return 0;
    if (v > c[m - 1]) return m - 1;
    int32_t l = 0, u = m - 1;
    while (l <= u) {
        int32_t mid = l + (u - l) / 2;
        if (v >= c[mid]) l = mid + 1; else u = mid - 1;
    }
    u = max(0, u);
    int32_t idx = (c[u] == v) ? u : min(m - 1, l);
    if (idx >= 1) {
        float s = __fadd_rn(c[idx], c[idx - 1]);
        if ((double)v < __dmul_rn((double)s, 0.5)) idx -= 1;
    }
    return idx;
}

__global__ void quantize_rows_k(const int32_t* __restrict__ xbits, int64_t nrows, int32_t stride, int32_t F,
                                const float* __restrict__ cuts, const int32_t* __restrict__ start,
                                uint8_t* __restrict__ bins, int32_t S8) {
    int64_t total = nrows * S8;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i / S8;
        int32_t f = (int32_t)(i - r * S8);
        uint8_t b = 0;
        if (f < F) {
            float v = __int_as_float(xbits[r * stride + f]);
            int32_t s0 = start[f];
            b = (uint8_t)find_bin(cuts + s0, start[f + 1] - s0, v);
        }
        bins[i] = b;
    }
}

__global__ void unpack_rows_k(const uint8_t* __restrict__ bins, int32_t S8, int64_t nrows, int32_t F,
                              int32_t* __restrict__ out) {
    int64_t total = nrows * F;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i / F;
        int32_t f = (int32_t)(i - r * F);
        out[i] = bins[r * S8 + f];
    }
}

// ------------------------------------------------------------------------------------
// Gradient fixed point.  The reference adds float (g,h) into double bins
// (J/data/gbdt/HistogramBuilder.java:87-88).  Shared memory only has native 32-bit
// integer atomics, so each tree's gradients are scaled by a power of two chosen from
// max|g| (resp. max|h|) to 31-bit integers and every sum downstream is an exact 64-bit
// integer: histograms do not depend on row order, sharding or pool hits, and
// parent - child equals the sibling bit for bit.  Per-row rounding <= 2^-31 of max|g|.
// A non-zero value never rounds to zero, so "bin has samples" (hasSample,
// J/optimizer/gbdt/DataParallelTreeMaker.java:657-663) is preserved exactly.
// ------------------------------------------------------------------------------------
__global__ void grad_absmax_k(const float* __restrict__ gp, int64_t n, int32_t K, int32_t gid,
                              unsigned* __restrict__ absmax_bits) {
    float mg = 0.f, mh = 0.f;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        float2 v = *reinterpret_cast<const float2*>(gp + (i * K + gid) * 2);
        mg = fmaxf(mg, fabsf(v.x));
        mh = fmaxf(mh, fabsf(v.y));
    }
    for (int d = 16; d > 0; d >>= 1) {
        mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, d));
        mh = fmaxf(mh, __shfl_xor_sync(0xffffffffu, mh, d));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(absmax_bits, __float_as_uint(mg));
        atomicMax(absmax_bits + 1, __float_as_uint(mh));
    }
}

__device__ __forceinline__ int scale_exponent(float mx) {
    if (!(mx > 0.f)) return 0;
    int e;
    frexpf(mx, &e);          // mx = m * 2^e, m in [0.5,1)  =>  mx * 2^(30-e) < 2^30
    return 30 - e;
}

__device__ __forceinline__ int32_t to_fixed(float v, double mul) {
    long long q = __double2ll_rn((double)v * mul);
    if (q == 0 && v != 0.f) q = v > 0.f ? 1 : -1;
    return (int32_t)q;
}

__global__ void grad_quantize_k(const float* __restrict__ gp, int64_t n, int32_t K, int32_t gid,
                                int2* __restrict__ gq, const unsigned* __restrict__ absmax_bits,
                                GradScale* __restrict__ scale) {
    int kg = scale_exponent(__uint_as_float(absmax_bits[0]));
    int kh = scale_exponent(__uint_as_float(absmax_bits[1]));
    double mg = ldexp(1.0, kg), mh = ldexp(1.0, kh);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        scale->mul_g = mg; scale->mul_h = mh;
        scale->inv_g = ldexp(1.0, -kg); scale->inv_h = ldexp(1.0, -kh);
    }
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        float2 v = *reinterpret_cast<const float2*>(gp + (i * K + gid) * 2);
        gq[i] = make_int2(to_fixed(v.x, mg), to_fixed(v.y, mh));
    }
}

// ------------------------------------------------------------------------------------
// K2: per-node (g,h) histogram -- HistogramBuilder.buildHist's accumulation loop
//     (J/data/gbdt/HistogramBuilder.java:69-90).
//
// One CTA = 32 warps sharing one 128 KB shared histogram of a 32-feature group.  Lane l
// of every warp owns feature 32*group + l, and the histogram is laid out
// [bin][component][lane], so the 32 atomics of one warp instruction always hit 32
// different banks whatever the bins are: no bank conflicts by construction.  A warp
// walks its rows one at a time: the row's 32 bin bytes are one coalesced 32 B sector
// (uint8 row-major rows, 32 B per group), its fixed-point (g,h) is broadcast from a
// per-warp staging slot.  Each update is a native 32-bit ATOMS.ADD on the low word whose
// returned value yields the carry into the high word (a second, rare atomic).
// ------------------------------------------------------------------------------------
__device__ __forceinline__ void smem_add64(uint32_t* p, int32_t v) {
    uint32_t old = atomicAdd(p, (uint32_t)v);
    uint32_t sum = old + (uint32_t)v;
    int32_t hi = (v >> 31) + (sum < old ? 1 : 0);
    if (hi != 0) atomicAdd(reinterpret_cast<int32_t*>(p + kGroup), hi);
}

__global__ void __launch_bounds__(kHistThreads, 1)
hist_build_k(const uint8_t* __restrict__ bins, int32_t S8, const int2* __restrict__ gq,
             const int32_t* __restrict__ rowidx, int64_t begin, int64_t end, int32_t F,
             const uint32_t* __restrict__ group_mask, const int32_t* __restrict__ nbins,
             const int64_t* __restrict__ feat_off, long long* __restrict__ hist, int64_t rows_per_cta) {
    extern __shared__ __align__(16) uint32_t sm[];
    uint32_t* sh = sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int4* stage = reinterpret_cast<int4*>(sm + kMaxBins * 4 * kGroup) + warp * 32;
    const int fg = blockIdx.y;

    {
        uint4* z = reinterpret_cast<uint4*>(sh);
        for (int i = tid; i < kMaxBins * 4 * kGroup / 4; i += kHistThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();

    const int f = fg * kGroup + lane;
    const bool active = (f < F) && ((group_mask[fg] >> lane) & 1u);
    const uint8_t* col = bins + fg * kGroup + lane;
    const int64_t r0 = begin + blockIdx.x * rows_per_cta;
    const int64_t r1 = min(end, r0 + rows_per_cta);

    for (int64_t c = r0 + warp * 32; c < r1; c += kHistWarps * 32) {
        int64_t i = c + lane;
        int32_t rid = 0;
        int2 q = make_int2(0, 0);
        if (i < r1) {
            rid = rowidx ? rowidx[i] : (int32_t)i;
            q = gq[rid];
        }
        __syncwarp();
        stage[lane] = make_int4(rid, q.x, q.y, 0);
        __syncwarp();
        const int cnt = (int)min((int64_t)32, r1 - c);
        for (int j0 = 0; j0 < cnt; j0 += 8) {
            int4 s[8];
            uint32_t b[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                if (j0 + u < cnt) {
                    s[u] = stage[j0 + u];
                    b[u] = col[(int64_t)s[u].x * S8];
                }
            }
#pragma unroll
            for (int u = 0; u < 8; u++) {
                if (j0 + u < cnt && active) {
                    uint32_t* p = sh + (b[u] * 4) * kGroup + lane;
                    smem_add64(p, s[u].y);
                    smem_add64(p + 2 * kGroup, s[u].z);
                }
            }
        }
    }
    __syncthreads();

    // flush: one 64-bit global reduction per non-empty (feature, bin, component)
    const uint32_t mask = group_mask[fg];
    for (int i = tid; i < kMaxBins * kGroup; i += kHistThreads) {
        int bin = i >> 5, ln = i & 31;
        int ff = fg * kGroup + ln;
        if (ff >= F || !((mask >> ln) & 1u) || bin >= nbins[ff]) continue;
        const uint32_t* p = sh + (bin * 4) * kGroup + ln;
        long long g = (long long)(((unsigned long long)p[kGroup] << 32) | p[0]);
        long long h = (long long)(((unsigned long long)p[3 * kGroup] << 32) | p[2 * kGroup]);
        long long* dst = hist + feat_off[ff] + 2 * bin;
        if (g != 0) atomicAdd(reinterpret_cast<unsigned long long*>(dst), (unsigned long long)g);
        if (h != 0) atomicAdd(reinterpret_cast<unsigned long long*>(dst + 1), (unsigned long long)h);
    }
}

// K3: sibling = parent - smaller child (HistogramBuilder.substract,
//     J/data/gbdt/HistogramBuilder.java:101-107), exact in the integer domain.
__global__ void hist_subtract_k(const long long* __restrict__ parent, const long long* __restrict__ small,
                                long long* __restrict__ out, int64_t len) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < len;
         i += (int64_t)gridDim.x * blockDim.x)
        out[i] = parent[i] - small[i];
}

__global__ void hist_to_double_k(const long long* __restrict__ hist, int64_t len,
                                 const GradScale* __restrict__ scale, double* __restrict__ out) {
    double ig = scale->inv_g, ih = scale->inv_h;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < len;
         i += (int64_t)gridDim.x * blockDim.x)
        out[i] = (double)hist[i] * ((i & 1) ? ih : ig);
}

__global__ void hist_from_double_k(const double* __restrict__ in, int64_t len,
                                   const GradScale* __restrict__ scale, long long* __restrict__ hist) {
    double mg = scale->mul_g, mh = scale->mul_h;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < len;
         i += (int64_t)gridDim.x * blockDim.x)
        hist[i] = __double2ll_rn(in[i] * ((i & 1) ? mh : mg));
}

// ------------------------------------------------------------------------------------
// K4/K5: node totals, split scan and arg-max.
//   updateTreeMakerNodeStats (J/optimizer/gbdt/DataParallelTreeMaker.java:543-573),
//   enumerateSplit (:598-637), SplitInfo.needReplace (J/data/gbdt/SplitInfo.java:99-104),
//   UpdateStrategy.calcGain / calcNodeValue (J/optimizer/gbdt/UpdateStrategy.java:55-100).
// One warp per feature: a segmented inclusive scan over the feature's bins in exact
// 64-bit integers (so the prefix does not depend on scan order), gains in IEEE double
// with explicit _rn intrinsics (never fused), loss change rounded to float as the
// reference does.
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

constexpr int kScanWarps = 8;

__global__ void __launch_bounds__(kScanWarps * 32)
split_scan_k(const long long* __restrict__ hist, const int32_t* __restrict__ start,
             const int64_t* __restrict__ feat_off, const uint8_t* __restrict__ fea_selected,
             int32_t f_from, int32_t f_to, int32_t first_fea, const GradScale* __restrict__ scale,
             GainParams gp, long long sample_cnt, const double* __restrict__ totals_in,
             ytk_split* __restrict__ out) {
    __shared__ double s_tot[2];
    __shared__ float s_lc[kScanWarps];
    __shared__ int s_fid[kScanWarps], s_lo[kScanWarps], s_hi[kScanWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const double ig = scale->inv_g, ih = scale->inv_h;

    if (warp == 0) {
        double G, H;
        if (totals_in) {
            G = totals_in[0]; H = totals_in[1];
        } else if (first_fea >= 0) {
            long long g = 0, h = 0;
            const long long* hp = hist + feat_off[first_fea];
            int nb = start[first_fea + 1] - start[first_fea];
            for (int j = lane; j < nb; j += 32) { g += hp[2 * j]; h += hp[2 * j + 1]; }
            for (int d = 16; d > 0; d >>= 1) {
                g += __shfl_xor_sync(0xffffffffu, g, d);
                h += __shfl_xor_sync(0xffffffffu, h, d);
            }
            G = (double)g * ig; H = (double)h * ih;
        } else {
            G = -CUDART_INF; H = -CUDART_INF;   // no feature here: neutral for the MAX allreduce (:549)
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
    if (can_split) {
        for (int f = f_from + warp; f < f_to; f += kScanWarps) {
            if (!fea_selected[f]) continue;
            const long long* hp = hist + feat_off[f];
            const int nb = start[f + 1] - start[f];
            long long carryG = 0, carryH = 0;
            int carry_last = -1;
            float b_lc = -CUDART_INF_F;
            int b_lo = -1, b_hi = 0x7fffffff;
            for (int c0 = 0; c0 < nb; c0 += 32) {
                int j = c0 + lane;
                long long g = 0, h = 0;
                if (j < nb) { g = hp[2 * j]; h = hp[2 * j + 1]; }
                bool ne = (g | h) != 0;                                   // hasSample
                long long pg = g, ph = h;
                for (int d = 1; d < 32; d <<= 1) {
                    long long tg = __shfl_up_sync(0xffffffffu, pg, d);
                    long long th = __shfl_up_sync(0xffffffffu, ph, d);
                    if (lane >= d) { pg += tg; ph += th; }
                }
                long long exG = carryG + pg - g, exH = carryH + ph - h;   // left branch before bin j
                unsigned m = __ballot_sync(0xffffffffu, ne);
                unsigned lower = m & ((1u << lane) - 1u);
                int prev = lower ? (c0 + 31 - __clz(lower)) : carry_last; // lastFeaValue
                if (ne && exH != 0) {                                      // !leftBranchGradStats.empty()
                    double lG = (double)exG * ig, lH = (double)exH * ih;
                    if (lH >= gp.min_child_weight) {
                        double rG = __dsub_rn(totG, lG), rH = __dsub_rn(totH, lH);
                        if (rH >= gp.min_child_weight) {
                            float lc = (float)__dsub_rn(__dadd_rn(calc_gain(gp, lG, lH), calc_gain(gp, rG, rH)),
                                                        (double)root_gain);
                            if (lc > b_lc) { b_lc = lc; b_lo = prev; b_hi = j; }
                        }
                    }
                }
                carryG += __shfl_sync(0xffffffffu, pg, 31);
                carryH += __shfl_sync(0xffffffffu, ph, 31);
                if (m) carry_last = c0 + 31 - __clz(m);
            }
            // first (lowest bin) maximum inside the feature
            for (int d = 16; d > 0; d >>= 1) {
                float o_lc = __shfl_xor_sync(0xffffffffu, b_lc, d);
                int o_lo = __shfl_xor_sync(0xffffffffu, b_lo, d);
                int o_hi = __shfl_xor_sync(0xffffffffu, b_hi, d);
                if (o_lc > b_lc || (o_lc == b_lc && o_hi < b_hi)) { b_lc = o_lc; b_lo = o_lo; b_hi = o_hi; }
            }
            // features are visited in increasing fid by this warp: strict > keeps the smaller fid
            if (b_lc > w_lc) { w_lc = b_lc; w_fid = f; w_lo = b_lo; w_hi = b_hi; }
        }
    }
    if (lane == 0) { s_lc[warp] = w_lc; s_fid[warp] = w_fid; s_lo[warp] = w_lo; s_hi[warp] = w_hi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float lc = -CUDART_INF_F;
        int fid = -1, lo = -1, hi = -1;
        for (int w = 0; w < kScanWarps; w++) {
            if (s_fid[w] < 0) continue;
            if (s_lc[w] > lc || (s_lc[w] == lc && s_fid[w] < fid)) {
                lc = s_lc[w]; fid = s_fid[w]; lo = s_lo[w]; hi = s_hi[w];
            }
        }
        out->loss_chg = lc; out->fid = fid; out->lo = lo; out->hi = hi;
        out->sum_grad = totG; out->sum_hess = totH;
        out->root_gain = root_gain;
        out->node_value = (float)calc_node_value(gp, totG, totH);
    }
}

// ------------------------------------------------------------------------------------
// K6: row partition.  SamplePositionData.resetPosition
//     (J/data/gbdt/SamplePositionData.java:115-133): row goes left iff
//     (float)slot(row, fid) < cond.  The reference then swaps in place (unstable); only
//     the left/right sets are contractual, so this is a stable out-of-place scatter:
//     flags + per-block counts, a scan of the block counts, then the scatter.
// ------------------------------------------------------------------------------------
constexpr int kPartThreads = 256;
constexpr int kPartItems = 8;
constexpr int kPartTile = kPartThreads * kPartItems;

template <int MODE>  // 0: bin compare, 1: byte mask (row sampling)
__global__ void __launch_bounds__(kPartThreads)
partition_count_k(const uint8_t* __restrict__ src, int32_t S8, const int32_t* __restrict__ rowidx,
                  int64_t begin, int64_t end, int32_t fid, float cond, uint8_t* __restrict__ flags,
                  int32_t* __restrict__ block_cnt) {
    __shared__ int s_cnt[kPartThreads / 32];
    int64_t base = begin + (int64_t)blockIdx.x * kPartTile;
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < kPartItems; k++) {
        int64_t i = base + k * kPartThreads + threadIdx.x;
        if (i < end) {
            int32_t rid = rowidx ? rowidx[i] : (int32_t)i;
            bool left;
            if (MODE == 0) left = (float)src[(int64_t)rid * S8 + fid] < cond;
            else left = src[rid] != 0;
            flags[i - begin] = left ? 1 : 0;
            cnt += left ? 1 : 0;
        }
    }
    for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
    if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < kPartThreads / 32; w++) t += s_cnt[w];
        block_cnt[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(1024)
partition_scan_k(int32_t* __restrict__ block_cnt, int32_t nblocks, int64_t n, PartitionCounts* counts) {
    __shared__ long long s_warp[32];
    __shared__ long long s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int c0 = 0; c0 < nblocks; c0 += 1024) {
        int i = c0 + threadIdx.x;
        long long v = (i < nblocks) ? block_cnt[i] : 0;
        long long p = v;
        for (int d = 1; d < 32; d <<= 1) {
            long long t = __shfl_up_sync(0xffffffffu, p, d);
            if (lane >= d) p += t;
        }
        if (lane == 31) s_warp[warp] = p;
        __syncthreads();
        if (warp == 0) {
            long long w = s_warp[lane];
            long long q = w;
            for (int d = 1; d < 32; d <<= 1) {
                long long t = __shfl_up_sync(0xffffffffu, q, d);
                if (lane >= d) q += t;
            }
            s_warp[lane] = q - w;  // exclusive
        }
        __syncthreads();
        long long excl = s_carry + s_warp[warp] + p - v;
        if (i < nblocks) block_cnt[i] = (int32_t)excl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) { counts->left = s_carry; counts->right = n - s_carry; }
}

__global__ void __launch_bounds__(kPartThreads)
partition_scatter_k(const int32_t* __restrict__ rowidx, int32_t* __restrict__ out, int64_t begin, int64_t end,
                    const uint8_t* __restrict__ flags, const int32_t* __restrict__ block_off,
                    const PartitionCounts* __restrict__ counts) {
    __shared__ int s_warp[kPartThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // thread t owns items [t*8, t*8+8) of the tile so that order is preserved
    int64_t base = begin + (int64_t)blockIdx.x * kPartTile + threadIdx.x * kPartItems;
    int32_t rid[kPartItems];
    bool fl[kPartItems];
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < kPartItems; k++) {
        int64_t i = base + k;
        fl[k] = false;
        if (i < end) {
            rid[k] = rowidx ? rowidx[i] : (int32_t)i;
            fl[k] = flags[i - begin] != 0;
            cnt += fl[k] ? 1 : 0;
        }
    }
    int p = cnt;
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, p, d);
        if (lane >= d) p += t;
    }
    if (lane == 31) s_warp[warp] = p;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < warp; w++) woff += s_warp[w];
    int64_t lefts_before = (int64_t)block_off[blockIdx.x] + woff + p - cnt;  // lefts before my first item
    const int64_t left_total = counts->left;
    int64_t lpos = begin + lefts_before;
    int64_t rpos = begin + left_total + ((base - begin) - lefts_before);
#pragma unroll
    for (int k = 0; k < kPartItems; k++) {
        int64_t i = base + k;
        if (i < end) {
            if (fl[k]) out[lpos++] = rid[k];
            else out[rpos++] = rid[k];
        }
    }
}

// K7: exact sums of the fixed-point (g,h) over a node's rows -- computeLeafStats
//     (J/optimizer/gbdt/DataParallelTreeMaker.java:352-372).
__global__ void leaf_sums_k(const int2* __restrict__ gq, const int32_t* __restrict__ rowidx, int64_t begin,
                            int64_t end, long long* __restrict__ out) {
    long long g = 0, h = 0;
    for (int64_t i = begin + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < end;
         i += (int64_t)gridDim.x * blockDim.x) {
        int2 q = gq[rowidx ? rowidx[i] : (int32_t)i];
        g += q.x; h += q.y;
    }
    for (int d = 16; d > 0; d >>= 1) {
        g += __shfl_xor_sync(0xffffffffu, g, d);
        h += __shfl_xor_sync(0xffffffffu, h, d);
    }
    if ((threadIdx.x & 31) == 0) {
        if (g) atomicAdd(reinterpret_cast<unsigned long long*>(out), (unsigned long long)g);
        if (h) atomicAdd(reinterpret_cast<unsigned long long*>(out + 1), (unsigned long long)h);
    }
}

__global__ void iota_k(int32_t* p, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = (int32_t)i;
}
__global__ void fill_i32_k(int32_t* p, int64_t n, int32_t v) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = v;
}
__global__ void fill_positions_k(const int32_t* __restrict__ rowidx, int64_t begin, int64_t end, int32_t nid,
                                 int32_t* __restrict__ node_of_row) {
    for (int64_t i = begin + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < end;
         i += (int64_t)gridDim.x * blockDim.x)
        node_of_row[rowidx ? rowidx[i] : (int32_t)i] = nid;
}

// ------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------
static inline int grid_for(int64_t n, int threads, int max_blocks = kNumSMs * 8) {
    int64_t b = (n + threads - 1) / threads;
    if (b < 1) b = 1;
    if (b > max_blocks) b = max_blocks;
    return (int)b;
}

cudaError_t launch_repack_rows(const int32_t* x, int64_t nrows, int32_t stride, int32_t F, uint8_t* bins,
                               int32_t S8, int* bad_flag, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    repack_rows_k<<<grid_for(nrows * S8, 256), 256, 0, st>>>(x, nrows, stride, F, bins, S8, bad_flag);
    return cudaGetLastError();
}
cudaError_t launch_quantize_rows(const int32_t* xbits, int64_t nrows, int32_t stride, int32_t F,
                                 const float* cuts, const int32_t* start, uint8_t* bins, int32_t S8,
                                 cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    quantize_rows_k<<<grid_for(nrows * S8, 256), 256, 0, st>>>(xbits, nrows, stride, F, cuts, start, bins, S8);
    return cudaGetLastError();
}
cudaError_t launch_unpack_rows(const uint8_t* bins, int32_t S8, int64_t nrows, int32_t F, int32_t* out,
                               cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    unpack_rows_k<<<grid_for(nrows * F, 256), 256, 0, st>>>(bins, S8, nrows, F, out);
    return cudaGetLastError();
}

cudaError_t launch_grad_quantize(const float* gp, int64_t n, int32_t K, int32_t gid, int2* gq,
                                 unsigned* absmax_bits, GradScale* scale, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(absmax_bits, 0, 2 * sizeof(unsigned), st);
    if (e != cudaSuccess) return e;
    int g = grid_for(n, 256, kNumSMs * 4);
    grad_absmax_k<<<g, 256, 0, st>>>(gp, n, K, gid, absmax_bits);
    grad_quantize_k<<<g, 256, 0, st>>>(gp, n, K, gid, gq, absmax_bits, scale);
    return cudaGetLastError();
}

cudaError_t launch_hist_build(const uint8_t* bins, int32_t S8, const int2* gq, const int32_t* rowidx,
                              int64_t begin, int64_t end, int32_t F, const uint32_t* group_mask,
                              const int32_t* nbins, const int64_t* feat_off, long long* hist, int variant,
                              cudaStream_t st, int* launches) {
    (void)variant;
    int64_t n = end - begin;
    if (n <= 0) return cudaSuccess;
    static bool attr_set = false;
    const int smem = kHistSmemBytes + kStageBytes;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(hist_build_k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    const int groups = S8 / kGroup;
    const int64_t chunk = kHistWarps * 32;  // rows one CTA consumes per sweep
    // at least 4 sweeps per CTA before another CTA is worth its 128 KB zero + flush
    int64_t gx = (n + 4 * chunk - 1) / (4 * chunk);
    int64_t cap = kNumSMs;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    int64_t rows_per_cta = ((n + gx - 1) / gx + chunk - 1) / chunk * chunk;
    gx = (n + rows_per_cta - 1) / rows_per_cta;
    dim3 grid((unsigned)gx, (unsigned)groups);
    hist_build_k<<<grid, kHistThreads, smem, st>>>(bins, S8, gq, rowidx, begin, end, F, group_mask, nbins,
                                                   feat_off, hist, rows_per_cta);
    if (launches) (*launches)++;
    return cudaGetLastError();
}

cudaError_t launch_hist_subtract(const long long* parent, const long long* small, long long* out, int64_t len,
                                 cudaStream_t st) {
    hist_subtract_k<<<grid_for(len, 256), 256, 0, st>>>(parent, small, out, len);
    return cudaGetLastError();
}
cudaError_t launch_hist_to_double(const long long* hist, int64_t len, const GradScale* scale, double* out,
                                  cudaStream_t st) {
    hist_to_double_k<<<grid_for(len, 256), 256, 0, st>>>(hist, len, scale, out);
    return cudaGetLastError();
}
cudaError_t launch_hist_from_double(const double* in, int64_t len, const GradScale* scale, long long* hist,
                                    cudaStream_t st) {
    hist_from_double_k<<<grid_for(len, 256), 256, 0, st>>>(in, len, scale, hist);
    return cudaGetLastError();
}

cudaError_t launch_split_scan(const long long* hist, const int32_t* start, const int64_t* feat_off,
                              const uint8_t* fea_selected, int32_t f_from, int32_t f_to, int32_t first_fea,
                              const GradScale* scale, GainParams gp, long long sample_cnt,
                              const double* totals_in, ytk_split* out, cudaStream_t st) {
    split_scan_k<<<1, kScanWarps * 32, 0, st>>>(hist, start, feat_off, fea_selected, f_from, f_to, first_fea,
                                                scale, gp, sample_cnt, totals_in, out);
    return cudaGetLastError();
}

int64_t partition_scratch_ints(int64_t n) {
    int64_t nblocks = (n + kPartTile - 1) / kPartTile;
    return nblocks + 1 + (n + 3) / 4;  // block counts, then one flag byte per row
}

template <int MODE>
static cudaError_t partition_impl(const uint8_t* src, int32_t S8, const int32_t* rowidx_in, int32_t* rowidx_out,
                                  int64_t begin, int64_t end, int32_t fid, float cond, int32_t* scratch,
                                  PartitionCounts* counts, cudaStream_t st, int* launches) {
    int64_t n = end - begin;
    if (n <= 0) {
        cudaError_t e = cudaMemsetAsync(counts, 0, sizeof(PartitionCounts), st);
        return e;
    }
    int nblocks = (int)((n + kPartTile - 1) / kPartTile);
    int32_t* block_cnt = scratch;
    uint8_t* flags = reinterpret_cast<uint8_t*>(scratch + nblocks + 1);
    partition_count_k<MODE><<<nblocks, kPartThreads, 0, st>>>(src, S8, rowidx_in, begin, end, fid, cond, flags,
                                                              block_cnt);
    partition_scan_k<<<1, 1024, 0, st>>>(block_cnt, nblocks, n, counts);
    partition_scatter_k<<<nblocks, kPartThreads, 0, st>>>(rowidx_in, rowidx_out, begin, end, flags, block_cnt,
                                                          counts);
    if (launches) *launches += 3;
    return cudaGetLastError();
}

cudaError_t launch_partition(const uint8_t* bins, int32_t S8, const int32_t* rowidx_in, int32_t* rowidx_out,
                             int64_t begin, int64_t end, int32_t fid, float cond, int32_t* scratch,
                             PartitionCounts* counts, cudaStream_t st, int* launches) {
    return partition_impl<0>(bins, S8, rowidx_in, rowidx_out, begin, end, fid, cond, scratch, counts, st,
                             launches);
}

cudaError_t launch_compact_mask(const uint8_t* mask, int64_t n, int32_t* rowidx_out, int32_t* scratch,
                                PartitionCounts* counts, cudaStream_t st, int* launches) {
    return partition_impl<1>(mask, 0, nullptr, rowidx_out, 0, n, 0, 0.f, scratch, counts, st, launches);
}

cudaError_t launch_leaf_sums(const int2* gq, const int32_t* rowidx, int64_t begin, int64_t end, long long* out,
                             cudaStream_t st) {
    if (end <= begin) return cudaSuccess;
    leaf_sums_k<<<grid_for(end - begin, 256, kNumSMs * 4), 256, 0, st>>>(gq, rowidx, begin, end, out);
    return cudaGetLastError();
}

cudaError_t launch_iota(int32_t* p, int64_t n, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    iota_k<<<grid_for(n, 256), 256, 0, st>>>(p, n);
    return cudaGetLastError();
}
cudaError_t launch_fill_i32(int32_t* p, int64_t n, int32_t v, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    fill_i32_k<<<grid_for(n, 256), 256, 0, st>>>(p, n, v);
    return cudaGetLastError();
}
cudaError_t launch_fill_positions(const int32_t* rowidx, int64_t begin, int64_t end, int32_t nid,
                                  int32_t* node_of_row, cudaStream_t st) {
    if (end <= begin) return cudaSuccess;
    fill_positions_k<<<grid_for(end - begin, 256), 256, 0, st>>>(rowidx, begin, end, nid, node_of_row);
    return cudaGetLastError();
}

}  // namespace ytk
