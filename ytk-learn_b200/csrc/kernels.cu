// CUDA kernels of the GBDT hot path (sm_100a).  Nothing here is a dense contraction:
// the work is HBM-/shared-memory-bound byte and integer traffic, so there are no tensor
// cores; what matters is sector-aligned row reads, conflict-free shared-memory atomics
// and grids sized to the 148 SMs.
//
// Reference semantics restated by each kernel are cited as J/... =
// /root/reference/src/main/java/com/fenbi/ytklearn/...
#include "common.cuh"

#include <math_constants.h>

#include <algorithm>
#include <cstdlib>

namespace ytk {

// Programmatic dependent launch for the chain of short kernels that ends every level (zero ->
// histogram -> subtract -> scan -> merge -> plan -> partition): a kernel launched with the attribute
// may start being scheduled while its predecessor in the stream drains; pdl_sync() -- the first
// statement of each of them -- lets the dependents go and then waits until the predecessor has
// completed and its writes are visible, so the kernels see exactly the serial order.
__device__ __forceinline__ void pdl_sync() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
static bool g_pdl = [] { const char* e = getenv("YTK_GBDT_NO_PDL"); return !(e && atoi(e) != 0); }();
template <typename... KArgs, typename... Args>
static cudaError_t launch_chain(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = g_pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ------------------------------------------------------------------------------------
// K0: int32 slot matrix (GBDTCoreData.x chunk, J/dataflow/GBDTCoreData.java:164-168)
//     -> uint8 row-major, rows padded to a multiple of 32 features (one 32 B sector per
//     feature group and row).
// ------------------------------------------------------------------------------------
__global__ void repack_rows_k(const int32_t* __restrict__ x, int64_t nrows, int32_t stride, int32_t F,
                              const int32_t* __restrict__ nbins, uint8_t* __restrict__ bins, int32_t S8,
                              uint8_t* __restrict__ col, int64_t col_stride, WideCols wide, int64_t row_base,
                              int* bad_flag) {
    int64_t total = nrows * S8;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i / S8;
        int32_t f = (int32_t)(i - r * S8);
        // padding columns get row-varying filler: the histogram kernel accumulates them too (and drops
        // them at flush), and identical bins in consecutive rows would make consecutive atomics hit
        // the same shared-memory word
        uint8_t b = (uint8_t)((r + row_base) * 37 + f * 11);
        if (f < F) {
            int32_t v = x[r * stride + f];
            if (v < 0 || v >= nbins[f]) { *bad_flag = 1; v = 0; }
            const int wi = wide.index ? wide.index[f] : -1;
            if (wi >= 0) {
                wide.data[wi * wide.stride + r] = (uint16_t)v;     // its byte in the uint8 matrices stays filler
            } else {
                b = (uint8_t)v;
                col[f * col_stride + r] = b;
            }
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
                                uint8_t* __restrict__ bins, int32_t S8, uint8_t* __restrict__ col, int64_t col_stride,
                                WideCols wide, int64_t row_base) {
    int64_t total = nrows * S8;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i / S8;
        int32_t f = (int32_t)(i - r * S8);
        uint8_t b = (uint8_t)((r + row_base) * 37 + f * 11);     // see repack_rows_k
        if (f < F) {
            float v = __int_as_float(xbits[r * stride + f]);
            int32_t s0 = start[f];
            const int32_t slot = find_bin(cuts + s0, start[f + 1] - s0, v);
            const int wi = wide.index ? wide.index[f] : -1;
            if (wi >= 0) {
                wide.data[wi * wide.stride + r] = (uint16_t)slot;
            } else {
                b = (uint8_t)slot;
                col[f * col_stride + r] = b;
            }
        }
        bins[i] = b;
    }
}

__global__ void unpack_rows_k(const uint8_t* __restrict__ bins, int32_t S8, int64_t nrows, int32_t F, WideCols wide,
                              int32_t* __restrict__ out) {
    int64_t total = nrows * F;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i / F;
        int32_t f = (int32_t)(i - r * F);
        const int wi = wide.index ? wide.index[f] : -1;
        out[i] = wi >= 0 ? (int32_t)wide.data[wi * wide.stride + r] : (int32_t)bins[r * S8 + f];
    }
}

// ------------------------------------------------------------------------------------
// Gradient fixed point.  The reference adds float (g,h) into double bins
// (J/data/gbdt/HistogramBuilder.java:87-88).  Shared memory only has native 32-bit
// integer atomics (a float or 64-bit atomicAdd is a CAS loop, measured 2.9x slower:
// profiles/r01_microbench_smem.txt), so each tree's gradients are scaled by a power of
// two chosen from max|g| (resp. max|h|) to signed integers below 2^bits and every sum
// downstream is an exact 64-bit integer: histograms do not depend on row order, sharding
// or pool hits, and parent - child equals the sibling bit for bit.  Per-row rounding is
// at most 2^-(bits+1) of max|g| (bits = 30 by default).
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

__device__ __forceinline__ int scale_exponent(float mx, int bits) {
    if (!(mx > 0.f)) return 0;
    int e;
    frexpf(mx, &e);          // mx = m * 2^e, m in [0.5,1)  =>  mx * 2^(bits-e) < 2^bits
    return bits - e;
}

__device__ __forceinline__ int32_t to_fixed(float v, double mul) {
    long long q = __double2ll_rn((double)v * mul);
    if (q == 0 && v != 0.f) q = v > 0.f ? 1 : -1;
    return (int32_t)q;
}

__global__ void grad_quantize_k(const float* __restrict__ gp, int64_t n, int32_t K, int32_t gid,
                                int2* __restrict__ gq, const unsigned* __restrict__ absmax_bits,
                                GradScale* __restrict__ scale, int bits) {
    int kg = scale_exponent(__uint_as_float(absmax_bits[0]), bits);
    int kh = scale_exponent(__uint_as_float(absmax_bits[1]), bits);
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
// What bounds this kernel on B200 is not HBM but the shared-memory atomic pipe
// (profiles/r01_microbench_smem.txt), so it is built to issue little else:
//   * rows are uint8, row-major, one 32 B sector per 32-feature group, so a row reached
//     through the row-id list costs exactly one sector; a warp fetches 32 rows with two
//     128-bit loads, adjacent lanes taking the two 16-feature halves of a row, plus that
//     row's fixed-point (g,h) -- no staging, no per-row broadcast;
//   * lane L works on row L>>1 and half L&1: at step k (0..15) it updates feature
//     16*(L&1) + (k ^ (L>>1)).  The 32 lanes therefore always touch 32 different
//     features, and the shared histogram is laid out
//     [bin][component][feature], i.e. bank = feature: the 32 atomics of a warp
//     instruction hit 32 different banks whatever the bins are (skewed or binary
//     features cannot serialise it); the address is one PRMT (bin*256 from the data
//     word) and one LOP3 ((bin256 | lane*4) ^ k*4);
//   * exact 64-bit sums WITHOUT carries: each fixed-point value (|v| < 2^30) is split into
//     its low 16 bits (unsigned) and its high part (signed, |.| <= 2^14), and the two
//     limbs are accumulated in separate 32-bit words by fire-and-forget atomics
//     (RED).  A work item covers at most 65536 rows, so neither limb sum can leave 32
//     bits; the flush recombines hi*65536 + lo in 64-bit.  Detecting carries of a single
//     64-bit accumulator instead costs more than the atomics themselves (7.5 vs 4.6
//     SM-cycles per row, same file).
// One CTA (32 warps, 128 KB of shared memory, 1 per SM) handles one work item: a run of
// rows of one node and one feature group; it flushes non-empty bins with 64-bit global
// reductions into that node's histogram.
// ------------------------------------------------------------------------------------
// smaller child = strictly fewer GLOBAL rows on the left, else the right
// (J/optimizer/gbdt/DataParallelTreeMaker.java:344-350, :493-499)
__device__ __forceinline__ RowItem resolve_item(RowItem it, const PartitionCounts* __restrict__ cnt_local,
                                                const PartitionCounts* __restrict__ cnt_global) {
    if (it.cseg < 0) return it;
    const PartitionCounts l = cnt_local[it.cseg], g = cnt_global[it.cseg];
    const bool small_left = g.left < g.right;
    const long long cb = small_left ? it.begin : it.begin + l.left;
    const long long ce = small_left ? it.begin + l.left : it.end;
    const long long b = cb + it.chunk * it.rows_per_chunk;
    it.begin = min(b, ce);
    it.end = min(ce, b + it.rows_per_chunk);
    return it;
}

__global__ void __launch_bounds__(kHistThreads, 1)
hist_build_k(const uint8_t* __restrict__ bins, int32_t S8, const int2* __restrict__ gq,
             const int32_t* __restrict__ rowidx0, const int32_t* __restrict__ rowidx1,
             const RowItem* __restrict__ items, int32_t F, const uint32_t* __restrict__ group_mask,
             const int32_t* __restrict__ nbins, const int32_t* __restrict__ feat_owner,
             const int64_t* __restrict__ feat_inner, long long owner_stride,
             long long* const* __restrict__ seg_out, const PartitionCounts* __restrict__ cnt_local,
             const PartitionCounts* __restrict__ cnt_global) {
    pdl_sync();
    // g limbs [bin][{lo,hi}][feature] (64 KB), then the h limbs in the same layout (64 KB)
    extern __shared__ __align__(128) uint32_t sh[];
    constexpr uint32_t kHOff = kMaxBins * 2 * kGroup * 4;        // bytes
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int fg = blockIdx.y;
    const RowItem it = resolve_item(items[blockIdx.x], cnt_local, cnt_global);
    if (it.begin >= it.end) return;          // a deferred chunk beyond the end of its child
    const int32_t* __restrict__ ridx = it.buf < 0 ? nullptr : (it.buf == 0 ? rowidx0 : rowidx1);

    {
        uint4* z = reinterpret_cast<uint4*>(sh);
        for (int i = tid; i < kMaxBins * 4 * kGroup / 4; i += kHistThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();

    const uint32_t gmask = group_mask[fg];
    // Lane L works on row (L >> 1) of a 16-row block and on the 16 features of half (L & 1) of the
    // group: at step k (0..15) it updates feature 16*(L&1) + (k ^ (L>>1)).  The 32 lanes of a warp
    // instruction therefore always touch 32 different features = 32 different banks.
    const int half = lane & 1, r16 = lane >> 1;
    const int m = r16 >> 2, bsel = r16 & 3;
    // PRMT selectors: byte (kk ^ bsel) of a data word moved to byte 1, i.e. bin * 256 = the byte
    // offset of that bin's 256 B row of limbs
    uint32_t selk[4];
#pragma unroll
    for (int kk = 0; kk < 4; kk++) selk[kk] = 0x4404u | ((uint32_t)(kk ^ bsel) << 4);
    // byte offset of feature 16*half + r16 inside a bin row; feature 16*half + (k ^ r16) is reached
    // by XOR-ing k*4 into it: (bin256 | lane_off) ^ k4 is one LOP3
    uint32_t lane_off = (uint32_t)(half * 16 + r16) << 2;
    asm volatile("" : "+r"(lane_off));
    const uint8_t* __restrict__ gbase = bins + fg * kGroup + (half << 4);
    char* const shb = reinterpret_cast<char*>(sh);

    // Three-stage software pipeline per warp: row ids of block t+2, row data of block t+1 and the
    // atomics of block t are in flight together, so no load waits on another load of the same
    // iteration (loads share the LSU queue with thousands of cycles of queued atomics).
    // A block is 32 rows = two 128-bit loads: adjacent lanes fetch the two halves of a row, so every
    // 32 B sector is requested once; for contiguous rows the warp reads 1 KB in 8 fully used lines.
    struct Fetch { uint4 a, b; int2 qa, qb; };
    auto fetch_rid = [&](long long base) -> int32_t {
        const long long i = base + lane;
        return i < it.end ? (ridx ? ridx[i] : (int32_t)i) : -1;
    };
    auto fetch = [&](int32_t rid, Fetch& f) {
        const int32_t ra = __shfl_sync(0xffffffffu, rid, r16);
        const int32_t rb = __shfl_sync(0xffffffffu, rid, 16 + r16);
        // past the end of the run: zeros, i.e. "add zero to bin 0", which changes nothing
        f.a = make_uint4(0, 0, 0, 0); f.b = f.a;
        f.qa = make_int2(0, 0); f.qb = f.qa;
        if (ra >= 0) { f.a = *reinterpret_cast<const uint4*>(gbase + (long long)ra * S8); f.qa = gq[ra]; }
        if (rb >= 0) { f.b = *reinterpret_cast<const uint4*>(gbase + (long long)rb * S8); f.qb = gq[rb]; }
    };

    // 16 steps on one 16-row half block.  Register j must hold data word j ^ m (m = r16 >> 2), the
    // word with the features this lane visits at steps 4j..4j+3: two conditional swaps.
    // Unsampled features and padding lanes are accumulated too (the instruction is issued for the
    // whole warp anyway) and dropped at flush time.
    auto accumulate16 = [&](const uint4 d, const int2 q) {
        uint32_t w[4] = {d.x, d.y, d.z, d.w};
        {
            const bool c0 = m & 1, c1 = m & 2;
            uint32_t t;
            t = c0 ? w[1] : w[0]; w[1] = c0 ? w[0] : w[1]; w[0] = t;
            t = c0 ? w[3] : w[2]; w[3] = c0 ? w[2] : w[3]; w[2] = t;
            t = c1 ? w[2] : w[0]; w[2] = c1 ? w[0] : w[2]; w[0] = t;
            t = c1 ? w[3] : w[1]; w[3] = c1 ? w[1] : w[3]; w[1] = t;
        }
        const uint32_t g_lo = (uint32_t)q.x & 0xffffu, h_lo = (uint32_t)q.y & 0xffffu;
        const int32_t g_hi = q.x >> 16, h_hi = q.y >> 16;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t word = w[j];
#pragma unroll
            for (int kk = 0; kk < 4; kk++) {
                const uint32_t off = (__byte_perm(word, 0u, selk[kk]) | lane_off) ^ (uint32_t)((4 * j + kk) << 2);
                uint32_t* p = reinterpret_cast<uint32_t*>(shb + off);
                atomicAdd(p, g_lo);
                atomicAdd(reinterpret_cast<int32_t*>(p + kGroup), g_hi);
                atomicAdd(p + kHOff / 4, h_lo);
                atomicAdd(reinterpret_cast<int32_t*>(p + kHOff / 4 + kGroup), h_hi);
            }
        }
    };

    // (two register sets in explicit ping-pong: a copy `cur = nxt` would be scheduled right behind
    // the loads and stall the warp on them)
    long long base = it.begin + warp * 32;
    Fetch fa, fb;
    int32_t r1 = fetch_rid(base + kHistThreads);
    fetch(fetch_rid(base), fa);
    while (base < it.end) {
        const int32_t r2 = fetch_rid(base + 2 * kHistThreads);
        fetch(r1, fb);
        accumulate16(fa.a, fa.qa);
        accumulate16(fa.b, fa.qb);
        base += kHistThreads;
        if (base >= it.end) break;
        r1 = fetch_rid(base + 2 * kHistThreads);
        fetch(r2, fa);
        accumulate16(fb.a, fb.qa);
        accumulate16(fb.b, fb.qb);
        base += kHistThreads;
    }
    __syncthreads();

    // flush: recombine the limbs, one 64-bit global reduction per non-empty (feature, bin, component)
    long long* __restrict__ out = seg_out[it.seg];
    for (int i = tid; i < kMaxBins * kGroup; i += kHistThreads) {
        int bin = i >> 5, ln = i & 31;
        int ff = fg * kGroup + ln;
        if (ff >= F || !((gmask >> ln) & 1u) || bin >= nbins[ff]) continue;
        const uint32_t* pg = sh + bin * (2 * kGroup) + ln;
        const uint32_t* ph = pg + kHOff / 4;
        long long g = (long long)(int32_t)pg[kGroup] * 65536ll + (long long)pg[0];
        long long h = (long long)(int32_t)ph[kGroup] * 65536ll + (long long)ph[0];
        long long* dst = out + feat_owner[ff] * owner_stride + feat_inner[ff] + 2 * bin;
        if (g != 0) atomicAdd(reinterpret_cast<unsigned long long*>(dst), (unsigned long long)g);
        if (h != 0) atomicAdd(reinterpret_cast<unsigned long long*>(dst + 1), (unsigned long long)h);
    }
}

// ------------------------------------------------------------------------------------
// Sparse rows (BASELINE config 4: 10 000 features at ~1 % density).  The reference's matrix is dense
// (J/dataflow/GBDTCoreData.java:164-168) and cannot hold that shape; here the rows are CSR, an
// absent feature has its default slot, and the histogram is what the dense path would produce.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ int sparse_slot(const SparseRows& sp, int32_t rid, int fid) {
    long long lo = sp.row_ptr[rid];
    const long long end = sp.row_ptr[rid + 1];
    long long hi = end;
    const unsigned key = (unsigned)fid;
    while (lo < hi) {                                    // entries are sorted by feature
        const long long mid = (lo + hi) >> 1;
        if ((sp.nz[mid] >> 16) < key) lo = mid + 1; else hi = mid;
    }
    if (lo < end) {
        const unsigned e = sp.nz[lo];
        if ((e >> 16) == key) return (int)(e & 0xffffu);
    }
    return sp.default_slot[fid];
}

// One CTA per work item (same items as hist_build_k), one warp per row at a time, lanes over the
// row's stored entries.  Thousands of bins per node: global 64-bit reductions, little contention.
__global__ void __launch_bounds__(256)
hist_csr_k(SparseRows sp, const uint8_t* __restrict__ fea_selected, const int2* __restrict__ gq,
           const int32_t* __restrict__ rowidx0, const int32_t* __restrict__ rowidx1, const RowItem* __restrict__ items,
           const int32_t* __restrict__ feat_owner, const int64_t* __restrict__ feat_inner, long long owner_stride,
           long long* const* __restrict__ seg_out, long long* __restrict__ tot,
           const PartitionCounts* __restrict__ cnt_local, const PartitionCounts* __restrict__ cnt_global) {
    pdl_sync();
    const RowItem it = resolve_item(items[blockIdx.x], cnt_local, cnt_global);
    const int32_t* __restrict__ rows = it.buf < 0 ? nullptr : (it.buf == 0 ? rowidx0 : rowidx1);
    long long* const base = seg_out[it.seg];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    long long tg = 0, th = 0;
    // a warp walks its rows one after the other; the next row's id, extent and (g,h) are fetched while
    // the current row's entries are being added, so the dependent loads of a row overlap the previous row
    long long i = it.begin + warp;
    int32_t rid = 0;
    long long p0 = 0, p1 = 0;
    int2 q = make_int2(0, 0);
    if (i < it.end) {
        rid = rows ? rows[i] : (int32_t)i;
        p0 = sp.row_ptr[rid]; p1 = sp.row_ptr[rid + 1];
        q = gq[rid];
    }
    while (i < it.end) {
        const long long ni = i + nwarps;
        int32_t nrid = 0;
        long long np0 = 0, np1 = 0;
        int2 nq = make_int2(0, 0);
        if (ni < it.end) {
            nrid = rows ? rows[ni] : (int32_t)ni;
            np0 = sp.row_ptr[nrid]; np1 = sp.row_ptr[nrid + 1];
            nq = gq[nrid];
        }
        tg += q.x; th += q.y;
        for (long long p = p0 + lane; p < p1; p += 32) {
            const unsigned e = sp.nz[p];
            const int f = (int)(e >> 16), slot = (int)(e & 0xffffu);
            if (fea_selected[f]) {
                unsigned long long* out = reinterpret_cast<unsigned long long*>(base + feat_owner[f] * owner_stride + feat_inner[f]) + 2 * slot;
                atomicAdd(out, (unsigned long long)(long long)q.x);
                atomicAdd(out + 1, (unsigned long long)(long long)q.y);
            }
        }
        i = ni; rid = nrid; p0 = np0; p1 = np1; q = nq;
    }
    if (lane == 0 && (tg | th) != 0) {
        atomicAdd(reinterpret_cast<unsigned long long*>(tot + 2 * it.seg), (unsigned long long)tg);
        atomicAdd(reinterpret_cast<unsigned long long*>(tot + 2 * it.seg + 1), (unsigned long long)th);
    }
}

// "Fill-value bin by subtraction": per node and selected feature, default slot += node totals - (sum of
// what the stored entries put into the feature's bins).  Exact (integers); linear, so with several
// GPUs it is applied to the local sums before the reduce-scatter.  grid (feature blocks, nodes), one
// warp per feature.
__global__ void __launch_bounds__(256)
csr_default_fix_k(SparseRows sp, const uint8_t* __restrict__ fea_selected, int32_t F, const int32_t* __restrict__ start,
                  const int32_t* __restrict__ feat_owner, const int64_t* __restrict__ feat_inner, long long owner_stride,
                  long long* const* __restrict__ seg_out, const long long* __restrict__ tot) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (f >= F || !fea_selected[f]) return;
    const int seg = blockIdx.y;
    long long* hp = seg_out[seg] + feat_owner[f] * owner_stride + feat_inner[f];
    const int nb = start[f + 1] - start[f];
    long long g = 0, h = 0;
    for (int j = lane; j < nb; j += 32) {
        const longlong2 v = *reinterpret_cast<const longlong2*>(hp + 2 * j);
        g += v.x; h += v.y;
    }
    for (int d = 16; d > 0; d >>= 1) {
        g += __shfl_xor_sync(0xffffffffu, g, d);
        h += __shfl_xor_sync(0xffffffffu, h, d);
    }
    if (lane == 0) {
        const int dslot = sp.default_slot[f];
        hp[2 * dslot] += tot[2 * seg] - g;
        hp[2 * dslot + 1] += tot[2 * seg + 1] - h;
    }
}

// upload: range and order checks, (feature, slot) packed into one word
__global__ void csr_pack_k(const int64_t* __restrict__ row_ptr, const int32_t* __restrict__ fids,
                           const int32_t* __restrict__ slots, int64_t nrows, int32_t F, const int32_t* __restrict__ nbins,
                           uint32_t* __restrict__ nz_out, int* bad_flag) {
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < nrows; r += (int64_t)gridDim.x * blockDim.x) {
        const int64_t p0 = row_ptr[r] - row_ptr[0], p1 = row_ptr[r + 1] - row_ptr[0];
        int prev = -1;
        for (int64_t p = p0; p < p1; p++) {
            const int f = fids[p], sl = slots[p];
            if (f <= prev || f >= F || sl < 0 || sl >= (f >= 0 && f < F ? nbins[f] : 0)) { *bad_flag = 1; nz_out[p] = 0; continue; }
            prev = f;
            nz_out[p] = ((unsigned)f << 16) | (unsigned)sl;
        }
    }
}

__global__ void csr_to_dense_k(SparseRows sp, int64_t row0, int64_t nrows, int32_t F, int32_t* __restrict__ out) {
    const int64_t total = nrows * F;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / F;
        const int f = (int)(i - r * F);
        out[i] = sparse_slot(sp, (int32_t)(row0 + r), f);
    }
}

// K3: sibling = parent - smaller child (HistogramBuilder.substract,
//     J/data/gbdt/HistogramBuilder.java:101-107), exact in the integer domain.
// Features with more than 256 slots ("wide", e.g. un-sampled continuous columns) do not fit the
// shared-memory layout above; their (g,h) go straight into the node's histogram with 64-bit global
// reductions -- thousands of bins, so little contention.  Same work items as hist_build_k; one
// CTA per (item, wide feature).
__global__ void __launch_bounds__(256)
hist_wide_k(WideCols wide, const int32_t* __restrict__ wide_fids, const int2* __restrict__ gq,
            const int32_t* __restrict__ rowidx0, const int32_t* __restrict__ rowidx1,
            const RowItem* __restrict__ items, const int32_t* __restrict__ feat_owner,
            const int64_t* __restrict__ feat_inner, long long owner_stride, long long* const* __restrict__ seg_out,
            const PartitionCounts* __restrict__ cnt_local, const PartitionCounts* __restrict__ cnt_global) {
    pdl_sync();
    const RowItem it = resolve_item(items[blockIdx.x], cnt_local, cnt_global);
    const int f = wide_fids[blockIdx.y];
    const uint16_t* __restrict__ colp = wide.data + (long long)wide.index[f] * wide.stride;
    const int32_t* __restrict__ rows = it.buf < 0 ? nullptr : (it.buf == 0 ? rowidx0 : rowidx1);
    unsigned long long* out = reinterpret_cast<unsigned long long*>(seg_out[it.seg] + feat_owner[f] * owner_stride + feat_inner[f]);
    for (long long i = it.begin + threadIdx.x; i < it.end; i += blockDim.x) {
        const int32_t rid = rows ? rows[i] : (int32_t)i;
        const int2 q = gq[rid];
        const int slot = colp[rid];
        atomicAdd(out + 2 * slot, (unsigned long long)(long long)q.x);
        atomicAdd(out + 2 * slot + 1, (unsigned long long)(long long)q.y);
    }
}

__global__ void hist_subtract_k(const SubItem* __restrict__ items, int64_t len) {
    pdl_sync();
    const SubItem it = items[blockIdx.y];
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < len;
         i += (int64_t)gridDim.x * blockDim.x)
        it.out[i] = it.parent[i] - it.small[i];
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

__global__ void zero_segments_k(long long* const* __restrict__ dst, int64_t len) {
    pdl_sync();
    long long* d = dst[blockIdx.y];
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < len;
         i += (int64_t)gridDim.x * blockDim.x)
        d[i] = 0;
}

__global__ void copy_segments_k(const long long* __restrict__ src, long long* const* __restrict__ dst, int64_t len) {
    const long long* s = src + blockIdx.y * len;
    long long* d = dst[blockIdx.y];
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < len;
         i += (int64_t)gridDim.x * blockDim.x)
        d[i] = s[i];
}

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

__global__ void __launch_bounds__(kScanWarps * 32)
split_scan_k(const ScanNode* __restrict__ nodes, const int32_t* __restrict__ start,
             const int64_t* __restrict__ feat_off, const uint8_t* __restrict__ fea_selected,
             int32_t f_from, int32_t f_to, int32_t first_fea, const GradScale* __restrict__ scale,
             GainParams gp, const PartitionCounts* __restrict__ cnt_global, SplitPart* __restrict__ parts,
             ytk_split* __restrict__ outs) {
    pdl_sync();
    __shared__ double s_tot[2];
    __shared__ float s_lc[kScanWarps];
    __shared__ int s_fid[kScanWarps], s_lo[kScanWarps], s_hi[kScanWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int node = blockIdx.x, slice = blockIdx.y;
    const double ig = scale->inv_g, ih = scale->inv_h;
    const long long* __restrict__ hist = nodes[node].hist;
    long long sample_cnt = nodes[node].sample_cnt;
    if (nodes[node].cseg >= 0) {
        const PartitionCounts g = cnt_global[nodes[node].cseg];
        const long long small = g.left < g.right ? g.left : g.right;
        sample_cnt = nodes[node].which == 0 ? small : (g.left + g.right - small);
    }

    if (warp == 0) {
        double G, H;
        if (first_fea >= 0) {
            long long g = 0, h = 0;
            const long long* hp = hist + feat_off[first_fea];
            int nb = start[first_fea + 1] - start[first_fea];
            for (int j = lane; j < nb; j += 32) {
                longlong2 gh = *reinterpret_cast<const longlong2*>(hp + 2 * j);
                g += gh.x; h += gh.y;
            }
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
    const int f = f_from + slice * kScanWarps + warp;       // one feature per warp
    if (can_split && f < f_to && fea_selected[f]) {
        const long long* hp = hist + feat_off[f];
        const int nb = start[f + 1] - start[f];
        long long carryG = 0, carryH = 0;
        int carry_last = -1;
        float b_lc = -CUDART_INF_F;
        int b_lo = -1, b_hi = 0x7fffffff;
        // 256 bins at a time (one pass for a feature with uint8 slots) are fetched up front -- 8 loads
        // in flight instead of one dependent memory round trip per 32-bin chunk
        for (int blk = 0; blk < nb; blk += kMaxBins) {
        longlong2 pre[kMaxBins / 32];
#pragma unroll
        for (int c = 0; c < kMaxBins / 32; c++) {
            const int j = blk + c * 32 + lane;
            pre[c] = make_longlong2(0, 0);
            if (j < nb) pre[c] = *reinterpret_cast<const longlong2*>(hp + 2 * j);
        }
#pragma unroll
        for (int c = 0; c < kMaxBins / 32; c++) {
            const int c0 = blk + c * 32;
            if (c0 >= nb) break;
            int j = c0 + lane;
            long long g = pre[c].x, h = pre[c].y;
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
        }
        // first (lowest bin) maximum inside the feature
        for (int d = 16; d > 0; d >>= 1) {
            float o_lc = __shfl_xor_sync(0xffffffffu, b_lc, d);
            int o_lo = __shfl_xor_sync(0xffffffffu, b_lo, d);
            int o_hi = __shfl_xor_sync(0xffffffffu, b_hi, d);
            if (o_lc > b_lc || (o_lc == b_lc && o_hi < b_hi)) { b_lc = o_lc; b_lo = o_lo; b_hi = o_hi; }
        }
        if (b_lc > w_lc) { w_lc = b_lc; w_fid = f; w_lo = b_lo; w_hi = b_hi; }
    }
    if (lane == 0) { s_lc[warp] = w_lc; s_fid[warp] = w_fid; s_lo[warp] = w_lo; s_hi[warp] = w_hi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        // warps hold increasing feature ids: strict > keeps the smaller fid on ties
        SplitPart best = {-CUDART_INF_F, -1, -1, -1};
        for (int w = 0; w < kScanWarps; w++)
            if (s_fid[w] >= 0 && s_lc[w] > best.lc) best = {s_lc[w], s_fid[w], s_lo[w], s_hi[w]};
        parts[(size_t)node * gridDim.y + slice] = best;
        if (slice == 0) {
            ytk_split* out = outs + node;
            out->sum_grad = totG; out->sum_hess = totH;
            out->root_gain = root_gain;
            out->node_value = (float)calc_node_value(gp, totG, totH);
        }
    }
}

// folds the slices of every node (increasing feature ids: strict > = smallest fid wins ties)
// (mirror: the finished record is also stored there -- pinned host memory on one GPU, so that the host
// reads it without a copy while device-side consumers keep reading `outs`)
__global__ void split_merge_k(const SplitPart* __restrict__ parts, int n_nodes, int n_slices,
                              ytk_split* __restrict__ outs, ytk_split* __restrict__ mirror) {
    pdl_sync();
    int node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= n_nodes) return;
    SplitPart best = {-CUDART_INF_F, -1, -1, -1};
    for (int s = 0; s < n_slices; s++) {
        SplitPart p = parts[(size_t)node * n_slices + s];
        if (p.fid >= 0 && p.lc > best.lc) best = p;
    }
    ytk_split r = outs[node];            // totals, root gain and node value were written by the scan
    r.loss_chg = best.lc; r.fid = best.fid; r.lo = best.lo; r.hi = best.hi;
    outs[node] = r;
    if (mirror) mirror[node] = r;
}

// ------------------------------------------------------------------------------------
// K6: row partition.  SamplePositionData.resetPosition
//     (J/data/gbdt/SamplePositionData.java:115-133): row goes left iff
//     (float)slot(row, fid) < cond.  The reference then swaps in place (unstable); only
//     the left/right sets are contractual, so this is a stable out-of-place scatter over
//     every node of the batch at once: flags + per-tile counts, a per-node scan of the
//     tile counts, then the scatter into the other row-id buffer.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ const int32_t* seg_src(const SplitSeg& s, const int32_t* r0, const int32_t* r1) {
    return s.buf < 0 ? nullptr : (s.buf == 0 ? r0 : r1);
}

__global__ void __launch_bounds__(kPartThreads)
partition_count_k(const uint8_t* __restrict__ src, int32_t S8, int64_t col_stride, WideCols wide, SparseRows sparse,
                  const int32_t* __restrict__ rowidx0, const int32_t* __restrict__ rowidx1, const SplitSeg* __restrict__ segs,
                  const int* __restrict__ tile_seg, uint8_t* __restrict__ flags, int32_t* __restrict__ block_cnt) {
    __shared__ int s_cnt[kPartThreads / 32];
    const SplitSeg sg = segs[tile_seg[blockIdx.x]];
    const int32_t* __restrict__ rowidx = seg_src(sg, rowidx0, rowidx1);
    const long long base = sg.begin + (long long)(blockIdx.x - sg.tile0) * kPartTile;
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < kPartItems; k++) {
        long long i = base + k * kPartThreads + threadIdx.x;
        if (i < sg.end) {
            int32_t rid = rowidx ? rowidx[i] : (int32_t)i;
            bool left;
            if (sg.fid < 0) left = src[rid] != 0;
            else if (sparse.row_ptr) left = (float)sparse_slot(sparse, rid, sg.fid) < sg.cond;
            else if (sg.wide > 0) left = (float)wide.data[(sg.wide - 1) * wide.stride + rid] < sg.cond;   // uint16 column
            else if (col_stride > 0) left = (float)src[sg.fid * col_stride + rid] < sg.cond;   // feature-major copy
            else left = (float)src[(long long)rid * S8 + sg.fid] < sg.cond;
            flags[i] = left ? 1 : 0;
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

// one CTA per segment: exclusive scan of its tiles' counts, total -> counts[seg]
__global__ void __launch_bounds__(1024)
partition_scan_k(const SplitSeg* __restrict__ segs, int32_t* __restrict__ block_cnt, PartitionCounts* counts) {
    __shared__ long long s_warp[32];
    __shared__ long long s_carry;
    const SplitSeg sg = segs[blockIdx.x];
    int32_t* bc = block_cnt + sg.tile0;
    const int nblocks = sg.ntiles;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int c0 = 0; c0 < nblocks; c0 += 1024) {
        int i = c0 + threadIdx.x;
        long long v = (i < nblocks) ? bc[i] : 0;
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
        if (i < nblocks) bc[i] = (int32_t)excl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        counts[blockIdx.x].left = s_carry;
        counts[blockIdx.x].right = (sg.end - sg.begin) - s_carry;
    }
}

__global__ void __launch_bounds__(kPartThreads)
partition_scatter_k(int32_t* __restrict__ rowidx0, int32_t* __restrict__ rowidx1,
                    const SplitSeg* __restrict__ segs, const int* __restrict__ tile_seg,
                    const uint8_t* __restrict__ flags, const int32_t* __restrict__ block_off,
                    const PartitionCounts* __restrict__ counts) {
    __shared__ int s_warp[kPartThreads / 32];
    __shared__ int32_t s_rows[kPartTile];      // the tile, left rows first, both sides in input order
    const int seg = tile_seg[blockIdx.x];
    const SplitSeg sg = segs[seg];
    const int32_t* __restrict__ rowidx = seg_src(sg, rowidx0, rowidx1);
    int32_t* __restrict__ out = (sg.buf == 0) ? rowidx1 : rowidx0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long tile_begin = sg.begin + (long long)(blockIdx.x - sg.tile0) * kPartTile;
    const int tile_n = (int)min((long long)kPartTile, sg.end - tile_begin);
    // thread t owns items [t*8, t*8+8) of the tile so that order is preserved
    const int first = threadIdx.x * kPartItems;
    int32_t rid[kPartItems];
    bool fl[kPartItems];
    int cnt = 0;
    if (first + kPartItems <= tile_n && rowidx &&
        ((reinterpret_cast<unsigned long long>(rowidx + tile_begin + first) & 15ull) == 0) &&
        ((reinterpret_cast<unsigned long long>(flags + tile_begin + first) & 7ull) == 0)) {
        // full, aligned run: two 128-bit loads of row ids and one 64-bit load of flags
        const int4* rp = reinterpret_cast<const int4*>(rowidx + tile_begin + first);
        int4 r0 = rp[0], r1 = rp[1];
        rid[0] = r0.x; rid[1] = r0.y; rid[2] = r0.z; rid[3] = r0.w;
        rid[4] = r1.x; rid[5] = r1.y; rid[6] = r1.z; rid[7] = r1.w;
        unsigned long long f8 = *reinterpret_cast<const unsigned long long*>(flags + tile_begin + first);
#pragma unroll
        for (int k = 0; k < kPartItems; k++) { fl[k] = (f8 >> (8 * k)) & 1ull; cnt += fl[k] ? 1 : 0; }
    } else {
#pragma unroll
        for (int k = 0; k < kPartItems; k++) {
            long long i = tile_begin + first + k;
            fl[k] = false;
            rid[k] = 0;
            if (first + k < tile_n) {
                rid[k] = rowidx ? rowidx[i] : (int32_t)i;
                fl[k] = flags[i] != 0;
                cnt += fl[k] ? 1 : 0;
            }
        }
    }
    int p = cnt;
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, p, d);
        if (lane >= d) p += t;
    }
    if (lane == 31) s_warp[warp] = p;
    __syncthreads();
    int woff = 0, tile_left = 0;
    for (int w = 0; w < kPartThreads / 32; w++) {
        if (w < warp) woff += s_warp[w];
        tile_left += s_warp[w];
    }
    int lpos = woff + p - cnt;                           // lefts of this tile before my first item
    int rpos = tile_left + (first - lpos);               // rights go after all lefts of the tile
#pragma unroll
    for (int k = 0; k < kPartItems; k++) {
        if (first + k < tile_n) {
            if (fl[k]) s_rows[lpos++] = rid[k];
            else s_rows[rpos++] = rid[k];
        }
    }
    __syncthreads();
    // coalesced write-out of the two runs
    const long long left_total = counts[seg].left;
    const long long lefts_before_tile = block_off[blockIdx.x];
    const long long ldst = sg.begin + lefts_before_tile;
    const long long rdst = sg.begin + left_total + ((tile_begin - sg.begin) - lefts_before_tile);
    for (int i = threadIdx.x; i < tile_n; i += kPartThreads) {
        if (i < tile_left) out[ldst + i] = s_rows[i];
        else out[rdst + (i - tile_left)] = s_rows[i];
    }
}

// K7: exact sums of the fixed-point (g,h) over a node's rows -- computeLeafStats
//     (J/optimizer/gbdt/DataParallelTreeMaker.java:352-372).
__global__ void leaf_sums_k(const int2* __restrict__ gq, const int32_t* __restrict__ rowidx0,
                            const int32_t* __restrict__ rowidx1, const RowItem* __restrict__ items,
                            long long* __restrict__ out, const PartitionCounts* __restrict__ cnt_local,
                            const PartitionCounts* __restrict__ cnt_global) {
    pdl_sync();
    const RowItem it = resolve_item(items[blockIdx.x], cnt_local, cnt_global);
    const int32_t* __restrict__ rowidx = it.buf < 0 ? nullptr : (it.buf == 0 ? rowidx0 : rowidx1);
    long long g = 0, h = 0;
    for (long long i = it.begin + threadIdx.x; i < it.end; i += blockDim.x) {
        int2 q = gq[rowidx ? rowidx[i] : (int32_t)i];
        g += q.x; h += q.y;
    }
    for (int d = 16; d > 0; d >>= 1) {
        g += __shfl_xor_sync(0xffffffffu, g, d);
        h += __shfl_xor_sync(0xffffffffu, h, d);
    }
    if ((threadIdx.x & 31) == 0) {
        if (g) atomicAdd(reinterpret_cast<unsigned long long*>(out + 2 * it.seg), (unsigned long long)g);
        if (h) atomicAdd(reinterpret_cast<unsigned long long*>(out + 2 * it.seg + 1), (unsigned long long)h);
    }
}

__global__ void fill_i32_k(int32_t* p, int64_t n, int32_t v) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = v;
}
__global__ void fill_positions_k(const int32_t* __restrict__ rowidx0, const int32_t* __restrict__ rowidx1,
                                 const PosItem* __restrict__ items, int32_t* __restrict__ node_of_row) {
    const PosItem it = items[blockIdx.y];
    const int32_t* __restrict__ rowidx = it.buf < 0 ? nullptr : (it.buf == 0 ? rowidx0 : rowidx1);
    for (long long i = it.begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < it.end;
         i += (long long)gridDim.x * blockDim.x)
        node_of_row[rowidx ? rowidx[i] : (int32_t)i] = it.nid;
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

cudaError_t launch_repack_rows(const int32_t* x, int64_t nrows, int32_t stride, int32_t F, const int32_t* nbins,
                               uint8_t* bins, int32_t S8, uint8_t* col, int64_t col_stride, WideCols wide,
                               int64_t row_base, int* bad_flag, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    repack_rows_k<<<grid_for(nrows * S8, 256), 256, 0, st>>>(x, nrows, stride, F, nbins, bins, S8, col, col_stride, wide,
                                                             row_base, bad_flag);
    return cudaGetLastError();
}
cudaError_t launch_quantize_rows(const int32_t* xbits, int64_t nrows, int32_t stride, int32_t F,
                                 const float* cuts, const int32_t* start, uint8_t* bins, int32_t S8,
                                 uint8_t* col, int64_t col_stride, WideCols wide, int64_t row_base, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    quantize_rows_k<<<grid_for(nrows * S8, 256), 256, 0, st>>>(xbits, nrows, stride, F, cuts, start, bins, S8, col,
                                                               col_stride, wide, row_base);
    return cudaGetLastError();
}
cudaError_t launch_unpack_rows(const uint8_t* bins, int32_t S8, int64_t nrows, int32_t F, WideCols wide, int32_t* out,
                               cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    unpack_rows_k<<<grid_for(nrows * F, 256), 256, 0, st>>>(bins, S8, nrows, F, wide, out);
    return cudaGetLastError();
}

cudaError_t launch_grad_quantize(const float* gp, int64_t n, int32_t K, int32_t gid, int2* gq,
                                 unsigned* absmax_bits, GradScale* scale, int bits, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(absmax_bits, 0, 2 * sizeof(unsigned), st);
    if (e != cudaSuccess) return e;
    int g = grid_for(n, 256, kNumSMs * 8);
    grad_absmax_k<<<g, 256, 0, st>>>(gp, n, K, gid, absmax_bits);
    grad_quantize_k<<<g, 256, 0, st>>>(gp, n, K, gid, gq, absmax_bits, scale, bits);
    return cudaGetLastError();
}

int64_t hist_rows_per_item(int64_t total_rows, int groups) {
    // one CTA per SM (128 KB of shared memory); aim at about two waves over all feature groups,
    // never less than 4 sweeps of the CTA (4096 rows) per item so zeroing + flushing stays small
    int64_t items = std::max<int64_t>(1, (2 * kNumSMs) / std::max(1, groups));
    int64_t rows = (total_rows + items - 1) / items;
    rows = std::max<int64_t>(rows, 4 * kHistThreads);
    rows = (rows + kHistThreads - 1) / kHistThreads * kHistThreads;
    return std::min<int64_t>(rows, kHistMaxRowsPerItem);    // the 16-bit limb sums must stay inside 32 bits
}

cudaError_t launch_hist_build(const uint8_t* bins, int32_t S8, const int2* gq, const int32_t* rowidx0,
                              const int32_t* rowidx1, const RowItem* items, int n_items, int32_t F,
                              const uint32_t* group_mask, const int32_t* nbins, const int32_t* feat_owner,
                              const int64_t* feat_inner, long long owner_stride, long long* const* seg_out,
                              const PartitionCounts* cnt_local, const PartitionCounts* cnt_global, cudaStream_t st) {
    if (n_items <= 0) return cudaSuccess;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(hist_build_k, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             kHistSmemBytes);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    dim3 grid((unsigned)n_items, (unsigned)(S8 / kGroup));
    return launch_chain(hist_build_k, grid, dim3(kHistThreads), kHistSmemBytes, st, bins, S8, gq, rowidx0, rowidx1, items, F, group_mask,
                                                             nbins, feat_owner, feat_inner, owner_stride, seg_out,
                                                             cnt_local, cnt_global);
}

cudaError_t launch_hist_wide(WideCols wide, const int32_t* wide_fids, int n_wide, const int2* gq, const int32_t* rowidx0,
                             const int32_t* rowidx1, const RowItem* items, int n_items, const int32_t* feat_owner,
                             const int64_t* feat_inner, long long owner_stride, long long* const* seg_out,
                             const PartitionCounts* cnt_local, const PartitionCounts* cnt_global, cudaStream_t st) {
    if (n_items <= 0 || n_wide <= 0) return cudaSuccess;
    dim3 grid((unsigned)n_items, (unsigned)n_wide);
    return launch_chain(hist_wide_k, grid, dim3(256), 0, st, wide, wide_fids, gq, rowidx0, rowidx1, items, feat_owner, feat_inner, owner_stride,
                                      seg_out, cnt_local, cnt_global);
}

cudaError_t launch_hist_csr(SparseRows sparse, const uint8_t* fea_selected, const int2* gq, const int32_t* rowidx0,
                            const int32_t* rowidx1, const RowItem* items, int n_items, const int32_t* feat_owner,
                            const int64_t* feat_inner, long long owner_stride, long long* const* seg_out, long long* tot,
                            const PartitionCounts* cnt_local, const PartitionCounts* cnt_global, cudaStream_t st) {
    if (n_items <= 0) return cudaSuccess;
    return launch_chain(hist_csr_k, dim3(n_items), dim3(256), 0, st, sparse, fea_selected, gq, rowidx0, rowidx1, items,
                        feat_owner, feat_inner, owner_stride, seg_out, tot, cnt_local, cnt_global);
}
cudaError_t launch_csr_default_fix(SparseRows sparse, const uint8_t* fea_selected, int32_t F, const int32_t* start,
                                   const int32_t* feat_owner, const int64_t* feat_inner, long long owner_stride,
                                   long long* const* seg_out, int n_segs, const long long* tot, cudaStream_t st) {
    if (n_segs <= 0) return cudaSuccess;
    dim3 grid((unsigned)((F + 7) / 8), (unsigned)n_segs);
    return launch_chain(csr_default_fix_k, grid, dim3(256), 0, st, sparse, fea_selected, F, start, feat_owner, feat_inner,
                        owner_stride, seg_out, tot);
}
cudaError_t launch_csr_pack(const int64_t* row_ptr, const int32_t* fids, const int32_t* slots, int64_t nrows, int32_t F,
                            const int32_t* nbins, uint32_t* nz_out, int* bad_flag, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    csr_pack_k<<<grid_for(nrows, 128), 128, 0, st>>>(row_ptr, fids, slots, nrows, F, nbins, nz_out, bad_flag);
    return cudaGetLastError();
}
cudaError_t launch_csr_to_dense(SparseRows sparse, int64_t row0, int64_t nrows, int32_t F, int32_t* out, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    csr_to_dense_k<<<grid_for(nrows * F, 256), 256, 0, st>>>(sparse, row0, nrows, F, out);
    return cudaGetLastError();
}

cudaError_t launch_hist_subtract(const SubItem* items, int n_items, int64_t len, cudaStream_t st) {
    if (n_items <= 0) return cudaSuccess;
    dim3 grid((unsigned)grid_for(len, 256, 64), (unsigned)n_items);
    return launch_chain(hist_subtract_k, grid, dim3(256), 0, st, items, len);
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
cudaError_t launch_zero_segments(long long* const* dst, int n, int64_t len, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    dim3 grid((unsigned)grid_for(len, 256, 64), (unsigned)n);
    return launch_chain(zero_segments_k, grid, dim3(256), 0, st, dst, len);
}
cudaError_t launch_copy_segments(const long long* src, long long* const* dst, int n, int64_t len, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    dim3 grid((unsigned)grid_for(len, 256, 64), (unsigned)n);
    copy_segments_k<<<grid, 256, 0, st>>>(src, dst, len);
    return cudaGetLastError();
}

int split_scan_slices(int32_t f_from, int32_t f_to) {
    int n = f_to - f_from;
    return n <= 0 ? 1 : (n + kScanWarps - 1) / kScanWarps;
}

cudaError_t launch_split_scan(const ScanNode* nodes, int n_nodes, const int32_t* start, const int64_t* feat_off,
                              const uint8_t* fea_selected, int32_t f_from, int32_t f_to, int32_t first_fea,
                              const GradScale* scale, GainParams gp, const PartitionCounts* cnt_global,
                              SplitPart* parts, ytk_split* out, ytk_split* mirror, cudaStream_t st) {
    if (n_nodes <= 0) return cudaSuccess;
    const int slices = split_scan_slices(f_from, f_to);
    dim3 grid((unsigned)n_nodes, (unsigned)slices);
    cudaError_t e = launch_chain(split_scan_k, grid, dim3(kScanWarps * 32), 0, st, nodes, start, feat_off, fea_selected, f_from,
                                 f_to, first_fea, scale, gp, cnt_global, parts, out);
    if (e != cudaSuccess) return e;
    return launch_chain(split_merge_k, dim3((n_nodes + 127) / 128), dim3(128), 0, st, (const SplitPart*)parts, n_nodes, slices,
                        out, mirror);
}

// Single-pass variant used inside make(): one kernel reads every row id once, gathers its slot,
// learns where the tile goes from a chained scan over the tiles of its node (decoupled look-back)
// and writes left rows forward from the node's begin, right rows BACKWARD from its end.  Both
// sides stay monotonic in input order (the right side reversed), which is all the gathers of the
// next level care about; the child intervals are the same as the stable version's.
//   tile_state[t] = epoch << 34 | status << 32 | count;  status 1: the tile's own left count,
//   2: left count of the node's tiles up to and including t.  Entries of another epoch are "not
//   yet written", so the array is never cleared.  Tiles are claimed in ticket order, hence every
//   tile a CTA waits for is already running.
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__global__ void __launch_bounds__(kFusedThreads)
partition_fused_k(const uint8_t* __restrict__ src, int32_t S8, int64_t col_stride, WideCols wide, SparseRows sparse,
                  int32_t* __restrict__ rowidx0,
                  int32_t* __restrict__ rowidx1, const SplitSeg* __restrict__ segs, int n_segs,
                  const int* __restrict__ tile_seg, int n_tiles, const int* __restrict__ n_tiles_dev,
                  unsigned long long* __restrict__ tile_state, unsigned int* __restrict__ ticket, unsigned long long epoch,
                  PartitionCounts* __restrict__ counts) {
    pdl_sync();
    __shared__ int s_tile, s_prefix, s_found;
    __shared__ int s_warp[kFusedThreads / 32], s_lb_sum[kFusedThreads / 32], s_lb_fp[kFusedThreads / 32];
    __shared__ int32_t s_rows[kFusedTile];      // the tile, left rows first, both sides in input order
    // tables planned on the device (plan_next_partition_k): the grid is an upper bound, the real
    // number of tiles is read here, and the planner has reset the ticket and the empty nodes' counts
    if (n_tiles_dev) n_tiles = *n_tiles_dev;
    if (threadIdx.x == 0)
        s_tile = n_tiles_dev ? (int)atomicAdd(ticket, 1u)
                             : (int)atomicInc(ticket, (unsigned)n_tiles - 1u);        // back to 0 after the last tile
    __syncthreads();
    const int tile = s_tile;
    if (tile >= n_tiles) return;
    if (tile == 0 && !n_tiles_dev)
        for (int k = threadIdx.x; k < n_segs; k += kFusedThreads)
            if (segs[k].ntiles == 0) counts[k] = {0, 0};
    const int seg = tile_seg[tile];
    const SplitSeg sg = segs[seg];
    const int32_t* __restrict__ rowidx = seg_src(sg, rowidx0, rowidx1);
    int32_t* __restrict__ out = (sg.buf == 0) ? rowidx1 : rowidx0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int j = tile - sg.tile0;
    const long long tile_begin = sg.begin + (long long)j * kFusedTile;
    const int tile_n = (int)min((long long)kFusedTile, sg.end - tile_begin);
    // thread t owns kFusedItems consecutive items of the tile so that order is preserved
    const int first = threadIdx.x * kFusedItems;
    const int mine = max(0, min(kFusedItems, tile_n - first));      // items of this thread that exist
    int32_t rid[kFusedItems];
    if (!rowidx) {
#pragma unroll
        for (int k = 0; k < kFusedItems; k++) rid[k] = (int32_t)(tile_begin + first + k);
    } else if (mine == kFusedItems && ((reinterpret_cast<unsigned long long>(rowidx + tile_begin + first) & 15ull) == 0)) {
        const int4* rp = reinterpret_cast<const int4*>(rowidx + tile_begin + first);
#pragma unroll
        for (int q = 0; q < kFusedItems / 4; q++) {
            const int4 r = rp[q];
            rid[4 * q] = r.x; rid[4 * q + 1] = r.y; rid[4 * q + 2] = r.z; rid[4 * q + 3] = r.w;
        }
    } else {
#pragma unroll
        for (int k = 0; k < kFusedItems; k++) rid[k] = k < mine ? rowidx[tile_begin + first + k] : 0;
    }
    if (!rowidx && mine < kFusedItems) {
#pragma unroll
        for (int k = 0; k < kFusedItems; k++) if (k >= mine) rid[k] = 0;       // never read past the last row
    }
    // one addressing form for the three sources: a byte mask over rows (left iff non-zero, i.e. NOT
    // below 0.5), the feature-major copy, row-major slots
    const bool mask = sg.fid < 0;
    const uint8_t* __restrict__ base = mask ? src : (col_stride > 0 ? src + (long long)sg.fid * col_stride : src + sg.fid);
    const int mul = (mask || col_stride > 0) ? 1 : S8;
    const float cond = mask ? 0.5f : sg.cond;
    unsigned short slot[kFusedItems];
    if (sparse.row_ptr && !mask) {                       // sparse rows: look the feature up in the row
#pragma unroll
        for (int k = 0; k < kFusedItems; k++) slot[k] = (unsigned short)sparse_slot(sparse, rid[k], sg.fid);
    } else if (sg.wide > 0) {                            // a feature with more than 256 slots: its uint16 column
        const uint16_t* __restrict__ wb = wide.data + (long long)(sg.wide - 1) * wide.stride;
#pragma unroll
        for (int k = 0; k < kFusedItems; k++) slot[k] = wb[rid[k]];
    } else {
#pragma unroll
        for (int k = 0; k < kFusedItems; k++) slot[k] = base[(long long)rid[k] * mul];     // independent gathers in flight
    }
    unsigned fl = 0;                                     // bit k: item k goes left
#pragma unroll
    for (int k = 0; k < kFusedItems; k++)
        fl |= (unsigned)((k < mine) && (((float)slot[k] < cond) != mask)) << k;
    const int cnt = __popc(fl);
    int p = cnt;
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, p, d);
        if (lane >= d) p += t;
    }
    if (lane == 31) s_warp[warp] = p;
    __syncthreads();
    int woff = 0, tile_left = 0;
    for (int w = 0; w < kFusedThreads / 32; w++) {
        if (w < warp) woff += s_warp[w];
        tile_left += s_warp[w];
    }
    // ---- chained scan: publish this tile's left count, then add up the tiles before it back to
    // the nearest one that already knows its prefix.  Every warp inspects its own window of 32
    // predecessors, so one round trip to L2 covers 512 tiles.
    const unsigned long long tag = epoch << 34;
    if (threadIdx.x == 0) st_volatile_u64(tile_state + tile, tag | ((j == 0 ? 2ull : 1ull) << 32) | (unsigned)tile_left);
    int excl = 0;                                        // thread 0's
    if (j > 0) {                                         // (uniform over the CTA)
        int look = tile - 1;
        for (;;) {
            const int t = look - warp * 32 - lane;
            unsigned long long v;
            int fp;
            for (;;) {
                v = tag | (2ull << 32);                  // before the node's first tile: a virtual "prefix 0"
                if (t >= sg.tile0) v = ld_volatile_u64(tile_state + t);
                const int status = (v >> 34) == epoch ? (int)((v >> 32) & 3ull) : 0;
                const unsigned not_ready = __ballot_sync(0xffffffffu, status == 0);
                const unsigned is_prefix = __ballot_sync(0xffffffffu, status == 2);
                fp = is_prefix ? __ffs(is_prefix) - 1 : 32;                    // nearest tile of the window that knows its prefix
                const unsigned need = fp >= 31 ? 0xffffffffu : ((2u << fp) - 1u);
                if (!(not_ready & need)) break;                                 // else: a tile in between has not published yet
            }
            int c = (lane <= fp) ? (int)(unsigned)v : 0;
            for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
            if (lane == 0) { s_lb_sum[warp] = c; s_lb_fp[warp] = fp; }
            __syncthreads();
            if (threadIdx.x == 0) {
                int found = 0;
                for (int w = 0; w < kFusedThreads / 32 && !found; w++) {
                    excl += s_lb_sum[w];
                    found = s_lb_fp[w] < 32;
                }
                s_found = found;
            }
            __syncthreads();
            if (s_found) break;
            look -= kFusedThreads;
        }
    }
    if (threadIdx.x == 0) {
        if (j > 0) st_volatile_u64(tile_state + tile, tag | (2ull << 32) | (unsigned)(excl + tile_left));
        s_prefix = excl;
        if (j == sg.ntiles - 1) {
            const long long left = (long long)excl + tile_left;
            counts[seg] = {left, (sg.end - sg.begin) - left};
        }
    }
    int lpos = woff + p - cnt;                           // lefts of this tile before my first item
    int rpos = tile_left + (first - lpos);               // rights go after all lefts of the tile
#pragma unroll
    for (int k = 0; k < kFusedItems; k++) {
        if (k < mine) {
            const bool l = (fl >> k) & 1u;
            s_rows[l ? lpos : rpos] = rid[k];
            lpos += l ? 1 : 0;
            rpos += l ? 0 : 1;
        }
    }
    __syncthreads();
    const long long left_base = sg.begin + s_prefix;
    const long long right_last = sg.end - 1 - ((long long)j * kFusedTile - s_prefix) + tile_left;   // minus (i - tile_left) below
    for (int i = threadIdx.x; i < tile_n; i += kFusedThreads)
        out[i < tile_left ? left_base + i : right_last - i] = s_rows[i];
}

cudaError_t launch_partition_fused(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0, int32_t* rowidx1,
                                   const SplitSeg* segs, int n_segs, const int* tile_seg, int n_tiles,
                                   unsigned long long* tile_state, unsigned int* ticket, unsigned long long epoch,
                                   PartitionCounts* counts, cudaStream_t st, SparseRows sparse) {
    if (n_segs <= 0) return cudaSuccess;
    if (n_tiles <= 0) return cudaMemsetAsync(counts, 0, (size_t)n_segs * sizeof(PartitionCounts), st);
    return launch_chain(partition_fused_k, dim3(n_tiles), dim3(kFusedThreads), 0, st, src, S8, col_stride, wide, sparse,
                        rowidx0, rowidx1, segs, n_segs, tile_seg, n_tiles, (const int*)nullptr, tile_state, ticket, epoch, counts);
}

// ------------------------------------------------------------------------------------
// Level-wise growth: the partition of the NEXT level, planned on the device.  What the next level
// splits is decided by the records the split search has just written: a child of this level's job k
// is split iff its lossChg > min_split_loss (the other stop rules are settled by the host before it
// asks for this).  Queued right behind the split search, it lets the next partition run while the
// host is still being woken up, reading the records and building the other tables.
//   slot 2k / 2k+1 = left / right child of job k;  records: recs[2*scan], recs[2*scan+1] = the
//   smaller / larger child of the job with scan index `scan` (-1: its children are leaves).
//   Slots that are not split get counts {-1,-1}: the host checks them against its own decision.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512)
plan_next_partition_k(const SplitSeg* __restrict__ cur_segs, const PartitionCounts* __restrict__ cur_counts,
                      const PartitionCounts* __restrict__ cur_counts_global, const SpecJob* __restrict__ jobs, int m, const ytk_split* recs, double min_split_loss,
                      const int32_t* __restrict__ wide_index, SplitSeg* __restrict__ out_segs,
                      int* __restrict__ out_tile_seg, int* __restrict__ n_tiles_out, unsigned int* __restrict__ ticket,
                      PartitionCounts* __restrict__ next_counts) {
    pdl_sync();
    __shared__ int s_warp[16];
    __shared__ int s_carry;
    const int nslots = 2 * m;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { s_carry = 0; *ticket = 0; }
    __syncthreads();
    for (int s0 = 0; s0 < nslots; s0 += 512) {
        const int s = s0 + threadIdx.x;
        SplitSeg sg = {0, 0, 0, 0, 0.f, 0, 0, 0};
        bool split = false;
        if (s < nslots) {
            const int k = s >> 1, side = s & 1;
            const SpecJob jb = jobs[k];
            if (jb.scan >= 0) {
                const SplitSeg pr = cur_segs[jb.seg];
                const PartitionCounts c = cur_counts[jb.cseg];
                const PartitionCounts cgl = cur_counts_global[jb.cseg];
                const bool small_is_left = cgl.left < cgl.right;     // by global rows (:344-350)
                const bool is_small = (side == 0) == small_is_left;
                // the host copy may be in mapped host memory; volatile: written by the kernel before this one
                const ytk_split r = recs[2 * jb.scan + (is_small ? 0 : 1)];
                if ((double)r.loss_chg > min_split_loss) {
                    split = true;
                    sg.begin = side == 0 ? pr.begin : pr.begin + c.left;
                    sg.end = side == 0 ? pr.begin + c.left : pr.end;
                    sg.buf = jb.child_buf;
                    sg.fid = r.fid;
                    sg.cond = 0.5f * (float)(r.lo + r.hi);            // TreeNode.setSplit (:144-149)
                    sg.wide = wide_index ? wide_index[r.fid] + 1 : 0;
                    sg.ntiles = (int)((sg.end - sg.begin + kFusedTile - 1) / kFusedTile);
                }
            }
        }
        // exclusive scan of ntiles over the slots
        int p = sg.ntiles;
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, p, d);
            if (lane >= d) p += t;
        }
        if (lane == 31) s_warp[warp] = p;
        __syncthreads();
        int woff = 0, total = 0;
        for (int w = 0; w < 16; w++) {
            if (w < warp) woff += s_warp[w];
            total += s_warp[w];
        }
        sg.tile0 = s_carry + woff + p - sg.ntiles;
        if (s < nslots) {
            out_segs[s] = sg;
            // not split: marked; split but without rows on this GPU: counted here, no tile will
            if (!split) next_counts[s] = {-1, -1};
            else if (sg.ntiles == 0) next_counts[s] = {0, 0};
        }
        // tile -> slot table: the lanes of a warp fill the tiles of the warp's 32 slots, one slot after the other
        for (int j = 0; j < 32; j++) {
            const int t0 = __shfl_sync(0xffffffffu, sg.tile0, j), nt = __shfl_sync(0xffffffffu, sg.ntiles, j);
            for (int t = lane; t < nt; t += 32) out_tile_seg[t0 + t] = s0 + warp * 32 + j;
        }
        __syncthreads();
        if (threadIdx.x == 0) s_carry += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_tiles_out = s_carry;
}

// The root's partition, planned from the root's own record: one slot.
__global__ void __launch_bounds__(256)
plan_root_partition_k(long long begin, long long end, int buf, const ytk_split* rec, double min_split_loss,
                      const int32_t* __restrict__ wide_index, SplitSeg* __restrict__ out_segs,
                      int* __restrict__ out_tile_seg, int* __restrict__ n_tiles_out, unsigned int* __restrict__ ticket,
                      PartitionCounts* __restrict__ next_counts) {
    pdl_sync();
    const ytk_split r = rec[0];
    const bool split = (double)r.loss_chg > min_split_loss;
    const int ntiles = split ? (int)((end - begin + kFusedTile - 1) / kFusedTile) : 0;
    if (threadIdx.x == 0) {
        SplitSeg sg = {0, 0, 0, 0, 0.f, 0, 0, 0};
        if (split) sg = {begin, end, buf, r.fid, 0.5f * (float)(r.lo + r.hi), 0, ntiles, wide_index ? wide_index[r.fid] + 1 : 0};
        out_segs[0] = sg;
        if (!split) next_counts[0] = {-1, -1};
        else if (ntiles == 0) next_counts[0] = {0, 0};
        *n_tiles_out = ntiles;
        *ticket = 0;
    }
    for (int t = threadIdx.x; t < ntiles; t += blockDim.x) out_tile_seg[t] = 0;
}

cudaError_t launch_partition_planned_root(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0,
                                          int32_t* rowidx1, long long begin, long long end, int buf, const ytk_split* rec,
                                          double min_split_loss, SplitSeg* out_segs, int* out_tile_seg, int* n_tiles_dev,
                                          int grid_bound, unsigned long long* tile_state, unsigned int* ticket,
                                          unsigned long long epoch, PartitionCounts* next_counts, cudaStream_t st,
                                          SparseRows sparse) {
    cudaError_t e = launch_chain(plan_root_partition_k, dim3(1), dim3(256), 0, st, begin, end, buf, rec, min_split_loss,
                                 wide.index, out_segs, out_tile_seg, n_tiles_dev, ticket, next_counts);
    if (e != cudaSuccess || grid_bound <= 0) return e;
    return launch_chain(partition_fused_k, dim3(grid_bound), dim3(kFusedThreads), 0, st, src, S8, col_stride, wide, sparse,
                        rowidx0, rowidx1, (const SplitSeg*)out_segs, 1, (const int*)out_tile_seg, 0, (const int*)n_tiles_dev,
                        tile_state, ticket, epoch, next_counts);
}

cudaError_t launch_partition_planned(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0,
                                     int32_t* rowidx1, const SplitSeg* cur_segs, const PartitionCounts* cur_counts,
                                     const PartitionCounts* cur_counts_global, const SpecJob* jobs, int m,
                                     const ytk_split* recs, double min_split_loss,
                                     SplitSeg* out_segs, int* out_tile_seg, int* n_tiles_dev, int grid_bound,
                                     unsigned long long* tile_state, unsigned int* ticket, unsigned long long epoch,
                                     PartitionCounts* next_counts, cudaStream_t st, SparseRows sparse) {
    if (m <= 0) return cudaSuccess;
    cudaError_t e = launch_chain(plan_next_partition_k, dim3(1), dim3(512), 0, st, cur_segs, cur_counts, cur_counts_global, jobs, m, recs,
                                 min_split_loss, wide.index, out_segs, out_tile_seg, n_tiles_dev, ticket, next_counts);
    if (e != cudaSuccess || grid_bound <= 0) return e;
    return launch_chain(partition_fused_k, dim3(grid_bound), dim3(kFusedThreads), 0, st, src, S8, col_stride, wide, sparse,
                        rowidx0, rowidx1, (const SplitSeg*)out_segs, 2 * m, (const int*)out_tile_seg, 0, (const int*)n_tiles_dev,
                        tile_state, ticket, epoch, next_counts);
}

cudaError_t launch_partition(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0, int32_t* rowidx1,
                             const SplitSeg* segs, int n_segs, const int* tile_seg, int n_tiles,
                             uint8_t* flags, int32_t* block_cnt, PartitionCounts* counts, cudaStream_t st, SparseRows sparse) {
    if (n_segs <= 0) return cudaSuccess;
    if (n_tiles > 0)
        partition_count_k<<<n_tiles, kPartThreads, 0, st>>>(src, S8, col_stride, wide, sparse, rowidx0, rowidx1, segs, tile_seg,
                                                            flags, block_cnt);
    partition_scan_k<<<n_segs, 1024, 0, st>>>(segs, block_cnt, counts);
    if (n_tiles > 0)
        partition_scatter_k<<<n_tiles, kPartThreads, 0, st>>>(rowidx0, rowidx1, segs, tile_seg, flags, block_cnt,
                                                              counts);
    return cudaGetLastError();
}

cudaError_t launch_leaf_sums(const int2* gq, const int32_t* rowidx0, const int32_t* rowidx1, const RowItem* items,
                             int n_items, long long* out, const PartitionCounts* cnt_local,
                             const PartitionCounts* cnt_global, cudaStream_t st) {
    if (n_items <= 0) return cudaSuccess;
    return launch_chain(leaf_sums_k, dim3(n_items), dim3(256), 0, st, gq, rowidx0, rowidx1, items, out, cnt_local, cnt_global);
}

cudaError_t launch_fill_i32(int32_t* p, int64_t n, int32_t v, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    fill_i32_k<<<grid_for(n, 256), 256, 0, st>>>(p, n, v);
    return cudaGetLastError();
}
cudaError_t launch_fill_positions(const int32_t* rowidx0, const int32_t* rowidx1, const PosItem* items, int n_items,
                                  int32_t* node_of_row, cudaStream_t st) {
    if (n_items <= 0) return cudaSuccess;
    dim3 grid(64, (unsigned)n_items);
    fill_positions_k<<<grid, 256, 0, st>>>(rowidx0, rowidx1, items, node_of_row);
    return cudaGetLastError();
}

}  // namespace ytk

// ====================================================================================
// score update, loss and gradient pairs on the device --
// GBDTOptimizer.predictAndCalcLossGrad, train branch (J/optimizer/GBDTOptimizer.java:513-609).
// ====================================================================================
namespace ytk {

// score[(row*K + gid)] += leaf value, for every row of the listed leaves (float add, as :651-655).
// The row's leaf is known from the final partition, so no tree walk is needed.
__global__ void add_leaf_values_k(const int32_t* __restrict__ rowidx0, const int32_t* __restrict__ rowidx1,
                                  const LeafItem* __restrict__ items, float* __restrict__ score, int32_t K, int32_t gid) {
    const LeafItem it = items[blockIdx.y];
    const int32_t* __restrict__ rowidx = it.buf < 0 ? nullptr : (it.buf == 0 ? rowidx0 : rowidx1);
    for (long long i = it.begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < it.end;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = rowidx ? rowidx[i] : i;
        score[r * K + gid] += it.value;
    }
}

// Rows that were not sampled for the tree are in no leaf interval: walk the tree on their slots
// with Tree.getNext's rule, slot <= cond goes left (J/data/gbdt/Tree.java:156-168).
__global__ void walk_rows_k(const uint8_t* __restrict__ bins, int32_t S8, WideCols wide, SparseRows sparse,
                            const int32_t* __restrict__ rows,
                            long long begin, long long end, const TreeNodeDev* __restrict__ nodes,
                            float* __restrict__ score, int32_t K, int32_t gid) {
    for (long long i = begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < end;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = rows[i];
        int pid = 0;
        TreeNodeDev nd = nodes[0];
        while (nd.left != -1) {
            const int wi = wide.index ? wide.index[nd.fid] : -1;
            const float v = sparse.row_ptr ? (float)sparse_slot(sparse, (int32_t)r, nd.fid)
                            : wi >= 0 ? (float)wide.data[wi * wide.stride + r] : (float)bins[r * S8 + nd.fid];
            pid = (v <= nd.info) ? nd.left : nd.right;
            nd = nodes[pid];
        }
        score[r * K + gid] += nd.info;
    }
}

// Test-set scoring on ORIGINAL feature values (GBDTOptimizer.score :630-650 with isOriginTree = true;
// Tree.getLeafIndex / getNext, J/data/gbdt/Tree.java:142-168): value <= threshold goes left, a value
// whose bits are Float.NaN's raw pattern (Constants.INT_MISSING_VALUE) takes the node's default child.
__global__ void score_values_k(const int32_t* __restrict__ xbits, long long n, int32_t stride,
                               const TreeNodeDev* __restrict__ nodes, const int32_t* __restrict__ default_child,
                               float* __restrict__ score, int32_t K, int32_t gid) {
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
        int pid = 0;
        TreeNodeDev nd = nodes[0];
        while (nd.left != -1) {
            const int32_t bits = xbits[r * stride + nd.fid];
            if (bits == 0x7fc00000) pid = default_child ? default_child[pid] : nd.left;
            else pid = (__int_as_float(bits) <= nd.info) ? nd.left : nd.right;
            nd = nodes[pid];
        }
        score[r * K + gid] += nd.info;
    }
}

__device__ __forceinline__ double block_sum(double v, double* s_part) {
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += s_part[w];
    return t;      // valid in thread 0
}

// numTreeInGroup == 1 (:536-557): SigmoidFunction (J/loss/SigmoidFunction.java:39-93, zMax = 0) or
// L2Function (J/loss/L2Function.java:30-55).  Float casts where the reference has them: the combined
// score is a float sum, predict is rounded to float before the derivatives, (g,h) are rounded to float.
__global__ void loss_grad_scalar_k(int loss_kind, const float* __restrict__ score, const float* __restrict__ init,
                                   const float* __restrict__ y, const float* __restrict__ w, long long n,
                                   float* __restrict__ gp, double* __restrict__ block_loss) {
    __shared__ double s_part[32];
    double loss = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double wei = w ? (double)w[i] : 1.0;
        const double s = (double)__fadd_rn(score[i], init ? init[i] : 0.0f);
        const double label = y[i];
        double l, pred;
        if (loss_kind == 0) {
            if (s >= 0.0) {
                const double e = exp(-s);
                l = __dadd_rn(log(__dadd_rn(1.0, e)), __dmul_rn(s, __dsub_rn(1.0, label)));
                pred = __ddiv_rn(1.0, __dadd_rn(1.0, e));
            } else {
                const double e = exp(s);
                l = __dsub_rn(log(__dadd_rn(1.0, e)), __dmul_rn(s, label));
                pred = __ddiv_rn(e, __dadd_rn(1.0, e));
            }
        } else {
            const double d = __dsub_rn(label, s);
            l = __dmul_rn(__dmul_rn(0.5, d), d);
            pred = s;
        }
        loss = __dadd_rn(loss, __dmul_rn(wei, l));
        const double p = (double)(float)pred;
        const double d0 = __dsub_rn(p, label);
        const double d1 = loss_kind == 0 ? __dmul_rn(p, __dsub_rn(1.0, p)) : 1.0;
        *reinterpret_cast<float2*>(gp + 2 * i) = make_float2((float)__dmul_rn(d0, wei), (float)__dmul_rn(d1, wei));
    }
    const double t = block_sum(loss, s_part);
    if (threadIdx.x == 0) block_loss[blockIdx.x] = t;
}

// softmax branch (:559-607, J/loss/SoftmaxFunction.java:44-113); y holds the class index of each row
constexpr int kMaxClasses = 64;
__global__ void loss_grad_softmax_k(const float* __restrict__ score, const float* __restrict__ init,
                                    const float* __restrict__ y_onehot, const float* __restrict__ w, long long n,
                                    int32_t K, float* __restrict__ gp, double* __restrict__ block_loss) {
    __shared__ double s_part[32];
    double loss = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double wei = w ? (double)w[i] : 1.0;
        double ls[kMaxClasses];
        for (int k = 0; k < K; k++) ls[k] = (double)__fadd_rn(score[i * K + k], init ? init[i * K + k] : 0.0f);
        double mx = ls[K - 1];
        for (int k = 0; k < K - 1; k++) if (ls[k] > mx) mx = ls[k];
        double esum = 0.0, sig = 0.0;
        for (int k = 0; k < K; k++) {
            const double nw = __dsub_rn(ls[k], mx);
            sig = __dadd_rn(sig, __dmul_rn(nw, (double)y_onehot[i * K + k]));
            ls[k] = exp(nw);
            esum = __dadd_rn(esum, ls[k]);
        }
        loss = __dadd_rn(loss, __dmul_rn(wei, __dsub_rn(log(esum), sig)));
        const double inv = __ddiv_rn(1.0, esum);
        for (int k = 0; k < K; k++) {
            const double pk = __dmul_rn(ls[k], inv);
            const double d0 = __dsub_rn(pk, (double)y_onehot[i * K + k]);
            const double d1 = __dmul_rn(2.0, __dmul_rn(pk, __dsub_rn(1.0, pk)));
            *reinterpret_cast<float2*>(gp + (i * K + k) * 2) = make_float2((float)__dmul_rn(d0, wei), (float)__dmul_rn(d1, wei));
        }
    }
    const double t = block_sum(loss, s_part);
    if (threadIdx.x == 0) block_loss[blockIdx.x] = t;
}

// fixed-order sum of the per-block partial losses (run-to-run reproducible)
__global__ void sum_blocks_k(const double* __restrict__ block_loss, int n, double* __restrict__ out) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < n; i++) t = __dadd_rn(t, block_loss[i]);
        *out = t;
    }
}

cudaError_t launch_add_leaf_values(const int32_t* rowidx0, const int32_t* rowidx1, const LeafItem* items, int n_items,
                                   float* score, int32_t K, int32_t gid, cudaStream_t st) {
    if (n_items <= 0) return cudaSuccess;
    dim3 grid(64, (unsigned)n_items);
    add_leaf_values_k<<<grid, 256, 0, st>>>(rowidx0, rowidx1, items, score, K, gid);
    return cudaGetLastError();
}
cudaError_t launch_walk_rows(const uint8_t* bins, int32_t S8, WideCols wide, const int32_t* rows, long long begin, long long end,
                             const TreeNodeDev* nodes, float* score, int32_t K, int32_t gid, cudaStream_t st,
                             SparseRows sparse) {
    if (end <= begin) return cudaSuccess;
    long long b = (end - begin + 255) / 256;
    walk_rows_k<<<(unsigned)std::min<long long>(b, kNumSMs * 8), 256, 0, st>>>(bins, S8, wide, sparse, rows, begin, end, nodes, score, K, gid);
    return cudaGetLastError();
}
cudaError_t launch_score_values(const int32_t* xbits, long long n, int32_t stride, const TreeNodeDev* nodes,
                                const int32_t* default_child, float* score, int32_t K, int32_t gid, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    long long b = (n + 255) / 256;
    score_values_k<<<(unsigned)std::min<long long>(b, kNumSMs * 16), 256, 0, st>>>(xbits, n, stride, nodes, default_child, score,
                                                                                 K, gid);
    return cudaGetLastError();
}
int loss_grad_blocks() { return kNumSMs * 4; }
cudaError_t launch_loss_grad(int loss_kind, const float* score, const float* init, const float* y, const float* w,
                             long long n, int32_t K, float* gp, double* block_loss, double* loss_out, cudaStream_t st) {
    const int blocks = loss_grad_blocks();
    if (loss_kind == 2) {
        if (K > kMaxClasses) return cudaErrorInvalidValue;
        loss_grad_softmax_k<<<blocks, 128, 0, st>>>(score, init, y, w, n, K, gp, block_loss);
    } else {
        loss_grad_scalar_k<<<blocks, 256, 0, st>>>(loss_kind, score, init, y, w, n, gp, block_loss);
    }
    sum_blocks_k<<<1, 32, 0, st>>>(block_loss, blocks, loss_out);
    return cudaGetLastError();
}

}  // namespace ytk

// ====================================================================================
// Collectives over NVLink peer memory (one node, one process per GPU).  Every rank owns an
// "exchange" buffer and a histogram staging buffer that all peers have mapped (CUDA IPC).  The
// small per-level exchanges (child row counts, split records, leaf sums) are pushes of a few
// hundred bytes into every peer's buffer followed by a flag; the histogram reduce-scatter is a
// pull: each rank sums the segments it owns straight out of its peers' staging buffers.  These
// replace ncclAllReduce / ncclAllGather / ncclReduceScatter on the per-level path, whose launch and
// protocol latency (3 collectives x ~30 us per level) is what limits scaling of a 3 ms tree.
// ====================================================================================
namespace ytk {

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// Copies `bytes` (multiple of 8) of this rank's payload into slot [rank] of every rank's exchange
// area, then raises flag[rank] = epoch there.
__device__ __forceinline__ void peer_push(const PeerTable& pt, const unsigned long long* __restrict__ src, size_t area_off,
                                          size_t slot_bytes, size_t bytes, int flag_set, unsigned long long epoch) {
    const size_t words = bytes / 8;
    for (int r = 0; r < pt.world; r++) {
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(pt.xchg[r] + area_off + (size_t)pt.rank * slot_bytes);
        for (size_t i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < pt.world) {
        unsigned long long* f = reinterpret_cast<unsigned long long*>(pt.xchg[threadIdx.x]) + flag_set * kMaxPeers + pt.rank;
        st_release_sys(f, epoch);
    }
}

// Waits (bounded) until every rank's flag of the set has reached `epoch`.  Returns false on timeout.
__device__ __forceinline__ bool peer_wait(const PeerTable& pt, int flag_set, unsigned long long epoch, int* err) {
    __shared__ int s_ok;
    if (threadIdx.x == 0) s_ok = 1;
    __syncthreads();
    if (threadIdx.x < pt.world) {
        const unsigned long long* f = reinterpret_cast<const unsigned long long*>(pt.xchg[pt.rank]) + flag_set * kMaxPeers + threadIdx.x;
        const long long t0 = clock64();
        while (ld_acquire_sys(f) < epoch) {
            if (clock64() - t0 > 6000000000ll) { s_ok = 0; *err = 1; break; }     // ~3 s: a peer died
            __nanosleep(100);
        }
    }
    __syncthreads();
    return s_ok != 0;
}

// All-reduce (sum, int64) of a small vector in one kernel: push my contribution to every rank, wait
// for everybody's, add them up.  out may alias src.  One CTA.
__global__ void peer_allreduce_k(PeerTable pt, const unsigned long long* __restrict__ src, size_t area_off, size_t slot_bytes,
                                 size_t words, int flag_set, unsigned long long epoch, long long* __restrict__ out, int* err) {
    peer_push(pt, src, area_off, slot_bytes, words * 8, flag_set, epoch);
    if (!peer_wait(pt, flag_set, epoch, err)) return;
    const char* base = pt.xchg[pt.rank] + area_off;
    for (size_t i = threadIdx.x; i < words; i += blockDim.x) {
        long long s = 0;
        for (int r = 0; r < pt.world; r++) s += reinterpret_cast<const long long*>(base + (size_t)r * slot_bytes)[i];
        out[i] = s;
    }
}

// All-gather of a small record block in one kernel: afterwards slot [r] of this rank's area holds rank r's.
__global__ void peer_allgather_k(PeerTable pt, const unsigned long long* __restrict__ src, size_t area_off, size_t slot_bytes,
                                 size_t bytes, int flag_set, unsigned long long epoch, int* err) {
    peer_push(pt, src, area_off, slot_bytes, bytes, flag_set, epoch);
    peer_wait(pt, flag_set, epoch, err);
}

// The same all-gather for the split records, followed by the merge every rank would otherwise do on
// its host (ytk_gbdt_merge_splits: needReplace in rank order, element-wise MAX of the node totals,
// DataParallelTreeMaker.java:560,640-653): merged record k -> out_dev[k] and out_host[k] (pinned).
__global__ void peer_allgather_merge_k(PeerTable pt, const ytk_split* __restrict__ src, size_t area_off, size_t slot_bytes,
                                       int m, int flag_set, unsigned long long epoch, ytk_split* __restrict__ out_dev,
                                       ytk_split* __restrict__ out_host, int* err) {
    peer_push(pt, reinterpret_cast<const unsigned long long*>(src), area_off, slot_bytes, (size_t)m * sizeof(ytk_split),
              flag_set, epoch);
    if (!peer_wait(pt, flag_set, epoch, err)) return;
    const char* base = pt.xchg[pt.rank] + area_off;
    for (int k = threadIdx.x; k < m; k += blockDim.x) {
        const ytk_split first = reinterpret_cast<const ytk_split*>(base)[k];
        ytk_split best = first;
        double G = first.sum_grad, H = first.sum_hess;
        for (int r = 1; r < pt.world; r++) {
            const ytk_split t2 = reinterpret_cast<const ytk_split*>(base + (size_t)r * slot_bytes)[k];
            const bool rep = (best.fid <= t2.fid) ? (t2.loss_chg > best.loss_chg) : (t2.loss_chg >= best.loss_chg);
            if (rep) best = t2;
            if (t2.sum_grad > G) G = t2.sum_grad;
            if (t2.sum_hess > H) H = t2.sum_hess;
        }
        // root_gain / node_value are functions of (G,H): from the first rank that had the totals
        for (int r = 0; r < pt.world; r++) {
            const ytk_split t2 = reinterpret_cast<const ytk_split*>(base + (size_t)r * slot_bytes)[k];
            if (t2.sum_grad == G && t2.sum_hess == H) { best.root_gain = t2.root_gain; best.node_value = t2.node_value; break; }
        }
        best.sum_grad = G; best.sum_hess = H;
        out_dev[k] = best;
        out_host[k] = best;
    }
}

// Reduce-scatter by pulling: this rank owns segment [rank] of every node; staging layout of every
// rank is [owner][node][seg_max].  grid = (chunks, nodes); block 0 of each launch waits for the flags
// ... every block waits (cheap: the flags are already up for all but the first to arrive).
__global__ void peer_pull_reduce_k(PeerTable pt, int m, long long seg_max, int flag_set, unsigned long long epoch,
                                   long long* const* __restrict__ dst, int* err) {
    // the histogram kernel ahead of this one in the stream has finished: tell everybody
    if (blockIdx.x == 0 && blockIdx.y == 0) {
        __threadfence_system();
        if (threadIdx.x < pt.world) {
            unsigned long long* f = reinterpret_cast<unsigned long long*>(pt.xchg[threadIdx.x]) + flag_set * kMaxPeers + pt.rank;
            st_release_sys(f, epoch);
        }
    }
    if (!peer_wait(pt, flag_set, epoch, err)) return;
    const int node = blockIdx.y;
    long long* __restrict__ out = dst[node];
    const size_t off = ((size_t)pt.rank * m + node) * (size_t)seg_max;
    for (long long i = (blockIdx.x * (long long)blockDim.x + threadIdx.x) * 2; i < seg_max;
         i += (long long)gridDim.x * blockDim.x * 2) {
        longlong2 s = make_longlong2(0, 0);
        for (int r = 0; r < pt.world; r++) {
            const longlong2 v = *reinterpret_cast<const longlong2*>(reinterpret_cast<const long long*>(pt.hist[r]) + off + i);
            s.x += v.x; s.y += v.y;
        }
        *reinterpret_cast<longlong2*>(out + i) = s;
    }
}

cudaError_t launch_peer_allreduce(const PeerTable& pt, const void* src, size_t area_off, size_t slot_bytes, size_t words,
                                  int flag_set, unsigned long long epoch, long long* out, int* err, cudaStream_t st) {
    peer_allreduce_k<<<1, 256, 0, st>>>(pt, static_cast<const unsigned long long*>(src), area_off, slot_bytes, words, flag_set,
                                        epoch, out, err);
    return cudaGetLastError();
}
cudaError_t launch_peer_allgather(const PeerTable& pt, const void* src, size_t area_off, size_t slot_bytes, size_t bytes,
                                  int flag_set, unsigned long long epoch, int* err, cudaStream_t st) {
    peer_allgather_k<<<1, 256, 0, st>>>(pt, static_cast<const unsigned long long*>(src), area_off, slot_bytes, bytes, flag_set,
                                        epoch, err);
    return cudaGetLastError();
}
cudaError_t launch_peer_allgather_merge(const PeerTable& pt, const ytk_split* src, size_t area_off, size_t slot_bytes, int m,
                                        int flag_set, unsigned long long epoch, ytk_split* out_dev, ytk_split* out_host,
                                        int* err, cudaStream_t st) {
    peer_allgather_merge_k<<<1, 256, 0, st>>>(pt, src, area_off, slot_bytes, m, flag_set, epoch, out_dev, out_host, err);
    return cudaGetLastError();
}
cudaError_t launch_peer_pull_reduce(const PeerTable& pt, int m, long long seg_max, int flag_set, unsigned long long epoch,
                                    long long* const* dst, int* err, cudaStream_t st) {
    if (m <= 0) return cudaSuccess;
    dim3 grid((unsigned)std::max<long long>(1, std::min<long long>(16, (seg_max / 2 + 255) / 256)), (unsigned)m);
    peer_pull_reduce_k<<<grid, 256, 0, st>>>(pt, m, seg_max, flag_set, epoch, dst, err);
    return cudaGetLastError();
}

}  // namespace ytk
