// Histograms: gradient fixed point, the per-node (g,h) histogram kernels, subtraction.
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
__global__ void grad_absmax_k(const float* __restrict__ gp, int64_t n, int32_t K, int32_t gid, const int* __restrict__ gid_dev,
                              unsigned* __restrict__ absmax_bits) {
    if (gid_dev) gid = *gid_dev;
    float mg = 0.f, mh = 0.f;
    const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
    int64_t i0 = 0;
    if (K == 1 && (reinterpret_cast<unsigned long long>(gp) & 15ull) == 0) {
        // one class: the pairs are contiguous -- two rows per 128-bit load, four loads in flight per thread
        const float4* __restrict__ gp4 = reinterpret_cast<const float4*>(gp);
        const int64_t n2 = n >> 1;
        int64_t j = tid;
        for (; j + 3 * nthr < n2; j += 4 * nthr) {
            const float4 a = gp4[j], b = gp4[j + nthr], c = gp4[j + 2 * nthr], d = gp4[j + 3 * nthr];
            mg = fmaxf(fmaxf(fmaxf(mg, fabsf(a.x)), fmaxf(fabsf(a.z), fabsf(b.x))), fmaxf(fmaxf(fabsf(b.z), fabsf(c.x)), fmaxf(fabsf(c.z), fmaxf(fabsf(d.x), fabsf(d.z)))));
            mh = fmaxf(fmaxf(fmaxf(mh, fabsf(a.y)), fmaxf(fabsf(a.w), fabsf(b.y))), fmaxf(fmaxf(fabsf(b.w), fabsf(c.y)), fmaxf(fabsf(c.w), fmaxf(fabsf(d.y), fabsf(d.w)))));
        }
        for (; j < n2; j += nthr) {
            const float4 a = gp4[j];
            mg = fmaxf(mg, fmaxf(fabsf(a.x), fabsf(a.z)));
            mh = fmaxf(mh, fmaxf(fabsf(a.y), fabsf(a.w)));
        }
        i0 = n2 << 1;                      // (an odd last row is left to the loop below)
    }
    for (int64_t i = i0 + tid; i < n; i += nthr) {
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

__global__ void grad_quantize_k(const float* __restrict__ gp, int64_t n, int32_t K, int32_t gid, const int* __restrict__ gid_dev,
                                int2* __restrict__ gq, const unsigned* __restrict__ absmax_bits,
                                GradScale* __restrict__ scale, int bits) {
    if (gid_dev) gid = *gid_dev;
    int kg = scale_exponent(__uint_as_float(absmax_bits[0]), bits);
    int kh = scale_exponent(__uint_as_float(absmax_bits[1]), bits);
    double mg = ldexp(1.0, kg), mh = ldexp(1.0, kh);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        scale->mul_g = mg; scale->mul_h = mh;
        scale->inv_g = ldexp(1.0, -kg); scale->inv_h = ldexp(1.0, -kh);
    }
    const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
    int64_t i0 = 0;
    if (K == 1 && ((reinterpret_cast<unsigned long long>(gp) | reinterpret_cast<unsigned long long>(gq)) & 15ull) == 0) {
        // one class: two rows per 128-bit load and store, two in flight
        const float4* __restrict__ gp4 = reinterpret_cast<const float4*>(gp);
        int4* __restrict__ gq4 = reinterpret_cast<int4*>(gq);
        const int64_t n2 = n >> 1;
        int64_t j = tid;
        for (; j + nthr < n2; j += 2 * nthr) {
            const float4 a = gp4[j], b = gp4[j + nthr];
            gq4[j] = make_int4(to_fixed(a.x, mg), to_fixed(a.y, mh), to_fixed(a.z, mg), to_fixed(a.w, mh));
            gq4[j + nthr] = make_int4(to_fixed(b.x, mg), to_fixed(b.y, mh), to_fixed(b.z, mg), to_fixed(b.w, mh));
        }
        for (; j < n2; j += nthr) {
            const float4 a = gp4[j];
            gq4[j] = make_int4(to_fixed(a.x, mg), to_fixed(a.y, mh), to_fixed(a.z, mg), to_fixed(a.w, mh));
        }
        i0 = n2 << 1;
    }
    for (int64_t i = i0 + tid; i < n; i += nthr) {
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
// ---- the core both histogram kernels share --------------------------------------------------------
// Per-thread constants of the lane -> (row, feature) mapping.
struct HistLane {
    int half, r16, m;
    uint32_t selk[4];
    uint32_t lane_off;
};
__device__ __forceinline__ HistLane hist_lane(int lane) {
    HistLane L;
    // Lane L works on row (L >> 1) of a 16-row block and on the 16 features of half (L & 1) of the
    // group: at step k (0..15) it updates feature 16*(L&1) + (k ^ (L>>1)).  The 32 lanes of a warp
    // instruction therefore always touch 32 different features = 32 different banks.
    L.half = lane & 1; L.r16 = lane >> 1;
    L.m = L.r16 >> 2;
    const int bsel = L.r16 & 3;
    // PRMT selectors: byte (kk ^ bsel) of a data word moved to byte 1, i.e. bin * 256 = the byte
    // offset of that bin's 256 B row of limbs
#pragma unroll
    for (int kk = 0; kk < 4; kk++) L.selk[kk] = 0x4404u | ((uint32_t)(kk ^ bsel) << 4);
    // byte offset of feature 16*half + r16 inside a bin row; feature 16*half + (k ^ r16) is reached
    // by XOR-ing k*4 into it: (bin256 | lane_off) ^ k4 is one LOP3
    L.lane_off = (uint32_t)(L.half * 16 + L.r16) << 2;
    asm volatile("" : "+r"(L.lane_off));
    return L;
}

constexpr uint32_t kHOff = kMaxBins * 2 * kGroup * 4;        // bytes: the h limbs follow the g limbs

// Accumulates rows [begin, end) (through ridx, or the rows themselves when ridx is null) of feature
// group `fg` into the CTA's shared histogram `sh` (zero, or holding earlier rows of the same node;
// at most kHistMaxRowsPerItem rows between two flushes).  Called by all threads of the CTA.
__device__ __forceinline__ void hist_rows(uint32_t* sh, const HistLane& L, const uint8_t* __restrict__ bins, int32_t S8,
                                          const int2* __restrict__ gq, const int32_t* __restrict__ ridx, int fg,
                                          long long begin, long long end) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int half = L.half, r16 = L.r16, m = L.m;
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
        return i < end ? (ridx ? ridx[i] : (int32_t)i) : -1;
    };
    auto fetch = [&](int32_t rid, Fetch& f) {
        const int32_t ra = __shfl_sync(0xffffffffu, rid, r16);
        const int32_t rb = __shfl_sync(0xffffffffu, rid, 16 + r16);
        // past the end of the run: zeros, i.e. "add zero to bin 0", which changes nothing
        f.a = make_uint4(0, 0, 0, 0); f.b = f.a;
        f.qa = make_int2(0, 0); f.qb = f.qa;
        // (ld_gather_*: no more than 64 B of HBM per miss -- a gathered row uses 32 + 8 bytes of the lines it touches)
        if (ra >= 0) { f.a = ld_gather_u4(gbase + (long long)ra * S8); f.qa = ld_gather_int2(gq + ra); }
        if (rb >= 0) { f.b = ld_gather_u4(gbase + (long long)rb * S8); f.qb = ld_gather_int2(gq + rb); }
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
                const uint32_t off = (__byte_perm(word, 0u, L.selk[kk]) | L.lane_off) ^ (uint32_t)((4 * j + kk) << 2);
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
    long long base = begin + warp * 32;
    Fetch fa, fb;
    int32_t r1 = fetch_rid(base + kHistThreads);
    fetch(fetch_rid(base), fa);
    while (base < end) {
        const int32_t r2 = fetch_rid(base + 2 * kHistThreads);
        fetch(r1, fb);
        accumulate16(fa.a, fa.qa);
        accumulate16(fa.b, fa.qb);
        base += kHistThreads;
        if (base >= end) break;
        r1 = fetch_rid(base + 2 * kHistThreads);
        fetch(r2, fa);
        accumulate16(fb.a, fb.qa);
        accumulate16(fb.b, fb.qb);
        base += kHistThreads;
    }
}

// Experiment (north_star: "TMA-staged tiles where that helps"; YTK_GBDT_HIST_TMA=1): the same rows for a run of
// CONTIGUOUS rows of a single feature group (S8 == 32: the root of the Higgs shape), with the row data brought into
// shared memory by the bulk-copy engine instead of LDG: a warp's block of 32 rows is 1 KB of slots + 256 B of
// gradient pairs, both contiguous, copied by two cp.async.bulk into a per-warp ring (2 stages) and signalled on an
// mbarrier; the lanes then read their 16 + 8 bytes with LDS.  begin must be a multiple of 32 rows (16-byte source
// alignment of the pairs).  `ring`: kHistThreads/32 warps x 2 stages x 1280 bytes, `bars`: 2 mbarriers per warp.
// Measured on the 11 M-row root (profiles/r02_tma_experiment.txt): no faster -- the LDS reads cost the LSU data pipe
// what the coalesced LDG did, and that pipe (shared-memory atomics) is what bounds the kernel.
constexpr int kTmaStageBytes = 32 * kGroup + 32 * 8;
constexpr int kHistSpanSmemBytes = kHistSmemBytes + kHistWarps * 2 * kTmaStageBytes + kHistWarps * 2 * 8;     // 128 KB + 80 KB + 512 B
__device__ __forceinline__ void hist_rows_tma(uint32_t* sh, const HistLane& L, const uint8_t* __restrict__ bins,
                                              const int2* __restrict__ gq, long long begin, long long end, char* ring,
                                              unsigned long long* bars, uint32_t (&phases)[2]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int half = L.half, r16 = L.r16, m = L.m;
    char* const shb = reinterpret_cast<char*>(sh);
    char* my_ring = ring + warp * 2 * kTmaStageBytes;
    const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(my_ring);
    const uint32_t bar_s = (uint32_t)__cvta_generic_to_shared(bars + 2 * warp);
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
                const uint32_t off = (__byte_perm(word, 0u, L.selk[kk]) | L.lane_off) ^ (uint32_t)((4 * j + kk) << 2);
                uint32_t* p = reinterpret_cast<uint32_t*>(shb + off);
                atomicAdd(p, g_lo);
                atomicAdd(reinterpret_cast<int32_t*>(p + kGroup), g_hi);
                atomicAdd(p + kHOff / 4, h_lo);
                atomicAdd(reinterpret_cast<int32_t*>(p + kHOff / 4 + kGroup), h_hi);
            }
        }
    };
    // blocks of this warp: rows [begin + warp*32 + t*1024, +32); only whole blocks go through the ring
    const long long n_full = end > begin + warp * 32 ? (end - begin - warp * 32 + kHistThreads - 32) / kHistThreads : 0;
    auto issue = [&](long long t, int stage) {           // lane 0
        const long long row = begin + warp * 32 + t * kHistThreads;
        const uint32_t dst = ring_s + stage * kTmaStageBytes, bar = bar_s + stage * 8;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)kTmaStageBytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(bins + row * kGroup), "r"(32u * kGroup), "r"(bar) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst + 32u * kGroup), "l"(gq + row), "r"(256u), "r"(bar) : "memory");
    };
    long long whole = 0;                                  // blocks of 32 rows that exist entirely
    for (long long t = 0; t < n_full; t++)
        if (begin + warp * 32 + t * kHistThreads + 32 <= end) whole = t + 1;
    if (lane == 0) {
        if (whole > 0) issue(0, 0);
        if (whole > 1) issue(1, 1);
    }
    for (long long t = 0; t < whole; t++) {
        const int stage = (int)(t & 1);
        const uint32_t parity = phases[stage]++ & 1u;    // (completed phases of the stage's barrier so far, over all calls)
        const uint32_t bar = bar_s + stage * 8;
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        const char* src = my_ring + stage * kTmaStageBytes;
        const uint4 a = *reinterpret_cast<const uint4*>(src + r16 * kGroup + (half << 4));
        const uint4 b = *reinterpret_cast<const uint4*>(src + (16 + r16) * kGroup + (half << 4));
        const int2 qa = *reinterpret_cast<const int2*>(src + 32 * kGroup + r16 * 8);
        const int2 qb = *reinterpret_cast<const int2*>(src + 32 * kGroup + (16 + r16) * 8);
        __syncwarp();                                     // everybody has read the stage: it may be refilled
        if (lane == 0 && t + 2 < whole) issue(t + 2, stage);
        accumulate16(a, qa);
        accumulate16(b, qb);
    }
    // a last partial block: plain loads
    const long long tail = begin + warp * 32 + whole * kHistThreads;
    if (tail < end) {
        const long long ia = tail + r16, ib = tail + 16 + r16;
        uint4 a = make_uint4(0, 0, 0, 0), b = a;
        int2 qa = make_int2(0, 0), qb = qa;
        if (ia < end) { a = *reinterpret_cast<const uint4*>(bins + ia * kGroup + (half << 4)); qa = gq[ia]; }
        if (ib < end) { b = *reinterpret_cast<const uint4*>(bins + ib * kGroup + (half << 4)); qb = gq[ib]; }
        accumulate16(a, qa);
        accumulate16(b, qb);
    }
}

// Flush: recombines the limbs, one 64-bit global reduction per non-empty (feature, bin, component) into
// `out` (the node's histogram), and -- kRezero -- leaves the table zero for the next node's rows.
// What bounds a flush is the number of 32 B sectors its reductions touch (one L1 tag cycle and one L2 request
// each): the four values of a sector of `out` -- (g,h) of two neighbouring bins of one feature -- are therefore
// sent by four neighbouring lanes of ONE instruction, which then covers 8 features x 2 bins in 8 sectors (a
// lane per feature would touch 32).  The price is a 4-way bank conflict on the limb reads (bank = feature),
// cheaper than the sectors it saves.  CTAs start at different bins so that they do not all hit the same
// addresses of a shared node at the same time.
template <bool kRezero>
__device__ __forceinline__ void hist_flush(uint32_t* sh, int fg, int32_t F, uint32_t gmask, const int32_t* __restrict__ nbins,
                                           const int32_t* __restrict__ feat_owner, const int64_t* __restrict__ feat_inner,
                                           long long owner_stride, long long* __restrict__ out) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int q = lane >> 2, s = (lane >> 1) & 1, c = lane & 1;
    constexpr int kPairs = kMaxBins / 2 + 1;             // bin pairs of a feature, one more when its first cell is odd
    // this lane's four features (one per quarter of the group): where their cells go, how many bins they have
    long long* dst0[4];
    int nb[4];
#pragma unroll
    for (int g = 0; g < 4; g++) {
        const int ln = g * 8 + q, ff = fg * kGroup + ln;
        const bool on = ff < F && ((gmask >> ln) & 1u);
        nb[g] = on ? nbins[ff] : 0;
        dst0[g] = on ? out + feat_owner[ff] * owner_stride + feat_inner[ff] : out;
    }
    const int rot = (int)(blockIdx.x * 37u % (unsigned)kPairs);
    for (int w = warp; w < 4 * kPairs; w += kHistWarps) {
        const int g = w & 3;
        int j = (w >> 2) + rot;
        j = j >= kPairs ? j - kPairs : j;
        long long* const d0 = g == 0 ? dst0[0] : (g == 1 ? dst0[1] : (g == 2 ? dst0[2] : dst0[3]));
        const int nbf = g == 0 ? nb[0] : (g == 1 ? nb[1] : (g == 2 ? nb[2] : nb[3]));
        // pairs are aligned to the sectors of `out`: a feature whose first cell sits in the upper half of one starts
        // with a single bin
        const int odd = (int)((reinterpret_cast<unsigned long long>(d0) >> 4) & 1ull);
        const int bin = 2 * j + s - odd;
        if (bin < 0 || bin >= nbf) continue;
        const uint32_t* p = sh + (c ? kHOff / 4 : 0) + bin * (2 * kGroup) + g * 8 + q;
        const uint32_t lo = p[0], hi = p[kGroup];
        if ((lo | hi) == 0) continue;
        const long long v = (long long)(int32_t)hi * 65536ll + (long long)lo;
        if (v != 0) atomicAdd(reinterpret_cast<unsigned long long*>(d0 + 2 * bin + c), (unsigned long long)v);
    }
    if (kRezero) {
        __syncthreads();
        uint4* z = reinterpret_cast<uint4*>(sh);
        for (int i = threadIdx.x; i < kMaxBins * 4 * kGroup / 4; i += kHistThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
}

// Between two runs of rows on the same table: the low limbs' carries move into the high limbs (lo >> 16 added to hi,
// lo kept below 2^16), after which a second run of up to kHistSecondRunRows rows fits -- |hi| <= 2^14 * (65536 + 65000)
// + 65535 < 2^31 and lo <= 65535 + 65000 * 65535 < 2^32 -- and the node is flushed once instead of twice.
constexpr int kHistSecondRunRows = 65000;
__device__ __forceinline__ void hist_carry(uint32_t* sh) {
    uint4* t = reinterpret_cast<uint4*>(sh);
    for (int i = threadIdx.x; i < 2 * kMaxBins * (kGroup / 4); i += kHistThreads) {
        const int region = i / (kMaxBins * (kGroup / 4)), r = i % (kMaxBins * (kGroup / 4));
        const int bin = r / (kGroup / 4), q = r % (kGroup / 4);
        uint4* lo = t + region * (int)(kHOff / 16) + bin * (2 * kGroup / 4) + q;
        uint4* hi = lo + kGroup / 4;
        uint4 l = *lo, h = *hi;
        h.x += l.x >> 16; h.y += l.y >> 16; h.z += l.z >> 16; h.w += l.w >> 16;
        l.x &= 0xffffu; l.y &= 0xffffu; l.z &= 0xffffu; l.w &= 0xffffu;
        *lo = l; *hi = h;
    }
}

__device__ __forceinline__ void hist_zero(uint32_t* sh) {
    uint4* z = reinterpret_cast<uint4*>(sh);
    for (int i = threadIdx.x; i < kMaxBins * 4 * kGroup / 4; i += kHistThreads) z[i] = make_uint4(0, 0, 0, 0);
}

// One CTA (32 warps, 128 KB of shared memory, 1 per SM) per work item: a run of rows of one node and one
// feature group -- the tables of the host-planned path (grow.cu) and the hooks.
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
    const int fg = blockIdx.y;
    const RowItem it = resolve_item(items[blockIdx.x], cnt_local, cnt_global);
    if (it.begin >= it.end) return;          // a deferred chunk beyond the end of its child
    const int32_t* __restrict__ ridx = it.buf < 0 ? nullptr : (it.buf == 0 ? rowidx0 : rowidx1);
    hist_zero(sh);
    __syncthreads();
    const HistLane L = hist_lane(threadIdx.x & 31);
    hist_rows(sh, L, bins, S8, gq, ridx, fg, it.begin, it.end);
    __syncthreads();
    hist_flush<false>(sh, fg, F, group_mask[fg], nbins, feat_owner, feat_inner, owner_stride, seg_out[it.seg]);
}

// ------------------------------------------------------------------------------------
// The same histograms for a whole tree LEVEL whose work is planned on the device (level.cu): the rows
// to be visited are the concatenation of the level's build nodes (the smaller child of every node
// split at this level), `nodes[j]` with a running prefix of COST = rows + a flush's worth of rows.
// Every CTA takes one contiguous span of equal cost -- no tickets, no tail, one set of dependent
// start-up loads per CTA -- walks the nodes in it and flushes at every node boundary (and every
// kHistMaxRowsPerItem rows), re-zeroing while it flushes.  grid = (CTAs per group, groups).
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kHistThreads, 1)
hist_span_k(const uint8_t* __restrict__ bins, int32_t S8, const int2* __restrict__ gq,
            const int32_t* __restrict__ rowidx0, const int32_t* __restrict__ rowidx1,
            const BuildNode* __restrict__ nodes, const int* __restrict__ n_nodes_dev, int32_t F,
            const uint32_t* __restrict__ group_mask, const int32_t* __restrict__ nbins,
            const int32_t* __restrict__ feat_owner, const int64_t* __restrict__ feat_inner, long long owner_stride,
            long long* const* __restrict__ seg_out, unsigned long long* __restrict__ stamp, int use_tma) {
    pdl_sync();
    extern __shared__ __align__(128) uint32_t sh[];
    const int nb = *n_nodes_dev;
    if (nb <= 0) return;
    stamp_begin(stamp);
    // (experiment) per-warp rings + mbarriers of the bulk-copy variant sit behind the histogram
    uint32_t tma_phases[2] = {0, 0};
    char* const ring = reinterpret_cast<char*>(sh) + kHistSmemBytes;
    unsigned long long* const bars = reinterpret_cast<unsigned long long*>(ring + kHistWarps * 2 * kTmaStageBytes);
    if (use_tma) {
        if ((threadIdx.x & 31) == 0) {
            const uint32_t b = (uint32_t)__cvta_generic_to_shared(bars + 2 * (threadIdx.x >> 5));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b + 8));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
    }
    const int fg = blockIdx.y;
    const long long total = nodes[nb - 1].cost_end;
    const long long c0 = total * blockIdx.x / gridDim.x, c1 = total * (blockIdx.x + 1) / gridDim.x;
    if (c0 >= c1) return;
    // first node whose cost range ends beyond c0
    int lo = 0, hi = nb - 1;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (nodes[mid].cost_end > c0) hi = mid; else lo = mid + 1;
    }
    hist_zero(sh);
    __syncthreads();
    const HistLane L = hist_lane(threadIdx.x & 31);
    const uint32_t gmask = group_mask[fg];
    for (int j = lo; j < nb; j++) {
        const BuildNode nd = nodes[j];
        const long long n = nd.end - nd.begin;
        const long long cost_begin = nd.cost_end - n - nd.cost;
        if (cost_begin >= c1) break;
        long long o0 = c0 - cost_begin, o1 = c1 - cost_begin;       // my share of the node, in rows from its begin
        o0 = o0 < 0 ? 0 : o0;
        // shares begin on multiples of 32 rows (the neighbouring span rounds the same boundary the same way)
        o0 = (o0 + 31) & ~31ll;
        o1 = o1 >= n ? n : ((o1 + 31) & ~31ll);
        o0 = o0 > n ? n : o0;
        o1 = o1 > n ? n : o1;
        if (o0 >= o1) continue;                                     // only (part of) its flush cost falls to me
        const int32_t* __restrict__ ridx = nd.buf < 0 ? nullptr : (nd.buf == 0 ? rowidx0 : rowidx1);
        long long* __restrict__ out = seg_out[nd.seg];
        const long long share_end = nd.begin + o1;
        long long b = nd.begin + o0;
        int run = 0;                                     // runs of rows on the table since its last flush
        while (b < share_end) {
            const long long e = min(share_end, b + (long long)(run == 0 ? kHistMaxRowsPerItem : kHistSecondRunRows));
            if (use_tma && !ridx && S8 == kGroup && (b & 31) == 0) hist_rows_tma(sh, L, bins, gq, b, e, ring, bars, tma_phases);
            else hist_rows(sh, L, bins, S8, gq, ridx, fg, b, e);
            __syncthreads();
            b = e;
            if (run == 0 && b < share_end) {             // more rows of the node: a second run on the same table
                hist_carry(sh);
                run = 1;
            } else {
                hist_flush<true>(sh, fg, F, gmask, nbins, feat_owner, feat_inner, owner_stride, out);
                run = 0;
            }
            __syncthreads();
        }
    }
    stamp_end(stamp);
}

// ------------------------------------------------------------------------------------
// Sparse rows (BASELINE config 4: 10 000 features at ~1 % density).  The reference's matrix is dense
// (J/dataflow/GBDTCoreData.java:164-168) and cannot hold that shape; here the rows are CSR, an
// absent feature has its default slot, and the histogram is what the dense path would produce.
// ------------------------------------------------------------------------------------
// One CTA per work item (same items as hist_build_k), one warp per row at a time, lane pairs over the
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
        q = ld_gather_int2(gq + rid);
    }
    while (i < it.end) {
        const long long ni = i + nwarps;
        int32_t nrid = 0;
        long long np0 = 0, np1 = 0;
        int2 nq = make_int2(0, 0);
        if (ni < it.end) {
            nrid = rows ? rows[ni] : (int32_t)ni;
            np0 = sp.row_ptr[nrid]; np1 = sp.row_ptr[nrid + 1];
            nq = ld_gather_int2(gq + nrid);
        }
        tg += q.x; th += q.y;
        // two neighbouring lanes per entry, one for g and one for h: the pair lies in one 32 B sector of the
        // histogram, so a reduction instruction covers 16 entries in 16 sectors (a lane per entry took two
        // instructions of 32 sectors each -- the sectors, i.e. L1 tag cycles and L2 requests, are what bounds this kernel)
        const long long add = (lane & 1) ? (long long)q.y : (long long)q.x;
        for (long long p = p0 + (lane >> 1); p < p1; p += 16) {
            const unsigned e = sp.nz[p];
            const int f = (int)(e >> 16), slot = (int)(e & 0xffffu);
            if (fea_selected[f]) {
                const long long own = owner_stride ? feat_owner[f] * owner_stride : 0;
                unsigned long long* out = reinterpret_cast<unsigned long long*>(base + own + feat_inner[f]) + 2 * slot + (lane & 1);
                atomicAdd(out, (unsigned long long)add);
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
    // a lane pair per row, one lane for g and one for h (one sector per pair, see hist_csr_k)
    const int c = threadIdx.x & 1;
    for (long long i = it.begin + (threadIdx.x >> 1); i < it.end; i += blockDim.x >> 1) {
        const int32_t rid = rows ? rows[i] : (int32_t)i;
        const int2 q = ld_gather_int2(gq + rid);
        const int slot = (int)ld_gather_u16(colp + rid);
        atomicAdd(out + 2 * slot + c, (unsigned long long)(long long)(c ? q.y : q.x));
    }
}

// K3: sibling = parent - smaller child (HistogramBuilder.substract,
//     J/data/gbdt/HistogramBuilder.java:101-107), exact in the integer domain.
__global__ void hist_subtract_k(const SubItem* __restrict__ items, int64_t len) {
    pdl_sync();
    const SubItem it = items[blockIdx.y];
    if (!it.out) return;                    // an unused entry of a device-planned table
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

cudaError_t launch_grad_absmax(const float* gp, int64_t n, int32_t K, int32_t gid, unsigned* absmax_bits, cudaStream_t st,
                               const int* gid_dev) {
    cudaError_t e = cudaMemsetAsync(absmax_bits, 0, 2 * sizeof(unsigned), st);
    if (e != cudaSuccess) return e;
    grad_absmax_k<<<grid_for(n, 256, kNumSMs * 8), 256, 0, st>>>(gp, n, K, gid, gid_dev, absmax_bits);
    return cudaGetLastError();
}
cudaError_t launch_grad_quantize(const float* gp, int64_t n, int32_t K, int32_t gid, int2* gq,
                                 const unsigned* absmax_bits, GradScale* scale, int bits, cudaStream_t st, const int* gid_dev) {
    grad_quantize_k<<<grid_for(n, 256, kNumSMs * 8), 256, 0, st>>>(gp, n, K, gid, gid_dev, gq, absmax_bits, scale, bits);
    return cudaGetLastError();
}

// The 128 KB of dynamic shared memory are an opt-in PER DEVICE: called by set_bins with the handle's
// device current (one process may drive several GPUs, one thread each).
cudaError_t hist_kernels_init() {
    cudaError_t e = cudaFuncSetAttribute(hist_build_k, cudaFuncAttributeMaxDynamicSharedMemorySize, kHistSmemBytes);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(hist_span_k, cudaFuncAttributeMaxDynamicSharedMemorySize, kHistSpanSmemBytes);
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
    dim3 grid((unsigned)n_items, (unsigned)(S8 / kGroup));
    return launch_chain(hist_build_k, grid, dim3(kHistThreads), kHistSmemBytes, st, bins, S8, gq, rowidx0, rowidx1, items, F, group_mask,
                                                             nbins, feat_owner, feat_inner, owner_stride, seg_out,
                                                             cnt_local, cnt_global);
}

cudaError_t launch_hist_span(const uint8_t* bins, int32_t S8, const int2* gq, const int32_t* rowidx0, const int32_t* rowidx1,
                             const BuildNode* nodes, const int* n_nodes_dev, int groups, int32_t F, const uint32_t* group_mask,
                             const int32_t* nbins, const int32_t* feat_owner, const int64_t* feat_inner, long long owner_stride,
                             long long* const* seg_out, cudaStream_t st, unsigned long long* stamp) {
    const int per_group = std::max(1, kNumSMs / std::max(1, groups));
    dim3 grid((unsigned)per_group, (unsigned)groups);
    static const int use_tma = [] { const char* e = getenv("YTK_GBDT_HIST_TMA"); return e ? atoi(e) : 0; }();
    return launch_chain(hist_span_k, grid, dim3(kHistThreads), use_tma ? kHistSpanSmemBytes : kHistSmemBytes, st, bins, S8, gq, rowidx0, rowidx1, nodes, n_nodes_dev,
                        F, group_mask, nbins, feat_owner, feat_inner, owner_stride, seg_out, stamp, use_tma);
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

// CTAs per segment for the streaming kernels over histogram segments: 64 is plenty for a 114 KB
// histogram; a few nodes with 20 MB histograms (sparse rows) need the whole GPU
static inline int seg_blocks(int n_segments) { return std::max(64, kNumSMs * 8 / std::max(1, n_segments)); }

cudaError_t launch_hist_subtract(const SubItem* items, int n_items, int64_t len, cudaStream_t st) {
    if (n_items <= 0) return cudaSuccess;
    dim3 grid((unsigned)grid_for(len, 256, seg_blocks(n_items)), (unsigned)n_items);
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
    dim3 grid((unsigned)grid_for(len, 256, seg_blocks(n)), (unsigned)n);
    return launch_chain(zero_segments_k, grid, dim3(256), 0, st, dst, len);
}
cudaError_t launch_copy_segments(const long long* src, long long* const* dst, int n, int64_t len, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    dim3 grid((unsigned)grid_for(len, 256, seg_blocks(n)), (unsigned)n);
    copy_segments_k<<<grid, 256, 0, st>>>(src, dst, len);
    return cudaGetLastError();
}

}  // namespace ytk
