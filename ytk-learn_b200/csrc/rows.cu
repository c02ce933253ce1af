// Row storage: slot matrices in (repack, value->bin, CSR packing) and out.
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
// K0: int32 slot matrix (GBDTCoreData.x chunk, J/dataflow/GBDTCoreData.java:164-168)
//     -> uint8 row-major, rows padded to a multiple of 32 features (one 32 B sector per
//     feature group and row).
// ------------------------------------------------------------------------------------
// Both upload kernels (slots in, original values in) write two copies of a row's slots: the row-major matrix the
// histograms read and the feature-major copy the partition gathers from.  A CTA takes kRepackRows rows and one
// 32-feature group at a time: a warp reads one row's 32 values (one line), writes its 32 B sector, and the group
// goes through a transposing tile in shared memory so that the feature-major copy is written in runs of
// kRepackRows consecutive bytes (a byte per thread straight to `col` was one sector per byte: 170 GB/s).
constexpr int kRepackRows = 128;
template <typename SlotOf>    // SlotOf(row, feature) -> the (validated) slot of a real feature
__device__ __forceinline__ void repack_tile(SlotOf slot_of, int64_t nrows, int32_t F, uint8_t* __restrict__ bins, int32_t S8,
                                            uint8_t* __restrict__ col, int64_t col_stride, WideCols wide, int64_t row_base) {
    __shared__ uint8_t tile[kGroup][kRepackRows + 4];
    const int64_t r0 = blockIdx.x * (int64_t)kRepackRows;
    const int nr = (int)min((int64_t)kRepackRows, nrows - r0);
    for (int g0 = 0; g0 < S8; g0 += kGroup) {
        for (int e = threadIdx.x; e < nr * kGroup; e += blockDim.x) {
            const int rl = e / kGroup, fl = e % kGroup;
            const int64_t r = r0 + rl;
            const int32_t f = g0 + fl;
            // padding columns get row-varying filler: the histogram kernel accumulates them too (and drops
            // them at flush), and identical bins in consecutive rows would make consecutive atomics hit
            // the same shared-memory word
            uint8_t b = (uint8_t)((r + row_base) * 37 + f * 11);
            if (f < F) {
                const int32_t v = slot_of(r, f);
                const int wi = wide.index ? wide.index[f] : -1;
                if (wi >= 0) wide.data[wi * wide.stride + r] = (uint16_t)v;     // its byte in the uint8 matrices stays filler
                else b = (uint8_t)v;
            }
            bins[r * S8 + f] = b;
            tile[fl][rl] = b;
        }
        __syncthreads();
        for (int e = threadIdx.x; e < kGroup * kRepackRows; e += blockDim.x) {
            const int fl = e / kRepackRows, rl = e % kRepackRows;
            const int32_t f = g0 + fl;
            if (f < F && rl < nr && !(wide.index && wide.index[f] >= 0)) col[f * col_stride + r0 + rl] = tile[fl][rl];
        }
        __syncthreads();
    }
}

template <typename TIn>
__global__ void __launch_bounds__(256)
repack_rows_k(const TIn* __restrict__ x, int64_t nrows, int32_t stride, int32_t F,
              const int32_t* __restrict__ nbins, uint8_t* __restrict__ bins, int32_t S8,
              uint8_t* __restrict__ col, int64_t col_stride, WideCols wide, int64_t row_base,
              int* bad_flag) {
    repack_tile([&](int64_t r, int32_t f) -> int32_t {
        int32_t v = (int32_t)x[r * stride + f];
        if (v < 0 || v >= nbins[f]) { *bad_flag = 1; v = 0; }
        return v;
    }, nrows, F, bins, S8, col, col_stride, wide, row_base);
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

__global__ void __launch_bounds__(256)
quantize_rows_k(const int32_t* __restrict__ xbits, int64_t nrows, int32_t stride, int32_t F,
                const float* __restrict__ cuts, const int32_t* __restrict__ start,
                uint8_t* __restrict__ bins, int32_t S8, uint8_t* __restrict__ col, int64_t col_stride,
                WideCols wide, int64_t row_base) {
    repack_tile([&](int64_t r, int32_t f) -> int32_t {
        const float v = __int_as_float(xbits[r * stride + f]);
        const int32_t s0 = start[f];
        return find_bin(cuts + s0, start[f + 1] - s0, v);
    }, nrows, F, bins, S8, col, col_stride, wide, row_base);
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

// upload: range and order checks, (feature, slot) packed into one word
// (cuts != null: `slots` holds float bits of original values, binned here with find_bin -- A1 for sparse input)
__global__ void csr_pack_k(const int64_t* __restrict__ row_ptr, const int32_t* __restrict__ fids,
                           const int32_t* __restrict__ slots, int64_t nrows, int32_t F, const int32_t* __restrict__ nbins,
                           const float* __restrict__ cuts, const int32_t* __restrict__ start,
                           uint32_t* __restrict__ nz_out, int* bad_flag) {
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < nrows; r += (int64_t)gridDim.x * blockDim.x) {
        const int64_t p0 = row_ptr[r] - row_ptr[0], p1 = row_ptr[r + 1] - row_ptr[0];
        int prev = -1;
        for (int64_t p = p0; p < p1; p++) {
            const int f = fids[p];
            if (f <= prev || f >= F) { *bad_flag = 1; nz_out[p] = 0; continue; }
            const int sl = cuts ? find_bin(cuts + start[f], nbins[f], __int_as_float(slots[p])) : slots[p];
            if (sl < 0 || sl >= nbins[f]) { *bad_flag = 1; nz_out[p] = 0; continue; }
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

cudaError_t launch_repack_rows(const int32_t* x, int64_t nrows, int32_t stride, int32_t F, const int32_t* nbins,
                               uint8_t* bins, int32_t S8, uint8_t* col, int64_t col_stride, WideCols wide,
                               int64_t row_base, int* bad_flag, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    repack_rows_k<int32_t><<<(unsigned)((nrows + kRepackRows - 1) / kRepackRows), 256, 0, st>>>(x, nrows, stride, F, nbins, bins, S8, col, col_stride, wide,
                                                                      row_base, bad_flag);
    return cudaGetLastError();
}
cudaError_t launch_repack_rows_u8(const uint8_t* x, int64_t nrows, int32_t stride, int32_t F, const int32_t* nbins,
                                  uint8_t* bins, int32_t S8, uint8_t* col, int64_t col_stride, WideCols wide,
                                  int64_t row_base, int* bad_flag, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    repack_rows_k<uint8_t><<<(unsigned)((nrows + kRepackRows - 1) / kRepackRows), 256, 0, st>>>(x, nrows, stride, F, nbins, bins, S8, col, col_stride, wide,
                                                                      row_base, bad_flag);
    return cudaGetLastError();
}
cudaError_t launch_quantize_rows(const int32_t* xbits, int64_t nrows, int32_t stride, int32_t F,
                                 const float* cuts, const int32_t* start, uint8_t* bins, int32_t S8,
                                 uint8_t* col, int64_t col_stride, WideCols wide, int64_t row_base, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    quantize_rows_k<<<(unsigned)((nrows + kRepackRows - 1) / kRepackRows), 256, 0, st>>>(xbits, nrows, stride, F, cuts, start, bins, S8, col,
                                                               col_stride, wide, row_base);
    return cudaGetLastError();
}
cudaError_t launch_unpack_rows(const uint8_t* bins, int32_t S8, int64_t nrows, int32_t F, WideCols wide, int32_t* out,
                               cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    unpack_rows_k<<<grid_for(nrows * F, 256), 256, 0, st>>>(bins, S8, nrows, F, wide, out);
    return cudaGetLastError();
}
cudaError_t launch_csr_pack(const int64_t* row_ptr, const int32_t* fids, const int32_t* slots, int64_t nrows, int32_t F,
                            const int32_t* nbins, const float* cuts, const int32_t* start, uint32_t* nz_out, int* bad_flag,
                            cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    csr_pack_k<<<grid_for(nrows, 128), 128, 0, st>>>(row_ptr, fids, slots, nrows, F, nbins, cuts, start, nz_out, bad_flag);
    return cudaGetLastError();
}
cudaError_t launch_csr_to_dense(SparseRows sparse, int64_t row0, int64_t nrows, int32_t F, int32_t* out, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    csr_to_dense_k<<<grid_for(nrows * F, 256), 256, 0, st>>>(sparse, row0, nrows, F, out);
    return cudaGetLastError();
}

}  // namespace ytk
