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
template <typename TIn>
__global__ void repack_rows_k(const TIn* __restrict__ x, int64_t nrows, int32_t stride, int32_t F,
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
            int32_t v = (int32_t)x[r * stride + f];
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
    repack_rows_k<int32_t><<<grid_for(nrows * S8, 256), 256, 0, st>>>(x, nrows, stride, F, nbins, bins, S8, col, col_stride, wide,
                                                                      row_base, bad_flag);
    return cudaGetLastError();
}
cudaError_t launch_repack_rows_u8(const uint8_t* x, int64_t nrows, int32_t stride, int32_t F, const int32_t* nbins,
                                  uint8_t* bins, int32_t S8, uint8_t* col, int64_t col_stride, WideCols wide,
                                  int64_t row_base, int* bad_flag, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    repack_rows_k<uint8_t><<<grid_for(nrows * S8, 256), 256, 0, st>>>(x, nrows, stride, F, nbins, bins, S8, col, col_stride, wide,
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
