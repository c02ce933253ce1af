// Shared declarations of libytkgbdt's device side (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "ytk_gbdt.h"

namespace ytk {

constexpr int kGroup = 32;        // features per group = lanes of a warp
constexpr int kMaxBins = 256;     // uint8 slots
constexpr int kHistThreads = 1024;
constexpr int kHistWarps = kHistThreads / 32;
// per-CTA shared histogram: [bin][component][lane] 32-bit words, components
// {g_lo, g_hi, h_lo, h_hi} of two 64-bit fixed-point accumulators
constexpr int kHistSmemBytes = kMaxBins * 4 * kGroup * 4;                 // 128 KB
constexpr int kStageBytes = kHistWarps * 32 * 16;                         // per-warp row staging
constexpr int kNumSMs = 148;

// Gradient fixed-point scales of the current tree: value = integer * inv[.]
struct GradScale {
    double inv_g, inv_h;   // 2^-k
    double mul_g, mul_h;   // 2^k
};

// Gain parameters (UpdateStrategy's fields, J/optimizer/gbdt/UpdateStrategy.java:36-48)
struct GainParams {
    double min_child_weight, max_abs_leaf_val, l1, l2;
    double min_split_loss;
    long long min_split_samples;
};

struct PartitionCounts { long long left, right; };

// ---- launchers (kernels.cu); all asynchronous on `st`, return cudaGetLastError() ----

cudaError_t launch_repack_rows(const int32_t* x, int64_t nrows, int32_t stride, int32_t F,
                               uint8_t* bins, int32_t S8, int* bad_flag, cudaStream_t st);
cudaError_t launch_quantize_rows(const int32_t* xbits, int64_t nrows, int32_t stride, int32_t F,
                                 const float* cuts, const int32_t* start, uint8_t* bins, int32_t S8,
                                 cudaStream_t st);
cudaError_t launch_unpack_rows(const uint8_t* bins, int32_t S8, int64_t nrows, int32_t F, int32_t* out,
                               cudaStream_t st);

// (g,h) float pairs of class gid -> fixed-point int2 + scales
cudaError_t launch_grad_quantize(const float* gp, int64_t n, int32_t K, int32_t gid, int2* gq,
                                 unsigned* absmax_bits /*2, scratch*/, GradScale* scale, cudaStream_t st);

// K2: histogram of the rows rowidx[begin,end) (rowidx == nullptr: rows begin..end-1) into
// hist (int64, pre-zeroed), feature f's bins at hist[feat_off[f] + 2*bin + {0,1}].
cudaError_t launch_hist_build(const uint8_t* bins, int32_t S8, const int2* gq, const int32_t* rowidx,
                              int64_t begin, int64_t end, int32_t F, const uint32_t* group_mask,
                              const int32_t* nbins, const int64_t* feat_off, long long* hist,
                              int variant, cudaStream_t st, int* launches);

cudaError_t launch_hist_subtract(const long long* parent, const long long* small, long long* out,
                                 int64_t len, cudaStream_t st);
cudaError_t launch_hist_to_double(const long long* hist, int64_t len, const GradScale* scale, double* out,
                                  cudaStream_t st);
cudaError_t launch_hist_from_double(const double* in, int64_t len, const GradScale* scale, long long* hist,
                                    cudaStream_t st);

// K4/K5: node totals + split scan + arg-max over the features [f_from, f_to) this rank owns.
cudaError_t launch_split_scan(const long long* hist, const int32_t* start, const int64_t* feat_off,
                              const uint8_t* fea_selected, int32_t f_from, int32_t f_to,
                              int32_t first_fea, const GradScale* scale, GainParams gp,
                              long long sample_cnt, const double* totals_in /*nullable: G,H*/,
                              ytk_split* out, cudaStream_t st);

// K6: stable partition of rowidx_in[begin,end) by (float)bin(row,fid) < cond into rowidx_out
cudaError_t launch_partition(const uint8_t* bins, int32_t S8, const int32_t* rowidx_in,
                             int32_t* rowidx_out, int64_t begin, int64_t end, int32_t fid, float cond,
                             int32_t* block_cnt /*scratch*/, PartitionCounts* counts, cudaStream_t st,
                             int* launches);
int64_t partition_scratch_ints(int64_t n);

// K7: exact fixed-point sums of (g,h) over rowidx[begin,end) -> out[2] (pre-zeroed)
cudaError_t launch_leaf_sums(const int2* gq, const int32_t* rowidx, int64_t begin, int64_t end,
                             long long* out, cudaStream_t st);

cudaError_t launch_iota(int32_t* p, int64_t n, cudaStream_t st);
cudaError_t launch_compact_mask(const uint8_t* mask, int64_t n, int32_t* rowidx_out, int32_t* block_cnt,
                                PartitionCounts* counts, cudaStream_t st, int* launches);
cudaError_t launch_fill_positions(const int32_t* rowidx, int64_t begin, int64_t end, int32_t nid,
                                  int32_t* node_of_row, cudaStream_t st);
cudaError_t launch_fill_i32(int32_t* p, int64_t n, int32_t v, cudaStream_t st);

}  // namespace ytk
