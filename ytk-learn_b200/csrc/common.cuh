// Shared declarations of libytkgbdt's device side (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "ytk_gbdt.h"

namespace ytk {

constexpr int kGroup = 32;        // features per group = lanes of a warp
constexpr int kMaxBins = 256;     // uint8 slots
constexpr int kMaxWideBins = 65536;   // features with more than kMaxBins slots ("wide") keep uint16 slots
constexpr int kHistThreads = 1024;
constexpr int kHistWarps = kHistThreads / 32;
// per-CTA shared histogram: 32-bit words [bin][{lo,hi}][feature] for g, then the same for h:
// the low-16-bit and high limbs of the fixed-point values, accumulated carry-free
constexpr int kHistSmemBytes = kMaxBins * 4 * kGroup * 4;                 // 128 KB
constexpr int kNumSMs = 148;
constexpr int kDefaultFixedBits = 30;           // |fixed-point value| < 2^30
constexpr int kHistMaxRowsPerItem = 65536;      // so that 16-bit limb sums fit 32-bit words

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

// ---- work tables: one launch serves every node of a batch (a tree level, or one node) ----

// A run of rows rowidx[buf][begin,end) (buf < 0: rows begin..end-1) whose sums go to segment `seg`.
// cseg >= 0 defers the range to the device: [begin,end) is then the interval of split segment cseg
// (the parent) and the item means "rows [chunk*rows_per_chunk, +rows_per_chunk) of its SMALLER
// child", resolved from the partition counts still on the device -- so the host need not wait for
// the partition before it queues the histogram kernels.
struct RowItem { long long begin, end; int buf; int seg; int cseg; int chunk; long long rows_per_chunk; };

// one node being split: rows rowidx[buf][begin,end) move to rowidx[1-buf] (buf < 0 -> buffer 0)
struct SplitSeg {
    long long begin, end;
    int buf;          // source buffer (-1 identity)
    int fid;          // split feature; < 0: `src` is a byte mask over rows (row sampling)
    float cond;       // go left iff (float)slot < cond
    int tile0;        // first tile of this segment in the tile table
    int ntiles;
    int wide;         // split feature is wide: its column in WideCols + 1; else 0
};

// Slots of the wide features: uint16, feature-major, one column of `stride` rows per wide feature.
// index[f] = column of feature f, -1 if it has at most kMaxBins slots (null: no wide features).
struct WideCols { uint16_t* data; int64_t stride; const int32_t* index; };

// Sparse rows (CSR): row r holds nz[row_ptr[r] .. row_ptr[r+1]), each entry = feature << 16 | slot,
// sorted by feature; a feature that is absent from a row has the slot default_slot[feature] (the bin
// of the value the reference fills missing values with).  row_ptr == null: the dense layout.
struct SparseRows { const int64_t* row_ptr; const uint32_t* nz; const int32_t* default_slot; };
constexpr int kMaxSparseFeatures = 65536;

// cseg >= 0: sample_cnt is the global row count of the smaller (which = 0) / larger (which = 1)
// child of split segment cseg, read from the device counts
// sub != null: the node's histogram is hist - sub and the search writes it over hist (tot_parent: the exact totals
// {G, H} of the parent, which hist still holds); tot_out (nullable): where the node's own exact totals go
struct ScanNode {
    const long long* hist; long long sample_cnt; int cseg; int which;
    const long long* sub; const long long* tot_parent; long long* tot_out;
};
struct alignas(16) SplitPart { float lc; int fid, lo, hi; long long lh; long long pad; };   // best split of one slice of a node's features (lh: hessian left of it, fixed point)
struct SubItem { const long long* parent; const long long* small; long long* out; };
struct PosItem { long long begin, end; int buf; int nid; };
struct LeafItem { long long begin, end; int buf; float value; };      // rows of one leaf and its value
struct TreeNodeDev { int left, right, fid; float info; };             // device copy of a tree for row walks

// ---- launchers (rows.cu, hist.cu, split.cu, partition.cu); all asynchronous on `st`, return cudaGetLastError() ----

// both also write the feature-major copy col[f * col_stride + row] the partition gathers from
// (`wide.data` is passed already advanced to the chunk's first row, like `bins` and `col`)
cudaError_t launch_repack_rows(const int32_t* x, int64_t nrows, int32_t stride, int32_t F, const int32_t* nbins,
                               uint8_t* bins, int32_t S8, uint8_t* col, int64_t col_stride, WideCols wide,
                               int64_t row_base, int* bad_flag, cudaStream_t st);
cudaError_t launch_repack_rows_u8(const uint8_t* x, int64_t nrows, int32_t stride, int32_t F, const int32_t* nbins,
                                  uint8_t* bins, int32_t S8, uint8_t* col, int64_t col_stride, WideCols wide,
                                  int64_t row_base, int* bad_flag, cudaStream_t st);
cudaError_t launch_quantize_rows(const int32_t* xbits, int64_t nrows, int32_t stride, int32_t F,
                                 const float* cuts, const int32_t* start, uint8_t* bins, int32_t S8,
                                 uint8_t* col, int64_t col_stride, WideCols wide, int64_t row_base, cudaStream_t st);
cudaError_t launch_unpack_rows(const uint8_t* bins, int32_t S8, int64_t nrows, int32_t F, WideCols wide, int32_t* out,
                               cudaStream_t st);

// (g,h) float pairs of class gid -> fixed-point int2 (|value| < 2^bits) + scales, in two steps: the
// largest |g| and |h| of the shard as float bits in absmax_bits[2] (with several GPUs the host takes
// the maximum over the ranks in between: all ranks must use the same scales), then the conversion
// gid_dev != null: the class index is read on the device (launches recorded in a CUDA graph)
cudaError_t launch_grad_absmax(const float* gp, int64_t n, int32_t K, int32_t gid, unsigned* absmax_bits, cudaStream_t st,
                               const int* gid_dev = nullptr);
cudaError_t launch_grad_quantize(const float* gp, int64_t n, int32_t K, int32_t gid, int2* gq,
                                 const unsigned* absmax_bits, GradScale* scale, int bits, cudaStream_t st,
                                 const int* gid_dev = nullptr);

// K2: histograms of n_items row runs; item -> segment -> seg_out[seg] (int64, pre-zeroed); feature f's
// bins at seg_out[seg] + feat_owner[f]*owner_stride + feat_inner[f] + 2*bin + {0,1}.
cudaError_t launch_hist_build(const uint8_t* bins, int32_t S8, const int2* gq, const int32_t* rowidx0,
                              const int32_t* rowidx1, const RowItem* items, int n_items, int32_t F,
                              const uint32_t* group_mask, const int32_t* nbins, const int32_t* feat_owner,
                              const int64_t* feat_inner, long long owner_stride, long long* const* seg_out,
                              const PartitionCounts* cnt_local, const PartitionCounts* cnt_global, cudaStream_t st);
// Level planned on the device (level.cu): the build nodes of a level with a running cost prefix
// (cost of a node = its rows + kHistFlushCost), walked by hist_span_k in spans of equal cost.
// cost of a node = its rows + flush_cost: what a node boundary inside a span costs (flush + re-zero of the shared
// histogram, the dependent loads that start the next node's rows, two barriers), in rows' worth of time
constexpr long long kHistFlushCost = 6144;     // measured on the Higgs shape when a flush took 14-21 us: 512 -> 1.09 ms of histogram time per tree, 1536 -> 1.05, 3000 -> 1.01, 6000 and 12000 -> 0.97; with the 6-7 us flush 2048..8192 all give 0.85
struct BuildNode { long long begin, end; long long cost_end; long long cost; int buf; int seg; };
cudaError_t launch_hist_span(const uint8_t* bins, int32_t S8, const int2* gq, const int32_t* rowidx0, const int32_t* rowidx1,
                             const BuildNode* nodes, const int* n_nodes_dev, int groups, int32_t F, const uint32_t* group_mask,
                             const int32_t* nbins, const int32_t* feat_owner, const int64_t* feat_inner, long long owner_stride,
                             long long* const* seg_out, cudaStream_t st, unsigned long long* stamp = nullptr);
// the wide features' share of the same work items: 64-bit global reductions, one CTA per (item, feature)
cudaError_t launch_hist_wide(WideCols wide, const int32_t* wide_fids, int n_wide, const int2* gq, const int32_t* rowidx0,
                             const int32_t* rowidx1, const RowItem* items, int n_items, const int32_t* feat_owner,
                             const int64_t* feat_inner, long long owner_stride, long long* const* seg_out,
                             const PartitionCounts* cnt_local, const PartitionCounts* cnt_global, cudaStream_t st);
// Sparse rows: every stored entry of the items' rows goes into the node's histogram with 64-bit global
// reductions, the rows' (g,h) totals into tot[2*seg + {0,1}]; then, per node and selected feature, the
// default slot receives what the stored entries did not account for ("fill-value bin by subtraction").
cudaError_t launch_hist_csr(SparseRows sparse, const uint8_t* fea_selected, const int2* gq, const int32_t* rowidx0,
                            const int32_t* rowidx1, const RowItem* items, int n_items, const int32_t* feat_owner,
                            const int64_t* feat_inner, long long owner_stride, long long* const* seg_out, long long* tot,
                            const PartitionCounts* cnt_local, const PartitionCounts* cnt_global, cudaStream_t st);
cudaError_t launch_csr_default_fix(SparseRows sparse, const uint8_t* fea_selected, int32_t F, const int32_t* start,
                                   const int32_t* feat_owner, const int64_t* feat_inner, long long owner_stride,
                                   long long* const* seg_out, int n_segs, const long long* tot, cudaStream_t st);
// checks one uploaded chunk (entries sorted by feature inside a row, features / slots in range) and packs it;
// cuts != null: `slots` are float bits of original values, binned with the A1 rule first
cudaError_t launch_csr_pack(const int64_t* row_ptr, const int32_t* fids, const int32_t* slots, int64_t nrows, int32_t F,
                            const int32_t* nbins, const float* cuts, const int32_t* start, uint32_t* nz_out, int* bad_flag,
                            cudaStream_t st);
cudaError_t launch_csr_to_dense(SparseRows sparse, int64_t row0, int64_t nrows, int32_t F, int32_t* out, cudaStream_t st);
// per-device opt-ins of the histogram kernels (call with the device current, once per handle)
cudaError_t hist_kernels_init();
cudaError_t partition_kernels_init();
// rows one hist work item should cover so that `total_rows` split into about two waves of CTAs
int64_t hist_rows_per_item(int64_t total_rows, int groups);

cudaError_t launch_hist_subtract(const SubItem* items, int n_items, int64_t len, cudaStream_t st);
cudaError_t launch_hist_to_double(const long long* hist, int64_t len, const GradScale* scale, double* out,
                                  cudaStream_t st);
cudaError_t launch_hist_from_double(const double* in, int64_t len, const GradScale* scale, long long* hist,
                                    cudaStream_t st);

// K4/K5: node totals + split scan + arg-max over the features [f_from, f_to) this rank owns, one CTA per node.
cudaError_t launch_split_scan(const ScanNode* nodes, int n_nodes, const int32_t* start, const int64_t* feat_off,
                              const uint8_t* fea_selected, int32_t f_from, int32_t f_to, int32_t first_fea,
                              const GradScale* scale, GainParams gp, const PartitionCounts* cnt_global,
                              SplitPart* parts /* n_nodes * split_scan_slices() scratch */, ytk_split* out,
                              ytk_split* mirror /* nullable: a second copy of the records (pinned host memory) */,
                              unsigned int* done /* n_nodes zero-initialised counters; left zero */, cudaStream_t st,
                              long long* lh_out = nullptr /* [n_nodes][2]: hessian left of the best split, of the node (fixed point) */,
                              unsigned long long* stamp = nullptr /* {~start, end} on the global timer, see device_util.cuh */);
int split_scan_slices(int32_t f_from, int32_t f_to);

// K6: stable partition of every segment by (float)bin(row,fid) < cond (or by a byte mask)
constexpr int kPartThreads = 256;
constexpr int kPartItems = 8;
constexpr int kPartTile = kPartThreads * kPartItems;
// src: col_stride > 0 -> feature-major slots src[fid * col_stride + row]; else row-major with S8 bytes per row
cudaError_t launch_partition(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0, int32_t* rowidx1,
                             const SplitSeg* segs, int n_segs, const int* tile_seg, int n_tiles,
                             uint8_t* flags, int32_t* block_cnt, PartitionCounts* counts, cudaStream_t st,
                             SparseRows sparse = SparseRows{nullptr, nullptr, nullptr});

constexpr int kFusedThreads = 512;
constexpr int kFusedItems = 16;
constexpr int kFusedTile = kFusedThreads * kFusedItems;     // larger tiles: the chain over a node's tiles is the critical path
// single pass (chained scan over a node's tiles); right rows come out in reverse input order.
// tile_state: n_tiles words that need no clearing; ticket: one zero-initialised word; epoch: a value
// (< 2^30) no earlier launch on the same tile_state used
cudaError_t launch_partition_fused(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0, int32_t* rowidx1,
                                   const SplitSeg* segs, int n_segs, const int* tile_seg, int n_tiles,
                                   unsigned long long* tile_state, unsigned int* ticket, unsigned long long epoch,
                                   PartitionCounts* counts, cudaStream_t st,
                                   SparseRows sparse = SparseRows{nullptr, nullptr, nullptr});

// The next level's partition planned on the device (level-wise growth): slot 2k / 2k+1 =
// left / right child of this level's job k.  jobs[k]: which of the current segments / count
// entries the job is, the index of its children's record pair (-1: leaves), the children's buffer.
struct SpecJob { int seg; int cseg; int scan; int child_buf; };
cudaError_t launch_partition_planned(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0,
                                     int32_t* rowidx1, const SplitSeg* cur_segs, const PartitionCounts* cur_counts,
                                     const PartitionCounts* cur_counts_global, const SpecJob* jobs, int m,
                                     const ytk_split* recs, double min_split_loss,
                                     SplitSeg* out_segs, int* out_tile_seg, int* n_tiles_dev, int grid_bound,
                                     unsigned long long* tile_state, unsigned int* ticket, unsigned long long epoch,
                                     PartitionCounts* next_counts, cudaStream_t st,
                                     SparseRows sparse = SparseRows{nullptr, nullptr, nullptr});

// ... and the root's own partition, planned from its record (one slot)
cudaError_t launch_partition_planned_root(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0,
                                          int32_t* rowidx1, long long begin, long long end, int buf, const ytk_split* rec,
                                          double min_split_loss, SplitSeg* out_segs, int* out_tile_seg, int* n_tiles_dev,
                                          int grid_bound, unsigned long long* tile_state, unsigned int* ticket,
                                          unsigned long long epoch, PartitionCounts* next_counts, cudaStream_t st,
                                          SparseRows sparse = SparseRows{nullptr, nullptr, nullptr});

// K7: exact fixed-point sums of (g,h) over row runs -> out[2*seg + {0,1}] (pre-zeroed)
cudaError_t launch_leaf_sums(const int2* gq, const int32_t* rowidx0, const int32_t* rowidx1, const RowItem* items,
                             int n_items, long long* out, const PartitionCounts* cnt_local,
                             const PartitionCounts* cnt_global, cudaStream_t st);

cudaError_t launch_fill_positions(const int32_t* rowidx0, const int32_t* rowidx1, const PosItem* items, int n_items,
                                  int32_t* node_of_row, cudaStream_t st);
cudaError_t launch_fill_i32(int32_t* p, int64_t n, int32_t v, cudaStream_t st);
cudaError_t launch_zero_segments(long long* const* dst, int n, int64_t len, cudaStream_t st);
cudaError_t launch_copy_segments(const long long* src, long long* const* dst, int n, int64_t len, cudaStream_t st);

// ---- collectives over NVLink peer memory (peer.cu) ----
constexpr int kMaxPeers = 64;
constexpr int kPeerFlagSets = 8;                   // host-planned path: counts, histogram staging, split records, leaf sums; 4..7: the same for the device-planned level loop
constexpr size_t kPeerFlagBytes = kPeerFlagSets * kMaxPeers * 8;
struct PeerTable {
    int rank, world;
    long long timeout_cycles;   // a wait for a peer gives up after this many SM clocks (YTK_GBDT_PEER_TIMEOUT_MS)
    char* xchg[16];        // every rank's exchange buffer: flags, then the slot areas
    char* hist[16];        // every rank's histogram staging buffer
};
cudaError_t launch_peer_allreduce(const PeerTable& pt, const void* src, size_t area_off, size_t slot_bytes, size_t words,
                                  int flag_set, unsigned long long epoch, long long* out, int* err, cudaStream_t st);
// raises this rank's "staging complete" flag, waits for everybody's, sums the owned segments out of all staging buffers
// all-gather + the needReplace / MAX merge of the split records: merged record k -> out_dev[k], out_host[k]
cudaError_t launch_peer_allgather_merge(const PeerTable& pt, const ytk_split* src, size_t area_off, size_t slot_bytes, int m,
                                        int flag_set, unsigned long long epoch, ytk_split* out_dev, ytk_split* out_host,
                                        int* err, cudaStream_t st);
cudaError_t launch_peer_pull_reduce(const PeerTable& pt, int m, long long seg_max, int flag_set, unsigned long long epoch,
                                    long long* const* dst, int* err, cudaStream_t st);

// §8(f) row 1: score update + loss + gradient pairs on the device
cudaError_t launch_add_leaf_values(const int32_t* rowidx0, const int32_t* rowidx1, const LeafItem* items, int n_items,
                                   float* score, int32_t K, int32_t gid, cudaStream_t st);
cudaError_t launch_walk_rows(const uint8_t* bins, int32_t S8, WideCols wide, const int32_t* rows, long long begin, long long end,
                             const TreeNodeDev* nodes, float* score, int32_t K, int32_t gid, cudaStream_t st,
                             SparseRows sparse = SparseRows{nullptr, nullptr, nullptr});
// test rows: original values (float bits), real-valued thresholds in nodes[].info of inner nodes, leaf values in leaves
cudaError_t launch_score_values(const int32_t* xbits, long long n, int32_t stride, const TreeNodeDev* nodes,
                                const int32_t* default_child, float* score, int32_t K, int32_t gid, cudaStream_t st);
int loss_grad_blocks();
// loss_kind 0 sigmoid, 1 l2, 2 softmax (y one-hot [n][K]); block_loss: loss_grad_blocks() doubles of scratch
cudaError_t launch_loss_grad(int loss_kind, const float* score, const float* init, const float* y, const float* w,
                             long long n, int32_t K, float* gp, double* block_loss, double* loss_out, cudaStream_t st);

}  // namespace ytk
