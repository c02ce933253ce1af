// Level-wise growth driven by the device: the tables of every kernel of a level (partition segments and
// tiles, histogram build nodes, scan nodes (with their subtractions), leaf-sum items) are functions of the split
// records and child counts the device already holds, so the device plans them itself and the host
// enqueues a whole tree without waiting (engine.cu).  Shared by level.cu (kernels) and engine.cu (host).
//
// Nodes are addressed by HEAP position while the tree grows: the node at depth d, position s among the
// 2^d possible nodes of its level, is heap node 2^d - 1 + s; its children are 2i+1 and 2i+2.  The ids
// the reference gives its nodes (Tree.AllocTreeNode order, J/data/gbdt/Tree.java:58-73) are assigned by
// the host afterwards, when it walks the records in the reference's polling order.
#pragma once

#include "common.cuh"

namespace ytk {

constexpr int kLvActive = 1;        // the node exists in the tree
constexpr int kLvSplit = 2;         // it is split at its level (its children exist)
constexpr int kLvSmallLeft = 4;     // the child whose histogram is built from rows is the left one

struct LvNode {
    long long begin, end;           // its rows on this GPU: rowidx[buffer of its depth][begin, end)
    long long cnt_l, cnt_r;         // local rows of its children (valid once split)
    long long gcnt_l, gcnt_r;       // global rows of its children
    long long sample_cnt;           // global rows of the node
    int flags;
    int slot;                       // histogram slot of the node (-1: none)
};

// Everything a level's kernels read, in one struct passed by value.
struct LvTables {
    LvNode* nodes;                  // [2^(D+1) - 1] heap
    ytk_split* recs;                // [heap] best split + totals of every searched node
    long long* left_h;              // [heap][2] fixed-point hessian of the best split's left side, of the node
    long long* leaf;                // [heap][2] fixed-point (g,h) sums of the SMALLER child of a node whose children are leaves
    SplitSeg* segs;                 // [slots of a level] partition segments
    int* tile_seg;                  // partition tile -> segment
    int* n_tiles;
    unsigned int* ticket;           // of the partition kernel
    unsigned long long* epoch;      // of its chained-scan state: bumped by the planner before every partition
    PartitionCounts* counts;        // [slots] local child counts of the level's partition
    PartitionCounts* gcounts;       // [slots] global (== counts on one GPU)
    BuildNode* build;               // [slots] histogram build nodes of the level
    int* n_build;
    ScanNode* scans;                // [2 * slots] (the larger child's entry asks for parent - smaller sibling)
    long long* tot;                 // [heap][2] exact fixed-point totals {G, H} of every searched node
    long long** seg_out;            // [slots] where hist_span_k accumulates node `seg`
    long long** seg_final;          // [slots] world > 1: the slot the reduced histogram of node `seg` goes to (null: none)
    ytk_split* recs_local;          // [2 * slots] world > 1: this rank's records of the level's children, before the merge
    long long* lh_local;            // [2 * slots][2]
    long long* pool;                // histogram slots, seg_max words each
    long long* staging;             // world > 1: [owner][slot][seg_max]
    long long seg_max;
    int* gid;                       // the class whose tree is being built (read by the gradient kernels)
    long long flush_cost;           // cost of a node boundary in a span, in rows (kHistFlushCost; YTK_GBDT_FLUSH_COST)
    int world;
};

// ---- exchanges of the device-driven level loop over peer memory (peer.cu) ----------------------------
// `phase`: kXAll = one kernel per exchange with flags between GPUs (one process per GPU); a clique of
// handles on ONE device (tests: the ranks' kernels run one after the other on one stream) runs
// kXPush for every rank, then kXReduce for every rank, and needs no flags.
constexpr int kXPush = 1, kXWait = 2, kXReduce = 4, kXAll = 7;
// largest |g|, |h| of all ranks (float bits): absmax[2] in place
// `xctr`: the exchange kind's epoch counter in device memory ([0] the last completed epoch, [1] scratch): every
// rank runs the same sequence of exchanges, so the counters agree, and a launch recorded in a CUDA graph needs no
// new argument per tree.
cudaError_t launch_lv_xmax(const PeerTable& pt, unsigned* absmax, size_t area_off, size_t slot_bytes, int flag_set,
                           unsigned long long* xctr, int phase, int* err, cudaStream_t st);
// histogram reduce-scatter of the `m` slots of a level (m = 1, d = -1: the root) + the global child counts
cudaError_t launch_lv_xhist(const PeerTable& pt, LvTables t, int d, int m, size_t area_off, size_t slot_bytes, int flag_set,
                            unsigned long long* xctr, int phase, int* err, cudaStream_t st, unsigned long long* stamp = nullptr);
// split records of `n` searched nodes whose first heap node is `first_heap`: all-gather + needReplace / MAX merge
cudaError_t launch_lv_xrecs(const PeerTable& pt, LvTables t, int n, long long first_heap, size_t area_off, size_t slot_bytes,
                            size_t lh_off, size_t lh_slot_bytes, int flag_set, unsigned long long* xctr, int phase, int* err,
                            cudaStream_t st, unsigned long long* stamp = nullptr,
                            int plan_d = -1 /* >= 0: the reduce step also plans the partition of that level (lv_plan_split_k's work) */,
                            int plan_buf = 0, double min_split_loss = 0.0);
// leaf sums of the 2m children of the last level (summed over ranks, in place) + the global child counts
cudaError_t launch_lv_xleaf(const PeerTable& pt, LvTables t, int d, size_t area_off, size_t slot_bytes, size_t cnt_off,
                            size_t cnt_slot_bytes, int flag_set, unsigned long long* xctr, int phase, int* err, cudaStream_t st,
                            unsigned long long* stamp = nullptr);

// global child counts of split slot s of level d (d = -1: the root's own row count, in gc.left) into the nodes
__device__ __forceinline__ void lv_apply_gcounts(const LvTables& t, int d, int s, PartitionCounts gc) {
    if (d < 0) { t.nodes[0].sample_cnt = gc.left; return; }
    const int p = (1 << d) - 1 + s;
    t.nodes[p].gcnt_l = gc.left; t.nodes[p].gcnt_r = gc.right;
    t.nodes[2 * p + 1].sample_cnt = gc.left;
    t.nodes[2 * p + 2].sample_cnt = gc.right;
}

cudaError_t launch_lv_root(LvTables t, long long begin, long long end, int buf, const PartitionCounts* mask_counts, cudaStream_t st);
// plans the partition of level d (slots = 2^d): which nodes split, their segments and tiles
cudaError_t launch_lv_plan_split(LvTables t, int d, int buf, double min_split_loss, const int32_t* wide_index, int tile_rows,
                                 cudaStream_t st);
// after the partition of level d: children's intervals, smaller child, and the tables of the kernels that
// follow: the build list (smaller children) and -- leaf = false -- scan nodes (the larger child: parent - sibling)
cudaError_t launch_lv_plan_build(LvTables t, int d, int child_buf, bool leaf, cudaStream_t st);
// leaf sums of the last level's build list into t.leaf (zero before)
cudaError_t launch_lv_leaf_sums(LvTables t, int d, const int2* gq, const int32_t* rowidx0, const int32_t* rowidx1, int grid,
                                cudaStream_t st);
// partition_fused_k over device-planned tables: grid_bound >= the number of tiles the planner can produce
cudaError_t launch_partition_fused_planned(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0,
                                           int32_t* rowidx1, const SplitSeg* segs, int n_segs, const int* tile_seg,
                                           const int* n_tiles_dev, int grid_bound, unsigned long long* tile_state,
                                           unsigned int* ticket, const unsigned long long* epoch_dev, PartitionCounts* counts,
                                           cudaStream_t st, SparseRows sparse = SparseRows{nullptr, nullptr, nullptr},
                                           unsigned long long* stamp = nullptr);

}  // namespace ytk
