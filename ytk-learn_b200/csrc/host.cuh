// Host-side internals of libytkgbdt.so shared by its translation units (tree_maker.cu: handle and
// helpers; grow.cu: make(); boost.cu: the step either side of make(); hooks.cu: per-stage test hooks).
#pragma once

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <deque>
#include <limits>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "level.cuh"
#include "nccl_dyn.h"

namespace ytkh {

// sets the thread's error message (ytk_gbdt_last_error) and returns `code`
int fail(int code, const char* fmt, ...);

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return fail(YTK_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
    } while (0)

#define NK(call)                                                                                   \
    do {                                                                                           \
        ncclResult_t r__ = (call);                                                                 \
        if (r__ != ncclSuccess)                                                                    \
            return fail(YTK_ERR_NCCL, "%s:%d %s: %s", __FILE__, __LINE__, #call,                   \
                        ytk::nccl_api().GetErrorString(r__));                                      \
    } while (0)

#define RET(call)                 \
    do {                          \
        int r__ = (call);         \
        if (r__ != YTK_OK) return r__; \
    } while (0)



inline double now_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

// All work tables of a batch packed into one buffer: one host->device copy per batch instead of
// one per table.
struct Blob {
    std::vector<char> host;
    size_t add(const void* p, size_t bytes) {
        size_t off = (host.size() + 255) & ~size_t(255);
        host.resize(off + bytes);
        if (bytes) memcpy(&host[off], p, bytes);
        return off;
    }
    template <typename T>
    size_t add(const std::vector<T>& v) { return add(v.data(), v.size() * sizeof(T)); }
};

// HistogramPool (J/data/gbdt/HistogramPool.java:190-267): bounded cache nid -> histogram.
// Slots keep their device buffers when recycled, exactly like the Java arrays, so the
// (degenerate) case of a parent evicted by its own child's allocation behaves as there.
struct HistPool {
    struct Item { int nid; long long* buf; bool own_buf = true; };     // own_buf: buf is its own cudaMalloc
    int capacity = 0;
    std::vector<Item*> used;     // allocation order, oldest first
    std::deque<Item*> avail;     // FIFO
    std::vector<Item*> all;

    std::unordered_map<int, Item*> by_key;    // the entries of `used`, by key

    void clear() {               // :200-206
        for (Item* it : used) avail.push_back(it);
        used.clear();
        by_key.clear();
    }
    void rename(int from, int to) {   // in-place reuse of a slot (sibling written over its parent)
        auto f = by_key.find(from);
        if (f == by_key.end()) return;
        Item* it = f->second;
        by_key.erase(f);
        it->nid = to;
        by_key[to] = it;
    }
    long long* get(int nid) {    // getHist :238-245
        auto f = by_key.find(nid);
        return f == by_key.end() ? nullptr : f->second->buf;
    }
    void release(int nid) {      // releaseHist :256-267
        auto f = by_key.find(nid);
        if (f == by_key.end()) return;
        Item* it = f->second;
        by_key.erase(f);
        // `used` keeps allocation order for the eviction rule; released entries are usually recent
        for (size_t i = used.size(); i-- > 0;) {
            if (used[i] == it) { used.erase(used.begin() + i); break; }
        }
        avail.push_back(it);
    }
    // allocHist :209-236; *fresh is set when a new device buffer is needed.  `unbounded`: the
    // reference's capacity (max_leaf_cnt) can never be exceeded by its own allocation order, so
    // no eviction ever happens there; the batched path may hold a level's parents and children
    // at once and simply takes another slot instead.
    Item* alloc(int nid, bool* fresh, bool unbounded = false) {
        Item* it;
        *fresh = false;
        if (!avail.empty()) {
            it = avail.front(); avail.pop_front();
        } else if (unbounded || (int)all.size() < capacity) {
            it = new Item{nid, nullptr};
            all.push_back(it);
            *fresh = true;
        } else {
            it = used.front(); used.erase(used.begin());   // evict the oldest
            by_key.erase(it->nid);
        }
        it->nid = nid;
        used.push_back(it);
        by_key[nid] = it;
        return it;
    }
};

struct Interval { int64_t b = 0, e = 0; int buf = -1; };   // buf -1: identity (rows b..e-1)

struct NodeStats {                 // TreeMakerNodeStats (DataParallelTreeMaker.java:52-81)
    double sumG = 0, sumH = 0;
    float root_gain = 0, node_value = 0;
    int64_t sample_cnt = 0;
    float best_lc = -std::numeric_limits<float>::infinity();
    int best_fid = -1, best_lo = -1, best_hi = -1;
};

// hd: node handle (make_tree); slot: the node's segment in a partition the device planned ahead (-1: none)
struct ExpandNode { int hd; float loss_chg; int depth; int seq; int slot = -1; };

inline int java_float_compare(float a, float b) {   // java.lang.Float.compare
    if (a < b) return -1;
    if (a > b) return 1;
    int32_t ia, ib;
    if (a != a) ia = 0x7fc00000; else memcpy(&ia, &a, 4);
    if (b != b) ib = 0x7fc00000; else memcpy(&ib, &b, 4);
    return ia == ib ? 0 : (ia < ib ? -1 : 1);
}

// Pinned staging for the small work tables: a bump allocator that is rewound whenever the stream
// has been synchronised (twice per batch), so uploads are true async copies and never make the
// host wait for kernels already queued.
struct PinnedArena {
    char* h = nullptr;
    size_t cap = 0, off = 0;
    void* take(size_t bytes, cudaStream_t st) {
        bytes = (bytes + 255) & ~size_t(255);
        if (off + bytes > cap) {
            if (cudaStreamSynchronize(st) != cudaSuccess) return nullptr;     // everything staged so far is consumed
            off = 0;
            if (bytes > cap) {
                if (h) cudaFreeHost(h);
                h = nullptr; cap = 0;
                size_t nc = std::max<size_t>(bytes * 2, size_t(4) << 20);
                if (cudaHostAlloc((void**)&h, nc, cudaHostAllocMapped) != cudaSuccess) return nullptr;
                cap = nc;
            }
        }
        void* p = h + off;
        off += bytes;
        return p;
    }
    void rewind() { off = 0; }
    void release() { if (h) cudaFreeHost(h); h = nullptr; cap = off = 0; }
};

// growable device array + host->device upload of a work table
template <typename T>
struct DevTable {
    T* d = nullptr;        // what kernels are given (= own)
    T* own = nullptr;      // the device array
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n > cap) {
            size_t nc = std::max<size_t>(n, cap * 2);
            nc = std::max<size_t>(nc, 64);
            if (own) { cudaError_t e = cudaFree(own); if (e != cudaSuccess) return e; own = nullptr; cap = 0; }
            cudaError_t e = cudaMalloc(&own, nc * sizeof(T));
            if (e != cudaSuccess) return e;
            cap = nc;
        }
        d = own;
        return cudaSuccess;
    }
    cudaError_t upload(const std::vector<T>& v, cudaStream_t st, PinnedArena& arena) {
        if (v.empty()) return cudaSuccess;
        const size_t bytes = v.size() * sizeof(T);
        void* stage = arena.take(bytes, st);
        if (!stage) return cudaErrorMemoryAllocation;
        memcpy(stage, v.data(), bytes);
        cudaError_t e = reserve(v.size());
        if (e != cudaSuccess) return e;
        return cudaMemcpyAsync(own, stage, bytes, cudaMemcpyHostToDevice, st);
    }
    void release() { if (own) cudaFree(own); own = d = nullptr; cap = 0; }
};

// growable pinned host array for results coming back from the device
template <typename T>
struct PinnedBuf {
    T* h = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        size_t nc = std::max<size_t>(n, cap * 2);
        nc = std::max<size_t>(nc, 64);
        if (h) cudaFreeHost(h);
        h = nullptr; cap = 0;
        cudaError_t e = cudaHostAlloc((void**)&h, nc * sizeof(T), cudaHostAllocMapped);
        if (e == cudaSuccess) cap = nc;
        return e;
    }
    void release() { if (h) cudaFreeHost(h); h = nullptr; cap = 0; }
};


// a tree's launch sequence recorded as a CUDA graph (one GPU), and what the replay must restore on the host
struct EngineGraph {
    unsigned long long sig = 0;
    cudaGraphExec_t exec = nullptr;
    std::vector<int> marks;
    int ev_stage_used = 0, ev_used = 0, launches = 0, rendezvous = 0;
};

// device tables (and pinned mirrors of what comes back) of the device-driven level loop (engine.cu)
struct EngineBuffers {
    std::vector<EngineGraph> graphs;
    int depth = -1;
    size_t tile_cap = 0, mirror_bytes = 0, clear_bytes = 0, heap = 0;
    char* dev = nullptr;
    char* host = nullptr;
    ytk::LvTables t{};
    ytk::LvNode* h_nodes = nullptr;
    ytk_split* h_recs = nullptr;
    long long* h_leaf = nullptr;
    unsigned long long* d_stamps = nullptr;      // {~start, end} of the tree's main kernels on the global timer
    unsigned long long* h_stamps = nullptr;
    int gid_on_device = -1;                      // the class index t.gid holds
};

}  // namespace ytkh

// Host mirror of UpdateStrategy.calcGain / calcNodeValue (tree_maker.cu)
void ytk_host_node_value(const ytk_gbdt_params* p, double G, double H, float* root_gain, float* node_value);

using namespace ytkh;

struct ytk_gbdt {
    ytk_gbdt_params p{};
    int device = 0, rank = 0, world = 1;
    int fixed_bits = ytk::kDefaultFixedBits;
    bool force_sequential = false;
    bool single_pass_partition = true;     // YTK_GBDT_STABLE_PARTITION=1: the three-kernel stable version inside make() too
    unsigned long long* d_tile_state = nullptr;
    size_t tile_state_cap = 0;
    unsigned int* d_ticket = nullptr;
    unsigned long long part_epoch = 0;
    // level-wise growth: the next level's partition planned and started by the device
    bool plan_ahead = true;                 // YTK_GBDT_NO_PLAN_AHEAD=1 turns it off
    bool spin_wait = false;                 // YTK_GBDT_SPIN=1: poll the level's event from user space (measured: no gain)
    bool use_engine = true;                 // YTK_GBDT_NO_ENGINE=1: level-wise growth through the host-planned path too
    bool use_graph = true;                  // YTK_GBDT_NO_GRAPH=1: queue every tree launch by launch
    EngineBuffers eng;
    size_t pool_chunk_slots = 0;            // histogram slots in d_pool_chunk
    std::vector<cudaEvent_t> ev_stage;      // stage boundaries of a device-driven tree
    int ev_stage_used = 0;
    bool recs_merged = false;               // the last split exchange merged the records on the device
    DevTable<ytk::SplitSeg> t_spec_segs[2];
    DevTable<ytk::PartitionCounts> t_spec_counts[2];
    PinnedBuf<ytk::PartitionCounts> h_spec_counts[2];
    DevTable<int> t_spec_tiles;
    int* d_spec_ntiles = nullptr;
    cudaEvent_t ev_level = nullptr;
    bool trace = false;         // YTK_GBDT_TRACE=1: per-batch host/GPU timeline on stderr
    struct TraceRec { int m; double t_pop, t_jobs, t_part, t_tables, t_queued, t_done; cudaEvent_t e0, e1; };
    std::vector<TraceRec> trace_recs;
    int lookahead = 64;         // loss-wise: expansions computed per batch of launches (YTK_GBDT_LOOKAHEAD)
    cudaStream_t st = nullptr;
    ncclComm_t comm = nullptr;

    int F = 0, S8 = 0, B = 0, groups = 0;
    int64_t N = 0;
    std::vector<int32_t> start, nbins, assign;   // assign: avgAssign(F, world)
    int f_from = 0, f_to = 0;
    int64_t seg_max = 0;      // owned-segment length (int64 words), equal on every rank

    uint8_t* d_bins = nullptr;
    uint8_t* d_bins_col = nullptr;     // feature-major copy [F][N] for the partition's one-byte gathers
    float* d_gp = nullptr;
    int2* d_gq = nullptr;
    int32_t* d_rowidx[2] = {nullptr, nullptr};
    uint8_t* d_flags = nullptr;        // partition scratch: one byte per row
    int32_t* d_block_cnt = nullptr;    // partition scratch: one int per tile
    size_t block_cnt_cap = 0;
    uint8_t* d_mask = nullptr;
    int32_t* d_start = nullptr;
    int32_t* d_nbins = nullptr;
    int32_t* d_feat_owner = nullptr;   // rank owning feature f
    int32_t* d_feat_zero = nullptr;    // all zeros: "everything in one flat segment"
    int64_t* d_off_inner = nullptr;    // offset of feature f inside its owner's segment
    int64_t* d_off_flat = nullptr;     // reference layout 2*start[f]
    // sparse rows (CSR) instead of the dense matrices: ytk_gbdt_set_bins_sparse / ytk_gbdt_upload_csr
    bool sparse = false;
    int64_t* d_row_ptr = nullptr;          // [N+1], absolute offsets into d_nz
    uint32_t* d_nz = nullptr;              // feature << 16 | slot, rows sorted by feature
    int32_t* d_default_slot = nullptr;     // [F]
    int64_t nnz_cap = 0, csr_rows_done = 0, csr_nnz_done = 0;
    DevTable<long long> t_node_tot;        // (g,h) totals of the rows of each node of a batch
    DevTable<unsigned int> t_scan_done;    // per node: slices of the split scan that have finished (kept at zero)
    uint32_t* d_group_mask = nullptr;
    // features with more than 256 slots: uint16 columns beside the uint8 matrices
    int n_wide = 0;
    std::vector<int32_t> wide_index;       // per feature: its uint16 column, -1 if it has at most 256 slots
    std::vector<int32_t> wide_sel;         // wide features selected for the current tree
    std::vector<uint8_t> sel_on_device;    // the selection d_fea_sel / d_group_mask / d_wide_fids hold (empty: none yet)
    int sel_first_owned = -1, sel_count = 0;
    uint16_t* d_wide = nullptr;            // [n_wide][N]
    int32_t* d_wide_index = nullptr;       // device copy of wide_index (null when n_wide == 0)
    int32_t* d_wide_fids = nullptr;        // device copy of wide_sel
    uint8_t* d_fea_sel = nullptr;
    long long* d_hist_flat = nullptr;  // hooks
    long long* d_hist_flat2 = nullptr;
    double* d_hist_dbl = nullptr;
    ytk::GradScale* d_scale = nullptr;
    ytk::GradScale* d_scale_hook = nullptr;
    unsigned* d_absmax = nullptr;
    int* d_bad = nullptr;
    int32_t* d_node_of_row = nullptr;
    void* d_stage = nullptr;  size_t stage_bytes = 0;   // upload staging

    // per-batch work tables and result buffers
    DevTable<ytk::RowItem> t_items;        // work items of the build_hist hook
    DevTable<ytk::ScanNode> t_scan;
    DevTable<ytk::SubItem> t_sub;
    DevTable<ytk::PosItem> t_pos;
    DevTable<long long*> t_seg_out;
    DevTable<ytk::PartitionCounts> t_counts;     // [2*m]: local, then global (world > 1)
    DevTable<ytk_split> t_split_local, t_split_all;
    DevTable<ytk::SplitPart> t_parts;
    DevTable<long long> t_leaf, t_hist_local, t_rs_out;
    PinnedBuf<ytk::PartitionCounts> h_counts;
    PinnedBuf<ytk_split> h_split;
    PinnedBuf<long long> h_leaf;
    PinnedBuf<ytk::GradScale> h_scale;
    PinnedArena arena;
    char* d_blob = nullptr;  size_t blob_cap = 0;     // device side of Blob
    char* d_blob_part = nullptr;  size_t blob_part_cap = 0;     // ... of the partition tables, which go up (and start work) first

    // collectives over NVLink peer memory (world > 1, one node): see peer.cu
    bool peer_mode = false;
    ytk::PeerTable pt{};
    char* my_xchg = nullptr; char* my_hist = nullptr;
    size_t hist_cap_words = 0;               // capacity of my_hist in int64 words
    int peer_cap_nodes = 0;                  // nodes per batch the exchange slots can hold
    size_t off_counts = 0, off_recs = 0, off_leaf = 0, slot_counts = 0, slot_recs = 0, slot_leaf = 0;
    size_t off_lh = 0, slot_lh = 0, off_max = 0, slot_max = 0;
    // a clique of `world` handles on ONE device sharing one stream (ytk_gbdt_create_clique): the ranks' kernels run one
    // after the other, the exchanges read the other handles' buffers directly
    bool clique = false;
    bool clique_linked = false;
    ytk_gbdt** clique_members = nullptr;     // [world], owned by rank 0
    unsigned long long epoch[ytk::kPeerFlagSets] = {0, 0, 0, 0};
    unsigned long long* d_xctr = nullptr;    // device-planned level loop: epoch counters of its exchange kinds, {last epoch, scratch} x 4
    int* d_peer_err = nullptr;               // mapped host memory: set by a peer kernel that gave up waiting
    double peer_timeout_ms = 10000.0;        // YTK_GBDT_PEER_TIMEOUT_MS

    // §8(f) row 1 state
    float* d_score = nullptr; float* d_y = nullptr; float* d_w = nullptr; float* d_init = nullptr;
    double* d_block_loss = nullptr; double* d_loss = nullptr;
    int loss_kind = -1;
    std::vector<ytk::TreeNodeDev> last_tree;
    int last_gid = -1;
    int64_t unsampled_begin = -1;      // rows rowidx[0][unsampled_begin, N) were excluded by row_mask
    DevTable<ytk::LeafItem> t_leafval;
    DevTable<ytk::TreeNodeDev> t_tree;

    HistPool pool;
    long long* d_pool_chunk = nullptr;      // the histogram slots allocated up front (one allocation)
    bool grads_set = false, rows_set = false;
    std::vector<std::pair<int, Interval>> leaf_ranges;   // (leaf node id, a run of its rows) of the last tree
    bool tree_made = false;
    int launches = 0;
    std::vector<cudaEvent_t> ev;       // event pairs around histogram builds
    int ev_used = 0;
    std::vector<cudaEvent_t> evc;      // event pairs around collectives (world > 1)
    int evc_used = 0;
};

namespace ytkh {

// cseg < 0: the rows of interval v; else: the smaller child of split segment cseg, whose parent has interval v
struct HistJob { long long* dst; int cseg; Interval v; int child_buf; };

struct HistPlan {
    int m = 0;
    std::vector<ytk::RowItem> items;
    std::vector<long long*> outs_local;    // where the kernel accumulates (the slot, or the local pre-reduce buffer)
    std::vector<long long*> outs_final;    // world > 1: the slots the reduce-scatter result is copied to
    size_t o_items = 0, o_local = 0, o_final = 0;
};

struct PartPlan {
    int m = 0;
    std::vector<ytk::SplitSeg> segs;
    std::vector<int> tile_seg;
    size_t o_segs = 0, o_tiles = 0;
    bool single_pass = false;       // tiles sized for the chained-scan kernel
    const ytk::SplitSeg* d_segs = nullptr;      // set by partition_commit
    const int* d_tiles = nullptr;
};

// handles of the node and its children; k: position in its batch; slot: as ExpandNode's
struct SplitJob { int hd, l, r, depth; bool terminal; int k = 0; int slot = -1; };

// What splitting a node produces, kept until the split is applied (make_tree)
struct Expansion {
    bool have = false;          // computed and not yet applied
    bool by_hist = false;       // children's histograms + split records (else: leaf sums of the smaller child)
    bool placeholders = false;  // pool entries of the children still carry placeholder keys
    ytk::PartitionCounts cl{0, 0}, cg{0, 0};
    ytk_split rec[2];           // smaller, larger child
    long long leaf[2] = {0, 0};
    int hl = -1, hr = -1;       // handles of the children
};


template <typename T>
const T* blob_ptr(const ytk_gbdt* h, size_t off) { return reinterpret_cast<const T*>(h->d_blob + off); }

// ---- helpers defined in tree_maker.cu ------------------------------------------------------------
ytk::GainParams gain_params(const ytk_gbdt_params& p);
int ensure_stage(ytk_gbdt* h, size_t bytes);
void free_all(ytk_gbdt* h);
ytk::SparseRows sparse_rows(const ytk_gbdt* h);
ytk::WideCols wide_cols(const ytk_gbdt* h, int64_t row0 = 0);
int set_features(ytk_gbdt* h, const int32_t* fea_index, int32_t n_fea, int* first_owned, int* n_sel);
// fixed-point copies of class gid's gradient pairs, with scales common to all ranks
int quantize_grads(ytk_gbdt* h, int gid);
int hist_events(ytk_gbdt* h, cudaEvent_t* a, cudaEvent_t* b);
int comm_mark(ytk_gbdt* h);
void add_row_items(std::vector<ytk::RowItem>& items, const Interval& v, int seg, int64_t rows_per_item);
void add_deferred_items(std::vector<ytk::RowItem>& items, const Interval& parent, int child_buf, int seg, int cseg,
                        int64_t rows_per_item, int world);
int peer_setup(ytk_gbdt* h);
// after a host wait: did a peer kernel give up?  Reports (and clears) the flag, so the handle stays usable.
int peer_check(ytk_gbdt* h, const char* where);
bool peer_ok(const ytk_gbdt* h, int m);
int blob_commit_to(ytk_gbdt* h, const Blob& bl, char** d_buf, size_t* cap);
int blob_commit(ytk_gbdt* h, const Blob& bl);
int hist_plan(ytk_gbdt* h, const std::vector<HistJob>& jobs, ytk_time_stats* T, int n_sel, HistPlan& pl, Blob& bl);
int hist_enqueue(ytk_gbdt* h, const HistPlan& pl, ytk_time_stats* T, const ytk::PartitionCounts* cnt_local,
                 const ytk::PartitionCounts* cnt_global);
int build_hists(ytk_gbdt* h, const std::vector<HistJob>& jobs, ytk_time_stats* T, int n_sel,
                const ytk::PartitionCounts* cnt_local, const ytk::PartitionCounts* cnt_global);
// at least m zeroed counters in h->t_scan_done
int scan_done_counters(ytk_gbdt* h, int m);
int search_enqueue(ytk_gbdt* h, const ytk::ScanNode* d_nodes, int m, int first_owned,
                   const ytk::PartitionCounts* cnt_global);
void search_collect(ytk_gbdt* h, int m, std::vector<ytk_split>& recs);
void apply_record(NodeStats& s, const ytk_split& r);
int partition_commit(ytk_gbdt* h, const Blob& bl, PartPlan& pl);
void partition_plan(const std::vector<ytk::SplitSeg>& segs_in, PartPlan& pl, Blob& bl, bool single_pass = false);
int next_partition_epoch(ytk_gbdt* h, size_t n_tiles);
int partition_enqueue(ytk_gbdt* h, const uint8_t* src, int64_t col_stride, const PartPlan& pl);
int partition_launch(ytk_gbdt* h, const uint8_t* src, int64_t col_stride, const std::vector<ytk::SplitSeg>& segs_in);
void partition_collect(ytk_gbdt* h, int m, std::vector<ytk::PartitionCounts>& local,
                       std::vector<ytk::PartitionCounts>& global);
int partition_nodes(ytk_gbdt* h, const uint8_t* src, int64_t col_stride, const std::vector<ytk::SplitSeg>& segs_in,
                    std::vector<ytk::PartitionCounts>& local, std::vector<ytk::PartitionCounts>& global);
// ---- engine.cu: level-wise growth planned on the device --------------------------------------------
bool engine_applicable(const ytk_gbdt* h);
void engine_drop_graphs(ytk_gbdt* h);
int engine_make_tree(ytk_gbdt* h, int gid, const int32_t* fea_index, int32_t n_fea, const uint8_t* row_mask, ytk_tree_out* out,
                     ytk_time_stats* T);
int engine_make_tree_clique(ytk_gbdt** hs, int world, int gid, const int32_t* fea_index, int32_t n_fea,
                            const uint8_t* const* row_masks, ytk_tree_out* outs, ytk_time_stats* Ts);
// sizes and offsets of the exchange areas (equal on every rank); returns the bytes of an exchange buffer
size_t peer_layout(ytk_gbdt* h, size_t* hist_words);

}  // namespace ytkh
