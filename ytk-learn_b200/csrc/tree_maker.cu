// Host side of libytkgbdt.so: the tree maker behind the reference's ITreeMaker seam.
// Restates the control flow of DataParallelTreeMaker.make
// (J/optimizer/gbdt/DataParallelTreeMaker.java:229-295) -- expansion queue, stop rules,
// smaller-child / subtraction trick, HistogramPool policy -- and drives the kernels of
// kernels.cu.  J/ = /root/reference/src/main/java/com/fenbi/ytklearn/.
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
#include "nccl_dyn.h"

namespace {

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

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

}  // namespace
namespace ytk { void set_last_error(const char* msg) { g_err = msg ? msg : ""; } }
namespace {

double now_ms() {
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

int java_float_compare(float a, float b) {   // java.lang.Float.compare
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

}  // namespace

// Host mirror of UpdateStrategy.calcGain / calcNodeValue
// (J/optimizer/gbdt/UpdateStrategy.java:55-100) for the two leaf totals computeLeafStats
// derives on the host (large = parent - small, DataParallelTreeMaker.java:381-384).
// Compiled without FMA contraction (see the build flags) so it matches the kernel's
// _rn arithmetic.
static double host_threshold_l1(double w, double lambda) {
    if (w > lambda) return w - lambda;
    if (w < -lambda) return w + lambda;
    return 0.0;
}
static double host_node_value(const ytk_gbdt_params* p, double G, double H) {
    if (H < p->min_child_hessian_sum) return 0.0;
    double val = (p->l1 == 0.0) ? -G / (H + p->l2) : -host_threshold_l1(G, p->l1) / (H + p->l2);
    if (p->max_abs_leaf_val > 0) {
        if (val > p->max_abs_leaf_val) val = p->max_abs_leaf_val;
        else if (val < -p->max_abs_leaf_val) val = -p->max_abs_leaf_val;
    }
    return val;
}
static double host_gain(const ytk_gbdt_params* p, double G, double H) {
    if (H < p->min_child_hessian_sum) return 0.0;
    if (p->max_abs_leaf_val <= 0) {
        double t = (p->l1 == 0.0) ? G : host_threshold_l1(G, p->l1);
        volatile double num = t * t;
        return num / (H + p->l2);
    }
    double v = host_node_value(p, G, H);
    volatile double a = G * v;
    volatile double hl = 0.5 * (H + p->l2);
    volatile double vv = v * v;
    volatile double b = hl * vv;
    volatile double c = p->l1 * std::fabs(v);
    volatile double s = a + b;
    volatile double s2 = s + c;
    return -2 * s2;
}
static void ytk_host_node_value(const ytk_gbdt_params* p, double G, double H, float* root_gain, float* node_value) {
    *root_gain = (float)host_gain(p, G, H);
    *node_value = (float)host_node_value(p, G, H);
}

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
    // level-wise growth on one GPU: the next level's partition planned and started by the device
    bool plan_ahead = true;                 // YTK_GBDT_NO_PLAN_AHEAD=1 turns it off
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
    uint32_t* d_group_mask = nullptr;
    // features with more than 256 slots: uint16 columns beside the uint8 matrices
    int n_wide = 0;
    std::vector<int32_t> wide_index;       // per feature: its uint16 column, -1 if it has at most 256 slots
    std::vector<int32_t> wide_sel;         // wide features selected for the current tree
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
    DevTable<ytk::RowItem> t_items, t_leaf_items;
    DevTable<ytk::SplitSeg> t_segs;
    DevTable<int> t_tile_seg;
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

    // collectives over NVLink peer memory (world > 1, one node): see kernels.cu, last section
    bool peer_mode = false;
    ytk::PeerTable pt{};
    char* my_xchg = nullptr; char* my_hist = nullptr;
    size_t hist_cap_words = 0;               // capacity of my_hist in int64 words
    int peer_cap_nodes = 0;                  // nodes per batch the exchange slots can hold
    size_t off_counts = 0, off_recs = 0, off_leaf = 0, slot_counts = 0, slot_recs = 0, slot_leaf = 0;
    unsigned long long epoch[ytk::kPeerFlagSets] = {0, 0, 0, 0};
    int* d_peer_err = nullptr;

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

namespace {

using ytk::GainParams;

GainParams gain_params(const ytk_gbdt_params& p) {
    GainParams g;
    g.min_child_weight = p.min_child_hessian_sum;
    g.max_abs_leaf_val = p.max_abs_leaf_val;
    g.l1 = p.l1; g.l2 = p.l2;
    g.min_split_loss = p.min_split_loss;
    g.min_split_samples = p.min_split_samples;
    return g;
}

int ensure_stage(ytk_gbdt* h, size_t bytes) {
    if (h->stage_bytes >= bytes) return YTK_OK;
    if (h->d_stage) CK(cudaFree(h->d_stage));
    h->d_stage = nullptr; h->stage_bytes = 0;
    CK(cudaMalloc(&h->d_stage, bytes));
    h->stage_bytes = bytes;
    return YTK_OK;
}

void free_all(ytk_gbdt* h) {
    cudaSetDevice(h->device);
    auto F = [](void* p) { if (p) cudaFree(p); };
    F(h->d_bins); F(h->d_bins_col); F(h->d_gp); F(h->d_gq); F(h->d_rowidx[0]); F(h->d_rowidx[1]); F(h->d_flags); F(h->d_block_cnt); F(h->d_tile_state); F(h->d_ticket);
    F(h->d_mask); F(h->d_start); F(h->d_nbins); F(h->d_feat_owner); F(h->d_feat_zero); F(h->d_off_inner);
    F(h->d_off_flat); F(h->d_group_mask); F(h->d_row_ptr); F(h->d_nz); F(h->d_default_slot); F(h->d_wide); F(h->d_wide_index); F(h->d_wide_fids); F(h->d_fea_sel); F(h->d_hist_flat); F(h->d_hist_flat2);
    F(h->d_hist_dbl); F(h->d_scale); F(h->d_scale_hook); F(h->d_absmax); F(h->d_bad); F(h->d_node_of_row);
    F(h->d_stage); F(h->d_blob); F(h->d_blob_part); F(h->d_score); F(h->d_y); F(h->d_w); F(h->d_init); F(h->d_block_loss); F(h->d_loss);
    h->t_leafval.release(); h->t_tree.release();
    if (h->peer_mode)
        for (int r = 0; r < h->world; r++)
            if (r != h->rank) { cudaIpcCloseMemHandle(h->pt.xchg[r]); cudaIpcCloseMemHandle(h->pt.hist[r]); }
    F(h->my_xchg); F(h->my_hist); F(h->d_peer_err);
    h->t_items.release(); h->t_leaf_items.release(); h->t_segs.release(); h->t_tile_seg.release();
    h->t_scan.release(); h->t_sub.release(); h->t_pos.release(); h->t_seg_out.release(); h->t_counts.release();
    h->t_split_local.release(); h->t_split_all.release(); h->t_parts.release(); h->t_leaf.release(); h->t_hist_local.release();
    h->t_rs_out.release(); h->t_node_tot.release();
    for (int i = 0; i < 2; i++) { h->t_spec_segs[i].release(); h->t_spec_counts[i].release(); h->h_spec_counts[i].release(); }
    h->t_spec_tiles.release();
    F(h->d_spec_ntiles);
    if (h->ev_level) cudaEventDestroy(h->ev_level);
    h->h_counts.release(); h->h_split.release(); h->h_leaf.release(); h->h_scale.release(); h->arena.release();
    for (auto* it : h->pool.all) { if (it->own_buf) F(it->buf); delete it; }
    F(h->d_pool_chunk);
    for (auto e : h->ev) cudaEventDestroy(e);
    for (auto e : h->evc) cudaEventDestroy(e);
    if (h->comm) ytk::nccl_api().CommDestroy(h->comm);
    if (h->st) cudaStreamDestroy(h->st);
}

ytk::SparseRows sparse_rows(const ytk_gbdt* h) {
    return h->sparse ? ytk::SparseRows{h->d_row_ptr, h->d_nz, h->d_default_slot} : ytk::SparseRows{nullptr, nullptr, nullptr};
}

// the uint16 columns, seen from row `row0` on
ytk::WideCols wide_cols(const ytk_gbdt* h, int64_t row0 = 0) {
    return ytk::WideCols{h->d_wide ? h->d_wide + row0 : nullptr, h->N, h->d_wide_index};
}

// Selected-feature tables for one make()/hook call (initAssistData, :425-469).
int set_features(ytk_gbdt* h, const int32_t* fea_index, int32_t n_fea, int* first_owned, int* n_sel) {
    std::vector<uint8_t> sel(h->F, 0);
    if (!fea_index) {
        std::fill(sel.begin(), sel.end(), 1);
    } else {
        if (n_fea < 1) return fail(YTK_ERR_ARG, "n_fea must be >= 1");
        for (int i = 0; i < n_fea; i++) {
            if (fea_index[i] < 0 || fea_index[i] >= h->F)
                return fail(YTK_ERR_ARG, "fea_index[%d]=%d out of range [0,%d)", i, fea_index[i], h->F);
            sel[fea_index[i]] = 1;
        }
    }
    std::vector<uint32_t> gm(h->groups, 0);
    int cnt = 0;
    h->wide_sel.clear();
    for (int f = 0; f < h->F; f++) {
        if (!sel[f]) continue;
        cnt++;
        if (h->wide_index[f] >= 0) h->wide_sel.push_back(f);                  // hist_wide_k's share
        else gm[f / ytk::kGroup] |= 1u << (f % ytk::kGroup);                  // hist_build_k's
    }
    if (n_sel) *n_sel = cnt;
    *first_owned = -1;   // firstSampledFeaIndexInThread (:451-467)
    for (int f = h->f_from; f < h->f_to; f++) if (sel[f]) { *first_owned = f; break; }
    void* s1 = h->arena.take(h->F, h->st);
    void* s2 = h->arena.take(h->groups * sizeof(uint32_t), h->st);
    if (!s1 || !s2) return fail(YTK_ERR_CUDA, "pinned staging allocation failed");
    memcpy(s1, sel.data(), h->F);
    memcpy(s2, gm.data(), h->groups * sizeof(uint32_t));
    CK(cudaMemcpyAsync(h->d_fea_sel, s1, h->F, cudaMemcpyHostToDevice, h->st));
    CK(cudaMemcpyAsync(h->d_group_mask, s2, h->groups * sizeof(uint32_t), cudaMemcpyHostToDevice, h->st));
    if (!h->wide_sel.empty()) {
        void* s3 = h->arena.take(h->wide_sel.size() * sizeof(int32_t), h->st);
        if (!s3) return fail(YTK_ERR_CUDA, "pinned staging allocation failed");
        memcpy(s3, h->wide_sel.data(), h->wide_sel.size() * sizeof(int32_t));
        CK(cudaMemcpyAsync(h->d_wide_fids, s3, h->wide_sel.size() * sizeof(int32_t), cudaMemcpyHostToDevice, h->st));
    }
    return YTK_OK;
}

int hist_events(ytk_gbdt* h, cudaEvent_t* a, cudaEvent_t* b) {
    if (h->ev_used + 2 > (int)h->ev.size()) {
        for (int i = 0; i < 2; i++) {
            cudaEvent_t e;
            CK(cudaEventCreate(&e));
            h->ev.push_back(e);
        }
    }
    *a = h->ev[h->ev_used]; *b = h->ev[h->ev_used + 1];
    h->ev_used += 2;
    return YTK_OK;
}

// brackets a collective with an event pair (summed into communicate_hist at the end of make())
int comm_mark(ytk_gbdt* h) {
    if (h->evc_used + 1 > (int)h->evc.size()) {
        cudaEvent_t e;
        CK(cudaEventCreate(&e));
        h->evc.push_back(e);
    }
    CK(cudaEventRecord(h->evc[h->evc_used++], h->st));
    return YTK_OK;
}

// Splits node intervals into work items of about `rows_per_item` rows.
void add_row_items(std::vector<ytk::RowItem>& items, const Interval& v, int seg, int64_t rows_per_item) {
    if (v.e <= v.b) return;
    for (int64_t b = v.b; b < v.e; b += rows_per_item)
        items.push_back({(long long)b, (long long)std::min(v.e, b + rows_per_item), v.buf, seg, -1, 0, 0});
}
// Work items for "the smaller child of split segment cseg" whose size is still on the device:
// enough chunks for the largest it can be (half the parent on one GPU; the whole local parent when
// the smaller child is decided by global counts).
void add_deferred_items(std::vector<ytk::RowItem>& items, const Interval& parent, int child_buf, int seg, int cseg,
                        int64_t rows_per_item, int world) {
    const int64_t n = parent.e - parent.b;
    const int64_t max_rows = world == 1 ? n / 2 : n;
    const int chunks = (int)((max_rows + rows_per_item - 1) / rows_per_item);
    for (int c = 0; c < chunks; c++)
        items.push_back({(long long)parent.b, (long long)parent.e, child_buf, seg, cseg, c, (long long)rows_per_item});
}

// Build a node's histogram into dst (this rank's segment).  cseg < 0: rows = iv[nid]; cseg >= 0: rows =
// the smaller child of split segment cseg (parent interval `parent`, children in buffer child_buf).
// cseg < 0: the rows of interval v; else: the smaller child of split segment cseg, whose parent has interval v
struct HistJob { long long* dst; int cseg; Interval v; int child_buf; };

// Maps every rank's exchange and histogram staging buffers into this process (CUDA IPC) so that the
// per-level collectives can run as plain kernels over NVLink.  All ranks must be processes on one node;
// if any rank cannot set this up, all of them stay on NCCL (the decision itself is an NCCL allreduce).
int peer_setup(ytk_gbdt* h) {
    h->peer_mode = false;
    if (h->world < 2 || h->world > 16) return YTK_OK;
    if (const char* e = getenv("YTK_GBDT_NO_PEER")) if (atoi(e) != 0) return YTK_OK;
    const int cap = 4096;
    h->peer_cap_nodes = cap;
    h->slot_counts = (size_t)cap * sizeof(ytk::PartitionCounts);
    h->slot_recs = (size_t)2 * cap * sizeof(ytk_split);
    h->slot_leaf = (size_t)2 * cap * sizeof(long long);
    h->off_counts = ytk::kPeerFlagBytes;
    h->off_recs = h->off_counts + h->slot_counts * h->world;
    h->off_leaf = h->off_recs + h->slot_recs * h->world;
    const size_t xchg_bytes = h->off_leaf + h->slot_leaf * h->world;
    const size_t local_len = (size_t)h->seg_max * h->world;
    size_t hist_words = std::max<size_t>(local_len * 256, (size_t(128) << 20) / 8);      // >= 256 nodes, >= 128 MB
    hist_words = std::min<size_t>(hist_words, std::max<size_t>(local_len * 2, (size_t(1) << 30) / 8));
    int ok = 1;
    cudaIpcMemHandle_t mine[2];
    if (cudaMalloc(&h->my_xchg, xchg_bytes) != cudaSuccess || cudaMalloc(&h->my_hist, hist_words * 8) != cudaSuccess ||
        cudaMalloc(&h->d_peer_err, sizeof(int)) != cudaSuccess)
        ok = 0;
    if (ok) {
        cudaMemset(h->my_xchg, 0, xchg_bytes);
        cudaMemset(h->d_peer_err, 0, sizeof(int));
        if (cudaIpcGetMemHandle(&mine[0], h->my_xchg) != cudaSuccess || cudaIpcGetMemHandle(&mine[1], h->my_hist) != cudaSuccess)
            ok = 0;
    }
    cudaGetLastError();
    // exchange the handles (and learn whether everybody got this far) over NCCL, once
    const size_t rec = sizeof(mine) + 8;
    char* d_send = nullptr; char* d_recv = nullptr;
    CK(cudaMalloc(&d_send, rec));
    CK(cudaMalloc(&d_recv, rec * h->world));
    std::vector<char> send(rec, 0), recv(rec * h->world);
    if (ok) memcpy(send.data(), mine, sizeof(mine));
    long long okv = ok;
    memcpy(send.data() + sizeof(mine), &okv, 8);
    CK(cudaMemcpy(d_send, send.data(), rec, cudaMemcpyHostToDevice));
    NK(ytk::nccl_api().AllGather(d_send, d_recv, rec, ncclChar, h->comm, h->st));
    CK(cudaStreamSynchronize(h->st));
    CK(cudaMemcpy(recv.data(), d_recv, rec * h->world, cudaMemcpyDeviceToHost));
    for (int r = 0; r < h->world; r++) { long long v; memcpy(&v, recv.data() + r * rec + sizeof(mine), 8); if (!v) ok = 0; }
    h->pt.rank = h->rank; h->pt.world = h->world;
    int opened = 0;
    for (int r = 0; r < h->world && ok; r++) {
        if (r == h->rank) { h->pt.xchg[r] = h->my_xchg; h->pt.hist[r] = h->my_hist; continue; }
        cudaIpcMemHandle_t hd[2];
        memcpy(hd, recv.data() + r * rec, sizeof(hd));
        void *px = nullptr, *ph = nullptr;
        if (cudaIpcOpenMemHandle(&px, hd[0], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
            cudaIpcOpenMemHandle(&ph, hd[1], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            ok = 0; cudaGetLastError();
            if (px) cudaIpcCloseMemHandle(px);
            break;
        }
        h->pt.xchg[r] = (char*)px; h->pt.hist[r] = (char*)ph;
        opened = r + 1;
    }
    // second agreement: did every rank manage to open every peer?
    okv = ok;
    CK(cudaMemcpy(d_send, &okv, 8, cudaMemcpyHostToDevice));
    NK(ytk::nccl_api().AllReduce(d_send, d_send, 1, ncclInt64, ncclMin, h->comm, h->st));
    CK(cudaStreamSynchronize(h->st));
    CK(cudaMemcpy(&okv, d_send, 8, cudaMemcpyDeviceToHost));
    cudaFree(d_send); cudaFree(d_recv);
    if (getenv("YTK_GBDT_VERBOSE"))
        fprintf(stderr, "ytk_gbdt[rank %d]: peer-memory collectives %s (local setup %s)\n", h->rank, okv ? "on" : "off", ok ? "ok" : "failed");
    if (!okv) {
        for (int r = 0; r < opened; r++)
            if (r != h->rank) { cudaIpcCloseMemHandle(h->pt.xchg[r]); cudaIpcCloseMemHandle(h->pt.hist[r]); }
        return YTK_OK;      // NCCL path
    }
    h->hist_cap_words = hist_words;
    h->peer_mode = true;
    return YTK_OK;
}
// may this batch of m nodes use the peer-memory collectives? (same answer on every rank)
bool peer_ok(const ytk_gbdt* h, int m) {
    return h->peer_mode && m <= h->peer_cap_nodes && (size_t)m * h->seg_max * h->world <= h->hist_cap_words;
}

// one host->device copy of a Blob; afterwards table at offset o lives at h->d_blob + o
int blob_commit_to(ytk_gbdt* h, const Blob& bl, char** d_buf, size_t* cap) {
    const size_t bytes = bl.host.size();
    if (bytes == 0) return YTK_OK;
    if (bytes > *cap) {
        if (*d_buf) CK(cudaFree(*d_buf));                // (synchronises: nobody reads the old one any more)
        *d_buf = nullptr; *cap = 0;
        size_t nc = std::max<size_t>(bytes * 2, size_t(1) << 20);
        CK(cudaMalloc(d_buf, nc));
        *cap = nc;
    }
    void* stage = h->arena.take(bytes, h->st);
    if (!stage) return fail(YTK_ERR_CUDA, "pinned staging allocation failed");
    memcpy(stage, bl.host.data(), bytes);
    CK(cudaMemcpyAsync(*d_buf, stage, bytes, cudaMemcpyHostToDevice, h->st));
    return YTK_OK;
}
int blob_commit(ytk_gbdt* h, const Blob& bl) { return blob_commit_to(h, bl, &h->d_blob, &h->blob_cap); }
template <typename T>
const T* blob_ptr(const ytk_gbdt* h, size_t off) { return reinterpret_cast<const T*>(h->d_blob + off); }

// ---- histogram build: plan (host tables) / enqueue (launches) ----------------------------------
struct HistPlan {
    int m = 0;
    std::vector<ytk::RowItem> items;
    std::vector<long long*> outs_local;    // where the kernel accumulates (the slot, or the local pre-reduce buffer)
    std::vector<long long*> outs_final;    // world > 1: the slots the reduce-scatter result is copied to
    size_t o_items = 0, o_local = 0, o_final = 0;
};

// HistogramBuilder.buildHist (J/data/gbdt/HistogramBuilder.java:56-98) for a batch of nodes:
// one kernel launch over all their rows, then (world > 1) one reduce-scatter by feature range.
int hist_plan(ytk_gbdt* h, const std::vector<HistJob>& jobs, ytk_time_stats* T, int n_sel, HistPlan& pl, Blob& bl) {
    const int m = pl.m = (int)jobs.size();
    if (m == 0) return YTK_OK;
    // Work-item size: the smallest multiple of 1024 rows (4096..65536) for which the items of all
    // nodes -- each node rounds up to whole items -- fit in two waves of one CTA per SM.
    std::vector<int64_t> sizes(m);
    for (int k = 0; k < m; k++) {
        const HistJob& j = jobs[k];
        sizes[k] = j.cseg < 0 ? j.v.e - j.v.b : (h->world == 1 ? (j.v.e - j.v.b) / 2 : j.v.e - j.v.b);
    }
    const int64_t target = std::max<int64_t>(1, (2 * ytk::kNumSMs) / std::max(1, h->groups));
    auto items_at = [&](int64_t rpi) { int64_t c = 0; for (int64_t n : sizes) c += (n + rpi - 1) / rpi; return c; };
    int64_t lo_u = 4, hi_u = ytk::kHistMaxRowsPerItem / 1024;        // in units of 1024 rows
    while (lo_u < hi_u) {
        int64_t mid = (lo_u + hi_u) / 2;
        if (items_at(mid * 1024) <= target) hi_u = mid; else lo_u = mid + 1;
    }
    // sparse rows: a work item is a CTA of 8 warps walking one row each -- many small items instead
    const int64_t rpi = h->sparse ? 256 : lo_u * 1024;
    pl.outs_local.resize(m);
    const int64_t local_len = h->seg_max * h->world;
    const bool peer = h->world > 1 && peer_ok(h, m);
    if (h->world > 1 && !peer) CK(h->t_hist_local.reserve((size_t)m * local_len));
    long long* const staging = peer ? reinterpret_cast<long long*>(h->my_hist) : h->t_hist_local.d;
    for (int k = 0; k < m; k++) {
        if (jobs[k].cseg < 0) {
            const Interval& v = jobs[k].v;
            add_row_items(pl.items, v, k, rpi);
            T->hist_built++;
            T->hist_cells += (double)(v.e - v.b) * n_sel;
            T->hist_rows += (double)(v.e - v.b);
            if (v.buf >= 0) T->hist_rows_indexed += (double)(v.e - v.b);
        } else {
            add_deferred_items(pl.items, jobs[k].v, jobs[k].child_buf, k, jobs[k].cseg, rpi, h->world);
        }
        // world > 1: local layout [rank][node][segment] so that ONE reduce-scatter serves the batch
        pl.outs_local[k] = h->world == 1 ? jobs[k].dst : staging + (int64_t)k * h->seg_max;
    }
    pl.o_items = bl.add(pl.items);
    pl.o_local = bl.add(pl.outs_local);
    if (h->world > 1) {
        pl.outs_final.resize(m);
        for (int k = 0; k < m; k++) pl.outs_final[k] = jobs[k].dst;
        pl.o_final = bl.add(pl.outs_final);
    }
    return YTK_OK;
}

int hist_enqueue(ytk_gbdt* h, const HistPlan& pl, ytk_time_stats* T, const ytk::PartitionCounts* cnt_local,
                 const ytk::PartitionCounts* cnt_global) {
    const int m = pl.m;
    if (m == 0) return YTK_OK;
    const int64_t local_len = h->seg_max * h->world;
    const bool peer = h->world > 1 && peer_ok(h, m);
    long long* const* d_local = blob_ptr<long long*>(h, pl.o_local);
    cudaEvent_t e0, e1;
    RET(hist_events(h, &e0, &e1));
    if (h->world == 1) {
        CK(ytk::launch_zero_segments(d_local, m, h->seg_max, h->st));
        h->launches++;
    } else {
        CK(cudaMemsetAsync(peer ? (void*)h->my_hist : (void*)h->t_hist_local.d, 0,
                           (size_t)m * local_len * sizeof(long long), h->st));
    }
    const long long owner_stride = h->world == 1 ? 0 : (long long)m * h->seg_max;
    if (h->sparse) {
        CK(h->t_node_tot.reserve(2 * (size_t)m));
        CK(cudaMemsetAsync(h->t_node_tot.d, 0, 2 * (size_t)m * sizeof(long long), h->st));
    }
    CK(cudaEventRecord(e0, h->st));        // the pair brackets the histogram-build launch only
    if (h->sparse) {
        CK(ytk::launch_hist_csr(sparse_rows(h), h->d_fea_sel, h->d_gq, h->d_rowidx[0], h->d_rowidx[1],
                                blob_ptr<ytk::RowItem>(h, pl.o_items), (int)pl.items.size(), h->d_feat_owner, h->d_off_inner,
                                owner_stride, d_local, h->t_node_tot.d, cnt_local, cnt_global, h->st));
    } else {
        CK(ytk::launch_hist_build(h->d_bins, h->S8, h->d_gq, h->d_rowidx[0], h->d_rowidx[1],
                                  blob_ptr<ytk::RowItem>(h, pl.o_items), (int)pl.items.size(), h->F, h->d_group_mask,
                                  h->d_nbins, h->d_feat_owner, h->d_off_inner, owner_stride, d_local, cnt_local, cnt_global,
                                  h->st));
    }
    if (!pl.items.empty()) h->launches++;
    CK(cudaEventRecord(e1, h->st));
    if (h->sparse) {
        CK(ytk::launch_csr_default_fix(sparse_rows(h), h->d_fea_sel, h->F, h->d_start, h->d_feat_owner, h->d_off_inner,
                                       owner_stride, d_local, m, h->t_node_tot.d, h->st));
        h->launches++;
    }
    if (!h->wide_sel.empty() && !pl.items.empty()) {
        CK(ytk::launch_hist_wide(wide_cols(h), h->d_wide_fids, (int)h->wide_sel.size(), h->d_gq, h->d_rowidx[0],
                                 h->d_rowidx[1], blob_ptr<ytk::RowItem>(h, pl.o_items), (int)pl.items.size(),
                                 h->d_feat_owner, h->d_off_inner, owner_stride, d_local, cnt_local, cnt_global, h->st));
        h->launches++;
    }
    if (peer) {
        // reduce-scatter over peer memory: flag "my staging buffer is complete", then every rank
        // sums the segments it owns straight out of all staging buffers into the histogram slots
        RET(comm_mark(h));
        const unsigned long long ep = ++h->epoch[1];
        CK(ytk::launch_peer_pull_reduce(h->pt, m, h->seg_max, 1, ep, blob_ptr<long long*>(h, pl.o_final), h->d_peer_err, h->st));
        RET(comm_mark(h));
        h->launches++;
    } else if (h->world > 1) {
        CK(h->t_rs_out.reserve((size_t)m * h->seg_max));
        RET(comm_mark(h));
        NK(ytk::nccl_api().ReduceScatter(h->t_hist_local.d, h->t_rs_out.d, (size_t)m * h->seg_max, ncclInt64,
                                         ncclSum, h->comm, h->st));
        CK(ytk::launch_copy_segments(h->t_rs_out.d, blob_ptr<long long*>(h, pl.o_final), m, h->seg_max, h->st));
        RET(comm_mark(h));
        h->launches++;
    }
    return YTK_OK;
}

int build_hists(ytk_gbdt* h, const std::vector<HistJob>& jobs, ytk_time_stats* T, int n_sel,
                const ytk::PartitionCounts* cnt_local, const ytk::PartitionCounts* cnt_global) {
    HistPlan pl; Blob bl;
    RET(hist_plan(h, jobs, T, n_sel, pl, bl));
    RET(blob_commit(h, bl));
    return hist_enqueue(h, pl, T, cnt_local, cnt_global);
}

// ---- split search: enqueue / collect ---------------------------------------------------------------
// updateTreeMakerNodeStats + findBestSplit(nid, hist) + syncBestSplit (:543-595, :640-653) for a
// batch of nodes; after the stream has been synchronised the merged record of node k is recs[k].
int search_enqueue(ytk_gbdt* h, const ytk::ScanNode* d_nodes, int m, int first_owned,
                   const ytk::PartitionCounts* cnt_global) {
    if (m == 0) return YTK_OK;
    CK(h->t_split_local.reserve(m));
    CK(h->t_parts.reserve((size_t)m * ytk::split_scan_slices(h->f_from, h->f_to)));
    CK(h->h_split.reserve((size_t)m * h->world));
    // the records stay in device memory (the planner of the next partition and the exchange read them
    // there); on one GPU the merge kernel also stores them in pinned host memory, so the host needs no copy
    CK(ytk::launch_split_scan(d_nodes, m, h->d_start, h->d_off_inner, h->d_fea_sel, h->f_from, h->f_to,
                              first_owned, h->d_scale, gain_params(h->p), cnt_global, h->t_parts.d, h->t_split_local.d,
                              h->world == 1 ? h->h_split.h : nullptr, h->st));
    h->launches += 2;
    if (h->world > 1) {
        // "split allgather": every rank's record for each node, merged on the host with the
        // needReplace rule -- replaces allreduceRpc(SplitInfo) and allreduceArrayRpc(MAX)
        CK(h->t_split_all.reserve((size_t)m * h->world));
        RET(comm_mark(h));
        h->recs_merged = false;
        if (h->peer_mode && (size_t)m * sizeof(ytk_split) <= h->slot_recs) {
            // gathered and merged on the device; the merged records land in pinned host memory
            const unsigned long long ep = ++h->epoch[2];
            CK(ytk::launch_peer_allgather_merge(h->pt, h->t_split_local.d, h->off_recs, h->slot_recs, m, 2, ep,
                                                h->t_split_all.d, h->h_split.h, h->d_peer_err, h->st));
            h->launches++;
            h->recs_merged = true;
            RET(comm_mark(h));
        } else {
            NK(ytk::nccl_api().AllGather(h->t_split_local.d, h->t_split_all.d, m * sizeof(ytk_split), ncclChar, h->comm,
                                         h->st));
            RET(comm_mark(h));
            CK(cudaMemcpyAsync(h->h_split.h, h->t_split_all.d, (size_t)h->world * m * sizeof(ytk_split),
                               cudaMemcpyDeviceToHost, h->st));
        }
    }
    return YTK_OK;
}
void search_collect(ytk_gbdt* h, int m, std::vector<ytk_split>& recs) {
    recs.resize(m);
    if (h->world == 1 || h->recs_merged) {
        for (int k = 0; k < m; k++) recs[k] = h->h_split.h[k];
    } else {
        std::vector<ytk_split> tmp(h->world);
        for (int k = 0; k < m; k++) {
            for (int r = 0; r < h->world; r++) tmp[r] = h->h_split.h[(size_t)r * m + k];
            ytk_gbdt_merge_splits(tmp.data(), h->world, &recs[k]);
        }
    }
}
void apply_record(NodeStats& s, const ytk_split& r) {
    s.sumG = r.sum_grad; s.sumH = r.sum_hess;
    s.root_gain = r.root_gain; s.node_value = r.node_value;
    s.best_lc = r.loss_chg; s.best_fid = r.fid; s.best_lo = r.lo; s.best_hi = r.hi;
}

// position.resetPosition (SamplePositionData.java:115-133) for a batch of nodes, plus the global
// child counts of initTreeMakerNodeStats(l, r) (:522-541).
struct PartPlan {
    int m = 0;
    std::vector<ytk::SplitSeg> segs;
    std::vector<int> tile_seg;
    size_t o_segs = 0, o_tiles = 0;
    bool single_pass = false;       // tiles sized for the chained-scan kernel
    const ytk::SplitSeg* d_segs = nullptr;      // set by partition_commit
    const int* d_tiles = nullptr;
};
// The partition tables have their own device buffer: they go up first so that the partition is
// already running while the host builds the tables of the kernels behind it.
int partition_commit(ytk_gbdt* h, const Blob& bl, PartPlan& pl) {
    RET(blob_commit_to(h, bl, &h->d_blob_part, &h->blob_part_cap));
    pl.d_segs = reinterpret_cast<const ytk::SplitSeg*>(h->d_blob_part + pl.o_segs);
    pl.d_tiles = reinterpret_cast<const int*>(h->d_blob_part + pl.o_tiles);
    return YTK_OK;
}
void partition_plan(const std::vector<ytk::SplitSeg>& segs_in, PartPlan& pl, Blob& bl, bool single_pass = false) {
    pl.m = (int)segs_in.size();
    pl.segs = segs_in;
    pl.single_pass = single_pass;
    const int64_t tile = single_pass ? ytk::kFusedTile : ytk::kPartTile;
    for (int k = 0; k < pl.m; k++) {
        int64_t n = pl.segs[k].end - pl.segs[k].begin;
        int nt = (int)((n + tile - 1) / tile);
        pl.segs[k].tile0 = (int)pl.tile_seg.size();
        pl.segs[k].ntiles = nt;
        for (int t = 0; t < nt; t++) pl.tile_seg.push_back(k);
    }
    pl.o_segs = bl.add(pl.segs);
    pl.o_tiles = bl.add(pl.tile_seg);
}
// Queues the kernels, the count allreduce and the copy of the counts to the host; does not wait.
// Device counts: t_counts.d[0..m) local, [m..2m) global (world > 1).
// Work space of the chained-scan partition for a launch of up to n_tiles tiles, and a fresh epoch.
int next_partition_epoch(ytk_gbdt* h, size_t n_tiles) {
    if (n_tiles > h->tile_state_cap) {
        if (h->d_tile_state) CK(cudaFree(h->d_tile_state));
        h->d_tile_state = nullptr;
        h->tile_state_cap = std::max<size_t>(n_tiles, 2 * h->tile_state_cap);
        CK(cudaMalloc(&h->d_tile_state, h->tile_state_cap * sizeof(unsigned long long)));
        CK(cudaMemsetAsync(h->d_tile_state, 0, h->tile_state_cap * sizeof(unsigned long long), h->st));
        h->part_epoch = 0;
    }
    if (!h->d_ticket) {        // [0]: launches with host-built tables (wraps back to 0), [1]: device-planned ones
        CK(cudaMalloc(&h->d_ticket, 2 * sizeof(unsigned int)));
        CK(cudaMemsetAsync(h->d_ticket, 0, 2 * sizeof(unsigned int), h->st));
    }
    if (++h->part_epoch >= (1ull << 30)) {       // entries of other epochs read as "not written yet"
        CK(cudaMemsetAsync(h->d_tile_state, 0, h->tile_state_cap * sizeof(unsigned long long), h->st));
        h->part_epoch = 1;
    }
    return YTK_OK;
}

// single_pass: the chained-scan kernel (right rows come out reversed; make() only); else the stable
// three-kernel version the public entry points promise.
int partition_enqueue(ytk_gbdt* h, const uint8_t* src, int64_t col_stride, const PartPlan& pl) {
    const bool single_pass = pl.single_pass;
    const int m = pl.m;
    if (m == 0) return YTK_OK;
    const size_t n_tiles = pl.tile_seg.size();
    if (single_pass) {
        RET(next_partition_epoch(h, n_tiles));
        CK(h->t_counts.reserve(2 * (size_t)m));
        CK(ytk::launch_partition_fused(src, h->S8, col_stride, wide_cols(h), h->d_rowidx[0], h->d_rowidx[1],
                                       pl.d_segs, m, pl.d_tiles, (int)n_tiles,
                                       h->d_tile_state, h->d_ticket, h->part_epoch, h->t_counts.d, h->st, sparse_rows(h)));
        h->launches++;
    } else {
    if (n_tiles + 1 > h->block_cnt_cap) {
        if (h->d_block_cnt) CK(cudaFree(h->d_block_cnt));
        h->d_block_cnt = nullptr;
        h->block_cnt_cap = std::max<size_t>(n_tiles + 1, 2 * h->block_cnt_cap);
        CK(cudaMalloc(&h->d_block_cnt, h->block_cnt_cap * sizeof(int32_t)));
    }
    CK(h->t_counts.reserve(2 * (size_t)m));
    CK(ytk::launch_partition(src, h->S8, col_stride, wide_cols(h), h->d_rowidx[0], h->d_rowidx[1],
                             pl.d_segs, m, pl.d_tiles, (int)n_tiles,
                             h->d_flags, h->d_block_cnt, h->t_counts.d, h->st, sparse_rows(h)));
    h->launches += n_tiles == 0 ? 1 : 3;
    }
    const int nc = h->world > 1 ? 2 * m : m;
    if (h->world > 1) {   // allreduceArray(long[2], SUM) (:538), all nodes of the batch at once
        RET(comm_mark(h));
        if (peer_ok(h, m)) {
            const unsigned long long ep = ++h->epoch[0];
            CK(ytk::launch_peer_allreduce(h->pt, h->t_counts.d, h->off_counts, h->slot_counts, 2 * (size_t)m, 0, ep,
                                          reinterpret_cast<long long*>(h->t_counts.d + m), h->d_peer_err, h->st));
            h->launches++;
        } else {
            NK(ytk::nccl_api().AllReduce(h->t_counts.d, h->t_counts.d + m, 2 * (size_t)m, ncclInt64, ncclSum, h->comm,
                                         h->st));
        }
        RET(comm_mark(h));
    }
    CK(h->h_counts.reserve(nc));
    CK(cudaMemcpyAsync(h->h_counts.h, h->t_counts.d, nc * sizeof(ytk::PartitionCounts), cudaMemcpyDeviceToHost,
                       h->st));
    return YTK_OK;
}
int partition_launch(ytk_gbdt* h, const uint8_t* src, int64_t col_stride, const std::vector<ytk::SplitSeg>& segs_in) {
    if (segs_in.empty()) return YTK_OK;
    PartPlan pl; Blob bl;
    partition_plan(segs_in, pl, bl);
    RET(partition_commit(h, bl, pl));
    return partition_enqueue(h, src, col_stride, pl);
}
// after the stream has been synchronised
void partition_collect(ytk_gbdt* h, int m, std::vector<ytk::PartitionCounts>& local,
                       std::vector<ytk::PartitionCounts>& global) {
    local.resize(m); global.resize(m);
    for (int k = 0; k < m; k++) {
        local[k] = h->h_counts.h[k];
        global[k] = h->h_counts.h[h->world > 1 ? m + k : k];
    }
}
int partition_nodes(ytk_gbdt* h, const uint8_t* src, int64_t col_stride, const std::vector<ytk::SplitSeg>& segs_in,
                    std::vector<ytk::PartitionCounts>& local, std::vector<ytk::PartitionCounts>& global) {
    RET(partition_launch(h, src, col_stride, segs_in));
    CK(cudaStreamSynchronize(h->st));
    h->arena.rewind();
    partition_collect(h, (int)segs_in.size(), local, global);
    return YTK_OK;
}

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

}  // namespace

// =====================================================================================
// C ABI
// =====================================================================================
extern "C" {

const char* ytk_gbdt_last_error(void) { return g_err.c_str(); }

int32_t ytk_gbdt_normalize_max_leaf_cnt(int32_t max_depth, int32_t max_leaf_cnt) {
    if (max_depth > 0) {
        int32_t cap = (int32_t)std::pow(2.0, (double)max_depth);
        if (max_leaf_cnt < 0) return cap;
        return std::min(max_leaf_cnt, cap);
    }
    return max_leaf_cnt;
}

int ytk_gbdt_avg_assign(int32_t amount, int32_t workers, int32_t* out) {
    if (workers < 1 || amount < 0 || !out) return fail(YTK_ERR_ARG, "avg_assign: bad arguments");
    int32_t avg = amount / workers, mod = amount % workers, fidx = 0;
    for (int32_t t = 0; t < workers; t++) {
        int32_t tnum = t < mod ? avg + 1 : avg;
        out[2 * t] = fidx; out[2 * t + 1] = fidx + tnum;
        fidx += tnum;
    }
    return YTK_OK;
}

int ytk_gbdt_merge_splits(const ytk_split* recs, int32_t n, ytk_split* out) {
    if (!recs || !out || n < 1) return fail(YTK_ERR_ARG, "merge_splits: bad arguments");
    ytk_split best = recs[0];
    double G = recs[0].sum_grad, H = recs[0].sum_hess;
    for (int i = 1; i < n; i++) {
        const ytk_split& t2 = recs[i];
        // t1.needReplace(t2.lossChg, t2.splitIndex) ? t2 : t1   (DataParallelTreeMaker.java:645-649)
        bool rep = (best.fid <= t2.fid) ? (t2.loss_chg > best.loss_chg) : (t2.loss_chg >= best.loss_chg);
        if (rep) best = t2;
        if (t2.sum_grad > G) G = t2.sum_grad;      // allreduce MAX, element-wise (:560)
        if (t2.sum_hess > H) H = t2.sum_hess;
    }
    // root_gain / node_value are functions of (G,H): take them from a rank that had the totals
    for (int i = 0; i < n; i++) {
        if (recs[i].sum_grad == G && recs[i].sum_hess == H) {
            best.root_gain = recs[i].root_gain; best.node_value = recs[i].node_value;
            break;
        }
    }
    best.sum_grad = G; best.sum_hess = H;
    *out = best;
    return YTK_OK;
}

int ytk_gbdt_nccl_unique_id(void* out128) {
    const char* why = "";
    if (!out128) return fail(YTK_ERR_ARG, "nccl_unique_id: null argument");
    if (!ytk::nccl_api().load(&why)) return fail(YTK_ERR_NCCL, "NCCL unavailable: %s", why);
    ncclUniqueId id;
    NK(ytk::nccl_api().GetUniqueId(&id));
    static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
    memcpy(out128, &id, 128);
    return YTK_OK;
}

int ytk_gbdt_create(const ytk_gbdt_params* params, int32_t device, int32_t rank, int32_t world,
                    const void* nccl_id, ytk_gbdt** out) {
    if (!params || !out) return fail(YTK_ERR_ARG, "create: null argument");
    if (world < 1 || rank < 0 || rank >= world) return fail(YTK_ERR_ARG, "create: bad rank %d / world %d", rank, world);
    if (world > 64) return fail(YTK_ERR_ARG, "create: world %d > 64", world);
    if (params->class_num < 1) return fail(YTK_ERR_ARG, "class_num(%d) should >= 1", params->class_num);
    if (params->max_leaf_cnt == 0) return fail(YTK_ERR_ARG, "max_leaf_cnt(0) should not be 0");
    if (params->max_leaf_cnt < 0 && params->max_depth < 0)
        return fail(YTK_ERR_ARG, "max_leaf_cnt(%d) and max_depth(%d) should not be both negative",
                    params->max_leaf_cnt, params->max_depth);
    if (params->grow_policy != 0 && params->grow_policy != 1)
        return fail(YTK_ERR_ARG, "tree maker policy(%d) invalid!", params->grow_policy);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev < 1)
        return fail(YTK_ERR_CUDA, "no CUDA device (%s); libytkgbdt has no CPU path",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(YTK_ERR_ARG, "create: device %d not in [0,%d)", device, ndev);
    CK(cudaSetDevice(device));
    ytk_gbdt* h = new ytk_gbdt();
    h->p = *params; h->device = device; h->rank = rank; h->world = world;
    if (const char* s = getenv("YTK_GBDT_FIXED_BITS")) {
        int b = atoi(s);
        if (b < 8 || b > 30) { delete h; return fail(YTK_ERR_ARG, "YTK_GBDT_FIXED_BITS=%s not in [8,30]", s); }
        h->fixed_bits = b;
    }
    if (const char* s = getenv("YTK_GBDT_SEQUENTIAL")) h->force_sequential = atoi(s) != 0;
    if (const char* s = getenv("YTK_GBDT_STABLE_PARTITION")) h->single_pass_partition = atoi(s) == 0;
    if (const char* s = getenv("YTK_GBDT_NO_PLAN_AHEAD")) h->plan_ahead = atoi(s) == 0;
    if (const char* s = getenv("YTK_GBDT_TRACE")) h->trace = atoi(s) != 0;
    if (const char* s = getenv("YTK_GBDT_LOOKAHEAD")) h->lookahead = std::max(1, std::min(1024, atoi(s)));
    auto bail = [&](int code) { free_all(h); delete h; return code; };
    if (cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking) != cudaSuccess)
        return bail(fail(YTK_ERR_CUDA, "cudaStreamCreate failed"));
    if (world > 1) {
        const char* why = "";
        if (!nccl_id) return bail(fail(YTK_ERR_ARG, "create: world > 1 needs the NCCL unique id"));
        if (!ytk::nccl_api().load(&why)) return bail(fail(YTK_ERR_NCCL, "NCCL unavailable: %s", why));
        ncclUniqueId id;
        memcpy(&id, nccl_id, 128);
        ncclResult_t r = ytk::nccl_api().CommInitRank(&h->comm, world, id, rank);
        if (r != ncclSuccess)
            return bail(fail(YTK_ERR_NCCL, "ncclCommInitRank: %s", ytk::nccl_api().GetErrorString(r)));
    }
    *out = h;
    return YTK_OK;
}

int ytk_gbdt_destroy(ytk_gbdt* h) {
    if (!h) return YTK_OK;
    free_all(h);
    delete h;
    return YTK_OK;
}

static int set_bins_impl(ytk_gbdt* h, const int32_t* start, int32_t F, int64_t n_rows, const int32_t* default_slot,
                         int64_t nnz);
int ytk_gbdt_set_bins(ytk_gbdt* h, const int32_t* start, int32_t F, int64_t n_rows) {
    return set_bins_impl(h, start, F, n_rows, nullptr, 0);
}
int ytk_gbdt_set_bins_sparse(ytk_gbdt* h, const int32_t* start, int32_t F, int64_t n_rows, const int32_t* default_slot,
                             int64_t nnz) {
    if (!default_slot || nnz < 0) return fail(YTK_ERR_ARG, "set_bins_sparse: bad arguments");
    if (F > ytk::kMaxSparseFeatures) return fail(YTK_ERR_UNSUPPORTED, "set_bins_sparse: %d features, at most %d", F, ytk::kMaxSparseFeatures);
    return set_bins_impl(h, start, F, n_rows, default_slot, nnz);
}
static int set_bins_impl(ytk_gbdt* h, const int32_t* start, int32_t F, int64_t n_rows, const int32_t* default_slot,
                         int64_t nnz) {
    if (!h || !start || F < 1 || n_rows < 0) return fail(YTK_ERR_ARG, "set_bins: bad arguments");
    if (h->rows_set) return fail(YTK_ERR_ARG, "set_bins: already called");
    if (n_rows > 0x7fffffffLL) return fail(YTK_ERR_ARG, "set_bins: %lld rows exceed int32 row ids", (long long)n_rows);
    CK(cudaSetDevice(h->device));
    h->F = F; h->N = n_rows;
    h->start.assign(start, start + F + 1);
    h->nbins.resize(F);
    if (start[0] != 0) return fail(YTK_ERR_ARG, "set_bins: start[0] must be 0");
    for (int f = 0; f < F; f++) {
        h->nbins[f] = start[f + 1] - start[f];
        if (h->nbins[f] < 1) return fail(YTK_ERR_ARG, "[GBDT] fid:%d has no global feature vals!", f);
        if (h->nbins[f] > ytk::kMaxWideBins)
            return fail(YTK_ERR_UNSUPPORTED, "feature %d has %d bins; slots are stored as uint8 or uint16 (max %d)",
                        f, h->nbins[f], ytk::kMaxWideBins);
    }
    h->B = start[F];
    h->groups = (F + ytk::kGroup - 1) / ytk::kGroup;
    h->S8 = h->groups * ytk::kGroup;
    // GBDTDataFlow.handleOtherTrainInfo (J/dataflow/GBDTDataFlow.java:252-263) and
    // FeatureApprData.updateGlobalHistAssign (J/data/gbdt/FeatureApprData.java:98-111)
    h->assign.resize(2 * h->world);
    ytk_gbdt_avg_assign(F, h->world, h->assign.data());
    h->f_from = h->assign[2 * h->rank]; h->f_to = h->assign[2 * h->rank + 1];
    h->seg_max = 0;
    for (int r = 0; r < h->world; r++)
        h->seg_max = std::max<int64_t>(h->seg_max, 2LL * (start[h->assign[2 * r + 1]] - start[h->assign[2 * r]]));
    if (h->seg_max == 0) h->seg_max = 2;
    std::vector<int32_t> owner(F), zeros(F, 0);
    std::vector<int64_t> off_inner(F), off_flat(F);
    for (int r = 0; r < h->world; r++)
        for (int f = h->assign[2 * r]; f < h->assign[2 * r + 1]; f++) {
            owner[f] = r;
            off_inner[f] = 2LL * (start[f] - start[h->assign[2 * r]]);
        }
    for (int f = 0; f < F; f++) off_flat[f] = 2LL * start[f];

    const int64_t N1 = std::max<int64_t>(n_rows, 1);
    const int K = h->p.class_num;
    h->sparse = default_slot != nullptr;
    if (h->sparse) {
        for (int f = 0; f < F; f++)
            if (default_slot[f] < 0 || default_slot[f] >= h->nbins[f])
                return fail(YTK_ERR_ARG, "set_bins_sparse: default slot %d of feature %d outside its %d bins", default_slot[f], f, h->nbins[f]);
        h->nnz_cap = nnz;
        CK(cudaMalloc(&h->d_row_ptr, (N1 + 1) * sizeof(int64_t)));
        CK(cudaMemsetAsync(h->d_row_ptr, 0, (N1 + 1) * sizeof(int64_t), h->st));
        CK(cudaMalloc(&h->d_nz, std::max<int64_t>(nnz, 1) * sizeof(uint32_t)));
        CK(cudaMalloc(&h->d_default_slot, F * sizeof(int32_t)));
        CK(cudaMemcpyAsync(h->d_default_slot, default_slot, F * sizeof(int32_t), cudaMemcpyHostToDevice, h->st));
    } else {
        CK(cudaMalloc(&h->d_bins, N1 * h->S8));
        CK(cudaMemsetAsync(h->d_bins, 0, N1 * h->S8, h->st));
        CK(cudaMalloc(&h->d_bins_col, N1 * F));
        CK(cudaMemsetAsync(h->d_bins_col, 0, N1 * F, h->st));
    }
    CK(cudaMalloc(&h->d_gp, N1 * K * 2 * sizeof(float)));
    CK(cudaMalloc(&h->d_gq, N1 * sizeof(int2)));
    CK(cudaMalloc(&h->d_rowidx[0], N1 * sizeof(int32_t)));
    CK(cudaMalloc(&h->d_rowidx[1], N1 * sizeof(int32_t)));
    CK(cudaMalloc(&h->d_flags, N1));
    CK(cudaMalloc(&h->d_node_of_row, N1 * sizeof(int32_t)));
    CK(cudaMalloc(&h->d_start, (F + 1) * sizeof(int32_t)));
    CK(cudaMalloc(&h->d_nbins, F * sizeof(int32_t)));
    CK(cudaMalloc(&h->d_feat_owner, F * sizeof(int32_t)));
    CK(cudaMalloc(&h->d_feat_zero, F * sizeof(int32_t)));
    CK(cudaMalloc(&h->d_off_inner, F * sizeof(int64_t)));
    CK(cudaMalloc(&h->d_off_flat, F * sizeof(int64_t)));
    CK(cudaMalloc(&h->d_group_mask, h->groups * sizeof(uint32_t)));
    h->wide_index.assign(F, -1);
    h->n_wide = 0;
    for (int f = 0; f < F; f++) if (h->nbins[f] > ytk::kMaxBins) h->wide_index[f] = h->n_wide++;
    if (h->sparse) {                        // slots live in the CSR entries (16 bits each): no uint16 columns
        h->wide_index.assign(F, -1);
        h->n_wide = 0;
    }
    if (h->n_wide > 0) {
        CK(cudaMalloc(&h->d_wide, (size_t)h->n_wide * N1 * sizeof(uint16_t)));
        CK(cudaMemsetAsync(h->d_wide, 0, (size_t)h->n_wide * N1 * sizeof(uint16_t), h->st));
        CK(cudaMalloc(&h->d_wide_index, F * sizeof(int32_t)));
        CK(cudaMemcpyAsync(h->d_wide_index, h->wide_index.data(), F * sizeof(int32_t), cudaMemcpyHostToDevice, h->st));
        CK(cudaMalloc(&h->d_wide_fids, h->n_wide * sizeof(int32_t)));
    }
    CK(cudaMalloc(&h->d_fea_sel, F));
    CK(cudaMalloc(&h->d_hist_flat, 2LL * h->B * sizeof(long long)));
    CK(cudaMalloc(&h->d_hist_flat2, 2LL * h->B * sizeof(long long)));
    CK(cudaMalloc(&h->d_hist_dbl, 2LL * h->B * sizeof(double)));
    CK(cudaMalloc(&h->d_scale, sizeof(ytk::GradScale)));
    CK(cudaMalloc(&h->d_scale_hook, sizeof(ytk::GradScale)));
    CK(cudaMalloc(&h->d_absmax, 2 * sizeof(unsigned)));
    CK(cudaMalloc(&h->d_bad, sizeof(int)));
    CK(cudaMemsetAsync(h->d_bad, 0, sizeof(int), h->st));
    CK(cudaMemcpyAsync(h->d_start, start, (F + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, h->st));
    CK(cudaMemcpyAsync(h->d_nbins, h->nbins.data(), F * sizeof(int32_t), cudaMemcpyHostToDevice, h->st));
    CK(cudaMemcpyAsync(h->d_feat_owner, owner.data(), F * sizeof(int32_t), cudaMemcpyHostToDevice, h->st));
    CK(cudaMemcpyAsync(h->d_feat_zero, zeros.data(), F * sizeof(int32_t), cudaMemcpyHostToDevice, h->st));
    CK(cudaMemcpyAsync(h->d_off_inner, off_inner.data(), F * sizeof(int64_t), cudaMemcpyHostToDevice, h->st));
    CK(cudaMemcpyAsync(h->d_off_flat, off_flat.data(), F * sizeof(int64_t), cudaMemcpyHostToDevice, h->st));
    CK(h->h_scale.reserve(1));
    CK(cudaStreamSynchronize(h->st));

    // pool capacity (DataParallelTreeMaker.java:193-203); UNIT = 64, MB = 1<<20
    int capacity;
    if (h->p.histogram_pool_capacity > 0) {
        capacity = (int)(h->p.histogram_pool_capacity * 1048576.0 / 64 / h->B);
        capacity = std::max(2, capacity);
        capacity = std::min(capacity, h->p.max_leaf_cnt);
    } else {
        capacity = h->p.max_leaf_cnt;
    }
    if (capacity < 1) return fail(YTK_ERR_ARG, "histogram pool capacity %d < 1 (max_leaf_cnt %d)", capacity, h->p.max_leaf_cnt);
    h->pool.capacity = capacity;
    {   // Histogram slots up front, in one allocation: a cudaMalloc in the middle of a tree costs a
        // device-wide synchronisation (measured ~0.35 ms each).  A level holds parents and children at
        // once and expansions computed ahead of time add a few more, hence twice the capacity; anything
        // beyond that (or beyond 1 GiB here) is still allocated on demand.
        const size_t slot_bytes = (size_t)h->seg_max * sizeof(long long);
        size_t n_pre = 2 * (size_t)capacity + 2;
        n_pre = std::min(n_pre, std::max<size_t>(2, (size_t(1) << 30) / slot_bytes));
        if (h->p.histogram_pool_capacity > 0) n_pre = (size_t)capacity;      // a bounded pool never holds more
        CK(cudaMalloc(&h->d_pool_chunk, n_pre * slot_bytes));
        for (size_t i = 0; i < n_pre; i++) {
            HistPool::Item* it = new HistPool::Item{-1, h->d_pool_chunk + i * (size_t)h->seg_max, false};
            h->pool.all.push_back(it);
            h->pool.avail.push_back(it);
        }
    }
    if (h->world > 1) RET(peer_setup(h));
    h->rows_set = true;
    return YTK_OK;
}

static int check_rows(ytk_gbdt* h, const void* p, int64_t row0, int64_t nrows, const char* who) {
    if (!h || !h->rows_set) return fail(YTK_ERR_ARG, "%s: call set_bins first", who);
    if (nrows < 0 || row0 < 0 || row0 + nrows > h->N)
        return fail(YTK_ERR_ARG, "%s: rows [%lld,%lld) outside [0,%lld)", who, (long long)row0,
                    (long long)(row0 + nrows), (long long)h->N);
    if (nrows > 0 && !p) return fail(YTK_ERR_ARG, "%s: null buffer", who);
    return YTK_OK;
}

int ytk_gbdt_upload_rows(ytk_gbdt* h, const int32_t* x_chunk, int64_t row0, int64_t nrows, int32_t stride) {
    RET(check_rows(h, x_chunk, row0, nrows, "upload_rows"));
    if (h->sparse) return fail(YTK_ERR_ARG, "upload_rows: this handle holds sparse rows (ytk_gbdt_upload_csr)");
    if (stride < h->F) return fail(YTK_ERR_ARG, "upload_rows: stride %d < feature dim %d", stride, h->F);
    if (nrows == 0) return YTK_OK;
    CK(cudaSetDevice(h->device));
    const int64_t step = std::max<int64_t>(1, (64LL << 20) / ((int64_t)stride * 4));
    for (int64_t r = 0; r < nrows; r += step) {
        int64_t n = std::min(step, nrows - r);
        size_t bytes = (size_t)n * stride * sizeof(int32_t);
        RET(ensure_stage(h, bytes));
        CK(cudaMemcpyAsync(h->d_stage, x_chunk + r * stride, bytes, cudaMemcpyHostToDevice, h->st));
        CK(ytk::launch_repack_rows((const int32_t*)h->d_stage, n, stride, h->F, h->d_nbins, h->d_bins + (row0 + r) * h->S8,
                                   h->S8, h->d_bins_col + (row0 + r), h->N, wide_cols(h, row0 + r), row0 + r, h->d_bad,
                                   h->st));
        CK(cudaStreamSynchronize(h->st));
    }
    int bad = 0;
    CK(cudaMemcpy(&bad, h->d_bad, sizeof(int), cudaMemcpyDeviceToHost));
    if (bad) {
        CK(cudaMemset(h->d_bad, 0, sizeof(int)));
        return fail(YTK_ERR_ARG, "upload_rows: slot index outside [0, bins of its feature)");
    }
    return YTK_OK;
}

int ytk_gbdt_quantize_rows(ytk_gbdt* h, const int32_t* xbits_chunk, int64_t row0, int64_t nrows, int32_t stride,
                           const float* cuts) {
    RET(check_rows(h, xbits_chunk, row0, nrows, "quantize_rows"));
    if (h->sparse) return fail(YTK_ERR_ARG, "quantize_rows: this handle holds sparse rows (ytk_gbdt_upload_csr)");
    if (!cuts) return fail(YTK_ERR_ARG, "quantize_rows: null cuts");
    if (stride < h->F) return fail(YTK_ERR_ARG, "quantize_rows: stride %d < feature dim %d", stride, h->F);
    if (nrows == 0) return YTK_OK;
    CK(cudaSetDevice(h->device));
    float* d_cuts = nullptr;
    CK(cudaMalloc(&d_cuts, (size_t)h->B * sizeof(float)));
    cudaError_t e = cudaMemcpyAsync(d_cuts, cuts, (size_t)h->B * sizeof(float), cudaMemcpyHostToDevice, h->st);
    const int64_t step = std::max<int64_t>(1, (64LL << 20) / ((int64_t)stride * 4));
    for (int64_t r = 0; r < nrows && e == cudaSuccess; r += step) {
        int64_t n = std::min(step, nrows - r);
        size_t bytes = (size_t)n * stride * sizeof(int32_t);
        if (ensure_stage(h, bytes) != YTK_OK) { cudaFree(d_cuts); return YTK_ERR_CUDA; }
        e = cudaMemcpyAsync(h->d_stage, xbits_chunk + r * stride, bytes, cudaMemcpyHostToDevice, h->st);
        if (e == cudaSuccess)
            e = ytk::launch_quantize_rows((const int32_t*)h->d_stage, n, stride, h->F, d_cuts, h->d_start,
                                          h->d_bins + (row0 + r) * h->S8, h->S8, h->d_bins_col + (row0 + r), h->N,
                                          wide_cols(h, row0 + r), row0 + r, h->st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->st);
    }
    cudaFree(d_cuts);
    CK(e);
    return YTK_OK;
}

int ytk_gbdt_upload_csr(ytk_gbdt* h, const int64_t* row_ptr, const int32_t* fids, const int32_t* slots, int64_t row0,
                        int64_t nrows) {
    RET(check_rows(h, row_ptr, row0, nrows, "upload_csr"));
    if (!h->sparse) return fail(YTK_ERR_ARG, "upload_csr: call ytk_gbdt_set_bins_sparse first");
    if (row0 != h->csr_rows_done) return fail(YTK_ERR_ARG, "upload_csr: chunks go in row order (expected row %lld, got %lld)",
                                              (long long)h->csr_rows_done, (long long)row0);
    if (nrows == 0) return YTK_OK;
    const int64_t nnz = row_ptr[nrows] - row_ptr[0];
    for (int64_t r = 0; r < nrows; r++)
        if (row_ptr[r + 1] < row_ptr[r]) return fail(YTK_ERR_ARG, "upload_csr: row_ptr decreases at row %lld", (long long)(row0 + r));
    if (nnz > 0 && (!fids || !slots)) return fail(YTK_ERR_ARG, "upload_csr: null entry arrays");
    if (h->csr_nnz_done + nnz > h->nnz_cap)
        return fail(YTK_ERR_CAPACITY, "upload_csr: %lld entries exceed the %lld announced to set_bins_sparse",
                    (long long)(h->csr_nnz_done + nnz), (long long)h->nnz_cap);
    CK(cudaSetDevice(h->device));
    // absolute row offsets
    std::vector<int64_t> abs_ptr(nrows + 1);
    for (int64_t r = 0; r <= nrows; r++) abs_ptr[r] = h->csr_nnz_done + (row_ptr[r] - row_ptr[0]);
    CK(cudaMemcpyAsync(h->d_row_ptr + row0, abs_ptr.data(), (nrows + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, h->st));
    if (nnz > 0) {
        RET(ensure_stage(h, (size_t)nnz * 2 * sizeof(int32_t)));
        int32_t* d_f = (int32_t*)h->d_stage;
        int32_t* d_s = d_f + nnz;
        CK(cudaMemcpyAsync(d_f, fids + row_ptr[0], nnz * sizeof(int32_t), cudaMemcpyHostToDevice, h->st));
        CK(cudaMemcpyAsync(d_s, slots + row_ptr[0], nnz * sizeof(int32_t), cudaMemcpyHostToDevice, h->st));
        CK(ytk::launch_csr_pack(h->d_row_ptr + row0, d_f, d_s, nrows, h->F, h->d_nbins, h->d_nz + h->csr_nnz_done, h->d_bad, h->st));
    }
    CK(cudaStreamSynchronize(h->st));
    int bad = 0;
    CK(cudaMemcpy(&bad, h->d_bad, sizeof(int), cudaMemcpyDeviceToHost));
    if (bad) {
        CK(cudaMemset(h->d_bad, 0, sizeof(int)));
        return fail(YTK_ERR_ARG, "upload_csr: an entry has a feature or slot out of range, or a row is not sorted by feature");
    }
    h->csr_rows_done += nrows;
    h->csr_nnz_done += nnz;
    return YTK_OK;
}

int ytk_gbdt_download_rows(ytk_gbdt* h, int32_t* out, int64_t row0, int64_t nrows) {
    RET(check_rows(h, out, row0, nrows, "download_rows"));
    if (nrows == 0) return YTK_OK;
    CK(cudaSetDevice(h->device));
    const int64_t step = std::max<int64_t>(1, (64LL << 20) / ((int64_t)h->F * 4));
    for (int64_t r = 0; r < nrows; r += step) {
        int64_t n = std::min(step, nrows - r);
        size_t bytes = (size_t)n * h->F * sizeof(int32_t);
        RET(ensure_stage(h, bytes));
        if (h->sparse) CK(ytk::launch_csr_to_dense(sparse_rows(h), row0 + r, n, h->F, (int32_t*)h->d_stage, h->st));
        else
        CK(ytk::launch_unpack_rows(h->d_bins + (row0 + r) * h->S8, h->S8, n, h->F, wide_cols(h, row0 + r),
                                   (int32_t*)h->d_stage, h->st));
        CK(cudaMemcpyAsync(out + r * h->F, h->d_stage, bytes, cudaMemcpyDeviceToHost, h->st));
        CK(cudaStreamSynchronize(h->st));
    }
    return YTK_OK;
}

int ytk_gbdt_set_grads(ytk_gbdt* h, const float* gp_chunk, int64_t row0, int64_t nrows) {
    RET(check_rows(h, gp_chunk, row0, nrows, "set_grads"));
    CK(cudaSetDevice(h->device));
    const int K = h->p.class_num;
    if (nrows > 0) {
        CK(cudaMemcpyAsync(h->d_gp + row0 * K * 2, gp_chunk, (size_t)nrows * K * 2 * sizeof(float),
                           cudaMemcpyHostToDevice, h->st));
        CK(cudaStreamSynchronize(h->st));
    }
    h->grads_set = true;
    return YTK_OK;
}

// -------------------------------------------------------------------------------------
// ITreeMaker.make
// -------------------------------------------------------------------------------------
}  // extern "C"

// =====================================================================================
// One make() call: DataParallelTreeMaker.make(groupId) (:229-295)
// =====================================================================================
namespace {

// What one batch of launches works on: a level, a single node, or a set of splits computed ahead of time.
struct Batch {
    std::vector<SplitJob> jobs;
    bool use_planned = false;       // its partition was planned (and started) by the device behind the previous batch
    int par = 0, slots = 0;         // ... in this set of planned tables, with this many slots
    std::vector<int> cs;            // job k -> its entry in the segment / count tables (k, or its planned slot)
    const ytk::SplitSeg* d_segs = nullptr;
    const ytk::PartitionCounts *d_cl = nullptr, *d_cg = nullptr;
    PartPlan pp;
    // work tables of the kernels behind the partition
    std::vector<ytk::RowItem> leaf_items;
    std::vector<HistJob> hist_jobs;
    std::vector<ytk::SubItem> subs;
    std::vector<ytk::ScanNode> scans;
    std::vector<int> scan_job;      // scans[2i], scans[2i+1] = smaller, larger child of jobs[scan_job[i]]
    int n_leaf_jobs = 0;
    HistPlan hp;
    size_t o_leaf = 0, o_subs = 0, o_scans = 0, o_spec = 0;
    bool plan_now = false;          // the device plans the next level's partition behind this batch
    long long grid_bound = 0;
    double t_pop = 0;
    ytk_gbdt::TraceRec tr{};
};

class TreeGrower {
public:
    TreeGrower(ytk_gbdt* handle, ytk_tree_out* tree, ytk_time_stats* stats)
        : h(handle), p(handle->p), out(tree), T(stats) {
        loss_wise = p.grow_policy == 1;
        lr = p.learning_rate;
        // The reference's pool never evicts when it holds max_leaf_cnt histograms (its own allocation
        // order cannot exceed that), so slot management is then free and whole levels can be batched.
        unbounded = (h->pool.capacity >= p.max_leaf_cnt && p.max_leaf_cnt > 0) || (p.max_leaf_cnt <= 0);
        batch_levels = unbounded && !loss_wise && !h->force_sequential;
        // With an unbounded pool and no row-count stop rule nothing the host decides next depends on
        // the child counts, so the histogram/leaf kernels are queued right behind the partition with
        // "smaller child of segment k" work items resolved on the device: one host wait per batch
        // instead of two.
        deferred = unbounded && p.min_split_samples <= 0;
        // Loss-wise growth polls one node at a time, but what a split produces (child rows, the children's
        // histograms and best splits) does not depend on when it is applied.  With lookahead > 1 the
        // expansions of the best few queue entries are computed by one batch of launches and applied
        // later, strictly in the reference's order: same tree, one host wait per batch instead of per split.
        lookahead = (loss_wise && deferred && !h->force_sequential) ? h->lookahead : 1;
        // Level-wise growth: right behind a level's split search the device itself plans and starts the
        // partition of the next level (plan_next_partition_k), so that it runs while the host is woken
        // up, reads the records and builds the other tables.  The host takes the same decisions from
        // the same records and checks them against the device's (counts of unused slots are -1).
        plan_ahead_ok = batch_levels && deferred && (h->world == 1 || h->peer_mode) && h->single_pass_partition
                        && h->plan_ahead;
    }

    int grow(int gid, const int32_t* fea_index, int32_t n_fea, const uint8_t* row_mask) {
        RET(begin_tree(gid, fea_index, n_fea, row_mask));
        RET(search_root());
        queue.push_back({root_id, ns[root_id].best_lc, 0, seq++, planned_next ? 0 : -1});
        while (!queue.empty()) {
            Batch b;
            b.t_pop = now_ms();
            bool applied = false;
            if (lookahead > 1) RET(poll_ahead(b, &applied));
            else RET(poll_batch(b));
            if (applied) continue;
            // this batch's partition may already be running (planned by the device behind the previous one)
            b.use_planned = planned_next; b.par = planned_par; b.slots = planned_slots;
            planned_next = false;
            if (b.jobs.empty()) continue;
            RET(expand(b));
            // applied at once, in the order polled -- unless these were expansions computed ahead of time
            if (lookahead <= 1)
                for (const SplitJob& j : b.jobs) commit(j);
        }
        return finish(row_mask);
    }

private:
    ytk_gbdt* h;
    const ytk_gbdt_params& p;
    ytk_tree_out* out;
    ytk_time_stats* T;
    bool loss_wise = false, unbounded = false, batch_levels = false, deferred = false, plan_ahead_ok = false;
    int lookahead = 1;
    float lr = 0.f;
    int first_owned = -1, n_sel = 0;

    // Per-node working state is indexed by HANDLE.  A node gets its handle when the split that
    // creates it is computed, and its id in the tree (Tree.AllocTreeNode order) when that split is
    // applied; unless expansions are computed ahead of time (lookahead) both happen together and
    // handle == node id.
    std::vector<Interval> iv;
    std::vector<NodeStats> ns;
    std::vector<Expansion> exps;        // by parent handle
    std::vector<int> nid_of;            // handle -> node id, -1 while the node is not in the tree
    std::vector<int> leaf_handles;
    Interval root;
    int root_id = 0;

    std::vector<ExpandNode> queue;
    int seq = 0, num_leaf = 1;
    std::vector<ytk::PartitionCounts> cl, cg;
    std::vector<ytk_split> recs;

    bool planned_next = false;          // the partition of the batch about to be polled has been planned by the device
    int planned_par = 0;                // which of the two planned-table sets it uses
    int planned_slots = 0;

    // ---- small helpers -----------------------------------------------------------------
    int new_handle() {
        iv.emplace_back(); ns.emplace_back(); exps.emplace_back(); nid_of.push_back(-1);
        return (int)iv.size() - 1;
    }
    int alloc_node(int parent) {        // Tree.AllocTreeNode (J/data/gbdt/Tree.java:58-64)
        if (out->n_nodes >= out->capacity) return -1;
        int nid = out->n_nodes++;
        out->parent[nid] = parent; out->left[nid] = -1; out->right[nid] = -1;
        out->fid[nid] = 0; out->lo[nid] = 0; out->hi[nid] = 0; out->node_info[nid] = 0.f;
        out->loss_chg[nid] = 0.f; out->hess_sum[nid] = 0.f; out->sample_cnt[nid] = 0;
        return nid;
    }
    void set_split_if_good(int hd) {    // findBestSplit tail (:587-592), TreeNode.setSplit (:144-149)
        const NodeStats& s = ns[hd];
        const int nid = nid_of[hd];
        if ((double)s.best_lc > p.min_split_loss) {
            out->fid[nid] = s.best_fid; out->lo[nid] = s.best_lo; out->hi[nid] = s.best_hi;
            out->node_info[nid] = 0.5f * (float)(s.best_lo + s.best_hi);
        }
    }
    void set_leaf(int hd) {             // :253
        const int nid = nid_of[hd];
        out->node_info[nid] = ns[hd].node_value * lr;
        out->left[nid] = -1; out->right[nid] = -1;
        leaf_handles.push_back(hd);
        if (unbounded) h->pool.release(hd);        // only recycles memory; the reference keeps it unused
    }
    int pool_slot(int key, long long** buf) {      // pool entries are keyed by handle
        bool fresh = false;
        HistPool::Item* it = h->pool.alloc(key, &fresh, unbounded);
        if (fresh) { CK(cudaMalloc(&it->buf, h->seg_max * sizeof(long long))); T->device_allocs++; }
        *buf = it->buf;
        return YTK_OK;
    }
    // PriorityQueue order with lossChgComparator / sequenceComparator (:105-115): true if a is polled before b
    bool polled_before(const ExpandNode& a, const ExpandNode& b) const {
        int c = loss_wise ? java_float_compare(a.loss_chg, b.loss_chg) : 0;      // larger lossChg first
        if (c != 0) return c > 0;
        return a.seq < b.seq;
    }
    size_t head_index(const std::vector<ExpandNode>& q) const {
        size_t bi = 0;
        for (size_t i = 1; i < q.size(); i++)
            if (polled_before(q[i], q[bi])) bi = i;
        return bi;
    }
    // stop rules of the loop head that hold whenever the node is polled (:249-250) ...
    bool never_splits(const ExpandNode& e) const {
        return (double)e.loss_chg <= p.min_split_loss || (p.max_depth >= 0 && p.max_depth == e.depth);
    }
    // ... and all of them (:249-252)
    bool becomes_leaf(const ExpandNode& e) const {
        return never_splits(e)
            || (p.max_leaf_cnt > 0 && p.max_leaf_cnt == num_leaf)
            || (p.min_split_samples > 0 && ns[e.hd].sample_cnt < p.min_split_samples);
    }
    // tree.addChilds (:256): the children of a node being split get their ids, in the order splits are applied
    int add_children(int hd, int hl, int hr) {
        const int nid = nid_of[hd];
        const int l = alloc_node(nid), r = alloc_node(nid);
        if (l < 0 || r < 0) return fail(YTK_ERR_CAPACITY, "make_tree: tree needs more than %d nodes", out->capacity);
        out->left[nid] = l; out->right[nid] = r;
        nid_of[hl] = l; nid_of[hr] = r;
        num_leaf++;
        return YTK_OK;
    }
    // first two clauses of the children-are-leaves test (:266-267); the third needs the counts
    bool children_are_leaves(int depth) const {
        return (p.max_depth >= 0 && p.max_depth == depth + 1) || (p.max_leaf_cnt > 0 && p.max_leaf_cnt == num_leaf);
    }

    // ---- initAssistData (:395-470) and the root node ----------------------------------------
    int begin_tree(int gid, const int32_t* fea_index, int32_t n_fea, const uint8_t* row_mask) {
        RET(set_features(h, fea_index, n_fea, &first_owned, &n_sel));
        h->pool.clear();
        CK(ytk::launch_grad_quantize(h->d_gp, h->N, p.class_num, gid, h->d_gq, h->d_absmax, h->d_scale, h->fixed_bits,
                                     h->st));
        h->launches += 2;
        CK(cudaMemcpyAsync(h->h_scale.h, h->d_scale, sizeof(ytk::GradScale), cudaMemcpyDeviceToHost, h->st));
        if (row_mask) {
            if (!h->d_mask) CK(cudaMalloc(&h->d_mask, std::max<int64_t>(h->N, 1)));
            CK(cudaMemcpyAsync(h->d_mask, row_mask, h->N, cudaMemcpyHostToDevice, h->st));
            std::vector<ytk::SplitSeg> sg(1);
            sg[0] = {0, (long long)h->N, -1, -1, 0.f, 0, 0, 0};
            RET(partition_nodes(h, h->d_mask, 0, sg, cl, cg));
            root.b = 0; root.e = cl[0].left; root.buf = 0;     // sampled rows first
        } else {
            root.b = 0; root.e = h->N; root.buf = -1;
        }
        out->n_nodes = 0;
        root_id = new_handle();
        nid_of[root_id] = alloc_node(-1);
        iv[root_id] = root;
        // initTreeMakerNodeStats(0) (:510-520): allreduce(long, SUM)
        int64_t local = root.e - root.b, global = local;
        if (h->world > 1) {
            CK(h->t_counts.reserve(2));
            CK(h->h_counts.reserve(2));
            h->h_counts.h[0].left = local; h->h_counts.h[0].right = 0;
            CK(cudaMemcpyAsync(h->t_counts.d, h->h_counts.h, sizeof(ytk::PartitionCounts), cudaMemcpyHostToDevice, h->st));
            NK(ytk::nccl_api().AllReduce(h->t_counts.d, h->t_counts.d + 1, 2, ncclInt64, ncclSum, h->comm, h->st));
            CK(cudaMemcpyAsync(h->h_counts.h + 1, h->t_counts.d + 1, sizeof(ytk::PartitionCounts), cudaMemcpyDeviceToHost, h->st));
            CK(cudaStreamSynchronize(h->st));
            global = h->h_counts.h[1].left;
        }
        ns[root_id].sample_cnt = global;
        return YTK_OK;
    }

    // Buffers of a device-planned partition with `nslots` slots and at most `grid_bound` tiles, in table set `par`.
    int reserve_planned(int par, int nslots, long long grid_bound) {
        const int ncnt = h->world > 1 ? 2 * nslots : nslots;      // [0,nslots) local, [nslots,2 nslots) global
        CK(h->t_spec_segs[par].reserve((size_t)nslots));
        CK(h->t_spec_counts[par].reserve((size_t)ncnt));
        CK(h->h_spec_counts[par].reserve((size_t)ncnt));
        CK(h->t_spec_tiles.reserve((size_t)std::max<long long>(1, grid_bound)));
        if (!h->d_spec_ntiles) CK(cudaMalloc(&h->d_spec_ntiles, sizeof(int)));
        if (!h->ev_level) CK(cudaEventCreateWithFlags(&h->ev_level, cudaEventDisableTiming));
        return next_partition_epoch(h, (size_t)std::max<long long>(1, grid_bound));
    }
    // Behind a planned partition: allreduceArray(long[2], SUM) (:538) of its counts, and the copy the host reads.
    int finish_planned(int par, int nslots) {
        const int ncnt = h->world > 1 ? 2 * nslots : nslots;
        if (h->world > 1) {
            RET(comm_mark(h));
            const unsigned long long ep = ++h->epoch[0];
            CK(ytk::launch_peer_allreduce(h->pt, h->t_spec_counts[par].d, h->off_counts, h->slot_counts, 2 * (size_t)nslots, 0,
                                          ep, reinterpret_cast<long long*>(h->t_spec_counts[par].d + nslots), h->d_peer_err,
                                          h->st));
            h->launches++;
            RET(comm_mark(h));
        }
        CK(cudaMemcpyAsync(h->h_spec_counts[par].h, h->t_spec_counts[par].d, (size_t)ncnt * sizeof(ytk::PartitionCounts),
                           cudaMemcpyDeviceToHost, h->st));
        h->launches += 2;
        planned_next = true; planned_par = par; planned_slots = nslots;
        return YTK_OK;
    }
    const ytk_split* device_records() const { return h->world > 1 ? h->t_split_all.d : h->t_split_local.d; }

    // ---- findBestSplit(root) (:302-318) ------------------------------------------------------
    int search_root() {
        double t0 = now_ms();
        long long* slot = nullptr;
        RET(pool_slot(root_id, &slot));
        RET(build_hists(h, {{slot, -1, root, 0}}, T, n_sel, nullptr, nullptr));
        double t1 = now_ms();
        Blob bl;
        const size_t o_scan = bl.add(std::vector<ytk::ScanNode>{{slot, (long long)ns[root_id].sample_cnt, -1, 0}});
        RET(blob_commit(h, bl));
        RET(search_enqueue(h, blob_ptr<ytk::ScanNode>(h, o_scan), 1, first_owned, nullptr));
        // the root's own partition, planned by the device from its record (as for the levels below)
        const bool plan_root = plan_ahead_ok && p.max_depth != 0 && (p.max_leaf_cnt <= 0 || p.max_leaf_cnt > 1)
                               && (h->world == 1 || (h->recs_merged && 1 <= h->peer_cap_nodes));
        if (plan_root) {
            const long long grid_bound = (root.e - root.b + ytk::kFusedTile - 1) / ytk::kFusedTile;
            RET(reserve_planned(0, 1, grid_bound));
            CK(cudaEventRecord(h->ev_level, h->st));
            CK(ytk::launch_partition_planned_root(h->d_bins_col, h->S8, h->N, wide_cols(h), h->d_rowidx[0], h->d_rowidx[1],
                                                  root.b, root.e, root.buf, device_records(), p.min_split_loss,
                                                  h->t_spec_segs[0].d, h->t_spec_tiles.d, h->d_spec_ntiles, (int)grid_bound,
                                                  h->d_tile_state, h->d_ticket + 1, h->part_epoch, h->t_spec_counts[0].d,
                                                  h->st, sparse_rows(h)));
            RET(finish_planned(0, 1));
            CK(cudaEventSynchronize(h->ev_level));
        } else {
            CK(cudaStreamSynchronize(h->st));
        }
        h->arena.rewind();
        search_collect(h, 1, recs);
        apply_record(ns[root_id], recs[0]);
        set_split_if_good(root_id);
        T->build_hist += t1 - t0;
        T->find_best_split += now_ms() - t1;
        return YTK_OK;
    }

    // What a computed expansion says about the two children, written to their handles: row
    // intervals, global counts and -- unless they are leaves -- histograms and best splits.
    void materialise(const SplitJob& j) {
        Expansion& x = exps[j.hd];
        const Interval pv = iv[j.hd];
        const int cbuf = pv.buf == 0 ? 1 : 0;
        iv[j.l].b = pv.b; iv[j.l].e = pv.b + x.cl.left; iv[j.l].buf = cbuf;
        iv[j.r].b = pv.b + x.cl.left; iv[j.r].e = pv.e; iv[j.r].buf = cbuf;
        ns[j.l].sample_cnt = x.cg.left; ns[j.r].sample_cnt = x.cg.right;
        if (!x.by_hist) return;
        // smaller child: strictly fewer global rows on the left, else the right (:344-350, :493-499)
        const int small = x.cg.left < x.cg.right ? j.l : j.r, large = small == j.l ? j.r : j.l;
        if (x.placeholders) {
            h->pool.rename(-2 - 2 * j.hd, small);
            h->pool.rename(-3 - 2 * j.hd, large);
            x.placeholders = false;
        }
        apply_record(ns[small], x.rec[0]);
        apply_record(ns[large], x.rec[1]);
    }

    // Applies a split whose expansion is there and whose children have their ids: leaf statistics
    // (computeLeafStats :341-385), or the children's splits and two new queue entries.
    void commit(const SplitJob& j) {
        Expansion& x = exps[j.hd];
        const int small = x.cg.left < x.cg.right ? j.l : j.r, large = small == j.l ? j.r : j.l;
        if (j.terminal) {
            NodeStats &ss = ns[small], &ls = ns[large], &ps = ns[j.hd];
            double sG, sH;
            if (x.by_hist) {
                // computed ahead of time as an inner split, applied as the last split of the tree: the
                // smaller child's totals are the same integer sums, read from its histogram instead of
                // its rows; the children were never searched as far as the tree is concerned
                sG = x.rec[0].sum_grad; sH = x.rec[0].sum_hess;
                for (NodeStats* s2 : {&ss, &ls}) {
                    const int64_t cnt = s2->sample_cnt;
                    *s2 = NodeStats();
                    s2->sample_cnt = cnt;
                }
                h->pool.release(small); h->pool.release(large);
            } else {
                sG = (double)x.leaf[0] * h->h_scale.h->inv_g;
                sH = (double)x.leaf[1] * h->h_scale.h->inv_h;
            }
            ss.sumG = sG; ss.sumH = sH;
            ls.sumG = ps.sumG - ss.sumG; ls.sumH = ps.sumH - ss.sumH;      // setSubstract (:382)
            for (NodeStats* s2 : {&ss, &ls}) ytk_host_node_value(&p, s2->sumG, s2->sumH, &s2->root_gain, &s2->node_value);
            set_leaf(j.l); set_leaf(j.r);
        } else {
            set_split_if_good(small);
            set_split_if_good(large);
            queue.push_back({j.l, ns[j.l].best_lc, j.depth + 1, seq++, planned_next ? 2 * j.k : -1});
            queue.push_back({j.r, ns[j.r].best_lc, j.depth + 1, seq++, planned_next ? 2 * j.k + 1 : -1});
        }
        x.have = false;
    }

    // ---- poll (:246-258), loss-wise with expansions computed ahead of time: the head becomes a leaf, or
    // its split is applied if it has been computed (*applied = true); otherwise b.jobs = what to compute.
    int poll_ahead(Batch& b, bool* applied) {
        const size_t hi = head_index(queue);
        const ExpandNode e = queue[hi];
        if (becomes_leaf(e)) {
            queue.erase(queue.begin() + hi);
            set_leaf(e.hd);
            *applied = true;
            return YTK_OK;
        }
        if (exps[e.hd].have) {
            queue.erase(queue.begin() + hi);
            const int hl = exps[e.hd].hl, hr = exps[e.hd].hr;
            RET(add_children(e.hd, hl, hr));
            commit({e.hd, hl, hr, e.depth, children_are_leaves(e.depth)});
            *applied = true;
            return YTK_OK;
        }
        // The head's split has not been computed.  Play the loop forward on a copy of the queue -- splits
        // already computed put their children into the copy -- and collect the first `lookahead` nodes
        // that would be split and have no expansion yet.  The forecast only chooses what to compute;
        // the order in which splits are applied is decided above.
        std::vector<ExpandNode> sim(queue);
        int vseq = seq;
        long long budget = p.max_leaf_cnt > 0 ? (long long)p.max_leaf_cnt - num_leaf : (1LL << 40);
        while (!sim.empty() && (int)b.jobs.size() < lookahead && budget > 0) {
            const size_t si = head_index(sim);
            const ExpandNode q = sim[si];
            sim.erase(sim.begin() + si);
            if (never_splits(q)) continue;
            budget--;
            if (exps[q.hd].have) {
                const Expansion& x = exps[q.hd];
                if (x.by_hist) {
                    sim.push_back({x.hl, ns[x.hl].best_lc, q.depth + 1, vseq++});
                    sim.push_back({x.hr, ns[x.hr].best_lc, q.depth + 1, vseq++});
                }
                continue;
            }
            const int hl = new_handle(), hr = new_handle();
            exps[q.hd].hl = hl; exps[q.hd].hr = hr;
            b.jobs.push_back({q.hd, hl, hr, q.depth, p.max_depth >= 0 && p.max_depth == q.depth + 1});
        }
        return YTK_OK;
    }

    // ---- poll: one node (the reference's order), or -- level-wise with an unbounded pool -- every node of
    // the level, which the reference would poll back to back in the same (seq) order
    int poll_batch(Batch& b) {
        std::vector<ExpandNode> polled;
        if (batch_levels) {
            polled.swap(queue);
            std::sort(polled.begin(), polled.end(),
                      [this](const ExpandNode& x, const ExpandNode& y) { return polled_before(x, y); });
        } else {
            const size_t hi = head_index(queue);
            polled.push_back(queue[hi]);
            queue.erase(queue.begin() + hi);
        }
        for (const ExpandNode& e : polled) {
            if (becomes_leaf(e)) { set_leaf(e.hd); continue; }
            const int hl = new_handle(), hr = new_handle();
            exps[e.hd].hl = hl; exps[e.hd].hr = hr;
            RET(add_children(e.hd, hl, hr));
            b.jobs.push_back({e.hd, hl, hr, e.depth, children_are_leaves(e.depth), (int)b.jobs.size(), e.slot});
        }
        return YTK_OK;
    }

    // ---- one batch: partition, children's work, one host wait, results into the Expansions ----------
    int expand(Batch& b) {
        const int m = (int)b.jobs.size();
        double t0 = now_ms();
        for (int k = 0; k < m; k++) {               // a fresh Expansion per job (keeping the children's handles)
            const int hl = exps[b.jobs[k].hd].hl, hr = exps[b.jobs[k].hd].hr;
            exps[b.jobs[k].hd] = Expansion();
            exps[b.jobs[k].hd].hl = hl; exps[b.jobs[k].hd].hr = hr;
        }
        b.tr = ytk_gbdt::TraceRec{m, b.t_pop, now_ms(), 0, 0, 0, 0, nullptr, nullptr};
        if (h->trace) {
            CK(cudaEventCreate(&b.tr.e0)); CK(cudaEventCreate(&b.tr.e1));
            CK(cudaEventRecord(b.tr.e0, h->st));
        }
        RET(start_partition(b));
        T->partition += now_ms() - t0;

        t0 = now_ms();
        Blob bl;
        RET(build_tables(b, bl));
        RET(blob_commit(h, bl));
        b.tr.t_tables = now_ms();
        RET(enqueue_children_work(b));
        double t1 = now_ms();
        RET(search_enqueue(h, blob_ptr<ytk::ScanNode>(h, b.o_scans), (int)b.scans.size(), first_owned, b.d_cg));
        if (h->trace) CK(cudaEventRecord(b.tr.e1, h->st));
        if (b.plan_now) {
            RET(plan_next_level(b));             // everything the host waits for is in front of its event
            b.tr.t_queued = now_ms();
            CK(cudaEventSynchronize(h->ev_level));   // the one host wait of the batch
        } else {
            b.tr.t_queued = now_ms();
            CK(cudaStreamSynchronize(h->st));        // the one host wait of the batch
        }
        b.tr.t_done = now_ms();
        T->max_batch_ms = std::max(T->max_batch_ms, b.tr.t_done - b.tr.t_pop);
        if (h->trace) h->trace_recs.push_back(b.tr);
        h->arena.rewind();
        RET(collect(b));
        T->build_hist += t1 - t0;
        T->find_best_split += now_ms() - t1;
        return YTK_OK;
    }

    // position.resetPosition (:260) + initTreeMakerNodeStats(l, r) (:264).  The partition's tables go up on
    // their own and the partition starts at once; the tables of everything behind it are built while it
    // runs.  (Or it is running already: planned by the device behind the previous batch.)
    int start_partition(Batch& b) {
        const int m = (int)b.jobs.size();
        // where job k's segment and counts are: entry k of the tables built here, or the slot the
        // device gave the node when it planned this partition
        b.cs.resize(m);
        for (int k = 0; k < m; k++) b.cs[k] = b.use_planned ? b.jobs[k].slot : k;
        if (b.use_planned) {
            for (int k = 0; k < m; k++)
                if (b.cs[k] < 0 || b.cs[k] >= b.slots) return fail(YTK_ERR_ARG, "internal: node without a planned partition slot");
            b.d_segs = h->t_spec_segs[b.par].d;
            b.d_cl = h->t_spec_counts[b.par].d;                               // [0,slots) local, [slots,2 slots) global
            b.d_cg = h->world > 1 ? b.d_cl + b.slots : b.d_cl;
        } else {
            std::vector<ytk::SplitSeg> segs(m);
            for (int k = 0; k < m; k++) {
                const Interval& pv = iv[b.jobs[k].hd];
                const NodeStats& s = ns[b.jobs[k].hd];       // TreeNode.setSplit: cond = 0.5f * (lo + hi) (:144-149)
                segs[k] = {(long long)pv.b, (long long)pv.e, pv.buf, s.best_fid, 0.5f * (float)(s.best_lo + s.best_hi), 0, 0,
                           h->wide_index[s.best_fid] + 1};
            }
            Blob bl_part;
            partition_plan(segs, b.pp, bl_part, h->single_pass_partition);
            CK(h->t_counts.reserve(2 * (size_t)m));     // its address is handed to the kernels queued below
            RET(partition_commit(h, bl_part, b.pp));
            RET(partition_enqueue(h, h->d_bins_col, h->N, b.pp));
            b.d_segs = b.pp.d_segs;
            b.d_cl = h->t_counts.d;
            b.d_cg = h->world > 1 ? h->t_counts.d + m : h->t_counts.d;
        }
        b.tr.t_part = now_ms();
        if (!deferred) {                        // the host needs the counts before it can go on
            CK(cudaStreamSynchronize(h->st));
            h->arena.rewind();
            partition_collect(h, m, cl, cg);
            for (int k = 0; k < m; k++) {
                SplitJob& j = b.jobs[k];
                Expansion& x = exps[j.hd];
                x.cl = cl[k]; x.cg = cg[k];
                materialise(j);
                if (p.min_split_samples > 0 && ns[j.l].sample_cnt < p.min_split_samples
                    && ns[j.r].sample_cnt < p.min_split_samples)
                    j.terminal = true;
            }
        }
        return YTK_OK;
    }

    // The children's work tables: leaf statistics (computeLeafStats :341-385) or histograms + split search
    // (buildHist(parent, l, r) :489-508), and -- if the device is to plan the next level -- the planner's.
    int build_tables(Batch& b, Blob& bl) {
        const int m = (int)b.jobs.size();
        auto small_of = [&](const SplitJob& j) { return ns[j.l].sample_cnt < ns[j.r].sample_cnt ? j.l : j.r; };
        for (int k = 0; k < m; k++) {
            const SplitJob& j = b.jobs[k];
            const Interval pv = iv[j.hd];
            const int cbuf = pv.buf == 0 ? 1 : 0;
            Expansion& x = exps[j.hd];
            if (j.terminal) {
                if (unbounded) h->pool.release(j.hd);
                if (deferred) {
                    const int64_t n = pv.e - pv.b;
                    add_deferred_items(b.leaf_items, pv, cbuf, b.n_leaf_jobs, b.cs[k], std::max<int64_t>(4096, (n / 2 + 63) / 64),
                                       h->world);
                } else {
                    const int small = small_of(j);
                    const int64_t n = iv[small].e - iv[small].b;
                    add_row_items(b.leaf_items, iv[small], b.n_leaf_jobs, std::max<int64_t>(4096, (n + 63) / 64));
                }
                b.n_leaf_jobs++;
                continue;
            }
            // Pool entries are keyed by handle; while the children's roles are still unknown they are
            // keyed by placeholders derived from the parent's handle and renamed once the counts are known.
            x.by_hist = true;
            x.placeholders = deferred;
            const int small = deferred ? -2 - 2 * j.hd : small_of(j);
            const int large = deferred ? -3 - 2 * j.hd : (small == j.l ? j.r : j.l);
            long long* parent_h = h->pool.get(j.hd);
            long long *small_h = nullptr, *large_h = nullptr;
            RET(pool_slot(small, &small_h));
            if (deferred) b.hist_jobs.push_back({small_h, b.cs[k], pv, cbuf});
            else b.hist_jobs.push_back({small_h, -1, iv[small], 0});
            if (parent_h) {
                if (unbounded) {                       // sibling written over its parent
                    h->pool.rename(j.hd, large);
                    large_h = parent_h;
                } else {
                    h->pool.release(j.hd);
                    RET(pool_slot(large, &large_h));
                }
                b.subs.push_back({parent_h, small_h, large_h});      // substract (:101-107)
            } else {
                if (deferred) return fail(YTK_ERR_ARG, "internal: parent histogram missing with an unbounded pool");
                RET(pool_slot(large, &large_h));
                b.hist_jobs.push_back({large_h, -1, iv[large], 0});
                T->hist_miss++;
            }
            if (deferred) {
                b.scans.push_back({small_h, 0, b.cs[k], 0});
                b.scans.push_back({large_h, 0, b.cs[k], 1});
            } else {
                b.scans.push_back({small_h, (long long)ns[small].sample_cnt, -1, 0});
                b.scans.push_back({large_h, (long long)ns[large].sample_cnt, -1, 0});
            }
            b.scan_job.push_back(k);
        }
        RET(hist_plan(h, b.hist_jobs, T, n_sel, b.hp, bl));
        b.o_leaf = bl.add(b.leaf_items); b.o_subs = bl.add(b.subs); b.o_scans = bl.add(b.scans);
        // Can the device plan the next level's partition?  Its candidates are the children searched in
        // this batch; of the stop rules only "lossChg <= min_split_loss" is still open (the depth rule made
        // the job terminal, row counts do not matter here), provided the leaf budget cannot run out
        // while the next level is polled.
        b.plan_now = plan_ahead_ok && !b.scan_job.empty()
                     && (p.max_leaf_cnt <= 0 || num_leaf + 2 * (long long)b.scan_job.size() <= p.max_leaf_cnt)
                     && (h->world == 1 || (2 * m <= h->peer_cap_nodes && (size_t)b.scans.size() * sizeof(ytk_split) <= h->slot_recs));
        if (b.plan_now) {
            std::vector<ytk::SpecJob> spec_jobs(m);
            for (int k = 0; k < m; k++) {
                const Interval& pv = iv[b.jobs[k].hd];
                spec_jobs[k] = {b.cs[k], b.cs[k], -1, pv.buf == 0 ? 1 : 0};
            }
            for (size_t i = 0; i < b.scan_job.size(); i++) {
                const Interval& pv = iv[b.jobs[b.scan_job[i]].hd];
                spec_jobs[b.scan_job[i]].scan = (int)i;
                b.grid_bound += (pv.e - pv.b + ytk::kFusedTile - 1) / ytk::kFusedTile + 1;
            }
            if (b.grid_bound > 0x3fffffff) b.plan_now = false;
            else b.o_spec = bl.add(spec_jobs);
        }
        return YTK_OK;
    }

    // leaf sums, histograms (+ their reduce-scatter), subtraction -- queued behind the partition
    int enqueue_children_work(Batch& b) {
        if (b.n_leaf_jobs > 0) {
            const size_t nl = 2 * (size_t)b.n_leaf_jobs;
            CK(h->t_leaf.reserve(nl));
            CK(h->h_leaf.reserve(nl));
            CK(cudaMemsetAsync(h->t_leaf.d, 0, nl * sizeof(long long), h->st));
            CK(ytk::launch_leaf_sums(h->d_gq, h->d_rowidx[0], h->d_rowidx[1], blob_ptr<ytk::RowItem>(h, b.o_leaf),
                                     (int)b.leaf_items.size(), h->t_leaf.d, b.d_cl, b.d_cg, h->st));
            if (!b.leaf_items.empty()) h->launches++;
            if (h->world > 1) {    // allreduceArray(double[2], SUM) (:375), exact in the integer domain
                if (h->peer_mode && (size_t)b.n_leaf_jobs * 16 <= h->slot_leaf) {
                    const unsigned long long ep = ++h->epoch[3];
                    CK(ytk::launch_peer_allreduce(h->pt, h->t_leaf.d, h->off_leaf, h->slot_leaf, nl, 3, ep, h->t_leaf.d,
                                                  h->d_peer_err, h->st));
                    h->launches++;
                } else {
                    NK(ytk::nccl_api().AllReduce(h->t_leaf.d, h->t_leaf.d, nl, ncclInt64, ncclSum, h->comm, h->st));
                }
            }
            CK(cudaMemcpyAsync(h->h_leaf.h, h->t_leaf.d, nl * sizeof(long long), cudaMemcpyDeviceToHost, h->st));
        }
        RET(hist_enqueue(h, b.hp, T, b.d_cl, b.d_cg));
        if (!b.subs.empty()) {
            CK(ytk::launch_hist_subtract(blob_ptr<ytk::SubItem>(h, b.o_subs), (int)b.subs.size(), h->seg_max, h->st));
            h->launches++;
        }
        return YTK_OK;
    }

    // Behind the split search: the event the host waits on, then the device-planned partition of the
    // next level (slot 2k / 2k+1 = left / right child of job k) and the exchange of its counts.
    int plan_next_level(Batch& b) {
        const int m = (int)b.jobs.size();
        const int np = b.use_planned ? 1 - b.par : 0;
        RET(reserve_planned(np, 2 * m, b.grid_bound));
        CK(cudaEventRecord(h->ev_level, h->st));
        CK(ytk::launch_partition_planned(h->d_bins_col, h->S8, h->N, wide_cols(h), h->d_rowidx[0], h->d_rowidx[1],
                                         b.d_segs, b.d_cl, b.d_cg, blob_ptr<ytk::SpecJob>(h, b.o_spec), m, device_records(),
                                         p.min_split_loss, h->t_spec_segs[np].d, h->t_spec_tiles.d, h->d_spec_ntiles,
                                         (int)b.grid_bound, h->d_tile_state, h->d_ticket + 1, h->part_epoch,
                                         h->t_spec_counts[np].d, h->st, sparse_rows(h)));
        return finish_planned(np, 2 * m);
    }

    // After the host wait: records, counts and leaf sums into the jobs' Expansions, then into the children.
    int collect(Batch& b) {
        const int m = (int)b.jobs.size();
        search_collect(h, (int)b.scans.size(), recs);
        if (b.use_planned) {
            // the device planned this partition: its decisions must be the ones taken here
            const ytk::PartitionCounts* hc = h->h_spec_counts[b.par].h;
            std::vector<char> used(b.slots, 0);
            cl.resize(m); cg.resize(m);
            for (int k = 0; k < m; k++) {
                used[b.cs[k]] = 1;
                cl[k] = hc[b.cs[k]];
                cg[k] = hc[h->world > 1 ? b.slots + b.cs[k] : b.cs[k]];
            }
            for (int sl = 0; sl < b.slots; sl++)
                if ((hc[sl].left >= 0) != (used[sl] != 0))
                    return fail(YTK_ERR_ARG, "internal: the device planned a different partition than the host (slot %d)", sl);
        } else if (deferred) {
            partition_collect(h, m, cl, cg);
        }
        T->batches++;
        for (int k = 0; k < m; k++) {
            Expansion& x = exps[b.jobs[k].hd];
            x.have = true; x.cl = cl[k]; x.cg = cg[k];
        }
        for (size_t i = 0; i < b.scan_job.size(); i++) {
            Expansion& x = exps[b.jobs[b.scan_job[i]].hd];
            x.rec[0] = recs[2 * i]; x.rec[1] = recs[2 * i + 1];
            if (deferred) {
                const double rows = (double)(x.cg.left < x.cg.right ? x.cl.left : x.cl.right);
                T->hist_built++;
                T->hist_cells += rows * n_sel;
                T->hist_rows += rows;
                T->hist_rows_indexed += rows;
            }
        }
        int lj = 0;
        for (int k = 0; k < m; k++) {
            if (!b.jobs[k].terminal) continue;
            Expansion& x = exps[b.jobs[k].hd];
            x.leaf[0] = h->h_leaf.h[2 * lj]; x.leaf[1] = h->h_leaf.h[2 * lj + 1];
            lj++;
        }
        for (int k = 0; k < m; k++) materialise(b.jobs[k]);
        return YTK_OK;
    }

    // updateTreeNodeStat (:387-393), and what get_positions / update_scores need afterwards
    int finish(const uint8_t* row_mask) {
        for (size_t hd = 0; hd < nid_of.size(); hd++) {
            const int nid = nid_of[hd];
            if (nid < 0) continue;                   // computed ahead of time, never applied
            out->loss_chg[nid] = ns[hd].best_lc;
            out->hess_sum[nid] = (float)ns[hd].sumH;
            out->sample_cnt[nid] = ns[hd].sample_cnt;
        }
        // Where the rows of every leaf are.  A leaf whose split was computed ahead of time but never
        // applied has had its rows moved on to its would-be children (and its own range in the other
        // buffer may since have been reused by their children), so its rows are the ranges of its
        // deepest computed descendants.
        h->leaf_ranges.clear();
        h->tree_made = true;
        std::vector<int> stack;
        for (int leaf : leaf_handles) {
            stack.assign(1, leaf);
            while (!stack.empty()) {
                const int hd = stack.back();
                stack.pop_back();
                if (exps[hd].hl >= 0) { stack.push_back(exps[hd].hl); stack.push_back(exps[hd].hr); }
                else if (iv[hd].e > iv[hd].b) h->leaf_ranges.push_back({nid_of[leaf], iv[hd]});
            }
        }
        h->unsampled_begin = row_mask ? root.e : -1;
        return YTK_OK;
    }
};

}  // namespace

extern "C" {

int ytk_gbdt_make_tree(ytk_gbdt* h, int32_t gid, const int32_t* fea_index, int32_t n_fea, const uint8_t* row_mask,
                       ytk_tree_out* out, ytk_time_stats* ts) {
    if (!h || !out) return fail(YTK_ERR_ARG, "make_tree: null argument");
    if (!h->rows_set || !h->grads_set) return fail(YTK_ERR_ARG, "make_tree: set_bins/upload_rows/set_grads first");
    if (gid < 0 || gid >= h->p.class_num) return fail(YTK_ERR_ARG, "make_tree: groupId %d not in [0,%d)", gid, h->p.class_num);
    if (out->capacity < 1) return fail(YTK_ERR_CAPACITY, "make_tree: tree capacity %d", out->capacity);
    CK(cudaSetDevice(h->device));
    const double t_start = now_ms();
    ytk_time_stats local_ts;
    memset(&local_ts, 0, sizeof(local_ts));
    ytk_time_stats* T = &local_ts;
    h->launches = 0;
    h->ev_used = 0;
    h->evc_used = 0;

    TreeGrower grower(h, out, T);
    RET(grower.grow(gid, fea_index, n_fea, row_mask));

    if (h->peer_mode) {
        int perr = 0;
        CK(cudaMemcpyAsync(&perr, h->d_peer_err, sizeof(int), cudaMemcpyDeviceToHost, h->st));
        CK(cudaStreamSynchronize(h->st));
        if (perr) return fail(YTK_ERR_NCCL, "make_tree: a peer GPU did not reach a collective within 3 s");
    }
    if (h->trace) {
        double prev_done = 0;
        for (const auto& r : h->trace_recs) {
            float gpu = 0.f;
            cudaEventElapsedTime(&gpu, r.e0, r.e1);
            fprintf(stderr, "[ytk trace] batch m=%3d  since_prev_wait %6.1f us | pop %5.1f | partition queued %5.1f | tables %6.1f | enqueue %6.1f | wait %7.1f | gpu_span %7.1f us\n",
                    r.m, prev_done > 0 ? (r.t_pop - prev_done) * 1e3 : 0.0, (r.t_jobs - r.t_pop) * 1e3,
                    (r.t_part - r.t_jobs) * 1e3, (r.t_tables - r.t_part) * 1e3,
                    (r.t_queued - r.t_tables) * 1e3, (r.t_done - r.t_queued) * 1e3, gpu * 1e3);
            prev_done = r.t_done;
            cudaEventDestroy(r.e0); cudaEventDestroy(r.e1);
        }
        h->trace_recs.clear();
    }
    h->last_tree.resize(out->n_nodes);
    for (int nid = 0; nid < out->n_nodes; nid++)
        h->last_tree[nid] = {out->left[nid], out->right[nid], out->fid[nid], out->node_info[nid]};
    h->last_gid = gid;
    CK(cudaStreamSynchronize(h->st));
    float gpu_ms = 0.f;
    for (int i = 0; i + 1 < h->ev_used; i += 2) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, h->ev[i], h->ev[i + 1]));
        gpu_ms += ms;
    }
    float comm_ms = 0.f;
    for (int i = 0; i + 1 < h->evc_used; i += 2) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, h->evc[i], h->evc[i + 1]));
        comm_ms += ms;
    }
    T->communicate_hist = comm_ms;     // CUDA-event time of the collectives (stream time, includes waiting for peers)
    T->gpu_hist_ms = gpu_ms;
    T->compute_hist = gpu_ms;
    T->kernel_launches = h->launches;
    T->hist_launches = h->ev_used / 2;
    T->peer_collectives = h->peer_mode ? 1 : 0;
    T->total = now_ms() - t_start;
    if (ts) *ts = *T;
    return YTK_OK;
}

int ytk_gbdt_get_positions(ytk_gbdt* h, int32_t* node_of_row) {
    if (!h || !node_of_row) return fail(YTK_ERR_ARG, "get_positions: null argument");
    if (!h->tree_made) return fail(YTK_ERR_ARG, "get_positions: no tree has been made");
    CK(cudaSetDevice(h->device));
    CK(ytk::launch_fill_i32(h->d_node_of_row, h->N, -1, h->st));
    std::vector<ytk::PosItem> items;
    for (const auto& lr : h->leaf_ranges)
        items.push_back({(long long)lr.second.b, (long long)lr.second.e, lr.second.buf, lr.first});
    CK(h->t_pos.upload(items, h->st, h->arena));
    CK(ytk::launch_fill_positions(h->d_rowidx[0], h->d_rowidx[1], h->t_pos.d, (int)items.size(), h->d_node_of_row, h->st));
    CK(cudaMemcpyAsync(node_of_row, h->d_node_of_row, h->N * sizeof(int32_t), cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    return YTK_OK;
}

// -------------------------------------------------------------------------------------
// §8(f) row 1: score update, loss and gradient pairs on the device
// -------------------------------------------------------------------------------------
int ytk_gbdt_set_targets(ytk_gbdt* h, int32_t loss, const float* y, const float* weight, const float* init_score) {
    if (!h || !y) return fail(YTK_ERR_ARG, "set_targets: null argument");
    if (!h->rows_set) return fail(YTK_ERR_ARG, "set_targets: call set_bins first");
    const int K = h->p.class_num;
    if (loss < 0 || loss > 2) return fail(YTK_ERR_ARG, "set_targets: unknown loss %d", loss);
    // GBDTOptimizationParams.checkParams: only softmax may have class_num > 1 (:222-224)
    if (loss != YTK_LOSS_SOFTMAX && K > 1)
        return fail(YTK_ERR_ARG, "[GBDT] class_num(%d) and objective(%d) inconsistent", K, loss);
    if (K > 64) return fail(YTK_ERR_UNSUPPORTED, "set_targets: class_num %d > 64", K);
    CK(cudaSetDevice(h->device));
    const size_t nk = (size_t)std::max<int64_t>(h->N, 1) * K, n1 = (size_t)std::max<int64_t>(h->N, 1);
    if (!h->d_score) {
        CK(cudaMalloc(&h->d_score, nk * sizeof(float)));
        CK(cudaMalloc(&h->d_y, nk * sizeof(float)));
        CK(cudaMalloc(&h->d_block_loss, ytk::loss_grad_blocks() * sizeof(double)));
        CK(cudaMalloc(&h->d_loss, sizeof(double)));
    }
    CK(cudaMemsetAsync(h->d_score, 0, nk * sizeof(float), h->st));
    CK(cudaMemcpyAsync(h->d_y, y, (size_t)h->N * K * sizeof(float), cudaMemcpyHostToDevice, h->st));
    if (weight) {
        if (!h->d_w) CK(cudaMalloc(&h->d_w, n1 * sizeof(float)));
        CK(cudaMemcpyAsync(h->d_w, weight, (size_t)h->N * sizeof(float), cudaMemcpyHostToDevice, h->st));
    } else if (h->d_w) { CK(cudaFree(h->d_w)); h->d_w = nullptr; }
    if (init_score) {
        if (!h->d_init) CK(cudaMalloc(&h->d_init, nk * sizeof(float)));
        CK(cudaMemcpyAsync(h->d_init, init_score, (size_t)h->N * K * sizeof(float), cudaMemcpyHostToDevice, h->st));
    } else if (h->d_init) { CK(cudaFree(h->d_init)); h->d_init = nullptr; }
    CK(cudaStreamSynchronize(h->st));
    h->loss_kind = loss;
    return YTK_OK;
}

int ytk_gbdt_update_scores(ytk_gbdt* h, int32_t gid) {
    if (!h) return fail(YTK_ERR_ARG, "update_scores: null argument");
    if (h->loss_kind < 0) return fail(YTK_ERR_ARG, "update_scores: call set_targets first");
    if (h->last_tree.empty() || gid != h->last_gid)
        return fail(YTK_ERR_ARG, "update_scores: groupId %d is not the tree the last make() built (%d)", gid, h->last_gid);
    CK(cudaSetDevice(h->device));
    std::vector<ytk::LeafItem> items;
    for (const auto& lr : h->leaf_ranges)
        items.push_back({(long long)lr.second.b, (long long)lr.second.e, lr.second.buf, h->last_tree[lr.first].info});
    CK(h->t_leafval.upload(items, h->st, h->arena));
    CK(ytk::launch_add_leaf_values(h->d_rowidx[0], h->d_rowidx[1], h->t_leafval.d, (int)items.size(), h->d_score,
                                   h->p.class_num, gid, h->st));
    if (h->unsampled_begin >= 0 && h->unsampled_begin < h->N) {
        CK(h->t_tree.upload(h->last_tree, h->st, h->arena));
        // the row-mask compaction left the excluded rows behind the sampled ones in buffer 0, and
        // later partitions only ever rewrite the sampled range
        CK(ytk::launch_walk_rows(h->d_bins, h->S8, wide_cols(h), h->d_rowidx[0], h->unsampled_begin, h->N, h->t_tree.d, h->d_score,
                                 h->p.class_num, gid, h->st, sparse_rows(h)));
    }
    CK(cudaStreamSynchronize(h->st));
    h->arena.rewind();
    return YTK_OK;
}

int ytk_gbdt_score_values(ytk_gbdt* h, const int32_t* xbits, int64_t nrows, int32_t stride, const ytk_tree_out* tree,
                          const float* threshold, const int32_t* default_child, int32_t class_num, int32_t gid,
                          float* score) {
    if (!h || !tree || !threshold || !score) return fail(YTK_ERR_ARG, "score_values: null argument");
    if (nrows < 0 || (nrows > 0 && !xbits)) return fail(YTK_ERR_ARG, "score_values: bad row buffer");
    if (class_num < 1 || gid < 0 || gid >= class_num) return fail(YTK_ERR_ARG, "score_values: groupId %d not in [0,%d)", gid, class_num);
    const int nn = tree->n_nodes;
    if (nn < 1) return fail(YTK_ERR_ARG, "score_values: empty tree");
    std::vector<ytk::TreeNodeDev> nodes(nn);
    for (int i = 0; i < nn; i++) {
        const bool leaf = tree->left[i] == -1;
        if (!leaf) {
            if (tree->left[i] < 0 || tree->left[i] >= nn || tree->right[i] < 0 || tree->right[i] >= nn)
                return fail(YTK_ERR_ARG, "score_values: node %d has children outside the tree", i);
            if (tree->fid[i] < 0 || tree->fid[i] >= stride)
                return fail(YTK_ERR_ARG, "score_values: node %d splits feature %d, rows have %d", i, tree->fid[i], stride);
            if (default_child && default_child[i] != tree->left[i] && default_child[i] != tree->right[i])
                return fail(YTK_ERR_ARG, "score_values: default child of node %d is neither of its children", i);
        }
        nodes[i] = {tree->left[i], tree->right[i], tree->fid[i], leaf ? tree->node_info[i] : threshold[i]};
    }
    if (nrows == 0) return YTK_OK;
    CK(cudaSetDevice(h->device));
    CK(h->t_tree.upload(nodes, h->st, h->arena));
    int32_t* d_def = nullptr;
    float* d_sc = nullptr;
    auto done = [&](int rc) { if (d_def) cudaFree(d_def); if (d_sc) cudaFree(d_sc); h->arena.rewind(); return rc; };
    if (default_child) {
        if (cudaMalloc(&d_def, nn * sizeof(int32_t)) != cudaSuccess) return done(fail(YTK_ERR_CUDA, "score_values: out of device memory"));
        if (cudaMemcpyAsync(d_def, default_child, nn * sizeof(int32_t), cudaMemcpyHostToDevice, h->st) != cudaSuccess)
            return done(fail(YTK_ERR_CUDA, "score_values: copy failed"));
    }
    // rows go through the staging buffer in chunks, their scores beside them
    const int64_t step = std::max<int64_t>(1, (64LL << 20) / ((int64_t)stride * 4));
    if (cudaMalloc(&d_sc, (size_t)std::min(step, nrows) * class_num * sizeof(float)) != cudaSuccess)
        return done(fail(YTK_ERR_CUDA, "score_values: out of device memory"));
    for (int64_t r = 0; r < nrows; r += step) {
        const int64_t n = std::min(step, nrows - r);
        if (ensure_stage(h, (size_t)n * stride * sizeof(int32_t)) != YTK_OK) return done(YTK_ERR_CUDA);
        cudaError_t e = cudaMemcpyAsync(h->d_stage, xbits + r * stride, (size_t)n * stride * sizeof(int32_t), cudaMemcpyHostToDevice, h->st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_sc, score + r * class_num, (size_t)n * class_num * sizeof(float), cudaMemcpyHostToDevice, h->st);
        if (e == cudaSuccess) e = ytk::launch_score_values((const int32_t*)h->d_stage, n, stride, h->t_tree.d, d_def, d_sc, class_num, gid, h->st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(score + r * class_num, d_sc, (size_t)n * class_num * sizeof(float), cudaMemcpyDeviceToHost, h->st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->st);
        if (e != cudaSuccess) return done(fail(YTK_ERR_CUDA, "score_values: %s", cudaGetErrorString(e)));
    }
    return done(YTK_OK);
}

int ytk_gbdt_compute_grads(ytk_gbdt* h, double* loss_sum) {
    if (!h) return fail(YTK_ERR_ARG, "compute_grads: null argument");
    if (h->loss_kind < 0) return fail(YTK_ERR_ARG, "compute_grads: call set_targets first");
    CK(cudaSetDevice(h->device));
    CK(ytk::launch_loss_grad(h->loss_kind, h->d_score, h->d_init, h->d_y, h->d_w, h->N, h->p.class_num, h->d_gp,
                             h->d_block_loss, h->d_loss, h->st));
    double l = 0.0;
    CK(cudaMemcpyAsync(&l, h->d_loss, sizeof(double), cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    if (loss_sum) *loss_sum = l;
    h->grads_set = true;
    return YTK_OK;
}

int ytk_gbdt_get_scores(ytk_gbdt* h, float* out) {
    if (!h || !out) return fail(YTK_ERR_ARG, "get_scores: null argument");
    if (!h->d_score) return fail(YTK_ERR_ARG, "get_scores: call set_targets first");
    CK(cudaSetDevice(h->device));
    CK(cudaMemcpyAsync(out, h->d_score, (size_t)h->N * h->p.class_num * sizeof(float), cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    return YTK_OK;
}

int ytk_gbdt_get_grads(ytk_gbdt* h, float* out) {
    if (!h || !out) return fail(YTK_ERR_ARG, "get_grads: null argument");
    if (!h->rows_set || !h->grads_set) return fail(YTK_ERR_ARG, "get_grads: no gradients on the device yet");
    CK(cudaSetDevice(h->device));
    CK(cudaMemcpyAsync(out, h->d_gp, (size_t)h->N * h->p.class_num * 2 * sizeof(float), cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    return YTK_OK;
}

// -------------------------------------------------------------------------------------
// per-stage hooks
// -------------------------------------------------------------------------------------
int ytk_gbdt_build_hist(ytk_gbdt* h, int32_t gid, const int32_t* rows, int64_t n_rows, const int32_t* fea_index,
                        int32_t n_fea, double* out) {
    if (!h || !out) return fail(YTK_ERR_ARG, "build_hist: null argument");
    if (!h->rows_set || !h->grads_set) return fail(YTK_ERR_ARG, "build_hist: set_bins/upload_rows/set_grads first");
    if (gid < 0 || gid >= h->p.class_num) return fail(YTK_ERR_ARG, "build_hist: groupId %d", gid);
    if (rows && (n_rows < 0 || n_rows > h->N)) return fail(YTK_ERR_ARG, "build_hist: n_rows %lld", (long long)n_rows);
    CK(cudaSetDevice(h->device));
    int first_owned;
    RET(set_features(h, fea_index, n_fea, &first_owned, nullptr));
    CK(ytk::launch_grad_quantize(h->d_gp, h->N, h->p.class_num, gid, h->d_gq, h->d_absmax, h->d_scale, h->fixed_bits,
                                 h->st));
    int64_t n = rows ? n_rows : h->N;
    if (rows && n > 0) {
        for (int64_t i = 0; i < n; i++)
            if (rows[i] < 0 || rows[i] >= h->N) return fail(YTK_ERR_ARG, "build_hist: row %d out of range", rows[i]);
        CK(cudaMemcpyAsync(h->d_rowidx[0], rows, n * sizeof(int32_t), cudaMemcpyHostToDevice, h->st));
    }
    const int64_t len = 2LL * h->B;
    CK(cudaMemsetAsync(h->d_hist_flat, 0, len * sizeof(long long), h->st));
    std::vector<ytk::RowItem> items;
    Interval v; v.b = 0; v.e = n; v.buf = rows ? 0 : -1;
    add_row_items(items, v, 0, ytk::hist_rows_per_item(n, h->groups));
    std::vector<long long*> outs(1, h->d_hist_flat);
    CK(h->t_items.upload(items, h->st, h->arena));
    CK(h->t_seg_out.upload(outs, h->st, h->arena));
    if (h->sparse) {
        CK(h->t_node_tot.reserve(2));
        CK(cudaMemsetAsync(h->t_node_tot.d, 0, 2 * sizeof(long long), h->st));
        CK(ytk::launch_hist_csr(sparse_rows(h), h->d_fea_sel, h->d_gq, h->d_rowidx[0], h->d_rowidx[1], h->t_items.d,
                                (int)items.size(), h->d_feat_zero, h->d_off_flat, 0, h->t_seg_out.d, h->t_node_tot.d,
                                nullptr, nullptr, h->st));
        CK(ytk::launch_csr_default_fix(sparse_rows(h), h->d_fea_sel, h->F, h->d_start, h->d_feat_zero, h->d_off_flat, 0,
                                       h->t_seg_out.d, 1, h->t_node_tot.d, h->st));
    } else
    CK(ytk::launch_hist_build(h->d_bins, h->S8, h->d_gq, h->d_rowidx[0], h->d_rowidx[1], h->t_items.d,
                              (int)items.size(), h->F, h->d_group_mask, h->d_nbins, h->d_feat_zero, h->d_off_flat, 0,
                              h->t_seg_out.d, nullptr, nullptr, h->st));
    if (!h->wide_sel.empty())
        CK(ytk::launch_hist_wide(wide_cols(h), h->d_wide_fids, (int)h->wide_sel.size(), h->d_gq, h->d_rowidx[0],
                                 h->d_rowidx[1], h->t_items.d, (int)items.size(), h->d_feat_zero, h->d_off_flat, 0,
                                 h->t_seg_out.d, nullptr, nullptr, h->st));
    if (h->world > 1)
        NK(ytk::nccl_api().AllReduce(h->d_hist_flat, h->d_hist_flat, (size_t)len, ncclInt64, ncclSum, h->comm, h->st));
    CK(ytk::launch_hist_to_double(h->d_hist_flat, len, h->d_scale, h->d_hist_dbl, h->st));
    CK(cudaMemcpyAsync(out, h->d_hist_dbl, len * sizeof(double), cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    return YTK_OK;
}

static int hook_scale(ytk_gbdt* h, const double* a, const double* b, int64_t len) {
    // power-of-two scales that keep every value below 2^52 (less for very wide features) so that the
    // prefix sums over a feature's bins stay exact in int64
    double mg = 0, mh = 0;
    for (int64_t i = 0; i < len; i++) {
        double v = std::fabs(a[i]);
        if (b) v = std::max(v, std::fabs(b[i]));
        if (!(v < std::numeric_limits<double>::infinity())) return fail(YTK_ERR_ARG, "histogram holds a non-finite value");
        if (i & 1) mh = std::max(mh, v); else mg = std::max(mg, v);
    }
    int max_nb = 1, lg = 0;
    for (int nb : h->nbins) max_nb = std::max(max_nb, nb);
    while ((1 << lg) < max_nb) lg++;
    const int top = std::min(52, 62 - lg);       // (wide features: up to 2^16 bins per prefix sum)
    auto k_of = [top](double m) { int e = 0; if (m > 0) std::frexp(m, &e); return top - e; };
    ytk::GradScale s;
    int kg = k_of(mg), kh = k_of(mh);
    s.mul_g = std::ldexp(1.0, kg); s.inv_g = std::ldexp(1.0, -kg);
    s.mul_h = std::ldexp(1.0, kh); s.inv_h = std::ldexp(1.0, -kh);
    CK(cudaMemcpyAsync(h->d_scale_hook, &s, sizeof(s), cudaMemcpyHostToDevice, h->st));
    CK(cudaStreamSynchronize(h->st));
    return YTK_OK;
}

int ytk_gbdt_subtract_hist(ytk_gbdt* h, const double* parent, const double* small, double* out) {
    if (!h || !parent || !small || !out) return fail(YTK_ERR_ARG, "subtract_hist: null argument");
    if (!h->rows_set) return fail(YTK_ERR_ARG, "subtract_hist: set_bins first");
    CK(cudaSetDevice(h->device));
    const int64_t len = 2LL * h->B;
    RET(hook_scale(h, parent, small, len));
    CK(cudaMemcpyAsync(h->d_hist_dbl, parent, len * sizeof(double), cudaMemcpyHostToDevice, h->st));
    CK(ytk::launch_hist_from_double(h->d_hist_dbl, len, h->d_scale_hook, h->d_hist_flat, h->st));
    CK(cudaMemcpyAsync(h->d_hist_dbl, small, len * sizeof(double), cudaMemcpyHostToDevice, h->st));
    CK(ytk::launch_hist_from_double(h->d_hist_dbl, len, h->d_scale_hook, h->d_hist_flat2, h->st));
    std::vector<ytk::SubItem> subs(1, {h->d_hist_flat, h->d_hist_flat2, h->d_hist_flat});
    CK(h->t_sub.upload(subs, h->st, h->arena));
    CK(ytk::launch_hist_subtract(h->t_sub.d, 1, len, h->st));
    CK(ytk::launch_hist_to_double(h->d_hist_flat, len, h->d_scale_hook, h->d_hist_dbl, h->st));
    CK(cudaMemcpyAsync(out, h->d_hist_dbl, len * sizeof(double), cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    return YTK_OK;
}

int ytk_gbdt_find_split(ytk_gbdt* h, const double* hist, int64_t sample_cnt, const int32_t* fea_index,
                        int32_t n_fea, ytk_split* out) {
    if (!h || !hist || !out) return fail(YTK_ERR_ARG, "find_split: null argument");
    if (!h->rows_set) return fail(YTK_ERR_ARG, "find_split: set_bins first");
    CK(cudaSetDevice(h->device));
    int first_owned;
    RET(set_features(h, fea_index, n_fea, &first_owned, nullptr));
    const int64_t len = 2LL * h->B;
    RET(hook_scale(h, hist, nullptr, len));
    CK(cudaMemcpyAsync(h->d_hist_dbl, hist, len * sizeof(double), cudaMemcpyHostToDevice, h->st));
    CK(ytk::launch_hist_from_double(h->d_hist_dbl, len, h->d_scale_hook, h->d_hist_flat, h->st));
    // single-worker view: all features scanned here, totals from the first selected one
    int first = -1;
    {
        std::vector<uint8_t> sel(h->F, fea_index ? 0 : 1);
        if (fea_index) for (int i = 0; i < n_fea; i++) sel[fea_index[i]] = 1;
        for (int f = 0; f < h->F; f++) if (sel[f]) { first = f; break; }
    }
    std::vector<ytk::ScanNode> nodes(1, {h->d_hist_flat, (long long)sample_cnt, -1, 0});
    CK(h->t_scan.upload(nodes, h->st, h->arena));
    CK(h->t_split_local.reserve(1));
    CK(h->h_split.reserve(1));
    CK(h->t_parts.reserve((size_t)ytk::split_scan_slices(0, h->F)));
    CK(ytk::launch_split_scan(h->t_scan.d, 1, h->d_start, h->d_off_flat, h->d_fea_sel, 0, h->F, first,
                              h->d_scale_hook, gain_params(h->p), nullptr, h->t_parts.d, h->t_split_local.d, nullptr, h->st));
    CK(cudaMemcpyAsync(h->h_split.h, h->t_split_local.d, sizeof(ytk_split), cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    *out = h->h_split.h[0];
    return YTK_OK;
}

int ytk_gbdt_partition_rows(ytk_gbdt* h, const int32_t* rows, int64_t n_rows, int32_t fid, float cond,
                            int32_t* out_rows, int64_t* left_cnt) {
    if (!h || !left_cnt) return fail(YTK_ERR_ARG, "partition_rows: null argument");
    if (!h->rows_set) return fail(YTK_ERR_ARG, "partition_rows: set_bins/upload_rows first");
    if (n_rows < 0 || n_rows > h->N || (n_rows > 0 && (!rows || !out_rows)))
        return fail(YTK_ERR_ARG, "partition_rows: bad row list");
    if (fid < 0 || fid >= h->F) return fail(YTK_ERR_ARG, "partition_rows: fid %d", fid);
    for (int64_t i = 0; i < n_rows; i++)
        if (rows[i] < 0 || rows[i] >= h->N) return fail(YTK_ERR_ARG, "partition_rows: row %d out of range", rows[i]);
    CK(cudaSetDevice(h->device));
    if (n_rows > 0) CK(cudaMemcpyAsync(h->d_rowidx[0], rows, n_rows * sizeof(int32_t), cudaMemcpyHostToDevice, h->st));
    std::vector<ytk::SplitSeg> sg(1);
    sg[0] = {0, (long long)n_rows, 0, fid, cond, 0, 0, h->wide_index[fid] + 1};
    std::vector<ytk::PartitionCounts> cl, cg;
    const int world = h->world;
    h->world = 1;                       // a local operation: no collective in this hook
    int rc = partition_nodes(h, h->d_bins_col, h->N, sg, cl, cg);
    h->world = world;
    RET(rc);
    if (n_rows > 0) CK(cudaMemcpyAsync(out_rows, h->d_rowidx[1], n_rows * sizeof(int32_t), cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    *left_cnt = cl[0].left;
    return YTK_OK;
}

}  // extern "C"
