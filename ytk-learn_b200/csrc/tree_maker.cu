// Host side of libytkgbdt.so: the handle behind the reference's ITreeMaker seam -- creation, bin layout,
// row and gradient uploads -- and the helpers that turn batches of nodes into table-driven launches
// (histogram plan/enqueue, split search, partition, peer-memory set-up).  The growth loop itself
// (DataParallelTreeMaker.make) is in grow.cu.  J/ = /root/reference/src/main/java/com/fenbi/ytklearn/.
#include "host.cuh"

namespace ytkh {


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


}  // namespace ytkh
namespace ytk { void set_last_error(const char* msg) { g_err = msg ? msg : ""; } }

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
void ytk_host_node_value(const ytk_gbdt_params* p, double G, double H, float* root_gain, float* node_value) {
    *root_gain = (float)host_gain(p, G, H);
    *node_value = (float)host_node_value(p, G, H);
}


namespace ytkh {



ytk::GainParams gain_params(const ytk_gbdt_params& p) {
    ytk::GainParams g;
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
    F(h->my_xchg); F(h->my_hist); F(h->d_xctr);
    if (h->d_peer_err) cudaFreeHost(h->d_peer_err);
    h->t_items.release();
    h->t_scan.release(); h->t_sub.release(); h->t_pos.release(); h->t_seg_out.release(); h->t_counts.release();
    h->t_split_local.release(); h->t_split_all.release(); h->t_parts.release(); h->t_leaf.release(); h->t_hist_local.release();
    h->t_rs_out.release(); h->t_node_tot.release(); h->t_scan_done.release();
    for (int i = 0; i < 2; i++) { h->t_spec_segs[i].release(); h->t_spec_counts[i].release(); h->h_spec_counts[i].release(); }
    h->t_spec_tiles.release();
    F(h->d_spec_ntiles);
    if (h->ev_level) cudaEventDestroy(h->ev_level);
    h->h_counts.release(); h->h_split.release(); h->h_leaf.release(); h->h_scale.release(); h->arena.release();
    for (auto* it : h->pool.all) { if (it->own_buf) F(it->buf); delete it; }
    F(h->d_pool_chunk);
    engine_drop_graphs(h);
    F(h->eng.dev);
    if (h->eng.host) cudaFreeHost(h->eng.host);
    for (auto e : h->ev_stage) cudaEventDestroy(e);
    for (auto e : h->ev) cudaEventDestroy(e);
    for (auto e : h->evc) cudaEventDestroy(e);
    if (h->comm) ytk::nccl_api().CommDestroy(h->comm);
    if (h->st && (!h->clique || h->rank == 0)) cudaStreamDestroy(h->st);      // a clique shares rank 0's stream
    if (h->clique && h->rank == 0) delete[] h->clique_members;
}

ytk::SparseRows sparse_rows(const ytk_gbdt* h) {
    return h->sparse ? ytk::SparseRows{h->d_row_ptr, h->d_nz, h->d_default_slot} : ytk::SparseRows{nullptr, nullptr, nullptr};
}

// the uint16 columns, seen from row `row0` on
ytk::WideCols wide_cols(const ytk_gbdt* h, int64_t row0) {
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
    // the same selection as the last tree's (no feature sampling: every tree): the device copies stand
    if (!h->sel_on_device.empty() && h->sel_on_device == sel) {
        if (n_sel) *n_sel = h->sel_count;
        *first_owned = h->sel_first_owned;
        return YTK_OK;
    }
    h->sel_on_device.clear();
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
    h->sel_on_device = sel;
    h->sel_first_owned = *first_owned;
    h->sel_count = cnt;
    return YTK_OK;
}

int quantize_grads(ytk_gbdt* h, int gid) {
    CK(ytk::launch_grad_absmax(h->d_gp, h->N, h->p.class_num, gid, h->d_absmax, h->st));
    // The scales are powers of two chosen from max|g| and max|h|: of ALL rows, or the integer sums of
    // different ranks would be in different units (float bits of non-negative values order like integers)
    if (h->world > 1 && h->comm)
        NK(ytk::nccl_api().AllReduce(h->d_absmax, h->d_absmax, 2, ncclUint32, ncclMax, h->comm, h->st));
    CK(ytk::launch_grad_quantize(h->d_gp, h->N, h->p.class_num, gid, h->d_gq, h->d_absmax, h->d_scale, h->fixed_bits, h->st));
    h->launches += 2;
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

// Maps every rank's exchange and histogram staging buffers into this process (CUDA IPC) so that the
// per-level collectives can run as plain kernels over NVLink.  All ranks must be processes on one node;
// if any rank cannot set this up, all of them stay on NCCL (the decision itself is an NCCL allreduce).
size_t peer_layout(ytk_gbdt* h, size_t* hist_words) {
    const int cap = 4096;
    h->peer_cap_nodes = cap;
    h->slot_counts = (size_t)cap * sizeof(ytk::PartitionCounts);
    h->slot_recs = (size_t)2 * cap * sizeof(ytk_split);
    h->slot_leaf = (size_t)2 * cap * sizeof(long long);
    h->slot_lh = (size_t)2 * cap * 2 * sizeof(long long);
    h->slot_max = 64;
    h->off_counts = ytk::kPeerFlagBytes;
    h->off_recs = h->off_counts + h->slot_counts * h->world;
    h->off_leaf = h->off_recs + h->slot_recs * h->world;
    h->off_lh = h->off_leaf + h->slot_leaf * h->world;
    h->off_max = h->off_lh + h->slot_lh * h->world;
    const size_t local_len = (size_t)h->seg_max * h->world;
    size_t words = std::max<size_t>(local_len * 256, (size_t(128) << 20) / 8);      // >= 256 nodes, >= 128 MB
    words = std::min<size_t>(words, std::max<size_t>(local_len * 2, (size_t(1) << 30) / 8));
    *hist_words = words;
    return h->off_max + h->slot_max * h->world;
}

// A clique member's own buffers; the members are linked (PeerTable) at the first ytk_gbdt_make_tree_clique.
static int clique_buffers(ytk_gbdt* h) {
    size_t hist_words = 0;
    const size_t xchg_bytes = peer_layout(h, &hist_words);
    CK(cudaMalloc(&h->my_xchg, xchg_bytes));
    CK(cudaMemsetAsync(h->my_xchg, 0, xchg_bytes, h->st));
    CK(cudaMalloc(&h->my_hist, hist_words * 8));
    CK(cudaHostAlloc((void**)&h->d_peer_err, sizeof(int), cudaHostAllocMapped));
    *h->d_peer_err = 0;
    h->hist_cap_words = hist_words;
    h->pt.rank = h->rank; h->pt.world = h->world; h->pt.timeout_cycles = 0;
    return YTK_OK;
}

int peer_setup(ytk_gbdt* h) {
    h->peer_mode = false;
    if (h->world < 2 || h->world > 16) return YTK_OK;
    if (const char* e = getenv("YTK_GBDT_NO_PEER")) if (atoi(e) != 0) return YTK_OK;
    if (const char* e = getenv("YTK_GBDT_PEER_TIMEOUT_MS")) h->peer_timeout_ms = std::max(1.0, atof(e));
    size_t hist_words = 0;
    const size_t xchg_bytes = peer_layout(h, &hist_words);
    // Every rank goes through BOTH agreements below whatever fails locally: a rank that returned early
    // would leave the others hanging in the collective.  Local failures only clear `ok`.
    int ok = 1;
    cudaIpcMemHandle_t mine[2];
    memset(mine, 0, sizeof(mine));
    if (cudaMalloc(&h->my_xchg, xchg_bytes) != cudaSuccess || cudaMalloc(&h->my_hist, hist_words * 8) != cudaSuccess ||
        cudaHostAlloc((void**)&h->d_peer_err, sizeof(int), cudaHostAllocMapped) != cudaSuccess)
        ok = 0;
    if (ok) {
        *h->d_peer_err = 0;
        if (cudaMemset(h->my_xchg, 0, xchg_bytes) != cudaSuccess ||
            cudaIpcGetMemHandle(&mine[0], h->my_xchg) != cudaSuccess || cudaIpcGetMemHandle(&mine[1], h->my_hist) != cudaSuccess)
            ok = 0;
    }
    cudaGetLastError();
    // exchange the handles (and learn whether everybody got this far) over NCCL, once
    const size_t rec = sizeof(mine) + 8;
    char* d_send = nullptr; char* d_recv = nullptr;
    std::vector<char> send(rec, 0), recv(rec * h->world, 0);
    if (ok) memcpy(send.data(), mine, sizeof(mine));
    long long okv = ok;
    memcpy(send.data() + sizeof(mine), &okv, 8);
    // (these two small allocations are needed for the agreement itself: without them nothing can be agreed)
    CK(cudaMalloc(&d_send, rec));
    CK(cudaMalloc(&d_recv, rec * h->world));
    cudaError_t ce = cudaMemcpy(d_send, send.data(), rec, cudaMemcpyHostToDevice);
    ncclResult_t nr = ytk::nccl_api().AllGather(d_send, d_recv, rec, ncclChar, h->comm, h->st);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(h->st);
    if (ce == cudaSuccess) ce = cudaMemcpy(recv.data(), d_recv, rec * h->world, cudaMemcpyDeviceToHost);
    if (ce != cudaSuccess || nr != ncclSuccess) ok = 0;
    for (int r = 0; r < h->world; r++) { long long v; memcpy(&v, recv.data() + r * rec + sizeof(mine), 8); if (!v) ok = 0; }
    h->pt.rank = h->rank; h->pt.world = h->world;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, h->device);
    h->pt.timeout_cycles = (long long)(h->peer_timeout_ms * (khz > 0 ? khz : 2000000));
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
    ce = cudaMemcpy(d_send, &okv, 8, cudaMemcpyHostToDevice);
    nr = ytk::nccl_api().AllReduce(d_send, d_send, 1, ncclInt64, ncclMin, h->comm, h->st);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(h->st);
    if (ce == cudaSuccess) ce = cudaMemcpy(&okv, d_send, 8, cudaMemcpyDeviceToHost);
    cudaFree(d_send); cudaFree(d_recv);
    if (nr != ncclSuccess) return fail(YTK_ERR_NCCL, "peer set-up: %s", ytk::nccl_api().GetErrorString(nr));
    CK(ce);
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
int peer_check(ytk_gbdt* h, const char* where) {
    if (!h->peer_mode || !h->d_peer_err || !*reinterpret_cast<volatile int*>(h->d_peer_err)) return YTK_OK;
    *reinterpret_cast<volatile int*>(h->d_peer_err) = 0;
    return fail(YTK_ERR_NCCL, "%s: a peer GPU did not reach a collective within %.0f ms; this tree is abandoned", where,
                h->peer_timeout_ms);
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

// ---- histogram build: plan (host tables) / enqueue (launches) ----------------------------------

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
int scan_done_counters(ytk_gbdt* h, int m) {
    const size_t before = h->t_scan_done.cap;
    CK(h->t_scan_done.reserve((size_t)std::max(m, 256)));
    if (h->t_scan_done.cap != before)       // a new array: the kernels leave the counters at zero afterwards
        CK(cudaMemsetAsync(h->t_scan_done.d, 0, h->t_scan_done.cap * sizeof(unsigned int), h->st));
    return YTK_OK;
}

int search_enqueue(ytk_gbdt* h, const ytk::ScanNode* d_nodes, int m, int first_owned,
                   const ytk::PartitionCounts* cnt_global) {
    if (m == 0) return YTK_OK;
    CK(h->t_split_local.reserve(m));
    CK(h->t_parts.reserve((size_t)m * ytk::split_scan_slices(h->f_from, h->f_to)));
    CK(h->h_split.reserve((size_t)m * h->world));
    // the records stay in device memory (the planner of the next partition and the exchange read them
    // there); on one GPU the scan also stores them in pinned host memory, so the host needs no copy
    RET(scan_done_counters(h, m));
    CK(ytk::launch_split_scan(d_nodes, m, h->d_start, h->d_off_inner, h->d_fea_sel, h->f_from, h->f_to,
                              first_owned, h->d_scale, gain_params(h->p), cnt_global, h->t_parts.d, h->t_split_local.d,
                              h->world == 1 ? h->h_split.h : nullptr, h->t_scan_done.d, h->st));
    h->launches++;
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
// The partition tables have their own device buffer: they go up first so that the partition is
// already running while the host builds the tables of the kernels behind it.
int partition_commit(ytk_gbdt* h, const Blob& bl, PartPlan& pl) {
    RET(blob_commit_to(h, bl, &h->d_blob_part, &h->blob_part_cap));
    pl.d_segs = reinterpret_cast<const ytk::SplitSeg*>(h->d_blob_part + pl.o_segs);
    pl.d_tiles = reinterpret_cast<const int*>(h->d_blob_part + pl.o_tiles);
    return YTK_OK;
}
void partition_plan(const std::vector<ytk::SplitSeg>& segs_in, PartPlan& pl, Blob& bl, bool single_pass) {
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
    if (++h->part_epoch >= (1ull << 29)) {     // (epochs from 2^29 on belong to the device-planned level loop)       // entries of other epochs read as "not written yet"
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
    RET(peer_check(h, "partition"));
    h->arena.rewind();
    partition_collect(h, (int)segs_in.size(), local, global);
    return YTK_OK;
}


}  // namespace ytkh

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
    if (const char* s = getenv("YTK_GBDT_SPIN")) h->spin_wait = atoi(s) != 0;
    if (const char* s = getenv("YTK_GBDT_NO_ENGINE")) h->use_engine = atoi(s) == 0;
    if (const char* s = getenv("YTK_GBDT_NO_GRAPH")) h->use_graph = atoi(s) == 0;
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

int ytk_gbdt_create_clique(const ytk_gbdt_params* params, int32_t device, int32_t world, ytk_gbdt** out) {
    if (!params || !out) return fail(YTK_ERR_ARG, "create_clique: null argument");
    if (world < 2 || world > 16) return fail(YTK_ERR_ARG, "create_clique: world %d not in [2,16]", world);
    ytk_gbdt** members = new ytk_gbdt*[world];
    for (int r = 0; r < world; r++) members[r] = nullptr;
    for (int r = 0; r < world; r++) {
        ytk_gbdt* h = nullptr;
        int rc = ytk_gbdt_create(params, device, 0, 1, nullptr, &h);      // (as a world of one: no NCCL)
        if (rc != YTK_OK) {
            for (int q = 0; q < r; q++) { members[q]->clique = false; ytk_gbdt_destroy(members[q]); }
            delete[] members;
            return rc;
        }
        h->rank = r; h->world = world; h->clique = true;
        if (r > 0) { cudaStreamDestroy(h->st); h->st = members[0]->st; }
        h->clique_members = members;
        members[r] = h;
    }
    for (int r = 0; r < world; r++) out[r] = members[r];
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
    CK(ytk::hist_kernels_init());
    CK(ytk::partition_kernels_init());
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
    h->sel_on_device.clear();
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
        n_pre = std::min(n_pre, std::max<size_t>(2, (size_t(8) << 30) / slot_bytes));
        if (h->p.histogram_pool_capacity > 0) n_pre = (size_t)capacity;      // a bounded pool never holds more
        CK(cudaMalloc(&h->d_pool_chunk, n_pre * slot_bytes));
        h->pool_chunk_slots = n_pre;
        for (size_t i = 0; i < n_pre; i++) {
            HistPool::Item* it = new HistPool::Item{-1, h->d_pool_chunk + i * (size_t)h->seg_max, false};
            h->pool.all.push_back(it);
            h->pool.avail.push_back(it);
        }
    }
    if (h->clique) RET(clique_buffers(h));
    else if (h->world > 1) RET(peer_setup(h));
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

int ytk_gbdt_upload_rows_u8(ytk_gbdt* h, const uint8_t* x_chunk, int64_t row0, int64_t nrows, int32_t stride) {
    RET(check_rows(h, x_chunk, row0, nrows, "upload_rows_u8"));
    if (h->sparse) return fail(YTK_ERR_ARG, "upload_rows_u8: this handle holds sparse rows (ytk_gbdt_upload_csr)");
    if (stride < h->F) return fail(YTK_ERR_ARG, "upload_rows_u8: stride %d < feature dim %d", stride, h->F);
    if (h->n_wide > 0) return fail(YTK_ERR_ARG, "upload_rows_u8: a feature has more than 256 bins; its slots need ytk_gbdt_upload_rows");
    if (nrows == 0) return YTK_OK;
    CK(cudaSetDevice(h->device));
    const int64_t step = std::max<int64_t>(1, (64LL << 20) / (int64_t)stride);
    for (int64_t r = 0; r < nrows; r += step) {
        int64_t n = std::min(step, nrows - r);
        size_t bytes = (size_t)n * stride;
        RET(ensure_stage(h, bytes));
        CK(cudaMemcpyAsync(h->d_stage, x_chunk + r * stride, bytes, cudaMemcpyHostToDevice, h->st));
        CK(ytk::launch_repack_rows_u8((const uint8_t*)h->d_stage, n, stride, h->F, h->d_nbins, h->d_bins + (row0 + r) * h->S8,
                                      h->S8, h->d_bins_col + (row0 + r), h->N, wide_cols(h, row0 + r), row0 + r, h->d_bad,
                                      h->st));
        CK(cudaStreamSynchronize(h->st));
    }
    int bad = 0;
    CK(cudaMemcpy(&bad, h->d_bad, sizeof(int), cudaMemcpyDeviceToHost));
    if (bad) {
        CK(cudaMemset(h->d_bad, 0, sizeof(int)));
        return fail(YTK_ERR_ARG, "upload_rows_u8: slot index outside [0, bins of its feature)");
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

static int upload_csr_impl(ytk_gbdt* h, const int64_t* row_ptr, const int32_t* fids, const int32_t* slots, int64_t row0,
                           int64_t nrows, const float* cuts) {
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
        float* d_cuts = nullptr;
        if (cuts) {
            CK(cudaMalloc(&d_cuts, (size_t)h->B * sizeof(float)));
            cudaError_t e = cudaMemcpyAsync(d_cuts, cuts, (size_t)h->B * sizeof(float), cudaMemcpyHostToDevice, h->st);
            if (e != cudaSuccess) { cudaFree(d_cuts); CK(e); }
        }
        cudaError_t e = ytk::launch_csr_pack(h->d_row_ptr + row0, d_f, d_s, nrows, h->F, h->d_nbins, d_cuts, h->d_start,
                                             h->d_nz + h->csr_nnz_done, h->d_bad, h->st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->st);
        if (d_cuts) cudaFree(d_cuts);
        CK(e);
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

int ytk_gbdt_upload_csr(ytk_gbdt* h, const int64_t* row_ptr, const int32_t* fids, const int32_t* slots, int64_t row0,
                        int64_t nrows) {
    return upload_csr_impl(h, row_ptr, fids, slots, row0, nrows, nullptr);
}

int ytk_gbdt_quantize_csr(ytk_gbdt* h, const int64_t* row_ptr, const int32_t* fids, const int32_t* value_bits, int64_t row0,
                          int64_t nrows, const float* cuts) {
    if (!cuts) return fail(YTK_ERR_ARG, "quantize_csr: null cuts");
    return upload_csr_impl(h, row_ptr, fids, value_bits, row0, nrows, cuts);
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
