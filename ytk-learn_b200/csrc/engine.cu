// Level-wise growth with the host out of the loop: DataParallelTreeMaker.make
// (J/optimizer/gbdt/DataParallelTreeMaker.java:229-295) for grow policy "level" as ONE enqueue and ONE
// host wait per tree.  Every table a level's kernels need is planned on the device from the split records
// and child counts it already holds (level.cu); the host queues root + max_depth levels of kernels with
// grids that are upper bounds, waits once, reads the records back and assembles the Tree in the
// reference's node order (Tree.addChilds order = polling order, :246-258).  The host-planned path
// (grow.cu) remains for everything this one does not take: loss-wise growth, a bounded histogram pool,
// the row-count stop rule, a leaf budget that can bind, sparse rows, features with more than 256 bins.
// J/ = /root/reference/src/main/java/com/fenbi/ytklearn/.
#include "host.cuh"
#include "level.cuh"

namespace ytkh {

namespace {

size_t align256(size_t v) { return (v + 255) & ~size_t(255); }

// kernels of a tree that report their own time span (device_util.cuh): stamp slot = kind * kStampLevels + level
constexpr int kStampLevels = 20;
enum StampKind { kStHist = 0, kStPart = 1, kStScan = 2, kStComm = 3, kStSync = 4, kStKinds = 5 };
constexpr int kMaxStamps = kStKinds * kStampLevels;

// buffer the rows of level d live in: the root's (-1 identity / 0 after a row mask), then alternating
int level_buf(int root_buf, int d) {
    int b = root_buf;
    for (int i = 0; i < d; i++) b = (b == 0) ? 1 : 0;
    return b;
}

}  // namespace

bool engine_applicable(const ytk_gbdt* h) {
    const ytk_gbdt_params& p = h->p;
    if (!h->use_engine || h->force_sequential || !h->single_pass_partition || !h->plan_ahead) return false;
    if (p.grow_policy != 0 || p.max_depth < 1 || p.max_depth > 14) return false;
    if (p.min_split_samples > 0 || p.histogram_pool_capacity > 0) return false;
    if (p.max_leaf_cnt > 0 && (long long)p.max_leaf_cnt < (1LL << p.max_depth)) return false;      // the leaf budget could bind
    if (h->sparse || h->n_wide > 0) return false;
    if ((size_t)(1u << (p.max_depth - 1)) > h->pool_chunk_slots) return false;
    if (h->world > 1) {
        // exchanges over peer memory (one process per GPU), or a clique of handles on one device
        if (!h->peer_mode && !h->clique) return false;
        if (p.max_depth > 12) return false;
        const size_t slots = size_t(1) << (p.max_depth - 1);
        if ((int)slots > h->peer_cap_nodes / 2) return false;
        if (slots * (size_t)h->seg_max * h->world > h->hist_cap_words) return false;
    }
    return true;
}

void engine_drop_graphs(ytk_gbdt* h) {
    for (auto& g : h->eng.graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
    h->eng.graphs.clear();
}

// Device tables and their pinned mirrors for trees of depth D (kept across trees).
static int engine_reserve(ytk_gbdt* h, int D) {
    EngineBuffers& e = h->eng;
    const size_t heap = (size_t(1) << (D + 1)) - 1, slots = size_t(1) << (D - 1);
    const size_t tiles = (size_t)(h->N / ytk::kFusedTile) + (size_t(1) << D) + 2;
    if (e.depth == D && e.tile_cap >= tiles) return YTK_OK;
    engine_drop_graphs(h);                               // they hold the old tables' addresses
    if (e.dev) CK(cudaFree(e.dev));
    if (e.host) CK(cudaFreeHost(e.host));
    e = EngineBuffers();
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align256(off + bytes); return o; };
    const size_t o_nodes = take(heap * sizeof(ytk::LvNode)), o_recs = take(heap * sizeof(ytk_split)),
                 o_lh = take(heap * 2 * sizeof(long long)), o_leaf = take(heap * 2 * sizeof(long long)),
                 o_stamps = take(2 * kMaxStamps * sizeof(unsigned long long));
    const size_t mirror_bytes = off;                     // everything up to here is copied back after the tree
    const size_t o_segs = take(slots * sizeof(ytk::SplitSeg)), o_tiles = take(tiles * sizeof(int)), o_ntiles = take(sizeof(int)),
                 o_counts = take(slots * sizeof(ytk::PartitionCounts)), o_gcounts = take(slots * sizeof(ytk::PartitionCounts)),
                 o_build = take(2 * slots * sizeof(ytk::BuildNode)), o_nbuild = take(sizeof(int)),
                 o_scans = take(2 * slots * sizeof(ytk::ScanNode)),
                 o_segout = take(2 * slots * sizeof(long long*)), o_segfinal = take(2 * slots * sizeof(long long*)),
                 o_recs_local = take(2 * slots * sizeof(ytk_split)), o_lh_local = take(4 * slots * sizeof(long long)),
                 o_epoch = take(sizeof(unsigned long long)), o_gid = take(sizeof(int)), o_tot = take(heap * 2 * sizeof(long long));
    CK(cudaMalloc(&e.dev, off));
    CK(cudaMemsetAsync(e.dev, 0, off, h->st));
    CK(cudaHostAlloc((void**)&e.host, mirror_bytes, cudaHostAllocDefault));
    e.depth = D; e.tile_cap = tiles; e.mirror_bytes = mirror_bytes; e.heap = heap;
    ytk::LvTables& t = e.t;
    t.nodes = reinterpret_cast<ytk::LvNode*>(e.dev + o_nodes);
    t.recs = reinterpret_cast<ytk_split*>(e.dev + o_recs);
    t.left_h = reinterpret_cast<long long*>(e.dev + o_lh);
    t.leaf = reinterpret_cast<long long*>(e.dev + o_leaf);
    t.segs = reinterpret_cast<ytk::SplitSeg*>(e.dev + o_segs);
    t.tile_seg = reinterpret_cast<int*>(e.dev + o_tiles);
    t.n_tiles = reinterpret_cast<int*>(e.dev + o_ntiles);
    t.counts = reinterpret_cast<ytk::PartitionCounts*>(e.dev + o_counts);
    t.gcounts = reinterpret_cast<ytk::PartitionCounts*>(e.dev + o_gcounts);
    t.build = reinterpret_cast<ytk::BuildNode*>(e.dev + o_build);
    t.n_build = reinterpret_cast<int*>(e.dev + o_nbuild);
    t.tot = reinterpret_cast<long long*>(e.dev + o_tot);
    t.scans = reinterpret_cast<ytk::ScanNode*>(e.dev + o_scans);
    t.seg_out = reinterpret_cast<long long**>(e.dev + o_segout);
    t.seg_final = reinterpret_cast<long long**>(e.dev + o_segfinal);
    t.recs_local = reinterpret_cast<ytk_split*>(e.dev + o_recs_local);
    t.lh_local = reinterpret_cast<long long*>(e.dev + o_lh_local);
    t.epoch = reinterpret_cast<unsigned long long*>(e.dev + o_epoch);
    t.gid = reinterpret_cast<int*>(e.dev + o_gid);
    // the exchange epochs outlive these tables: the peers' flags keep counting
    if (!h->d_xctr) {
        CK(cudaMalloc(&h->d_xctr, 8 * 2 * sizeof(unsigned long long)));
        CK(cudaMemsetAsync(h->d_xctr, 0, 8 * 2 * sizeof(unsigned long long), h->st));
    }
    e.h_nodes = reinterpret_cast<ytk::LvNode*>(e.host + o_nodes);
    e.h_recs = reinterpret_cast<ytk_split*>(e.host + o_recs);
    e.h_leaf = reinterpret_cast<long long*>(e.host + o_leaf);
    e.d_stamps = reinterpret_cast<unsigned long long*>(e.dev + o_stamps);
    e.h_stamps = reinterpret_cast<unsigned long long*>(e.host + o_stamps);
    e.clear_bytes = o_segs;                              // nodes, records, left hessians, leaf sums: cleared per tree
    return YTK_OK;
}

namespace {

class LevelEngine {
public:
    LevelEngine(ytk_gbdt* handle, ytk_tree_out* tree, ytk_time_stats* stats)
        : h(handle), p(handle->p), out(tree), T(stats), D(handle->p.max_depth) {}

    // One rank on its own (one GPU, or one process per GPU with exchanges over peer memory).
    int run(int gid, const int32_t* fea_index, int32_t n_fea, const uint8_t* row_mask) {
        const double t0 = now_ms();
        RET(prepare(gid, fea_index, n_fea));
        // On one GPU the launch sequence of a tree is the same every time (grids are upper bounds, tables and the
        // class index live on the device): it is recorded once as a CUDA graph and replayed, which takes the
        // per-launch gaps out of the ~60 short kernels of a tree.  (Not with a row mask: its table upload is per tree.)
        bool launched = false;
        if (h->use_graph && (h->world == 1 || h->peer_mode) && !row_mask) {
            const unsigned long long sig = signature();
            EngineGraph* g = nullptr;
            for (auto& c : h->eng.graphs) if (c.sig == sig) g = &c;
            if (!g) RET(record_graph(sig, &g));
            if (g) {
                CK(cudaGraphLaunch(g->exec, h->st));
                marks = g->marks;
                h->ev_stage_used = g->ev_stage_used; h->ev_used = g->ev_used; h->launches += g->launches;
                T->rendezvous = g->rendezvous;
                root_buf = -1;
                launched = true;
                replayed = true;
                T->graph_replayed = 1;
            }
        }
        if (!launched) RET(enqueue_tree(row_mask));
        const double t1 = now_ms();
        CK(cudaStreamSynchronize(h->st));                 // the one host wait of the tree
        RET(peer_check(h, "make_tree"));
        const double t2 = now_ms();
        RET(after_wait(row_mask != nullptr, t0, t1, t2));
        return YTK_OK;
    }

    // everything a tree queues, from the gradient scales to the copy of the records back to the host
    int enqueue_tree(const uint8_t* row_mask) {
        const int x = h->world > 1 ? ytk::kXAll : 0;
        RET(begin_scale());
        if (x) RET(x_max(x));
        RET(begin_root(row_mask));
        RET(root_hist());
        if (x) RET(x_hist(-1, x));
        RET(root_scan());
        if (x) RET(x_recs(-1, x));
        for (int d = 0; d < D; d++) {
            RET(level_partition(d));
            if (d == D - 1) {
                RET(level_leaf(d));
                if (x) RET(x_leaf(d, x));
            } else {
                RET(level_hist(d));
                if (x) RET(x_hist(d, x));
                RET(level_scan(d));
                if (x) RET(x_recs(d, x));
            }
        }
        return finish_enqueue();
    }

    // what a recorded launch sequence depends on besides the contents of device memory
    unsigned long long signature() const {
        unsigned long long v = 1469598103934665603ull;
        auto mix = [&](unsigned long long x) { v = (v ^ x) * 1099511628211ull; };
        mix((unsigned long long)(uintptr_t)h->eng.dev); mix((unsigned long long)(uintptr_t)h->eng.host);
        mix((unsigned long long)(uintptr_t)h->d_tile_state); mix((unsigned long long)(uintptr_t)h->d_ticket);
        mix((unsigned long long)(uintptr_t)h->t_parts.d); mix((unsigned long long)(uintptr_t)h->t_scan_done.d);
        mix((unsigned long long)(uintptr_t)h->h_scale.h); mix((unsigned long long)(uintptr_t)h->d_gq);
        mix((unsigned long long)(first_owned + 7)); mix((unsigned long long)D); mix((unsigned long long)h->N);
        mix((unsigned long long)t.flush_cost); mix((unsigned long long)plan_in_xrecs());
        mix((unsigned long long)h->world);
        for (int r = 0; r < h->world && h->world > 1; r++) { mix((unsigned long long)(uintptr_t)h->pt.xchg[r]); mix((unsigned long long)(uintptr_t)h->pt.hist[r]); }
        return v;
    }

    int record_graph(unsigned long long sig, EngineGraph** out) {
        *out = nullptr;
        cudaGraph_t graph = nullptr;
        const int launches_before = h->launches;
        h->ev_used = 0;
        T->rendezvous = 0;
        if (cudaStreamBeginCapture(h->st, cudaStreamCaptureModeRelaxed) != cudaSuccess) { cudaGetLastError(); h->use_graph = false; return YTK_OK; }
        capturing = true;
        const int rc = enqueue_tree(nullptr);
        capturing = false;
        const cudaError_t ce = cudaStreamEndCapture(h->st, &graph);
        cudaGraphExec_t exec = nullptr;
        if (rc != YTK_OK || ce != cudaSuccess || !graph || cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) {
            // recording is an optimisation: without it the same calls are queued directly
            cudaGetLastError();
            if (graph) cudaGraphDestroy(graph);
            h->use_graph = false;
            h->launches = launches_before;
            if (getenv("YTK_GBDT_VERBOSE")) fprintf(stderr, "ytk_gbdt: CUDA graph recording failed (%s); trees are queued launch by launch\n", cudaGetErrorString(ce));
            return YTK_OK;
        }
        cudaGraphDestroy(graph);
        if (h->eng.graphs.size() >= 32) engine_drop_graphs(h);
        EngineGraph g;
        g.sig = sig; g.exec = exec; g.marks = marks; g.ev_stage_used = h->ev_stage_used; g.ev_used = h->ev_used;
        g.launches = h->launches - launches_before;
        g.rendezvous = T->rendezvous;
        h->launches = launches_before;
        h->eng.graphs.push_back(g);
        *out = &h->eng.graphs.back();
        return YTK_OK;
    }

    // ---- the phases, also driven rank by rank for a clique of handles on one device (engine_make_tree_clique) ----
    int after_wait(bool masked, double t0, double t1, double t2) {
        T->batches = 0;
        T->max_batch_ms = t2 - t0;
        h->arena.rewind();
        RET(assemble(masked));
        RET(stage_times());
        T->enqueue_ms = t1 - t0;
        T->host_wait_ms = t2 - t1;
        return YTK_OK;
    }

    ytk_gbdt* h;
    const ytk_gbdt_params& p;
    ytk_tree_out* out;
    ytk_time_stats* T;
    const int D;
    int first_owned = -1, n_sel = 0;
    int root_buf = -1;
    long long root_rows_bound = 0;
    ytk::LvTables t{};
    // CUDA events at the stage boundaries of the tree (TimeStats.getStats, J/data/gbdt/TimeStats.java:69-78)
    enum Stage { kInit = 0, kHist, kComm, kSplit, kSync, kPart, kStages };
    std::vector<int> marks;              // stage that ENDS at event i + 1

    bool capturing = false;            // a recorded sequence carries no CUDA events (they cannot be read back from a replay)
    int mark(Stage s) {
        if (capturing) return YTK_OK;
        if (h->ev_stage_used >= (int)h->ev_stage.size()) {
            cudaEvent_t e;
            CK(cudaEventCreate(&e));
            h->ev_stage.push_back(e);
        }
        CK(cudaEventRecord(h->ev_stage[h->ev_stage_used++], h->st));
        marks.push_back((int)s);
        return YTK_OK;
    }

    // Outside any recording: tables and scratch at their final sizes (nothing is allocated while a tree is queued),
    // the selected features and the class index on the device.
    int prepare(int gid, const int32_t* fea_index, int32_t n_fea) {
        RET(engine_reserve(h, D));
        t = h->eng.t;
        t.pool = h->d_pool_chunk;
        t.staging = h->world > 1 ? reinterpret_cast<long long*>(h->my_hist) : nullptr;
        t.seg_max = h->seg_max;
        t.world = h->world;
        static const long long flush_cost = [] { const char* e = getenv("YTK_GBDT_FLUSH_COST"); return e ? std::max(0, atoi(e)) : (int)ytk::kHistFlushCost; }();
        t.flush_cost = flush_cost;
        RET(set_features(h, fea_index, n_fea, &first_owned, &n_sel));
        h->pool.clear();
        RET(next_partition_epoch(h, h->eng.tile_cap));    // the partition's chained-scan state for the largest level
        t.ticket = h->d_ticket + 1;
        CK(h->t_parts.reserve((size_t)(2 << D) * ytk::split_scan_slices(h->f_from, h->f_to)));
        RET(scan_done_counters(h, 2 << D));
        while ((int)h->ev_stage.size() < 6 * D + 16) { cudaEvent_t e; CK(cudaEventCreate(&e)); h->ev_stage.push_back(e); }
        while ((int)h->ev.size() < 2 * D + 4) { cudaEvent_t e; CK(cudaEventCreate(&e)); h->ev.push_back(e); }
        if (h->eng.gid_on_device != gid) {
            int* stage = static_cast<int*>(h->arena.take(sizeof(int), h->st));
            if (!stage) return fail(YTK_ERR_CUDA, "pinned staging allocation failed");
            *stage = gid;
            CK(cudaMemcpyAsync(t.gid, stage, sizeof(int), cudaMemcpyHostToDevice, h->st));
            h->eng.gid_on_device = gid;
        }
        return YTK_OK;
    }

    int begin_scale() {
        h->ev_stage_used = 0;
        marks.clear();
        RET(mark(kInit));                                 // (event 0: the start; its stage label is unused)
        CK(ytk::launch_grad_absmax(h->d_gp, h->N, p.class_num, 0, h->d_absmax, h->st, t.gid));
        h->launches++;
        return YTK_OK;
    }

    int begin_root(const uint8_t* row_mask) {
        CK(ytk::launch_grad_quantize(h->d_gp, h->N, p.class_num, 0, h->d_gq, h->d_absmax, h->d_scale, h->fixed_bits, h->st, t.gid));
        h->launches++;
        CK(cudaMemcpyAsync(h->h_scale.h, h->d_scale, sizeof(ytk::GradScale), cudaMemcpyDeviceToHost, h->st));
        CK(cudaMemsetAsync(h->eng.dev, 0, h->eng.clear_bytes, h->st));
        root_buf = -1;
        root_rows_bound = h->N;
        const ytk::PartitionCounts* mask_counts = nullptr;
        if (row_mask) {
            // resetPositionAfterSample (SamplePositionData.java:84-87): the sampled rows first, in buffer 0
            if (!h->d_mask) CK(cudaMalloc(&h->d_mask, std::max<int64_t>(h->N, 1)));
            void* stage = h->arena.take((size_t)std::max<int64_t>(h->N, 1), h->st);
            if (!stage) return fail(YTK_ERR_CUDA, "pinned staging allocation failed");
            memcpy(stage, row_mask, (size_t)h->N);
            CK(cudaMemcpyAsync(h->d_mask, stage, h->N, cudaMemcpyHostToDevice, h->st));
            std::vector<ytk::SplitSeg> sg(1);
            sg[0] = {0, (long long)h->N, -1, -1, 0.f, 0, 0, 0};
            const int world = h->world;
            h->world = 1;                                 // a local operation: the count stays on the device, no exchange
            const int rc = partition_launch(h, h->d_mask, 0, sg);
            h->world = world;
            RET(rc);
            mask_counts = h->t_counts.d;
            root_buf = 0;
        }
        CK(ytk::launch_lv_root(t, 0, h->N, root_buf, mask_counts, h->st));
        h->launches++;
        RET(mark(kInit));
        return YTK_OK;
    }

    // ---- exchanges (world > 1) ------------------------------------------------------------------------------
    int x_max(int phase) {
        if (phase != ytk::kXReduce) T->rendezvous++;
        CK(ytk::launch_lv_xmax(h->pt, h->d_absmax, h->off_max, h->slot_max, 4, h->d_xctr + 0, phase, h->d_peer_err, h->st));
        h->launches++;
        return YTK_OK;
    }
    int x_hist(int d, int phase) {
        if (phase != ytk::kXReduce) T->rendezvous++;
        CK(ytk::launch_lv_xhist(h->pt, t, d, d < 0 ? 1 : (1 << d), h->off_counts, h->slot_counts, 5, h->d_xctr + 2, phase,
                                h->d_peer_err, h->st, phase == ytk::kXPush ? nullptr : stamp(kStComm, d + 1)));
        h->launches++;
        if (phase != ytk::kXPush) RET(mark(kComm));
        return YTK_OK;
    }
    int x_recs(int d, int phase) {
        if (phase != ytk::kXReduce) T->rendezvous++;
        const int n = d < 0 ? 1 : (2 << d);
        const long long first = d < 0 ? 0 : (2ll << d) - 1;
        // (the level after these records is d + 1: its partition is planned in the tail of the merge)
        CK(ytk::launch_lv_xrecs(h->pt, t, n, first, h->off_recs, h->slot_recs, h->off_lh, h->slot_lh, 6, h->d_xctr + 4, phase,
                                h->d_peer_err, h->st, phase == ytk::kXPush ? nullptr : stamp(kStSync, d + 1),
                                plan_in_xrecs() ? d + 1 : -1, level_buf(root_buf, d + 1), p.min_split_loss));
        h->launches++;
        if (phase != ytk::kXPush) RET(mark(kSync));
        return YTK_OK;
    }
    int x_leaf(int d, int phase) {
        if (phase != ytk::kXReduce) T->rendezvous++;
        CK(ytk::launch_lv_xleaf(h->pt, t, d, h->off_leaf, h->slot_leaf, h->off_counts, h->slot_counts, 7, h->d_xctr + 6, phase,
                                h->d_peer_err, h->st, phase == ytk::kXPush ? nullptr : stamp(kStComm, kStampLevels - 1)));
        h->launches++;
        if (phase != ytk::kXPush) RET(mark(kComm));
        return YTK_OK;
    }

    unsigned long long* stamp(StampKind k, int level) { return h->eng.d_stamps + 2 * (k * kStampLevels + level); }

    int hist_launch(int m_slots, int level) {
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (!capturing) {
            RET(hist_events(h, &e0, &e1));
            CK(cudaEventRecord(e0, h->st));
        }
        CK(ytk::launch_hist_span(h->d_bins, h->S8, h->d_gq, h->d_rowidx[0], h->d_rowidx[1], t.build, t.n_build, h->groups, h->F,
                                 h->d_group_mask, h->d_nbins, h->d_feat_owner, h->d_off_inner,
                                 h->world == 1 ? 0 : (long long)m_slots * h->seg_max, t.seg_out, h->st, stamp(kStHist, level)));
        if (!capturing) CK(cudaEventRecord(e1, h->st));
        h->launches++;
        return YTK_OK;
    }

    int scan_launch(const ytk::ScanNode* nodes, int n, size_t first_heap, int level) {
        // several GPUs: this rank's records first, the merged ones go to the heap arrays (x_recs)
        CK(ytk::launch_split_scan(nodes, n, h->d_start, h->d_off_inner, h->d_fea_sel, h->f_from, h->f_to, first_owned,
                                  h->d_scale, gain_params(p), nullptr, h->t_parts.d,
                                  h->world == 1 ? t.recs + first_heap : t.recs_local, nullptr, h->t_scan_done.d, h->st,
                                  h->world == 1 ? t.left_h + 2 * first_heap : t.lh_local, stamp(kStScan, level)));
        h->launches++;
        return YTK_OK;
    }

    // histogram slots that receive atomics start at zero: the slots themselves on one GPU, else the staging buffer
    int zero_build_targets(int first_slot, int m_slots) {
        if (h->world == 1) {
            // One GPU: level d builds into slots [2^d, 2^(d+1)) and the root into slot 0 -- all different, so the
            // whole tree's targets are cleared by ONE memset ahead of the root instead of one per level on the
            // tree's critical path (unless that is a lot of memory to touch at once: deep trees clear level by level)
            const size_t all = ((size_t)1 << (D - 1)) * h->seg_max * sizeof(long long);
            if (all <= ((size_t)64 << 20)) {
                if (first_slot == 0) CK(cudaMemsetAsync(t.pool, 0, all, h->st));
                return YTK_OK;
            }
            CK(cudaMemsetAsync(t.pool + (size_t)first_slot * h->seg_max, 0, (size_t)m_slots * h->seg_max * sizeof(long long), h->st));
        } else
            CK(cudaMemsetAsync(t.staging, 0, (size_t)m_slots * h->seg_max * h->world * sizeof(long long), h->st));
        return YTK_OK;
    }

    // findBestSplit(root) (:302-318)
    int root_hist() {
        RET(zero_build_targets(0, 1));
        RET(hist_launch(1, 0));
        RET(mark(kHist));
        return YTK_OK;
    }
    int root_scan() {
        RET(scan_launch(t.scans, 1, 0, 0));
        RET(mark(kSplit));
        return YTK_OK;
    }

    // One level of the loop (:244-281) for all its nodes at once: position.resetPosition (:260) ...
    int level_partition(int d) {
        const int m = 1 << d;
        const int buf = level_buf(root_buf, d), child_buf = buf == 0 ? 1 : 0;
        const bool leaf = d == D - 1;                     // the children are leaves (:266)
        if (!plan_in_xrecs()) CK(ytk::launch_lv_plan_split(t, d, buf, p.min_split_loss, nullptr, ytk::kFusedTile, h->st));
        const long long grid_bound = root_rows_bound / ytk::kFusedTile + m + 1;
        CK(ytk::launch_partition_fused_planned(h->d_bins_col, h->S8, h->N, wide_cols(h), h->d_rowidx[0], h->d_rowidx[1], t.segs, m,
                                               t.tile_seg, t.n_tiles, (int)grid_bound, h->d_tile_state, t.ticket, t.epoch,
                                               t.counts, h->st, sparse_rows(h), stamp(kStPart, d)));
        RET(mark(kPart));
        CK(ytk::launch_lv_plan_build(t, d, child_buf, leaf, h->st));
        h->launches += plan_in_xrecs() ? 2 : 3;
        return YTK_OK;
    }
    // several GPUs: the record exchange ahead of a level's partition plans it (lv_xrecs_k); YTK_GBDT_NO_FUSE_PLAN=1: a launch of its own
    bool plan_in_xrecs() const {
        static const bool off = [] { const char* e = getenv("YTK_GBDT_NO_FUSE_PLAN"); return e && atoi(e); }();
        return h->world > 1 && !off;
    }
    // ... computeLeafStats (:341-385) when the children are leaves ...
    int level_leaf(int d) {
        CK(ytk::launch_lv_leaf_sums(t, d, h->d_gq, h->d_rowidx[0], h->d_rowidx[1], ytk::kNumSMs * 4, h->st));
        h->launches++;
        RET(mark(kInit));
        return YTK_OK;
    }
    // ... else buildHist(parent, l, r) (:489-508): the smaller children from rows ...
    int level_hist(int d) {
        const int m = 1 << d;
        RET(zero_build_targets(m, m));
        RET(hist_launch(m, d + 1));
        RET(mark(kHist));
        return YTK_OK;
    }
    // ... the larger ones by subtraction (inside the search: ScanNode::sub), and findBestSplit of both
    int level_scan(int d) {
        const int m = 1 << d;
        RET(scan_launch(t.scans, 2 * m, (size_t)2 * m - 1, d + 1));
        RET(mark(kSplit));
        return YTK_OK;
    }

    int finish_enqueue() {
        CK(cudaMemcpyAsync(h->eng.host, h->eng.dev, h->eng.mirror_bytes, cudaMemcpyDeviceToHost, h->st));
        return YTK_OK;
    }

    // ---- after the wait: the Tree, in the reference's order ---------------------------------------------------
    int alloc_node(int parent) {        // Tree.AllocTreeNode (J/data/gbdt/Tree.java:58-64)
        if (out->n_nodes >= out->capacity) return -1;
        int nid = out->n_nodes++;
        out->parent[nid] = parent; out->left[nid] = -1; out->right[nid] = -1;
        out->fid[nid] = 0; out->lo[nid] = 0; out->hi[nid] = 0; out->node_info[nid] = 0.f;
        out->loss_chg[nid] = -std::numeric_limits<float>::infinity(); out->hess_sum[nid] = 0.f; out->sample_cnt[nid] = 0;
        return nid;
    }

    int assemble(bool masked) {
        const ytk::LvNode* nodes = h->eng.h_nodes;
        const ytk_split* recs = h->eng.h_recs;
        const long long* leaf = h->eng.h_leaf;
        const float lr = p.learning_rate;
        const ytk::GradScale sc = *h->h_scale.h;
        out->n_nodes = 0;
        h->leaf_ranges.clear();
        struct Polled { int heap, nid; };
        std::vector<Polled> cur, next;
        if (!(nodes[0].flags & ytk::kLvActive)) return fail(YTK_ERR_ARG, "internal: the device built no root");
        const int root_nid = alloc_node(-1);
        if (root_nid < 0) return fail(YTK_ERR_CAPACITY, "make_tree: tree capacity %d", out->capacity);
        cur.push_back({0, root_nid});
        auto searched = [&](int heap, int nid) {            // updateTreeNodeStat (:387-393) of a searched node + setSplit
            const ytk_split& r = recs[heap];
            out->loss_chg[nid] = r.loss_chg;
            out->hess_sum[nid] = (float)r.sum_hess;
            out->sample_cnt[nid] = nodes[heap].sample_cnt;
            if ((double)r.loss_chg > p.min_split_loss) {         // findBestSplit tail (:587-592)
                out->fid[nid] = r.fid; out->lo[nid] = r.lo; out->hi[nid] = r.hi;
                out->node_info[nid] = 0.5f * (float)(r.lo + r.hi);
            }
        };
        auto range_of = [&](int heap, int depth, int nid) {
            if (nodes[heap].end > nodes[heap].begin) {
                Interval v; v.b = nodes[heap].begin; v.e = nodes[heap].end; v.buf = level_buf(root_buf, depth);
                h->leaf_ranges.push_back({nid, v});
            }
        };
        searched(0, root_nid);
        T->hist_built = 1;
        double rows_root = (double)(nodes[0].end - nodes[0].begin);
        T->hist_rows = rows_root; T->hist_cells = rows_root * n_sel;
        T->hist_rows_indexed = root_buf >= 0 ? rows_root : 0.0;
        for (int d = 0; d <= D && !cur.empty(); d++) {
            next.clear();
            for (const Polled& e : cur) {
                const ytk_split& r = recs[e.heap];
                const bool is_leaf = (double)r.loss_chg <= p.min_split_loss || d == D;      // (:249-250)
                const bool dev_split = (nodes[e.heap].flags & ytk::kLvSplit) != 0;
                if (is_leaf == dev_split)
                    return fail(YTK_ERR_ARG, "internal: the device split a different set of nodes than the host (heap node %d)", e.heap);
                if (is_leaf) {                               // (:253)
                    out->node_info[e.nid] = r.node_value * lr;
                    range_of(e.heap, d, e.nid);
                    continue;
                }
                const int l = alloc_node(e.nid), rr = alloc_node(e.nid);      // tree.addChilds (:256)
                if (l < 0 || rr < 0) return fail(YTK_ERR_CAPACITY, "make_tree: tree needs more than %d nodes", out->capacity);
                out->left[e.nid] = l; out->right[e.nid] = rr;
                const int hl = 2 * e.heap + 1, hr = 2 * e.heap + 2;
                const ytk::LvNode& pn = nodes[e.heap];
                const bool small_left_dev = (pn.flags & ytk::kLvSmallLeft) != 0;
                const double rows_small = (double)(small_left_dev ? pn.cnt_l : pn.cnt_r);
                if (d + 1 == D) {
                    // computeLeafStats (:341-385): the smaller child's sums, the larger one by subtraction
                    const bool small_left = pn.gcnt_l < pn.gcnt_r;          // (:344-350)
                    if (h->world == 1 && small_left != small_left_dev)    // (several GPUs: both children were summed)
                        return fail(YTK_ERR_ARG, "internal: leaf sums were taken over the other child (heap node %d)", e.heap);
                    const int hs = small_left ? hl : hr;
                    const double sG = (double)leaf[2 * hs] * sc.inv_g, sH = (double)leaf[2 * hs + 1] * sc.inv_h;
                    const double lG = r.sum_grad - sG, lH = r.sum_hess - sH;                 // setSubstract (:382)
                    const int nid_s = small_left ? l : rr, nid_l = small_left ? rr : l;
                    float rg, nv;
                    ytk_host_node_value(&p, sG, sH, &rg, &nv);
                    out->node_info[nid_s] = nv * lr; out->hess_sum[nid_s] = (float)sH;
                    ytk_host_node_value(&p, lG, lH, &rg, &nv);
                    out->node_info[nid_l] = nv * lr; out->hess_sum[nid_l] = (float)lH;
                    out->sample_cnt[l] = pn.gcnt_l; out->sample_cnt[rr] = pn.gcnt_r;
                    range_of(hl, d + 1, l);
                    range_of(hr, d + 1, rr);
                } else {
                    searched(hl, l);
                    searched(hr, rr);
                    next.push_back({hl, l});
                    next.push_back({hr, rr});
                    T->hist_built++;
                    T->hist_rows += rows_small; T->hist_cells += rows_small * n_sel; T->hist_rows_indexed += rows_small;
                }
            }
            cur.swap(next);
        }
        h->tree_made = true;
        h->unsampled_begin = masked ? nodes[0].end : -1;
        return YTK_OK;
    }

    // TimeStats (J/data/gbdt/TimeStats.java:69-78) from the kernels' own time spans on the device: histogram build,
    // row partition, split search (with the sibling subtraction); everything else on the device (gradient fixed point, planners, leaf
    // sums, gaps) is the remainder of the tree's device time and goes to init_stats.
    int stage_times() {
        const unsigned long long* st = h->eng.h_stamps;
        double ms[kStKinds] = {0, 0, 0, 0, 0};
        unsigned long long first = ~0ull, last = 0;
        for (int k = 0; k < kStKinds; k++)
            for (int l = 0; l < kStampLevels; l++) {
                const unsigned long long a = ~st[2 * (k * kStampLevels + l)], b = st[2 * (k * kStampLevels + l) + 1];
                if (b == 0 || b < a) continue;
                ms[k] += (double)(b - a) * 1e-6;
                first = std::min(first, a); last = std::max(last, b);
            }
        T->compute_hist = ms[kStHist];
        T->gpu_hist_ms = ms[kStHist];
        T->hist_launches = 0;
        for (int l = 0; l < kStampLevels; l++) T->hist_launches += st[2 * (kStHist * kStampLevels + l) + 1] != 0;
        T->partition = ms[kStPart];
        T->find_best_split = ms[kStScan];
        T->communicate_hist = ms[kStComm];      // exchange kernels: includes waiting for the slowest rank to arrive
        T->sync_best_split = ms[kStSync];
        T->build_hist = T->compute_hist + T->communicate_hist;
        const double span = last > first ? (double)(last - first) * 1e-6 : 0.0;
        T->device_span_ms = span;
        T->init_stats = std::max(0.0, span - ms[kStHist] - ms[kStPart] - ms[kStScan] - T->communicate_hist - T->sync_best_split);
        T->stage_times_from_gpu = 1;
        // YTK_GBDT_TRACE_STAMPS=1: the tree's timeline on the GPU's own clock -- start (from the first stamp) and
        // duration of every stamped kernel, in time order; what lies between two lines is unstamped (planners,
        // gradient scales, memsets, leaf sums) or idle
        static const bool trace = [] { const char* e = getenv("YTK_GBDT_TRACE_STAMPS"); return e && atoi(e); }();
        if (trace) {
            static const char* const names[kStKinds] = {"hist", "partition", "scan", "xhist", "xrecs"};
            struct Ev { unsigned long long a, b; int k, l; };
            std::vector<Ev> ev;
            for (int k = 0; k < kStKinds; k++)
                for (int l = 0; l < kStampLevels; l++) {
                    const unsigned long long a = ~st[2 * (k * kStampLevels + l)], b = st[2 * (k * kStampLevels + l) + 1];
                    if (b != 0 && b >= a) ev.push_back({a, b, k, l});
                }
            std::sort(ev.begin(), ev.end(), [](const Ev& x, const Ev& y) { return x.a < y.a; });
            unsigned long long prev_end = first;
            fprintf(stderr, "ytk_gbdt timeline (us): span %.1f\n", span * 1e3);
            for (const Ev& e : ev) {
                fprintf(stderr, "  +%8.1f  %-9s level %2d  %7.1f us   (gap before: %.1f)\n", (double)(e.a - first) * 1e-3, names[e.k], e.l,
                        (double)(e.b - e.a) * 1e-3, (double)((long long)e.a - (long long)prev_end) * 1e-3);
                prev_end = std::max(prev_end, e.b);
            }
        }
        return YTK_OK;
    }
    bool replayed = false;
};

}  // namespace

int engine_make_tree(ytk_gbdt* h, int gid, const int32_t* fea_index, int32_t n_fea, const uint8_t* row_mask, ytk_tree_out* out,
                     ytk_time_stats* T) {
    LevelEngine e(h, out, T);
    return e.run(gid, fea_index, n_fea, row_mask);
}

// The ranks of a clique of handles on ONE device, phase by phase on their shared stream: every rank's
// kernels of a phase, then the exchange as "push" for every rank and "reduce" for every rank.  The same
// kernels, tables and staging layout as one process per GPU; only the flags between GPUs are not needed.
int engine_make_tree_clique(ytk_gbdt** hs, int world, int gid, const int32_t* fea_index, int32_t n_fea,
                            const uint8_t* const* row_masks, ytk_tree_out* outs, ytk_time_stats* Ts) {
    std::vector<LevelEngine> eng;
    eng.reserve(world);
    for (int r = 0; r < world; r++) eng.emplace_back(hs[r], &outs[r], &Ts[r]);
    const int D = hs[0]->p.max_depth;
    const double t0 = now_ms();
#define EACH(call) for (int r = 0; r < world; r++) RET(eng[r].call)
#define XCHG(fn, d) do { EACH(fn(d, ytk::kXPush)); EACH(fn(d, ytk::kXReduce)); } while (0)
    EACH(prepare(gid, fea_index, n_fea));
    EACH(begin_scale());
    EACH(x_max(ytk::kXPush)); EACH(x_max(ytk::kXReduce));
    for (int r = 0; r < world; r++) RET(eng[r].begin_root(row_masks ? row_masks[r] : nullptr));
    EACH(root_hist());
    XCHG(x_hist, -1);
    EACH(root_scan());
    XCHG(x_recs, -1);
    for (int d = 0; d < D; d++) {
        EACH(level_partition(d));
        if (d == D - 1) {
            EACH(level_leaf(d));
            XCHG(x_leaf, d);
        } else {
            EACH(level_hist(d));
            XCHG(x_hist, d);
            EACH(level_scan(d));
            XCHG(x_recs, d);
        }
    }
    EACH(finish_enqueue());
#undef XCHG
#undef EACH
    const double t1 = now_ms();
    CK(cudaStreamSynchronize(hs[0]->st));
    const double t2 = now_ms();
    for (int r = 0; r < world; r++) RET(eng[r].after_wait(row_masks && row_masks[r], t0, t1, t2));
    return YTK_OK;
}

}  // namespace ytkh
