// ytk_gbdt_make_tree: the control flow of DataParallelTreeMaker.make
// (J/optimizer/gbdt/DataParallelTreeMaker.java:229-295) -- expansion queue, stop rules, smaller-child /
// subtraction trick, HistogramPool policy -- driving the kernels (hist.cu, split.cu, partition.cu) in batches.
// J/ = /root/reference/src/main/java/com/fenbi/ytklearn/.
#include "host.cuh"

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
    // The host wait of a batch (optionally polled from user space; cudaEventSynchronize measured the same)
    int wait_event(cudaEvent_t ev) {
        if (!h->spin_wait) { CK(cudaEventSynchronize(ev)); return peer_check(h, "make_tree"); }
        for (;;) {
            cudaError_t e = cudaEventQuery(ev);
            if (e == cudaSuccess) return peer_check(h, "make_tree");
            if (e != cudaErrorNotReady) CK(e);
        }
    }
    int wait_stream() {
        CK(cudaStreamSynchronize(h->st));
        return peer_check(h, "make_tree");
    }
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
        RET(quantize_grads(h, gid));
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
            RET(wait_event(h->ev_level));
        } else {
            RET(wait_stream());
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
            RET(wait_event(h->ev_level));   // the one host wait of the batch
        } else {
            b.tr.t_queued = now_ms();
            RET(wait_stream());                      // the one host wait of the batch
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
            RET(wait_stream());
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
    if (h->clique) return fail(YTK_ERR_ARG, "make_tree: this handle is a member of a clique (ytk_gbdt_make_tree_clique)");
    CK(cudaSetDevice(h->device));
    const double t_start = now_ms();
    ytk_time_stats local_ts;
    memset(&local_ts, 0, sizeof(local_ts));
    ytk_time_stats* T = &local_ts;
    h->launches = 0;
    h->ev_used = 0;
    h->evc_used = 0;

    if (engine_applicable(h)) {
        RET(engine_make_tree(h, gid, fea_index, n_fea, row_mask, out, T));
    } else {
        TreeGrower grower(h, out, T);
        RET(grower.grow(gid, fea_index, n_fea, row_mask));
    }

    if (h->peer_mode) {
        CK(cudaStreamSynchronize(h->st));
        RET(peer_check(h, "make_tree"));
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
    T->hist_events_ms = gpu_ms;            // CUDA events around the histogram launches (0 when replayed from a graph)
    if (!T->stage_times_from_gpu) {
        T->gpu_hist_ms = gpu_ms;
        T->hist_launches = h->ev_used / 2;
        T->rendezvous = h->evc_used / 2;
        T->communicate_hist = comm_ms;     // CUDA-event time of the collectives (stream time, includes waiting for peers)
        T->compute_hist = gpu_ms;
    }
    T->kernel_launches = h->launches;
    T->peer_collectives = h->peer_mode ? 1 : 0;
    T->total = now_ms() - t_start;
    if (ts) *ts = *T;
    return YTK_OK;
}

int ytk_gbdt_make_tree_clique(ytk_gbdt** hs, int32_t world, int32_t gid, const int32_t* fea_index, int32_t n_fea,
                              const uint8_t* const* row_masks, ytk_tree_out* outs, ytk_time_stats* ts) {
    if (!hs || !outs || world < 2) return fail(YTK_ERR_ARG, "make_tree_clique: bad arguments");
    for (int r = 0; r < world; r++) {
        ytk_gbdt* h = hs[r];
        if (!h || !h->clique || h->world != world || h->rank != r || h->clique_members != hs[0]->clique_members)
            return fail(YTK_ERR_ARG, "make_tree_clique: handle %d is not rank %d of one clique of %d", r, r, world);
        if (!h->rows_set || !h->grads_set) return fail(YTK_ERR_ARG, "make_tree_clique: set_bins/upload_rows/set_grads first (rank %d)", r);
        if (gid < 0 || gid >= h->p.class_num) return fail(YTK_ERR_ARG, "make_tree_clique: groupId %d", gid);
        if (h->F != hs[0]->F || h->B != hs[0]->B) return fail(YTK_ERR_ARG, "make_tree_clique: the ranks' bin layouts differ");
        if (outs[r].capacity < 1) return fail(YTK_ERR_CAPACITY, "make_tree_clique: tree capacity %d", outs[r].capacity);
    }
    CK(cudaSetDevice(hs[0]->device));
    if (!hs[0]->clique_linked) {           // every member sees every member's exchange and staging buffers
        for (int r = 0; r < world; r++)
            for (int q = 0; q < world; q++) { hs[r]->pt.xchg[q] = hs[q]->my_xchg; hs[r]->pt.hist[q] = hs[q]->my_hist; }
        for (int r = 0; r < world; r++) hs[r]->clique_linked = true;
    }
    if (!engine_applicable(hs[0]))
        return fail(YTK_ERR_UNSUPPORTED, "make_tree_clique: only what the device-planned level loop takes (level-wise growth, "
                                         "max_depth 1..12, unbounded pool, no row-count rule, dense rows of at most 256 bins)");
    const double t_start = now_ms();
    std::vector<ytk_time_stats> T(world);
    memset(T.data(), 0, sizeof(ytk_time_stats) * world);
    for (int r = 0; r < world; r++) { hs[r]->launches = 0; hs[r]->ev_used = 0; hs[r]->evc_used = 0; }
    RET(engine_make_tree_clique(hs, world, gid, fea_index, n_fea, row_masks, outs, T.data()));
    for (int r = 0; r < world; r++) {
        ytk_gbdt* h = hs[r];
        h->last_tree.resize(outs[r].n_nodes);
        for (int nid = 0; nid < outs[r].n_nodes; nid++)
            h->last_tree[nid] = {outs[r].left[nid], outs[r].right[nid], outs[r].fid[nid], outs[r].node_info[nid]};
        h->last_gid = gid;
        T[r].kernel_launches = h->launches;
        T[r].total = now_ms() - t_start;
        if (ts) ts[r] = T[r];
    }
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


}  // extern "C"
