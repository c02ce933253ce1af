// Per-stage entry points for the parity tests and ncu benches.
#include "host.cuh"

extern "C" {

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
    RET(quantize_grads(h, gid));
    int64_t n = rows ? n_rows : h->N;
    // this hook overwrites the fixed-point gradients (and, with a row list, the row-id buffers) the last
    // tree's partition lives in: get_positions / update_scores must not use them afterwards
    h->tree_made = false; h->leaf_ranges.clear(); h->last_tree.clear(); h->last_gid = -1;
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
    if (h->world > 1 && h->comm)           // (a clique member has no communicator: its hook returns the local histogram)
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
    RET(scan_done_counters(h, 1));
    CK(ytk::launch_split_scan(h->t_scan.d, 1, h->d_start, h->d_off_flat, h->d_fea_sel, 0, h->F, first,
                              h->d_scale_hook, gain_params(h->p), nullptr, h->t_parts.d, h->t_split_local.d, nullptr,
                              h->t_scan_done.d, h->st));
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
    h->tree_made = false; h->leaf_ranges.clear(); h->last_tree.clear(); h->last_gid = -1;      // the row-id buffers are overwritten
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
