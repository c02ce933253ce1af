// the step either side of make(), kept on the device, and test-set scoring.
#include "host.cuh"

extern "C" {

// -------------------------------------------------------------------------------------
// §8(f) row 1: score update, loss and gradient pairs on the device
// -------------------------------------------------------------------------------------
int ytk_gbdt_set_targets(ytk_gbdt* h, int32_t loss, const float* y, const float* weight, const float* init_score) {
    if (!h || !y) return fail(YTK_ERR_ARG, "set_targets: null argument");
    if (!h->rows_set) return fail(YTK_ERR_ARG, "set_targets: call set_bins first");
    const int K = h->p.class_num;
    if (loss < 0 || loss > 3) return fail(YTK_ERR_ARG, "set_targets: unknown loss %d", loss);
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


}  // extern "C"
