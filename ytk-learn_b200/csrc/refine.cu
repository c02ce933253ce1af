// TreeRefiner.refine for loss_function = l1
// (J/optimizer/gbdt/TreeRefiner.java:209-255, exact data-parallel variant :104-123): after make(), every
// leaf's value becomes learning_rate x the WEIGHTED MEDIAN of (label - score) over the leaf's rows, as
// PreciseQuantile computes it (J/utils/PreciseQuantile.java:156-208 / :303-372): the (value, weight) pairs
// of the leaf from ALL workers, sorted by value, weights accumulated in double in that order, the first
// value whose accumulated weight reaches half the total (an empty leaf gives QUANTILE_MISSGING = +inf; the
// allreduceMap(MIN) at :118 only carries rank 0's answer to everybody).
//
// Here: one 64-bit key per row -- (leaf index << 32 | order-preserving float bits) -- from the final
// partition the tree maker still holds, one radix sort of all rows (CUB; with several GPUs every rank
// gathers all ranks' pairs first, as the reference concatenates its workers' lists), then one thread per
// leaf walks its sorted segment sequentially, so the double accumulation is in the reference's order.
// Not on the hot path (l1 only); J/ = /root/reference/src/main/java/com/fenbi/ytklearn/.
#include <cub/device/device_radix_sort.cuh>
#include <math_constants.h>

#include "host.cuh"

namespace ytk {

__device__ __forceinline__ unsigned ordered_bits(float v) {
    const unsigned b = __float_as_uint(v);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float from_ordered(unsigned o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

// keys/weights of the rows of the leaf ranges; entries [n_local, n_pad) are padding that sorts last
__global__ void refine_keys_k(const int32_t* __restrict__ rowidx0, const int32_t* __restrict__ rowidx1,
                              const PosItem* __restrict__ items, int n_items, const long long* __restrict__ item_off,
                              const float* __restrict__ residual, const float* __restrict__ y, const float* __restrict__ score,
                              const float* __restrict__ weight, unsigned long long* __restrict__ keys, float* __restrict__ wts) {
    const PosItem it = items[blockIdx.y];
    const int32_t* __restrict__ rowidx = it.buf < 0 ? nullptr : (it.buf == 0 ? rowidx0 : rowidx1);
    const long long base = item_off[blockIdx.y];
    for (long long i = it.begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < it.end;
         i += (long long)gridDim.x * blockDim.x) {
        const int32_t row = rowidx ? rowidx[i] : (int32_t)i;
        const float r = residual ? residual[row] : __fsub_rn(y[row], score[row]);      // labels[k][i] - scores[k][i] (:226)
        keys[base + (i - it.begin)] = ((unsigned long long)(unsigned)it.nid << 32) | ordered_bits(r);
        wts[base + (i - it.begin)] = weight ? weight[row] : 1.0f;
    }
}
__global__ void refine_pad_k(unsigned long long* keys, float* wts, long long from, long long to) {
    for (long long i = from + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < to; i += (long long)gridDim.x * blockDim.x) {
        keys[i] = ~0ull;
        wts[i] = 0.f;
    }
}

// one thread per leaf: its segment in the sorted keys by binary search, then PreciseQuantile's two passes
__global__ void refine_median_k(const unsigned long long* __restrict__ keys, const float* __restrict__ wts, long long n,
                                int n_leaves, float* __restrict__ median) {
    const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf >= n_leaves) return;
    auto lower = [&](unsigned long long k) {
        long long lo = 0, hi = n;
        while (lo < hi) {
            const long long mid = (lo + hi) >> 1;
            if (keys[mid] < k) lo = mid + 1; else hi = mid;
        }
        return lo;
    };
    const long long b = lower((unsigned long long)leaf << 32), e = lower((unsigned long long)(leaf + 1) << 32);
    if (b >= e) { median[leaf] = CUDART_INF_F; return; }           // Constants.QUANTILE_MISSGING
    double total = 0.0;
    for (long long i = b; i < e; i++) total = __dadd_rn(total, (double)wts[i]);
    const double target = __dmul_rn(total, 0.5);
    double acc = 0.0;
    long long at = e - 1;
    for (long long i = b; i < e; i++) {
        acc = __dadd_rn(acc, (double)wts[i]);
        if (acc >= target) { at = i; break; }
    }
    median[leaf] = from_ordered((unsigned)(keys[at] & 0xffffffffull));
}

}  // namespace ytk

extern "C" int ytk_gbdt_refine_leaves(ytk_gbdt* h, const float* residual, const float* weight, ytk_tree_out* tree) {
    if (!h || !tree) return fail(YTK_ERR_ARG, "refine_leaves: null argument");
    if (!h->tree_made) return fail(YTK_ERR_ARG, "refine_leaves: no tree has been made");
    if (h->clique) return fail(YTK_ERR_UNSUPPORTED, "refine_leaves: not for clique members");
    if (tree->n_nodes != (int)h->last_tree.size()) return fail(YTK_ERR_ARG, "refine_leaves: not the tree the last make() built");
    if (!residual && (h->loss_kind < 0 || h->p.class_num != 1))
        return fail(YTK_ERR_ARG, "refine_leaves: pass residuals, or call set_targets first (one class)");
    CK(cudaSetDevice(h->device));
    const int64_t N = h->N;
    // leaves in increasing node id: leaf index = rank among the leaves
    std::vector<int> leaf_index(tree->n_nodes, -1), leaf_nid;
    for (int nid = 0; nid < tree->n_nodes; nid++)
        if (tree->left[nid] == -1) { leaf_index[nid] = (int)leaf_nid.size(); leaf_nid.push_back(nid); }
    const int n_leaves = (int)leaf_nid.size();
    std::vector<ytk::PosItem> items;
    std::vector<long long> item_off;
    long long n_local = 0;
    for (const auto& lr : h->leaf_ranges) {
        if (lr.first < 0 || lr.first >= tree->n_nodes || leaf_index[lr.first] < 0)
            return fail(YTK_ERR_ARG, "refine_leaves: the tree's leaves are not the last make()'s");
        items.push_back({(long long)lr.second.b, (long long)lr.second.e, lr.second.buf, leaf_index[lr.first]});
        item_off.push_back(n_local);
        n_local += lr.second.e - lr.second.b;
    }
    // several GPUs: every rank sorts all ranks' pairs (equal block sizes for the all-gather: the largest shard)
    long long n_pad = n_local;
    if (h->world > 1) {
        long long* d_n = nullptr;
        CK(cudaMalloc(&d_n, sizeof(long long)));
        CK(cudaMemcpyAsync(d_n, &n_local, sizeof(long long), cudaMemcpyHostToDevice, h->st));
        NK(ytk::nccl_api().AllReduce(d_n, d_n, 1, ncclInt64, ncclMax, h->comm, h->st));
        CK(cudaMemcpyAsync(&n_pad, d_n, sizeof(long long), cudaMemcpyDeviceToHost, h->st));
        CK(cudaStreamSynchronize(h->st));
        CK(cudaFree(d_n));
    }
    const long long n_all = n_pad * h->world;
    const size_t cap = (size_t)std::max<long long>(n_all, 1);
    unsigned long long *d_keys = nullptr, *d_keys2 = nullptr;
    float *d_w = nullptr, *d_w2 = nullptr, *d_res = nullptr, *d_wt = nullptr, *d_med = nullptr;
    long long* d_off = nullptr;
    void* d_tmp = nullptr;
    auto done = [&](int rc) {
        for (void* p : {(void*)d_keys, (void*)d_keys2, (void*)d_w, (void*)d_w2, (void*)d_res, (void*)d_wt, (void*)d_med, (void*)d_off, d_tmp})
            if (p) cudaFree(p);
        h->arena.rewind();
        return rc;
    };
#define RCK(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return done(fail(YTK_ERR_CUDA, "refine_leaves: %s", cudaGetErrorString(e__))); } while (0)
    RCK(cudaMalloc(&d_keys, cap * 8)); RCK(cudaMalloc(&d_keys2, cap * 8));
    RCK(cudaMalloc(&d_w, cap * 4)); RCK(cudaMalloc(&d_w2, cap * 4));
    RCK(cudaMalloc(&d_med, (size_t)std::max(n_leaves, 1) * 4));
    if (residual) {
        RCK(cudaMalloc(&d_res, (size_t)std::max<int64_t>(N, 1) * 4));
        RCK(cudaMemcpyAsync(d_res, residual, (size_t)N * 4, cudaMemcpyHostToDevice, h->st));
    }
    const float* w_dev = nullptr;
    if (weight) {
        RCK(cudaMalloc(&d_wt, (size_t)std::max<int64_t>(N, 1) * 4));
        RCK(cudaMemcpyAsync(d_wt, weight, (size_t)N * 4, cudaMemcpyHostToDevice, h->st));
        w_dev = d_wt;
    } else if (!residual) {
        w_dev = h->d_w;                                    // GBDTCoreData.weight as set_targets holds it (null: all 1)
    }
    unsigned long long* mine = d_keys + (size_t)(h->world > 1 ? h->rank * n_pad : 0);
    float* mine_w = d_w + (size_t)(h->world > 1 ? h->rank * n_pad : 0);
    if (!items.empty()) {
        RCK(h->t_pos.upload(items, h->st, h->arena));
        RCK(cudaMalloc(&d_off, items.size() * sizeof(long long)));
        RCK(cudaMemcpyAsync(d_off, item_off.data(), items.size() * sizeof(long long), cudaMemcpyHostToDevice, h->st));
        dim3 grid(64, (unsigned)items.size());
        ytk::refine_keys_k<<<grid, 256, 0, h->st>>>(h->d_rowidx[0], h->d_rowidx[1], h->t_pos.d, (int)items.size(), d_off, d_res,
                                                    h->d_y, h->d_score, w_dev, mine, mine_w);
        RCK(cudaGetLastError());
    }
    if (n_pad > n_local) {
        ytk::refine_pad_k<<<64, 256, 0, h->st>>>(mine, mine_w, n_local, n_pad);
        RCK(cudaGetLastError());
    }
    if (h->world > 1) {
        ncclResult_t r1 = ytk::nccl_api().AllGather(mine, d_keys, (size_t)n_pad * 8, ncclChar, h->comm, h->st);
        ncclResult_t r2 = ytk::nccl_api().AllGather(mine_w, d_w, (size_t)n_pad * 4, ncclChar, h->comm, h->st);
        if (r1 != ncclSuccess || r2 != ncclSuccess) return done(fail(YTK_ERR_NCCL, "refine_leaves: all-gather failed"));
    }
    int end_bit = 32;
    while (end_bit < 64 && (1ll << (end_bit - 32)) <= (long long)n_leaves) end_bit++;
    if (h->world > 1) end_bit = 64;                        // the padding keys are all ones
    if (n_all > 0) {
        size_t tmp_bytes = 0;
        RCK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_keys, d_keys2, d_w, d_w2, (int64_t)n_all, 0, end_bit, h->st));
        RCK(cudaMalloc(&d_tmp, std::max<size_t>(tmp_bytes, 16)));
        RCK(cub::DeviceRadixSort::SortPairs(d_tmp, tmp_bytes, d_keys, d_keys2, d_w, d_w2, (int64_t)n_all, 0, end_bit, h->st));
    }
    ytk::refine_median_k<<<(n_leaves + 63) / 64, 64, 0, h->st>>>(d_keys2, d_w2, n_all, n_leaves, d_med);
    RCK(cudaGetLastError());
    std::vector<float> med(n_leaves);
    RCK(cudaMemcpyAsync(med.data(), d_med, (size_t)n_leaves * 4, cudaMemcpyDeviceToHost, h->st));
    RCK(cudaStreamSynchronize(h->st));
#undef RCK
    for (int k = 0; k < n_leaves; k++) {
        const float v = med[k] * h->p.learning_rate;       // tree.getNode(nodeId).setLeaf(value * learningRate) (:252)
        tree->node_info[leaf_nid[k]] = v;
        h->last_tree[leaf_nid[k]].info = v;                // update_scores adds the refined values
    }
    return done(YTK_OK);
}
