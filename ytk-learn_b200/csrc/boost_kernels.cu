// ====================================================================================
// score update, loss and gradient pairs on the device --
// GBDTOptimizer.predictAndCalcLossGrad, train branch (J/optimizer/GBDTOptimizer.java:513-609).
// ====================================================================================
#include "common.cuh"
#include "device_util.cuh"

#include <math_constants.h>

#include <algorithm>
#include <cstdlib>

namespace ytk {

// score[(row*K + gid)] += leaf value, for every row of the listed leaves (float add, as :651-655).
// The row's leaf is known from the final partition, so no tree walk is needed.
__global__ void add_leaf_values_k(const int32_t* __restrict__ rowidx0, const int32_t* __restrict__ rowidx1,
                                  const LeafItem* __restrict__ items, float* __restrict__ score, int32_t K, int32_t gid) {
    const LeafItem it = items[blockIdx.y];
    const int32_t* __restrict__ rowidx = it.buf < 0 ? nullptr : (it.buf == 0 ? rowidx0 : rowidx1);
    for (long long i = it.begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < it.end;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = rowidx ? rowidx[i] : i;
        float* s = score + r * K + gid;
        *s = ld_gather_f32_rw(s) + it.value;
    }
}

// Rows that were not sampled for the tree are in no leaf interval: walk the tree on their slots
// with Tree.getNext's rule, slot <= cond goes left (J/data/gbdt/Tree.java:156-168).
__global__ void walk_rows_k(const uint8_t* __restrict__ bins, int32_t S8, WideCols wide, SparseRows sparse,
                            const int32_t* __restrict__ rows,
                            long long begin, long long end, const TreeNodeDev* __restrict__ nodes,
                            float* __restrict__ score, int32_t K, int32_t gid) {
    for (long long i = begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < end;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = rows[i];
        int pid = 0;
        TreeNodeDev nd = nodes[0];
        while (nd.left != -1) {
            const int wi = wide.index ? wide.index[nd.fid] : -1;
            const float v = sparse.row_ptr ? (float)sparse_slot(sparse, (int32_t)r, nd.fid)
                            : wi >= 0 ? (float)wide.data[wi * wide.stride + r] : (float)bins[r * S8 + nd.fid];
            pid = (v <= nd.info) ? nd.left : nd.right;
            nd = nodes[pid];
        }
        score[r * K + gid] += nd.info;
    }
}

// Test-set scoring on ORIGINAL feature values (GBDTOptimizer.score :630-650 with isOriginTree = true;
// Tree.getLeafIndex / getNext, J/data/gbdt/Tree.java:142-168): value <= threshold goes left, a value
// whose bits are Float.NaN's raw pattern (Constants.INT_MISSING_VALUE) takes the node's default child.
__global__ void score_values_k(const int32_t* __restrict__ xbits, long long n, int32_t stride,
                               const TreeNodeDev* __restrict__ nodes, const int32_t* __restrict__ default_child,
                               float* __restrict__ score, int32_t K, int32_t gid) {
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
        int pid = 0;
        TreeNodeDev nd = nodes[0];
        while (nd.left != -1) {
            const int32_t bits = xbits[r * stride + nd.fid];
            if (bits == 0x7fc00000) pid = default_child ? default_child[pid] : nd.left;
            else pid = (__int_as_float(bits) <= nd.info) ? nd.left : nd.right;
            nd = nodes[pid];
        }
        score[r * K + gid] += nd.info;
    }
}

__device__ __forceinline__ double block_sum(double v, double* s_part) {
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += s_part[w];
    return t;      // valid in thread 0
}

// numTreeInGroup == 1 (:536-557): SigmoidFunction (J/loss/SigmoidFunction.java:39-93, zMax = 0) or
// L2Function (J/loss/L2Function.java:30-55).  Float casts where the reference has them: the combined
// score is a float sum, predict is rounded to float before the derivatives, (g,h) are rounded to float.
__global__ void loss_grad_scalar_k(int loss_kind, const float* __restrict__ score, const float* __restrict__ init,
                                   const float* __restrict__ y, const float* __restrict__ w, long long n,
                                   float* __restrict__ gp, double* __restrict__ block_loss) {
    __shared__ double s_part[32];
    double loss = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double wei = w ? (double)w[i] : 1.0;
        const double s = (double)__fadd_rn(score[i], init ? init[i] : 0.0f);
        const double label = y[i];
        double l, pred;
        if (loss_kind == 0) {
            if (s >= 0.0) {
                const double e = exp(-s);
                l = __dadd_rn(log(__dadd_rn(1.0, e)), __dmul_rn(s, __dsub_rn(1.0, label)));
                pred = __ddiv_rn(1.0, __dadd_rn(1.0, e));
            } else {
                const double e = exp(s);
                l = __dsub_rn(log(__dadd_rn(1.0, e)), __dmul_rn(s, label));
                pred = __ddiv_rn(e, __dadd_rn(1.0, e));
            }
        } else if (loss_kind == 3) {                        // L1Function (J/loss/L1Function.java:31-54)
            l = fabs(__dsub_rn(label, s));
            pred = s;
        } else {
            const double d = __dsub_rn(label, s);
            l = __dmul_rn(__dmul_rn(0.5, d), d);
            pred = s;
        }
        loss = __dadd_rn(loss, __dmul_rn(wei, l));
        const double p = (double)(float)pred;
        double d0 = __dsub_rn(p, label);
        if (loss_kind == 3) d0 = d0 > 0.0 ? 1.0 : (d0 < 0.0 ? -1.0 : d0);      // Math.signum(score - label)
        const double d1 = loss_kind == 0 ? __dmul_rn(p, __dsub_rn(1.0, p)) : 1.0;
        *reinterpret_cast<float2*>(gp + 2 * i) = make_float2((float)__dmul_rn(d0, wei), (float)__dmul_rn(d1, wei));
    }
    const double t = block_sum(loss, s_part);
    if (threadIdx.x == 0) block_loss[blockIdx.x] = t;
}

// softmax branch (:559-607, J/loss/SoftmaxFunction.java:44-113); y holds the class index of each row
constexpr int kMaxClasses = 64;
__global__ void loss_grad_softmax_k(const float* __restrict__ score, const float* __restrict__ init,
                                    const float* __restrict__ y_onehot, const float* __restrict__ w, long long n,
                                    int32_t K, float* __restrict__ gp, double* __restrict__ block_loss) {
    __shared__ double s_part[32];
    double loss = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double wei = w ? (double)w[i] : 1.0;
        double ls[kMaxClasses];
        for (int k = 0; k < K; k++) ls[k] = (double)__fadd_rn(score[i * K + k], init ? init[i * K + k] : 0.0f);
        double mx = ls[K - 1];
        for (int k = 0; k < K - 1; k++) if (ls[k] > mx) mx = ls[k];
        double esum = 0.0, sig = 0.0;
        for (int k = 0; k < K; k++) {
            const double nw = __dsub_rn(ls[k], mx);
            sig = __dadd_rn(sig, __dmul_rn(nw, (double)y_onehot[i * K + k]));
            ls[k] = exp(nw);
            esum = __dadd_rn(esum, ls[k]);
        }
        loss = __dadd_rn(loss, __dmul_rn(wei, __dsub_rn(log(esum), sig)));
        const double inv = __ddiv_rn(1.0, esum);
        for (int k = 0; k < K; k++) {
            const double pk = __dmul_rn(ls[k], inv);
            const double d0 = __dsub_rn(pk, (double)y_onehot[i * K + k]);
            const double d1 = __dmul_rn(2.0, __dmul_rn(pk, __dsub_rn(1.0, pk)));
            *reinterpret_cast<float2*>(gp + (i * K + k) * 2) = make_float2((float)__dmul_rn(d0, wei), (float)__dmul_rn(d1, wei));
        }
    }
    const double t = block_sum(loss, s_part);
    if (threadIdx.x == 0) block_loss[blockIdx.x] = t;
}

// fixed-order sum of the per-block partial losses (run-to-run reproducible)
__global__ void sum_blocks_k(const double* __restrict__ block_loss, int n, double* __restrict__ out) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < n; i++) t = __dadd_rn(t, block_loss[i]);
        *out = t;
    }
}

cudaError_t launch_add_leaf_values(const int32_t* rowidx0, const int32_t* rowidx1, const LeafItem* items, int n_items,
                                   float* score, int32_t K, int32_t gid, cudaStream_t st) {
    if (n_items <= 0) return cudaSuccess;
    dim3 grid(64, (unsigned)n_items);
    add_leaf_values_k<<<grid, 256, 0, st>>>(rowidx0, rowidx1, items, score, K, gid);
    return cudaGetLastError();
}
cudaError_t launch_walk_rows(const uint8_t* bins, int32_t S8, WideCols wide, const int32_t* rows, long long begin, long long end,
                             const TreeNodeDev* nodes, float* score, int32_t K, int32_t gid, cudaStream_t st,
                             SparseRows sparse) {
    if (end <= begin) return cudaSuccess;
    long long b = (end - begin + 255) / 256;
    walk_rows_k<<<(unsigned)std::min<long long>(b, kNumSMs * 8), 256, 0, st>>>(bins, S8, wide, sparse, rows, begin, end, nodes, score, K, gid);
    return cudaGetLastError();
}
cudaError_t launch_score_values(const int32_t* xbits, long long n, int32_t stride, const TreeNodeDev* nodes,
                                const int32_t* default_child, float* score, int32_t K, int32_t gid, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    long long b = (n + 255) / 256;
    score_values_k<<<(unsigned)std::min<long long>(b, kNumSMs * 16), 256, 0, st>>>(xbits, n, stride, nodes, default_child, score,
                                                                                 K, gid);
    return cudaGetLastError();
}
int loss_grad_blocks() { return kNumSMs * 4; }
cudaError_t launch_loss_grad(int loss_kind, const float* score, const float* init, const float* y, const float* w,
                             long long n, int32_t K, float* gp, double* block_loss, double* loss_out, cudaStream_t st) {
    const int blocks = loss_grad_blocks();
    if (loss_kind == 2) {
        if (K > kMaxClasses) return cudaErrorInvalidValue;
        loss_grad_softmax_k<<<blocks, 128, 0, st>>>(score, init, y, w, n, K, gp, block_loss);
    } else {
        loss_grad_scalar_k<<<blocks, 256, 0, st>>>(loss_kind, score, init, y, w, n, gp, block_loss);
    }
    sum_blocks_k<<<1, 32, 0, st>>>(block_loss, blocks, loss_out);
    return cudaGetLastError();
}
}  // namespace ytk
