// ====================================================================================
// Collectives over NVLink peer memory (one node, one process per GPU).  Every rank owns an
// "exchange" buffer and a histogram staging buffer that all peers have mapped (CUDA IPC).  The
// small per-level exchanges (child row counts, split records, leaf sums) are pushes of a few
// hundred bytes into every peer's buffer followed by a flag; the histogram reduce-scatter is a
// pull: each rank sums the segments it owns straight out of its peers' staging buffers.  These
// replace ncclAllReduce / ncclAllGather / ncclReduceScatter on the per-level path, whose launch and
// protocol latency (3 collectives x ~30 us per level) is what limits scaling of a 3 ms tree.
// ====================================================================================
#include "common.cuh"
#include "device_util.cuh"

#include <math_constants.h>

#include <algorithm>
#include <cstdlib>

namespace ytk {

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// Copies `bytes` (multiple of 8) of this rank's payload into slot [rank] of every rank's exchange
// area, then raises flag[rank] = epoch there.
__device__ __forceinline__ void peer_push(const PeerTable& pt, const unsigned long long* __restrict__ src, size_t area_off,
                                          size_t slot_bytes, size_t bytes, int flag_set, unsigned long long epoch) {
    const size_t words = bytes / 8;
    for (int r = 0; r < pt.world; r++) {
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(pt.xchg[r] + area_off + (size_t)pt.rank * slot_bytes);
        for (size_t i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < pt.world) {
        unsigned long long* f = reinterpret_cast<unsigned long long*>(pt.xchg[threadIdx.x]) + flag_set * kMaxPeers + pt.rank;
        st_release_sys(f, epoch);
    }
}

// Waits (bounded) until every rank's flag of the set has reached `epoch`.  Returns false on timeout.
__device__ __forceinline__ bool peer_wait(const PeerTable& pt, int flag_set, unsigned long long epoch, int* err) {
    __shared__ int s_ok;
    if (threadIdx.x == 0) s_ok = 1;
    __syncthreads();
    if (threadIdx.x < pt.world) {
        const unsigned long long* f = reinterpret_cast<const unsigned long long*>(pt.xchg[pt.rank]) + flag_set * kMaxPeers + threadIdx.x;
        const long long t0 = clock64();
        while (ld_acquire_sys(f) < epoch) {
            if (clock64() - t0 > pt.timeout_cycles) {                              // a peer died or hangs
                s_ok = 0;
                *reinterpret_cast<volatile int*>(err) = 1;                         // (mapped host memory: the host reads it at its next wait)
                break;
            }
            __nanosleep(100);
        }
    }
    __syncthreads();
    return s_ok != 0;
}

// All-reduce (sum, int64) of a small vector in one kernel: push my contribution to every rank, wait
// for everybody's, add them up.  out may alias src.  One CTA.
__global__ void peer_allreduce_k(PeerTable pt, const unsigned long long* __restrict__ src, size_t area_off, size_t slot_bytes,
                                 size_t words, int flag_set, unsigned long long epoch, long long* __restrict__ out, int* err) {
    peer_push(pt, src, area_off, slot_bytes, words * 8, flag_set, epoch);
    if (!peer_wait(pt, flag_set, epoch, err)) return;
    const char* base = pt.xchg[pt.rank] + area_off;
    for (size_t i = threadIdx.x; i < words; i += blockDim.x) {
        long long s = 0;
        for (int r = 0; r < pt.world; r++) s += reinterpret_cast<const long long*>(base + (size_t)r * slot_bytes)[i];
        out[i] = s;
    }
}

// All-gather of the split records (every rank pushes its block into slot [rank] of every peer's area),
// followed by the merge every rank would otherwise do on
// its host (ytk_gbdt_merge_splits: needReplace in rank order, element-wise MAX of the node totals,
// DataParallelTreeMaker.java:560,640-653): merged record k -> out_dev[k] and out_host[k] (pinned).
__global__ void peer_allgather_merge_k(PeerTable pt, const ytk_split* __restrict__ src, size_t area_off, size_t slot_bytes,
                                       int m, int flag_set, unsigned long long epoch, ytk_split* __restrict__ out_dev,
                                       ytk_split* __restrict__ out_host, int* err) {
    peer_push(pt, reinterpret_cast<const unsigned long long*>(src), area_off, slot_bytes, (size_t)m * sizeof(ytk_split),
              flag_set, epoch);
    if (!peer_wait(pt, flag_set, epoch, err)) return;
    const char* base = pt.xchg[pt.rank] + area_off;
    for (int k = threadIdx.x; k < m; k += blockDim.x) {
        const ytk_split first = reinterpret_cast<const ytk_split*>(base)[k];
        ytk_split best = first;
        double G = first.sum_grad, H = first.sum_hess;
        for (int r = 1; r < pt.world; r++) {
            const ytk_split t2 = reinterpret_cast<const ytk_split*>(base + (size_t)r * slot_bytes)[k];
            const bool rep = (best.fid <= t2.fid) ? (t2.loss_chg > best.loss_chg) : (t2.loss_chg >= best.loss_chg);
            if (rep) best = t2;
            if (t2.sum_grad > G) G = t2.sum_grad;
            if (t2.sum_hess > H) H = t2.sum_hess;
        }
        // root_gain / node_value are functions of (G,H): from the first rank that had the totals
        for (int r = 0; r < pt.world; r++) {
            const ytk_split t2 = reinterpret_cast<const ytk_split*>(base + (size_t)r * slot_bytes)[k];
            if (t2.sum_grad == G && t2.sum_hess == H) { best.root_gain = t2.root_gain; best.node_value = t2.node_value; break; }
        }
        best.sum_grad = G; best.sum_hess = H;
        out_dev[k] = best;
        out_host[k] = best;
    }
}

// Reduce-scatter by pulling: this rank owns segment [rank] of every node; staging layout of every
// rank is [owner][node][seg_max].  grid = (chunks, nodes); block 0 of each launch waits for the flags
// ... every block waits (cheap: the flags are already up for all but the first to arrive).
__global__ void peer_pull_reduce_k(PeerTable pt, int m, long long seg_max, int flag_set, unsigned long long epoch,
                                   long long* const* __restrict__ dst, int* err) {
    // the histogram kernel ahead of this one in the stream has finished: tell everybody
    if (blockIdx.x == 0 && blockIdx.y == 0) {
        __threadfence_system();
        if (threadIdx.x < pt.world) {
            unsigned long long* f = reinterpret_cast<unsigned long long*>(pt.xchg[threadIdx.x]) + flag_set * kMaxPeers + pt.rank;
            st_release_sys(f, epoch);
        }
    }
    if (!peer_wait(pt, flag_set, epoch, err)) return;
    const int node = blockIdx.y;
    long long* __restrict__ out = dst[node];
    const size_t off = ((size_t)pt.rank * m + node) * (size_t)seg_max;
    for (long long i = (blockIdx.x * (long long)blockDim.x + threadIdx.x) * 2; i < seg_max;
         i += (long long)gridDim.x * blockDim.x * 2) {
        longlong2 s = make_longlong2(0, 0);
        for (int r = 0; r < pt.world; r++) {
            const longlong2 v = *reinterpret_cast<const longlong2*>(reinterpret_cast<const long long*>(pt.hist[r]) + off + i);
            s.x += v.x; s.y += v.y;
        }
        *reinterpret_cast<longlong2*>(out + i) = s;
    }
}

cudaError_t launch_peer_allreduce(const PeerTable& pt, const void* src, size_t area_off, size_t slot_bytes, size_t words,
                                  int flag_set, unsigned long long epoch, long long* out, int* err, cudaStream_t st) {
    peer_allreduce_k<<<1, 256, 0, st>>>(pt, static_cast<const unsigned long long*>(src), area_off, slot_bytes, words, flag_set,
                                        epoch, out, err);
    return cudaGetLastError();
}
cudaError_t launch_peer_allgather_merge(const PeerTable& pt, const ytk_split* src, size_t area_off, size_t slot_bytes, int m,
                                        int flag_set, unsigned long long epoch, ytk_split* out_dev, ytk_split* out_host,
                                        int* err, cudaStream_t st) {
    peer_allgather_merge_k<<<1, 256, 0, st>>>(pt, src, area_off, slot_bytes, m, flag_set, epoch, out_dev, out_host, err);
    return cudaGetLastError();
}
cudaError_t launch_peer_pull_reduce(const PeerTable& pt, int m, long long seg_max, int flag_set, unsigned long long epoch,
                                    long long* const* dst, int* err, cudaStream_t st) {
    if (m <= 0) return cudaSuccess;
    dim3 grid((unsigned)std::max<long long>(1, std::min<long long>(16, (seg_max / 2 + 255) / 256)), (unsigned)m);
    peer_pull_reduce_k<<<grid, 256, 0, st>>>(pt, m, seg_max, flag_set, epoch, dst, err);
    return cudaGetLastError();
}
}  // namespace ytk
