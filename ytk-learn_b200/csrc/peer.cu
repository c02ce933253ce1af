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
#include "level_plan.cuh"

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

// ====================================================================================
// Exchanges of the device-driven level loop (engine.cu).  Same transport as above -- plain stores into the
// peers' exchange areas, a release flag per (exchange kind, rank), acquire-polling of one's own flags --
// but every kernel can also run its three steps separately (phase bits kXPush / kXWait / kXReduce): a
// clique of handles on ONE device runs "push" for every rank and then "reduce" for every rank on one
// stream, where kernels must not wait for one another (nothing guarantees they are co-scheduled).
// ====================================================================================
__device__ __forceinline__ void lv_push_words(const PeerTable& pt, const unsigned long long* __restrict__ src, size_t area_off,
                                              size_t slot_bytes, size_t words) {
    for (int r = 0; r < pt.world; r++) {
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(pt.xchg[r] + area_off + (size_t)pt.rank * slot_bytes);
        for (size_t i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
    }
}
__device__ __forceinline__ void lv_raise(const PeerTable& pt, int flag_set, unsigned long long epoch) {
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < pt.world) {
        unsigned long long* f = reinterpret_cast<unsigned long long*>(pt.xchg[threadIdx.x]) + flag_set * kMaxPeers + pt.rank;
        st_release_sys(f, epoch);
    }
}

// this exchange's epoch: one more than the kind's last completed one
__device__ __forceinline__ unsigned long long lv_epoch(const unsigned long long* xctr) {
    return *reinterpret_cast<const volatile unsigned long long*>(xctr) + 1;
}

__global__ void lv_xmax_k(PeerTable pt, unsigned* __restrict__ absmax, size_t area_off, size_t slot_bytes, int flag_set,
                          unsigned long long* xctr, int phase, int* err) {
    pdl_sync();                                   // launched with programmatic stream serialization, like its neighbours
    __shared__ unsigned long long s_v;
    const unsigned long long epoch = lv_epoch(xctr);
    if (phase & kXPush) {
        if (threadIdx.x == 0) s_v = (unsigned long long)absmax[0] | ((unsigned long long)absmax[1] << 32);
        __syncthreads();
        lv_push_words(pt, &s_v, area_off, slot_bytes, 1);
        if (phase & kXWait) lv_raise(pt, flag_set, epoch);
    }
    if ((phase & kXWait) && !peer_wait(pt, flag_set, epoch, err)) return;
    if ((phase & kXReduce) && threadIdx.x == 0) {
        // float bits of non-negative values order like unsigned integers
        unsigned g = 0, h = 0;
        for (int r = 0; r < pt.world; r++) {
            const unsigned long long v = *reinterpret_cast<const unsigned long long*>(pt.xchg[pt.rank] + area_off + (size_t)r * slot_bytes);
            g = max(g, (unsigned)v); h = max(h, (unsigned)(v >> 32));
        }
        absmax[0] = g; absmax[1] = h;
    }
    __syncthreads();
    if ((phase & kXReduce) && threadIdx.x == 0) xctr[0] = epoch;
}

// grid (chunks, m).  Raises "my staging buffer is complete" (the histogram kernel ahead in the stream has
// finished), waits for everybody's, and sums the segment this rank owns of every slot that is being built
// straight out of all staging buffers into the slot's histogram.  Block (0,0) also exchanges the child
// counts of the level's partition (d < 0: the root's row count).
__global__ void lv_xhist_k(PeerTable pt, LvTables t, int d, int m, size_t area_off, size_t slot_bytes, int flag_set,
                           unsigned long long* xctr, int phase, int* err, unsigned long long* stamp) {
    pdl_sync();                                   // launched with programmatic stream serialization, like its neighbours
    const bool lead = blockIdx.x == 0 && blockIdx.y == 0;
    const unsigned long long epoch = lv_epoch(xctr);
    stamp_begin(stamp);
    struct Done {      // the last block to leave completes the epoch
        unsigned long long* xctr; unsigned long long epoch; int phase; unsigned long long* stamp;
        __device__ ~Done() {
            stamp_end(stamp);
            if (!(phase & kXReduce)) return;
            __syncthreads();
            if (threadIdx.x == 0) {
                __threadfence();
                unsigned int* done = reinterpret_cast<unsigned int*>(xctr + 1);
                if (atomicAdd(done, 1u) == gridDim.x * gridDim.y - 1) { *done = 0; xctr[0] = epoch; }
            }
        }
    } on_exit{xctr, epoch, phase, stamp};
    if ((phase & kXPush) && lead) {
        lv_push_words(pt, reinterpret_cast<const unsigned long long*>(t.counts), area_off, slot_bytes, 2 * (size_t)m);
        if (phase & kXWait) lv_raise(pt, flag_set, epoch);
    }
    if ((phase & kXWait) && !peer_wait(pt, flag_set, epoch, err)) return;
    if (!(phase & kXReduce)) return;
    if (lead) {
        const char* base = pt.xchg[pt.rank] + area_off;
        for (int s = threadIdx.x; s < m; s += blockDim.x) {
            PartitionCounts g = {0, 0};
            bool unused = false;
            for (int r = 0; r < pt.world; r++) {
                const PartitionCounts c = reinterpret_cast<const PartitionCounts*>(base + (size_t)r * slot_bytes)[s];
                if (c.left < 0) unused = true;          // the slot is not split (every rank says so)
                g.left += c.left; g.right += c.right;
            }
            if (unused) continue;
            t.gcounts[s] = g;
            lv_apply_gcounts(t, d, s, g);
        }
    }
    const int slot = blockIdx.y;
    long long* __restrict__ out = t.seg_final[slot];
    if (!out) return;
    const size_t off = ((size_t)pt.rank * m + slot) * (size_t)t.seg_max;
    for (long long i = (blockIdx.x * (long long)blockDim.x + threadIdx.x) * 2; i < t.seg_max;
         i += (long long)gridDim.x * blockDim.x * 2) {
        // eight ranks' loads in flight at a time: a read over NVLink is a round trip of microseconds, and one after
        // the other they were what an exchange on 8 GPUs took
        longlong2 s2 = make_longlong2(0, 0);
        for (int r0 = 0; r0 < pt.world; r0 += 8) {
            longlong2 v[8];
#pragma unroll
            for (int q = 0; q < 8; q++) {
                v[q] = make_longlong2(0, 0);
                if (r0 + q < pt.world)
                    v[q] = *reinterpret_cast<const longlong2*>(reinterpret_cast<const long long*>(pt.hist[r0 + q]) + off + i);
            }
#pragma unroll
            for (int q = 0; q < 8; q++) { s2.x += v[q].x; s2.y += v[q].y; }
        }
        *reinterpret_cast<longlong2*>(out + i) = s2;
    }
}

// One CTA.  Every rank's records (and left / total hessians) of the n searched nodes go to everybody; merged
// as every rank's host would (ytk_gbdt_merge_splits: needReplace in rank order, element-wise MAX of the node
// totals, DataParallelTreeMaker.java:560,640-653) into the heap arrays.
__global__ void lv_xrecs_k(PeerTable pt, LvTables t, int n, long long first_heap, size_t area_off, size_t slot_bytes, size_t lh_off,
                           size_t lh_slot_bytes, int flag_set, unsigned long long* xctr, int phase, int* err,
                           unsigned long long* stamp, int plan_d, int plan_buf, double min_split_loss) {
    pdl_sync();                                   // launched with programmatic stream serialization, like its neighbours
    const unsigned long long epoch = lv_epoch(xctr);
    stamp_begin(stamp);
    struct Done {
        unsigned long long* xctr; unsigned long long epoch; int phase; unsigned long long* stamp;
        __device__ ~Done() {
            stamp_end(stamp);
            __syncthreads();
            if ((phase & kXReduce) && threadIdx.x == 0) xctr[0] = epoch;
        }
    } on_exit{xctr, epoch, phase, stamp};
    if (phase & kXPush) {
        lv_push_words(pt, reinterpret_cast<const unsigned long long*>(t.recs_local), area_off, slot_bytes,
                      (size_t)n * sizeof(ytk_split) / 8);
        lv_push_words(pt, reinterpret_cast<const unsigned long long*>(t.lh_local), lh_off, lh_slot_bytes, 2 * (size_t)n);
        if (phase & kXWait) lv_raise(pt, flag_set, epoch);
    }
    if ((phase & kXWait) && !peer_wait(pt, flag_set, epoch, err)) return;
    if (!(phase & kXReduce)) return;
    const char* base = pt.xchg[pt.rank] + area_off;
    const char* lbase = pt.xchg[pt.rank] + lh_off;
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
        if (!t.scans[k].hist) continue;                 // not searched (same on every rank)
        ytk_split best = reinterpret_cast<const ytk_split*>(base)[k];
        int best_r = 0;
        double G = best.sum_grad, H = best.sum_hess;
        for (int r = 1; r < pt.world; r++) {
            const ytk_split t2 = reinterpret_cast<const ytk_split*>(base + (size_t)r * slot_bytes)[k];
            const bool rep = (best.fid <= t2.fid) ? (t2.loss_chg > best.loss_chg) : (t2.loss_chg >= best.loss_chg);
            if (rep) { best = t2; best_r = r; }
            if (t2.sum_grad > G) G = t2.sum_grad;
            if (t2.sum_hess > H) H = t2.sum_hess;
        }
        long long tot_h = 0;
        for (int r = 0; r < pt.world; r++) {            // root_gain / node_value / total hessian: from the first rank that had the totals
            const ytk_split t2 = reinterpret_cast<const ytk_split*>(base + (size_t)r * slot_bytes)[k];
            if (t2.sum_grad == G && t2.sum_hess == H) {
                best.root_gain = t2.root_gain; best.node_value = t2.node_value;
                tot_h = reinterpret_cast<const long long*>(lbase + (size_t)r * lh_slot_bytes)[2 * k + 1];
                break;
            }
        }
        best.sum_grad = G; best.sum_hess = H;
        t.recs[first_heap + k] = best;
        t.left_h[2 * (first_heap + k)] = reinterpret_cast<const long long*>(lbase + (size_t)best_r * lh_slot_bytes)[2 * k];
        t.left_h[2 * (first_heap + k) + 1] = tot_h;
    }
    if (plan_d >= 0) {                                  // the merged records decide the next partition: it is planned here
        __syncthreads();
        lv_plan_split_dev(t, plan_d, plan_buf, min_split_loss, nullptr, kFusedTile);
    }
}

// One CTA.  allreduceArray(double[2], SUM) of computeLeafStats (:375) for the 2m children of the last
// level, exact in the integer domain, and the global child counts of the level's partition.
__global__ void lv_xleaf_k(PeerTable pt, LvTables t, int d, size_t area_off, size_t slot_bytes, size_t cnt_off, size_t cnt_slot_bytes,
                           int flag_set, unsigned long long* xctr, int phase, int* err, unsigned long long* stamp) {
    pdl_sync();                                   // launched with programmatic stream serialization, like its neighbours
    const unsigned long long epoch = lv_epoch(xctr);
    stamp_begin(stamp);
    struct Done {
        unsigned long long* xctr; unsigned long long epoch; int phase; unsigned long long* stamp;
        __device__ ~Done() {
            stamp_end(stamp);
            __syncthreads();
            if ((phase & kXReduce) && threadIdx.x == 0) xctr[0] = epoch;
        }
    } on_exit{xctr, epoch, phase, stamp};
    const int m = 1 << d;
    const long long C0 = 2ll * m - 1;
    if (phase & kXPush) {
        lv_push_words(pt, reinterpret_cast<const unsigned long long*>(t.leaf + 2 * C0), area_off, slot_bytes, 4 * (size_t)m);
        lv_push_words(pt, reinterpret_cast<const unsigned long long*>(t.counts), cnt_off, cnt_slot_bytes, 2 * (size_t)m);
        if (phase & kXWait) lv_raise(pt, flag_set, epoch);
    }
    if ((phase & kXWait) && !peer_wait(pt, flag_set, epoch, err)) return;
    if (!(phase & kXReduce)) return;
    const char* base = pt.xchg[pt.rank] + area_off;
    for (int i = threadIdx.x; i < 4 * m; i += blockDim.x) {
        long long sum = 0;
        for (int r = 0; r < pt.world; r++) sum += reinterpret_cast<const long long*>(base + (size_t)r * slot_bytes)[i];
        t.leaf[2 * C0 + i] = sum;
    }
    const char* cbase = pt.xchg[pt.rank] + cnt_off;
    for (int s = threadIdx.x; s < m; s += blockDim.x) {
        PartitionCounts g = {0, 0};
        bool unused = false;
        for (int r = 0; r < pt.world; r++) {
            const PartitionCounts c = reinterpret_cast<const PartitionCounts*>(cbase + (size_t)r * cnt_slot_bytes)[s];
            if (c.left < 0) unused = true;
            g.left += c.left; g.right += c.right;
        }
        if (unused) continue;
        t.gcounts[s] = g;
        lv_apply_gcounts(t, d, s, g);
    }
}

cudaError_t launch_lv_xmax(const PeerTable& pt, unsigned* absmax, size_t area_off, size_t slot_bytes, int flag_set,
                           unsigned long long* xctr, int phase, int* err, cudaStream_t st) {
    return launch_chain(lv_xmax_k, dim3(1), dim3(64), 0, st, pt, absmax, area_off, slot_bytes, flag_set, xctr, phase, err);
}
cudaError_t launch_lv_xhist(const PeerTable& pt, LvTables t, int d, int m, size_t area_off, size_t slot_bytes, int flag_set,
                            unsigned long long* xctr, int phase, int* err, cudaStream_t st, unsigned long long* stamp) {
    dim3 grid((unsigned)std::max<long long>(1, std::min<long long>(16, (t.seg_max / 2 + 255) / 256)), (unsigned)m);
    if (phase == kXPush) grid = dim3(1, 1);
    return launch_chain(lv_xhist_k, grid, dim3(256), 0, st, pt, t, d, m, area_off, slot_bytes, flag_set, xctr, phase, err, stamp);
}
cudaError_t launch_lv_xrecs(const PeerTable& pt, LvTables t, int n, long long first_heap, size_t area_off, size_t slot_bytes,
                            size_t lh_off, size_t lh_slot_bytes, int flag_set, unsigned long long* xctr, int phase, int* err,
                            cudaStream_t st, unsigned long long* stamp, int plan_d, int plan_buf, double min_split_loss) {
    if (!(phase & kXReduce)) plan_d = -1;
    return launch_chain(lv_xrecs_k, dim3(1), dim3(256), 0, st, pt, t, n, first_heap, area_off, slot_bytes, lh_off, lh_slot_bytes, flag_set,
                        xctr, phase, err, stamp, plan_d, plan_buf, min_split_loss);
}
cudaError_t launch_lv_xleaf(const PeerTable& pt, LvTables t, int d, size_t area_off, size_t slot_bytes, size_t cnt_off,
                            size_t cnt_slot_bytes, int flag_set, unsigned long long* xctr, int phase, int* err, cudaStream_t st,
                            unsigned long long* stamp) {
    return launch_chain(lv_xleaf_k, dim3(1), dim3(256), 0, st, pt, t, d, area_off, slot_bytes, cnt_off, cnt_slot_bytes, flag_set, xctr, phase, err, stamp);
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
