// The partition planner of the device-driven level loop as a device function: level.cu wraps it in a kernel of its
// own (one GPU: queued behind the split search), and with several GPUs it runs in the tail of the record exchange
// (peer.cu: lv_xrecs_k), which has just merged the records it reads -- one single-CTA launch less per level.
// Callable by all threads of one CTA whose size is a multiple of 32 (at most 1024).
#pragma once

#include "level.cuh"
#include "device_util.cuh"

namespace ytk {

constexpr int kPlanThreads = 512;
constexpr int kPlanMaxWarps = 32;

// exclusive scan of one value per thread over the CTA; *total = sum over the CTA.  s_warp: kPlanMaxWarps words
template <typename T>
__device__ __forceinline__ T block_excl_scan(T v, T* s_warp, T* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (int)blockDim.x >> 5;
    T p = v;
    for (int d = 1; d < 32; d <<= 1) {
        T t = __shfl_up_sync(0xffffffffu, p, d);
        if (lane >= d) p += t;
    }
    __syncthreads();                     // s_warp may still be read from an earlier call
    if (lane == 31) s_warp[warp] = p;
    __syncthreads();
    T woff = 0, tot = 0;
    for (int w = 0; w < nw; w++) {
        if (w < warp) woff += s_warp[w];
        tot += s_warp[w];
    }
    *total = tot;
    return woff + p - v;
}

// Which nodes of level d are split -- lossChg > min_split_loss (the depth rule is the host's: it does not
// ask for a level beyond max_depth) -- and the partition's segment and tile tables.
__device__ __forceinline__ void lv_plan_split_dev(const LvTables& t, int d, int buf, double min_split_loss,
                                                  const int32_t* __restrict__ wide_index, int tile_rows) {
    __shared__ int s_warp[kPlanMaxWarps];
    __shared__ int s_carry;
    const int m = 1 << d, L0 = m - 1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        s_carry = 0; *t.ticket = 0;
        // a fresh epoch for the partition's chained-scan state; the device's epochs live in [2^29, 2^30), the host's
        // (tables planned there) below
        unsigned long long e = *t.epoch + 1;
        if (e < (1ull << 29) || e >= (1ull << 30)) e = (1ull << 29) + 1;
        *t.epoch = e;
    }
    __syncthreads();
    for (int s0 = 0; s0 < m; s0 += (int)blockDim.x) {
        const int s = s0 + threadIdx.x;
        SplitSeg sg = {0, 0, 0, 0, 0.f, 0, 0, 0};
        bool split = false;
        if (s < m) {
            const LvNode nd = t.nodes[L0 + s];
            if (nd.flags & kLvActive) {
                const ytk_split r = t.recs[L0 + s];
                if (!((double)r.loss_chg <= min_split_loss)) {          // the negation of the leaf rule (:249)
                    split = true;
                    sg.begin = nd.begin; sg.end = nd.end; sg.buf = buf;
                    sg.fid = r.fid;
                    sg.cond = 0.5f * (float)(r.lo + r.hi);               // TreeNode.setSplit (:144-149)
                    sg.wide = wide_index ? wide_index[r.fid] + 1 : 0;
                    sg.ntiles = (int)((sg.end - sg.begin + tile_rows - 1) / tile_rows);
                    t.nodes[L0 + s].flags = nd.flags | kLvSplit;
                }
            }
        }
        int total;
        sg.tile0 = s_carry + block_excl_scan<int>(sg.ntiles, s_warp, &total);
        if (s < m) {
            t.segs[s] = sg;
            if (!split) t.counts[s] = {-1, -1};
            else if (sg.ntiles == 0) t.counts[s] = {0, 0};
        }
        // tile -> slot table: the lanes of a warp fill the tiles of the warp's 32 slots, one slot after the other
        for (int j = 0; j < 32; j++) {
            const int t0 = __shfl_sync(0xffffffffu, sg.tile0, j), nt = __shfl_sync(0xffffffffu, sg.ntiles, j);
            for (int q = lane; q < nt; q += 32) t.tile_seg[t0 + q] = s0 + warp * 32 + j;
        }
        __syncthreads();
        if (threadIdx.x == 0) s_carry += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *t.n_tiles = s_carry;
}

}  // namespace ytk
