// Row partition (stable and single-pass), its device-side planner, leaf sums, per-row node ids.
// Part of the CUDA side of libytkgbdt.so (sm_100a).  Nothing here is a dense contraction: the work is
// HBM-/shared-memory-bound byte and integer traffic, so there are no tensor cores; what matters is
// sector-aligned row reads, conflict-free shared-memory atomics and grids sized to the 148 SMs.
// Reference semantics restated by each kernel are cited as J/... =
// /root/reference/src/main/java/com/fenbi/ytklearn/...
#include "common.cuh"
#include "device_util.cuh"
#include "level.cuh"

#include <math_constants.h>

#include <algorithm>
#include <cstdlib>

#ifndef YTK_PART_MIN_BLOCKS
#define YTK_PART_MIN_BLOCKS 2
#endif
namespace ytk {

// ------------------------------------------------------------------------------------
// K6: row partition.  SamplePositionData.resetPosition
//     (J/data/gbdt/SamplePositionData.java:115-133): row goes left iff
//     (float)slot(row, fid) < cond.  The reference then swaps in place (unstable); only
//     the left/right sets are contractual, so this is a stable out-of-place scatter over
//     every node of the batch at once: flags + per-tile counts, a per-node scan of the
//     tile counts, then the scatter into the other row-id buffer.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kPartThreads)
partition_count_k(const uint8_t* __restrict__ src, int32_t S8, int64_t col_stride, WideCols wide, SparseRows sparse,
                  const int32_t* __restrict__ rowidx0, const int32_t* __restrict__ rowidx1, const SplitSeg* __restrict__ segs,
                  const int* __restrict__ tile_seg, uint8_t* __restrict__ flags, int32_t* __restrict__ block_cnt) {
    __shared__ int s_cnt[kPartThreads / 32];
    const SplitSeg sg = segs[tile_seg[blockIdx.x]];
    const int32_t* __restrict__ rowidx = seg_src(sg, rowidx0, rowidx1);
    const long long base = sg.begin + (long long)(blockIdx.x - sg.tile0) * kPartTile;
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < kPartItems; k++) {
        long long i = base + k * kPartThreads + threadIdx.x;
        if (i < sg.end) {
            int32_t rid = rowidx ? rowidx[i] : (int32_t)i;
            bool left;
            if (sg.fid < 0) left = src[rid] != 0;
            else if (sparse.row_ptr) left = (float)sparse_slot(sparse, rid, sg.fid) < sg.cond;
            else if (sg.wide > 0) left = (float)wide.data[(sg.wide - 1) * wide.stride + rid] < sg.cond;   // uint16 column
            else if (col_stride > 0) left = (float)src[sg.fid * col_stride + rid] < sg.cond;   // feature-major copy
            else left = (float)src[(long long)rid * S8 + sg.fid] < sg.cond;
            flags[i] = left ? 1 : 0;
            cnt += left ? 1 : 0;
        }
    }
    for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
    if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < kPartThreads / 32; w++) t += s_cnt[w];
        block_cnt[blockIdx.x] = t;
    }
}

// one CTA per segment: exclusive scan of its tiles' counts, total -> counts[seg]
__global__ void __launch_bounds__(1024)
partition_scan_k(const SplitSeg* __restrict__ segs, int32_t* __restrict__ block_cnt, PartitionCounts* counts) {
    __shared__ long long s_warp[32];
    __shared__ long long s_carry;
    const SplitSeg sg = segs[blockIdx.x];
    int32_t* bc = block_cnt + sg.tile0;
    const int nblocks = sg.ntiles;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int c0 = 0; c0 < nblocks; c0 += 1024) {
        int i = c0 + threadIdx.x;
        long long v = (i < nblocks) ? bc[i] : 0;
        long long p = v;
        for (int d = 1; d < 32; d <<= 1) {
            long long t = __shfl_up_sync(0xffffffffu, p, d);
            if (lane >= d) p += t;
        }
        if (lane == 31) s_warp[warp] = p;
        __syncthreads();
        if (warp == 0) {
            long long w = s_warp[lane];
            long long q = w;
            for (int d = 1; d < 32; d <<= 1) {
                long long t = __shfl_up_sync(0xffffffffu, q, d);
                if (lane >= d) q += t;
            }
            s_warp[lane] = q - w;  // exclusive
        }
        __syncthreads();
        long long excl = s_carry + s_warp[warp] + p - v;
        if (i < nblocks) bc[i] = (int32_t)excl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        counts[blockIdx.x].left = s_carry;
        counts[blockIdx.x].right = (sg.end - sg.begin) - s_carry;
    }
}

__global__ void __launch_bounds__(kPartThreads)
partition_scatter_k(int32_t* __restrict__ rowidx0, int32_t* __restrict__ rowidx1,
                    const SplitSeg* __restrict__ segs, const int* __restrict__ tile_seg,
                    const uint8_t* __restrict__ flags, const int32_t* __restrict__ block_off,
                    const PartitionCounts* __restrict__ counts) {
    __shared__ int s_warp[kPartThreads / 32];
    __shared__ int32_t s_rows[kPartTile];      // the tile, left rows first, both sides in input order
    const int seg = tile_seg[blockIdx.x];
    const SplitSeg sg = segs[seg];
    const int32_t* __restrict__ rowidx = seg_src(sg, rowidx0, rowidx1);
    int32_t* __restrict__ out = (sg.buf == 0) ? rowidx1 : rowidx0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long tile_begin = sg.begin + (long long)(blockIdx.x - sg.tile0) * kPartTile;
    const int tile_n = (int)min((long long)kPartTile, sg.end - tile_begin);
    // thread t owns items [t*8, t*8+8) of the tile so that order is preserved
    const int first = threadIdx.x * kPartItems;
    int32_t rid[kPartItems];
    bool fl[kPartItems];
    int cnt = 0;
    if (first + kPartItems <= tile_n && rowidx &&
        ((reinterpret_cast<unsigned long long>(rowidx + tile_begin + first) & 15ull) == 0) &&
        ((reinterpret_cast<unsigned long long>(flags + tile_begin + first) & 7ull) == 0)) {
        // full, aligned run: two 128-bit loads of row ids and one 64-bit load of flags
        const int4* rp = reinterpret_cast<const int4*>(rowidx + tile_begin + first);
        int4 r0 = rp[0], r1 = rp[1];
        rid[0] = r0.x; rid[1] = r0.y; rid[2] = r0.z; rid[3] = r0.w;
        rid[4] = r1.x; rid[5] = r1.y; rid[6] = r1.z; rid[7] = r1.w;
        unsigned long long f8 = *reinterpret_cast<const unsigned long long*>(flags + tile_begin + first);
#pragma unroll
        for (int k = 0; k < kPartItems; k++) { fl[k] = (f8 >> (8 * k)) & 1ull; cnt += fl[k] ? 1 : 0; }
    } else {
#pragma unroll
        for (int k = 0; k < kPartItems; k++) {
            long long i = tile_begin + first + k;
            fl[k] = false;
            rid[k] = 0;
            if (first + k < tile_n) {
                rid[k] = rowidx ? rowidx[i] : (int32_t)i;
                fl[k] = flags[i] != 0;
                cnt += fl[k] ? 1 : 0;
            }
        }
    }
    int p = cnt;
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, p, d);
        if (lane >= d) p += t;
    }
    if (lane == 31) s_warp[warp] = p;
    __syncthreads();
    int woff = 0, tile_left = 0;
    for (int w = 0; w < kPartThreads / 32; w++) {
        if (w < warp) woff += s_warp[w];
        tile_left += s_warp[w];
    }
    int lpos = woff + p - cnt;                           // lefts of this tile before my first item
    int rpos = tile_left + (first - lpos);               // rights go after all lefts of the tile
#pragma unroll
    for (int k = 0; k < kPartItems; k++) {
        if (first + k < tile_n) {
            if (fl[k]) s_rows[lpos++] = rid[k];
            else s_rows[rpos++] = rid[k];
        }
    }
    __syncthreads();
    // coalesced write-out of the two runs
    const long long left_total = counts[seg].left;
    const long long lefts_before_tile = block_off[blockIdx.x];
    const long long ldst = sg.begin + lefts_before_tile;
    const long long rdst = sg.begin + left_total + ((tile_begin - sg.begin) - lefts_before_tile);
    for (int i = threadIdx.x; i < tile_n; i += kPartThreads) {
        if (i < tile_left) out[ldst + i] = s_rows[i];
        else out[rdst + (i - tile_left)] = s_rows[i];
    }
}

// K7: exact sums of the fixed-point (g,h) over a node's rows -- computeLeafStats
//     (J/optimizer/gbdt/DataParallelTreeMaker.java:352-372).
__global__ void leaf_sums_k(const int2* __restrict__ gq, const int32_t* __restrict__ rowidx0,
                            const int32_t* __restrict__ rowidx1, const RowItem* __restrict__ items,
                            long long* __restrict__ out, const PartitionCounts* __restrict__ cnt_local,
                            const PartitionCounts* __restrict__ cnt_global) {
    pdl_sync();
    const RowItem it = resolve_item(items[blockIdx.x], cnt_local, cnt_global);
    const int32_t* __restrict__ rowidx = it.buf < 0 ? nullptr : (it.buf == 0 ? rowidx0 : rowidx1);
    long long g = 0, h = 0;
    for (long long i = it.begin + threadIdx.x; i < it.end; i += blockDim.x) {
        int2 q = ld_gather_int2(gq + (rowidx ? rowidx[i] : (int32_t)i));
        g += q.x; h += q.y;
    }
    for (int d = 16; d > 0; d >>= 1) {
        g += __shfl_xor_sync(0xffffffffu, g, d);
        h += __shfl_xor_sync(0xffffffffu, h, d);
    }
    if ((threadIdx.x & 31) == 0) {
        if (g) atomicAdd(reinterpret_cast<unsigned long long*>(out + 2 * it.seg), (unsigned long long)g);
        if (h) atomicAdd(reinterpret_cast<unsigned long long*>(out + 2 * it.seg + 1), (unsigned long long)h);
    }
}

__global__ void fill_i32_k(int32_t* p, int64_t n, int32_t v) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = v;
}
__global__ void fill_positions_k(const int32_t* __restrict__ rowidx0, const int32_t* __restrict__ rowidx1,
                                 const PosItem* __restrict__ items, int32_t* __restrict__ node_of_row) {
    const PosItem it = items[blockIdx.y];
    const int32_t* __restrict__ rowidx = it.buf < 0 ? nullptr : (it.buf == 0 ? rowidx0 : rowidx1);
    for (long long i = it.begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < it.end;
         i += (long long)gridDim.x * blockDim.x)
        node_of_row[rowidx ? rowidx[i] : (int32_t)i] = it.nid;
}

// Single-pass variant used inside make(): one kernel reads every row id once, gathers its slot,
// learns where the tile goes from a chained scan over the tiles of its node (decoupled look-back)
// and writes left rows forward from the node's begin, right rows BACKWARD from its end.  Both
// sides stay monotonic in input order (the right side reversed), which is all the gathers of the
// next level care about; the child intervals are the same as the stable version's.
//   tile_state[t] = epoch << 34 | status << 32 | count;  status 1: the tile's own left count,
//   2: left count of the node's tiles up to and including t.  Entries of another epoch are "not
//   yet written", so the array is never cleared.  Tiles are claimed in ticket order, hence every
//   tile a CTA waits for is already running.
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// kSrc: where a row's slot comes from -- 0: uint8 slots only, 1: some features have uint16 columns, 2: sparse rows.
// (One kernel per kind: the sparse look-up would cost the common case its registers.)
template <int kSrc>
__global__ void __launch_bounds__(kFusedThreads, YTK_PART_MIN_BLOCKS)
partition_fused_k(const uint8_t* __restrict__ src, int32_t S8, int64_t col_stride, WideCols wide, SparseRows sparse,
                  int32_t* __restrict__ rowidx0,
                  int32_t* __restrict__ rowidx1, const SplitSeg* __restrict__ segs, int n_segs,
                  const int* __restrict__ tile_seg, int n_tiles, const int* __restrict__ n_tiles_dev,
                  unsigned long long* __restrict__ tile_state, unsigned int* __restrict__ ticket, unsigned long long epoch,
                  PartitionCounts* __restrict__ counts, const unsigned long long* __restrict__ epoch_dev,
                  unsigned long long* __restrict__ stamp) {
    pdl_sync();
    if (epoch_dev) epoch = *epoch_dev;       // kept on the device (a launch recorded in a CUDA graph cannot take a new value)
    __shared__ int s_tile, s_prefix, s_found, s_tile_left, s_total;
    __shared__ int s_lb_sum[kFusedThreads / 32], s_lb_fp[kFusedThreads / 32];
    __shared__ int s_cnt[kFusedItems * (kFusedThreads / 32)];
    __shared__ int32_t s_rows[kFusedTile];      // the tile, left rows first, both sides in input order
    // tables planned on the device (plan_next_partition_k): the grid is an upper bound, the real
    // number of tiles is read here, and the planner has reset the ticket and the empty nodes' counts
    if (n_tiles_dev) n_tiles = *n_tiles_dev;
    if (threadIdx.x == 0) {
        s_total = 0;
        s_tile = n_tiles_dev ? (int)atomicAdd(ticket, 1u)
                             : (int)atomicInc(ticket, (unsigned)n_tiles - 1u);        // back to 0 after the last tile
    }
    __syncthreads();
    const int tile = s_tile;
    if (tile >= n_tiles) return;
    stamp_begin(stamp);
    if (tile == 0 && !n_tiles_dev)
        for (int k = threadIdx.x; k < n_segs; k += kFusedThreads)
            if (segs[k].ntiles == 0) counts[k] = {0, 0};
    const int seg = tile_seg[tile];
    const SplitSeg sg = segs[seg];
    const int32_t* __restrict__ rowidx = seg_src(sg, rowidx0, rowidx1);
    int32_t* __restrict__ out = (sg.buf == 0) ? rowidx1 : rowidx0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int j = tile - sg.tile0;
    const long long tile_begin = sg.begin + (long long)j * kFusedTile;
    const int tile_n = (int)min((long long)kFusedTile, sg.end - tile_begin);
    // Item i of the tile = k * kFusedThreads + thread: neighbouring lanes hold neighbouring rows, so that where the
    // row-id list is (nearly) contiguous -- the upper levels -- a warp's 32 one-byte gathers fall into one or two
    // sectors instead of sixteen (the gathers' L1 tag cycles were half of this kernel's time there).
    int32_t rid[kFusedItems];
#pragma unroll
    for (int k = 0; k < kFusedItems; k++) {
        const int i = k * kFusedThreads + threadIdx.x;
        rid[k] = i < tile_n ? (rowidx ? rowidx[tile_begin + i] : (int32_t)(tile_begin + i)) : 0;   // (never read past the last row)
    }
    // one addressing form for the three sources: a byte mask over rows (left iff non-zero, i.e. NOT
    // below 0.5), the feature-major copy, row-major slots
    const bool mask = sg.fid < 0;
    const uint8_t* __restrict__ base = mask ? src : (col_stride > 0 ? src + (long long)sg.fid * col_stride : src + sg.fid);
    const int mul = (mask || col_stride > 0) ? 1 : S8;
    const float cond = mask ? 0.5f : sg.cond;
    unsigned short slot[kFusedItems];
    if (kSrc == 2 && !mask) {                            // sparse rows: look the feature up in the row
#pragma unroll
        for (int k = 0; k < kFusedItems; k++) slot[k] = (unsigned short)sparse_slot(sparse, rid[k], sg.fid);
    } else if (kSrc == 1 && sg.wide > 0) {               // a feature with more than 256 slots: its uint16 column
        const uint16_t* __restrict__ wb = wide.data + (long long)(sg.wide - 1) * wide.stride;
#pragma unroll
        for (int k = 0; k < kFusedItems; k++) slot[k] = (unsigned short)ld_gather_u16(wb + rid[k]);
    } else {
#pragma unroll
        for (int k = 0; k < kFusedItems; k++) slot[k] = (unsigned short)ld_gather_u8(base + (long long)rid[k] * mul);     // independent gathers in flight
    }
    unsigned fl = 0;                                     // bit k: item k goes left
#pragma unroll
    for (int k = 0; k < kFusedItems; k++)
        fl |= (unsigned)((k * kFusedThreads + (int)threadIdx.x < tile_n) && (((float)slot[k] < cond) != mask)) << k;
    // lefts per (k, warp), in item order; their exclusive scan gives every item the number of lefts before it
    int warp_left = 0;
#pragma unroll
    for (int k = 0; k < kFusedItems; k++) {
        const unsigned m = __ballot_sync(0xffffffffu, (fl >> k) & 1u);
        if (lane == 0) s_cnt[k * (kFusedThreads / 32) + warp] = __popc(m);
        warp_left += __popc(m);
    }
    if (lane == 0) atomicAdd(&s_total, warp_left);
    __syncthreads();
    // the tile's own left count goes out as early as it is known (the tiles behind this one wait for it), by a warp
    // that is not busy with the scan of the counts
    const unsigned long long tag = epoch << 34;
    if (threadIdx.x == kFusedThreads - 32)
        st_volatile_u64(tile_state + tile, tag | ((j == 0 ? 2ull : 1ull) << 32) | (unsigned)s_total);
    if (warp == 0) {
        constexpr int kPerLane = kFusedItems * (kFusedThreads / 32) / 32;
        int v[kPerLane], sum = 0;
#pragma unroll
        for (int q = 0; q < kPerLane; q++) { v[q] = s_cnt[lane * kPerLane + q]; sum += v[q]; }
        int p = sum;
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, p, d);
            if (lane >= d) p += t;
        }
        int excl = p - sum;
#pragma unroll
        for (int q = 0; q < kPerLane; q++) { s_cnt[lane * kPerLane + q] = excl; excl += v[q]; }
        if (lane == 31) s_tile_left = p;
    }
    __syncthreads();
    const int tile_left = s_tile_left;
    // the rows take their places inside the tile (left rows first) while the tiles in front are still counting: ranks
    // inside a tile need nothing from outside it
#pragma unroll
    for (int k = 0; k < kFusedItems; k++) {
        const int i = k * kFusedThreads + threadIdx.x;
        const bool l = (fl >> k) & 1u;
        const unsigned m = __ballot_sync(0xffffffffu, l);
        const int before = s_cnt[k * (kFusedThreads / 32) + warp] + __popc(m & ((1u << lane) - 1u));   // lefts before item i
        if (i < tile_n) s_rows[l ? before : tile_left + (i - before)] = rid[k];           // rights go after all lefts of the tile
    }
    // ---- chained scan: publish this tile's left count, then add up the tiles before it back to
    // the nearest one that already knows its prefix.  Every warp inspects its own window of 32
    // predecessors, so one round trip to L2 covers 512 tiles.
    int excl = 0;                                        // thread 0's
    if (j > 0) {                                         // (uniform over the CTA)
        int look = tile - 1;
        for (;;) {
            const int t = look - warp * 32 - lane;
            unsigned long long v;
            int fp;
            for (;;) {
                v = tag | (2ull << 32);                  // before the node's first tile: a virtual "prefix 0"
                if (t >= sg.tile0) v = ld_volatile_u64(tile_state + t);
                const int status = (v >> 34) == epoch ? (int)((v >> 32) & 3ull) : 0;
                const unsigned not_ready = __ballot_sync(0xffffffffu, status == 0);
                const unsigned is_prefix = __ballot_sync(0xffffffffu, status == 2);
                fp = is_prefix ? __ffs(is_prefix) - 1 : 32;                    // nearest tile of the window that knows its prefix
                const unsigned need = fp >= 31 ? 0xffffffffu : ((2u << fp) - 1u);
                if (!(not_ready & need)) break;                                 // else: a tile in between has not published yet
            }
            int c = (lane <= fp) ? (int)(unsigned)v : 0;
            for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
            if (lane == 0) { s_lb_sum[warp] = c; s_lb_fp[warp] = fp; }
            __syncthreads();
            if (threadIdx.x == 0) {
                int found = 0;
                for (int w = 0; w < kFusedThreads / 32 && !found; w++) {
                    excl += s_lb_sum[w];
                    found = s_lb_fp[w] < 32;
                }
                s_found = found;
            }
            __syncthreads();
            if (s_found) break;
            look -= kFusedThreads;
        }
    }
    if (threadIdx.x == 0) {
        if (j > 0) st_volatile_u64(tile_state + tile, tag | (2ull << 32) | (unsigned)(excl + tile_left));
        s_prefix = excl;
        if (j == sg.ntiles - 1) {
            const long long left = (long long)excl + tile_left;
            counts[seg] = {left, (sg.end - sg.begin) - left};
        }
    }
    __syncthreads();
    const long long left_base = sg.begin + s_prefix;
    const long long right_last = sg.end - 1 - ((long long)j * kFusedTile - s_prefix) + tile_left;   // minus (i - tile_left) below
    for (int i = threadIdx.x; i < tile_n; i += kFusedThreads)
        out[i < tile_left ? left_base + i : right_last - i] = s_rows[i];
    stamp_end(stamp);
}

// the kernel for the handle's kind of rows
template <typename... Args>
static cudaError_t launch_fused(dim3 grid, cudaStream_t st, const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide,
                                SparseRows sparse, Args... args) {
    if (sparse.row_ptr)
        return launch_chain(partition_fused_k<2>, grid, dim3(kFusedThreads), 0, st, src, S8, col_stride, wide, sparse, args...);
    if (wide.data)
        return launch_chain(partition_fused_k<1>, grid, dim3(kFusedThreads), 0, st, src, S8, col_stride, wide, sparse, args...);
    return launch_chain(partition_fused_k<0>, grid, dim3(kFusedThreads), 0, st, src, S8, col_stride, wide, sparse, args...);
}

cudaError_t launch_partition_fused(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0, int32_t* rowidx1,
                                   const SplitSeg* segs, int n_segs, const int* tile_seg, int n_tiles,
                                   unsigned long long* tile_state, unsigned int* ticket, unsigned long long epoch,
                                   PartitionCounts* counts, cudaStream_t st, SparseRows sparse) {
    if (n_segs <= 0) return cudaSuccess;
    if (n_tiles <= 0) return cudaMemsetAsync(counts, 0, (size_t)n_segs * sizeof(PartitionCounts), st);
    return launch_fused(dim3(n_tiles), st, src, S8, col_stride, wide, sparse,
                        rowidx0, rowidx1, segs, n_segs, tile_seg, n_tiles, (const int*)nullptr, tile_state, ticket, epoch, counts,
                        (const unsigned long long*)nullptr, (unsigned long long*)nullptr);
}

// ------------------------------------------------------------------------------------
// Level-wise growth: the partition of the NEXT level, planned on the device.  What the next level
// splits is decided by the records the split search has just written: a child of this level's job k
// is split iff its lossChg > min_split_loss (the other stop rules are settled by the host before it
// asks for this).  Queued right behind the split search, it lets the next partition run while the
// host is still being woken up, reading the records and building the other tables.
//   slot 2k / 2k+1 = left / right child of job k;  records: recs[2*scan], recs[2*scan+1] = the
//   smaller / larger child of the job with scan index `scan` (-1: its children are leaves).
//   Slots that are not split get counts {-1,-1}: the host checks them against its own decision.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512)
plan_next_partition_k(const SplitSeg* __restrict__ cur_segs, const PartitionCounts* __restrict__ cur_counts,
                      const PartitionCounts* __restrict__ cur_counts_global, const SpecJob* __restrict__ jobs, int m, const ytk_split* recs, double min_split_loss,
                      const int32_t* __restrict__ wide_index, SplitSeg* __restrict__ out_segs,
                      int* __restrict__ out_tile_seg, int* __restrict__ n_tiles_out, unsigned int* __restrict__ ticket,
                      PartitionCounts* __restrict__ next_counts) {
    pdl_sync();
    __shared__ int s_warp[16];
    __shared__ int s_carry;
    const int nslots = 2 * m;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { s_carry = 0; *ticket = 0; }
    __syncthreads();
    for (int s0 = 0; s0 < nslots; s0 += 512) {
        const int s = s0 + threadIdx.x;
        SplitSeg sg = {0, 0, 0, 0, 0.f, 0, 0, 0};
        bool split = false;
        if (s < nslots) {
            const int k = s >> 1, side = s & 1;
            const SpecJob jb = jobs[k];
            if (jb.scan >= 0) {
                const SplitSeg pr = cur_segs[jb.seg];
                const PartitionCounts c = cur_counts[jb.cseg];
                const PartitionCounts cgl = cur_counts_global[jb.cseg];
                const bool small_is_left = cgl.left < cgl.right;     // by global rows (:344-350)
                const bool is_small = (side == 0) == small_is_left;
                // the host copy may be in mapped host memory; volatile: written by the kernel before this one
                const ytk_split r = recs[2 * jb.scan + (is_small ? 0 : 1)];
                if ((double)r.loss_chg > min_split_loss) {
                    split = true;
                    sg.begin = side == 0 ? pr.begin : pr.begin + c.left;
                    sg.end = side == 0 ? pr.begin + c.left : pr.end;
                    sg.buf = jb.child_buf;
                    sg.fid = r.fid;
                    sg.cond = 0.5f * (float)(r.lo + r.hi);            // TreeNode.setSplit (:144-149)
                    sg.wide = wide_index ? wide_index[r.fid] + 1 : 0;
                    sg.ntiles = (int)((sg.end - sg.begin + kFusedTile - 1) / kFusedTile);
                }
            }
        }
        // exclusive scan of ntiles over the slots
        int p = sg.ntiles;
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, p, d);
            if (lane >= d) p += t;
        }
        if (lane == 31) s_warp[warp] = p;
        __syncthreads();
        int woff = 0, total = 0;
        for (int w = 0; w < 16; w++) {
            if (w < warp) woff += s_warp[w];
            total += s_warp[w];
        }
        sg.tile0 = s_carry + woff + p - sg.ntiles;
        if (s < nslots) {
            out_segs[s] = sg;
            // not split: marked; split but without rows on this GPU: counted here, no tile will
            if (!split) next_counts[s] = {-1, -1};
            else if (sg.ntiles == 0) next_counts[s] = {0, 0};
        }
        // tile -> slot table: the lanes of a warp fill the tiles of the warp's 32 slots, one slot after the other
        for (int j = 0; j < 32; j++) {
            const int t0 = __shfl_sync(0xffffffffu, sg.tile0, j), nt = __shfl_sync(0xffffffffu, sg.ntiles, j);
            for (int t = lane; t < nt; t += 32) out_tile_seg[t0 + t] = s0 + warp * 32 + j;
        }
        __syncthreads();
        if (threadIdx.x == 0) s_carry += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_tiles_out = s_carry;
}

// The root's partition, planned from the root's own record: one slot.
__global__ void __launch_bounds__(256)
plan_root_partition_k(long long begin, long long end, int buf, const ytk_split* rec, double min_split_loss,
                      const int32_t* __restrict__ wide_index, SplitSeg* __restrict__ out_segs,
                      int* __restrict__ out_tile_seg, int* __restrict__ n_tiles_out, unsigned int* __restrict__ ticket,
                      PartitionCounts* __restrict__ next_counts) {
    pdl_sync();
    const ytk_split r = rec[0];
    const bool split = (double)r.loss_chg > min_split_loss;
    const int ntiles = split ? (int)((end - begin + kFusedTile - 1) / kFusedTile) : 0;
    if (threadIdx.x == 0) {
        SplitSeg sg = {0, 0, 0, 0, 0.f, 0, 0, 0};
        if (split) sg = {begin, end, buf, r.fid, 0.5f * (float)(r.lo + r.hi), 0, ntiles, wide_index ? wide_index[r.fid] + 1 : 0};
        out_segs[0] = sg;
        if (!split) next_counts[0] = {-1, -1};
        else if (ntiles == 0) next_counts[0] = {0, 0};
        *n_tiles_out = ntiles;
        *ticket = 0;
    }
    for (int t = threadIdx.x; t < ntiles; t += blockDim.x) out_tile_seg[t] = 0;
}

cudaError_t launch_partition_planned_root(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0,
                                          int32_t* rowidx1, long long begin, long long end, int buf, const ytk_split* rec,
                                          double min_split_loss, SplitSeg* out_segs, int* out_tile_seg, int* n_tiles_dev,
                                          int grid_bound, unsigned long long* tile_state, unsigned int* ticket,
                                          unsigned long long epoch, PartitionCounts* next_counts, cudaStream_t st,
                                          SparseRows sparse) {
    cudaError_t e = launch_chain(plan_root_partition_k, dim3(1), dim3(256), 0, st, begin, end, buf, rec, min_split_loss,
                                 wide.index, out_segs, out_tile_seg, n_tiles_dev, ticket, next_counts);
    if (e != cudaSuccess || grid_bound <= 0) return e;
    return launch_fused(dim3(grid_bound), st, src, S8, col_stride, wide, sparse,
                        rowidx0, rowidx1, (const SplitSeg*)out_segs, 1, (const int*)out_tile_seg, 0, (const int*)n_tiles_dev,
                        tile_state, ticket, epoch, next_counts, (const unsigned long long*)nullptr, (unsigned long long*)nullptr);
}

cudaError_t launch_partition_planned(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0,
                                     int32_t* rowidx1, const SplitSeg* cur_segs, const PartitionCounts* cur_counts,
                                     const PartitionCounts* cur_counts_global, const SpecJob* jobs, int m,
                                     const ytk_split* recs, double min_split_loss,
                                     SplitSeg* out_segs, int* out_tile_seg, int* n_tiles_dev, int grid_bound,
                                     unsigned long long* tile_state, unsigned int* ticket, unsigned long long epoch,
                                     PartitionCounts* next_counts, cudaStream_t st, SparseRows sparse) {
    if (m <= 0) return cudaSuccess;
    cudaError_t e = launch_chain(plan_next_partition_k, dim3(1), dim3(512), 0, st, cur_segs, cur_counts, cur_counts_global, jobs, m, recs,
                                 min_split_loss, wide.index, out_segs, out_tile_seg, n_tiles_dev, ticket, next_counts);
    if (e != cudaSuccess || grid_bound <= 0) return e;
    return launch_fused(dim3(grid_bound), st, src, S8, col_stride, wide, sparse,
                        rowidx0, rowidx1, (const SplitSeg*)out_segs, 2 * m, (const int*)out_tile_seg, 0, (const int*)n_tiles_dev,
                        tile_state, ticket, epoch, next_counts, (const unsigned long long*)nullptr, (unsigned long long*)nullptr);
}

// The shared-memory carve-out of the partition is left to the driver unless YTK_GBDT_PART_CARVEOUT (percent) says
// otherwise: its one-byte gathers live on L1 hits (a sector is asked for 16 times), and a maximal carve-out,
// which would allow more CTAs per SM, takes that cache away (measured: 0.64 -> 1.0 ms per tree).
cudaError_t partition_kernels_init() {
    if (const char* e = getenv("YTK_GBDT_PART_CARVEOUT"))
        for (const void* k : {(const void*)partition_fused_k<0>, (const void*)partition_fused_k<1>, (const void*)partition_fused_k<2>}) {
            const cudaError_t err = cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(e));
            if (err != cudaSuccess) return err;
        }
    return cudaSuccess;
}

cudaError_t launch_partition_fused_planned(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0,
                                           int32_t* rowidx1, const SplitSeg* segs, int n_segs, const int* tile_seg,
                                           const int* n_tiles_dev, int grid_bound, unsigned long long* tile_state,
                                           unsigned int* ticket, const unsigned long long* epoch_dev, PartitionCounts* counts,
                                           cudaStream_t st, SparseRows sparse, unsigned long long* stamp) {
    if (grid_bound <= 0) return cudaSuccess;
    return launch_fused(dim3(grid_bound), st, src, S8, col_stride, wide, sparse,
                        rowidx0, rowidx1, segs, n_segs, tile_seg, 0, n_tiles_dev, tile_state, ticket, 0ull, counts, epoch_dev, stamp);
}

cudaError_t launch_partition(const uint8_t* src, int32_t S8, int64_t col_stride, WideCols wide, int32_t* rowidx0, int32_t* rowidx1,
                             const SplitSeg* segs, int n_segs, const int* tile_seg, int n_tiles,
                             uint8_t* flags, int32_t* block_cnt, PartitionCounts* counts, cudaStream_t st, SparseRows sparse) {
    if (n_segs <= 0) return cudaSuccess;
    if (n_tiles > 0)
        partition_count_k<<<n_tiles, kPartThreads, 0, st>>>(src, S8, col_stride, wide, sparse, rowidx0, rowidx1, segs, tile_seg,
                                                            flags, block_cnt);
    partition_scan_k<<<n_segs, 1024, 0, st>>>(segs, block_cnt, counts);
    if (n_tiles > 0)
        partition_scatter_k<<<n_tiles, kPartThreads, 0, st>>>(rowidx0, rowidx1, segs, tile_seg, flags, block_cnt,
                                                              counts);
    return cudaGetLastError();
}

cudaError_t launch_leaf_sums(const int2* gq, const int32_t* rowidx0, const int32_t* rowidx1, const RowItem* items,
                             int n_items, long long* out, const PartitionCounts* cnt_local,
                             const PartitionCounts* cnt_global, cudaStream_t st) {
    if (n_items <= 0) return cudaSuccess;
    return launch_chain(leaf_sums_k, dim3(n_items), dim3(256), 0, st, gq, rowidx0, rowidx1, items, out, cnt_local, cnt_global);
}

cudaError_t launch_fill_i32(int32_t* p, int64_t n, int32_t v, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    fill_i32_k<<<grid_for(n, 256), 256, 0, st>>>(p, n, v);
    return cudaGetLastError();
}
cudaError_t launch_fill_positions(const int32_t* rowidx0, const int32_t* rowidx1, const PosItem* items, int n_items,
                                  int32_t* node_of_row, cudaStream_t st) {
    if (n_items <= 0) return cudaSuccess;
    dim3 grid(64, (unsigned)n_items);
    fill_positions_k<<<grid, 256, 0, st>>>(rowidx0, rowidx1, items, node_of_row);
    return cudaGetLastError();
}

}  // namespace ytk
