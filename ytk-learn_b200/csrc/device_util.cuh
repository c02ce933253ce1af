// Small device/launch helpers shared by the kernel files.
#pragma once

#include "common.cuh"

#include <cstdlib>

namespace ytk {

// Programmatic dependent launch for the chain of short kernels that ends every level (zero ->
// histogram -> subtract -> scan -> merge -> plan -> partition): a kernel launched with the attribute
// may start being scheduled while its predecessor in the stream drains; pdl_sync() -- the first
// statement of each of them -- lets the dependents go and then waits until the predecessor has
// completed and its writes are visible, so the kernels see exactly the serial order.
__device__ __forceinline__ void pdl_sync() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
static bool g_pdl = [] { const char* e = getenv("YTK_GBDT_NO_PDL"); return !(e && atoi(e) != 0); }();
template <typename... KArgs, typename... Args>
static cudaError_t launch_chain(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = g_pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// Loads of data reached through a row-id list (a row's slots, its gradient pair, one slot of a column): a miss in L2
// fetches 128 B from HBM by default, of which such a load uses 1..32; `.L2::64B` asks for 64 (SASS LDG.E.LTC64B).
// Measured on the Higgs shape: DRAM bytes of the gathered histogram launches -40 %, their time -12 %
// (cudaLimitMaxL2FetchGranularity, the device-wide hint, changes nothing: profiles/r02_hist_record_experiment.txt).
__device__ __forceinline__ uint4 ld_gather_u4(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ int2 ld_gather_int2(const int2* p) {
    int2 v;
    asm volatile("ld.global.nc.L2::64B.v2.s32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ unsigned ld_gather_u8(const uint8_t* p) {
    unsigned v;
    asm volatile("ld.global.nc.L2::64B.u8 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ float ld_gather_f32_rw(const float* p) {      // (of memory the same kernel writes: not .nc)
    float v;
    asm volatile("ld.global.L2::64B.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ unsigned ld_gather_u16(const uint16_t* p) {
    unsigned short v;
    asm volatile("ld.global.nc.L2::64B.u16 %0, [%1];" : "=h"(v) : "l"(p));
    return v;
}

// A kernel's own account of when it ran: the first CTA's start and the last CTA's end on the GPU's global timer
// (ns), kept as {max of ~start, max of end} so that an all-zero pair means "did not run".  Launches replayed
// from a CUDA graph cannot be bracketed with CUDA events (cudaEventElapsedTime refuses events recorded by a
// graph), so the device-driven level loop times its kernels this way.
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void stamp_begin(unsigned long long* s) {
    if (s && threadIdx.x == 0) atomicMax(s, ~global_ns());
}
__device__ __forceinline__ void stamp_end(unsigned long long* s) {
    if (s && threadIdx.x == 0) atomicMax(s + 1, global_ns());
}

// smaller child = strictly fewer GLOBAL rows on the left, else the right
// (J/optimizer/gbdt/DataParallelTreeMaker.java:344-350, :493-499)
__device__ __forceinline__ RowItem resolve_item(RowItem it, const PartitionCounts* __restrict__ cnt_local,
                                                const PartitionCounts* __restrict__ cnt_global) {
    if (it.cseg < 0) return it;
    const PartitionCounts l = cnt_local[it.cseg], g = cnt_global[it.cseg];
    const bool small_left = g.left < g.right;
    const long long cb = small_left ? it.begin : it.begin + l.left;
    const long long ce = small_left ? it.begin + l.left : it.end;
    const long long b = cb + it.chunk * it.rows_per_chunk;
    it.begin = min(b, ce);
    it.end = min(ce, b + it.rows_per_chunk);
    return it;
}

// slot of feature `fid` in sparse row `rid`: binary search among the row's entries, else the default slot
__device__ __forceinline__ int sparse_slot(const SparseRows& sp, int32_t rid, int fid) {
    long long lo = sp.row_ptr[rid];
    const long long end = sp.row_ptr[rid + 1];
    long long hi = end;
    const unsigned key = (unsigned)fid;
    while (lo < hi) {                                    // entries are sorted by feature
        const long long mid = (lo + hi) >> 1;
        if ((sp.nz[mid] >> 16) < key) lo = mid + 1; else hi = mid;
    }
    if (lo < end) {
        const unsigned e = sp.nz[lo];
        if ((e >> 16) == key) return (int)(e & 0xffffu);
    }
    return sp.default_slot[fid];
}

// the row-id buffer a split segment reads (null: rows begin..end-1 themselves)
__device__ __forceinline__ const int32_t* seg_src(const SplitSeg& s, const int32_t* r0, const int32_t* r1) {
    return s.buf < 0 ? nullptr : (s.buf == 0 ? r0 : r1);
}

static inline int grid_for(int64_t n, int threads, int max_blocks = kNumSMs * 8) {
    int64_t b = (n + threads - 1) / threads;
    if (b < 1) b = 1;
    if (b > max_blocks) b = max_blocks;
    return (int)b;
}

}  // namespace ytk
