// ytk_gbdt_partition_rows_static: the row partition (K6) for callers that must not allocate.
//
// Every other entry point of this library gets its work space from cudaMalloc.  Some hosts cannot
// afford that: a trainer that has handed the whole device to its own memory pool (a further
// cudaMalloc fails, or is against its rules), or one that splits rows between two steps of a loop
// where the device-wide synchronisation hidden in cudaMalloc/cudaFree is not acceptable.  For them
// this one entry point runs the same three partition kernels with the work space taken from a
// fixed array that is part of the library image.
//
// The price of that: the array exists once per GPU, is shared by every call in the process, and
// has a fixed size, so there is a hard row limit and calls that overlap on one GPU would corrupt
// each other.
#include <mutex>

#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace ytk { void set_last_error(const char* msg); }

namespace {

// ---- the work space: 4 MiB + 17 KiB, resident on every GPU that loads the library, until exit ----
constexpr long long kStaticMaxRows = 1LL << 22;                                   // 4,194,304 rows
constexpr int kStaticMaxTiles = (int)(kStaticMaxRows / ytk::kPartTile);           // 2048 tiles
constexpr size_t kOffFlags = 0;                                                    // one byte per row
constexpr size_t kOffBlockCnt = kOffFlags + (size_t)kStaticMaxRows;                // int per tile (+1)
constexpr size_t kOffTileSeg = kOffBlockCnt + 256 * ((sizeof(int) * (kStaticMaxTiles + 1) + 255) / 256);
constexpr size_t kOffSeg = kOffTileSeg + 256 * ((sizeof(int) * (kStaticMaxTiles + 1) + 255) / 256);
constexpr size_t kOffCounts = kOffSeg + 256;
constexpr size_t kScratchBytes = kOffCounts + 256;

// In host code this name is only a handle for the runtime, not a device address: handing it to a
// kernel would fault (or, where the GPU can reach host memory, quietly use the wrong bytes).  The
// device address comes from cudaGetSymbolAddress, below.
static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

struct DeviceScratch { bool ready = false; unsigned char* base = nullptr; };
DeviceScratch g_dev[64];
std::mutex g_mu;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    ytk::set_last_error(buf);
    return code;
}

__global__ void static_begin_k(ytk::SplitSeg* seg, ytk::SplitSeg v) { *seg = v; }
__global__ void static_end_k(const ytk::PartitionCounts* c, long long* left_out) { *left_out = c->left; }

}  // namespace

extern "C" int ytk_gbdt_partition_rows_static(const uint8_t* d_bins, int32_t stride, const int32_t* d_rows_in,
                                              int64_t n_rows, int32_t fid, float cond, int32_t* d_rows_out,
                                              int64_t* d_left_cnt, void* cuda_stream) {
    if (!d_bins || !d_left_cnt || (n_rows > 0 && (!d_rows_in || !d_rows_out)))
        return fail(YTK_ERR_ARG, "partition_rows_static: null device pointer");
    if (n_rows < 0) return fail(YTK_ERR_ARG, "partition_rows_static: n_rows %lld < 0", (long long)n_rows);
    if (n_rows > kStaticMaxRows)
        return fail(YTK_ERR_CAPACITY,
                    "partition_rows_static: %lld rows exceed the fixed work space of %lld rows; "
                    "use ytk_gbdt_partition_rows (which allocates) or split the list",
                    (long long)n_rows, kStaticMaxRows);
    if (stride < 1 || fid < 0 || fid >= stride)
        return fail(YTK_ERR_ARG, "partition_rows_static: fid %d not in [0, stride %d)", fid, stride);
    // the scatter reads the input list while other CTAs already write the output list
    if (n_rows > 0 && d_rows_in < d_rows_out + n_rows && d_rows_out < d_rows_in + n_rows)
        return fail(YTK_ERR_ARG, "partition_rows_static: the input and output row lists overlap");

    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(YTK_ERR_CUDA, "partition_rows_static: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= 64) return fail(YTK_ERR_ARG, "partition_rows_static: device %d not supported", dev);

    unsigned char* base;
    {
        // Set-up, once per device and with that device current: ask the runtime where this GPU's
        // copy of the array lives.  This first call on a device loads the module, so it belongs
        // before any stream capture the caller may do.
        std::lock_guard<std::mutex> lk(g_mu);
        if (!g_dev[dev].ready) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_static_scratch);
            if (e != cudaSuccess || !p)
                return fail(YTK_ERR_CUDA, "partition_rows_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            g_dev[dev].base = static_cast<unsigned char*>(p);
            g_dev[dev].ready = true;
        }
        base = g_dev[dev].base;
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    uint8_t* flags = base + kOffFlags;
    int32_t* block_cnt = reinterpret_cast<int32_t*>(base + kOffBlockCnt);
    int* tile_seg = reinterpret_cast<int*>(base + kOffTileSeg);
    ytk::SplitSeg* seg = reinterpret_cast<ytk::SplitSeg*>(base + kOffSeg);
    ytk::PartitionCounts* counts = reinterpret_cast<ytk::PartitionCounts*>(base + kOffCounts);

    const int n_tiles = (int)((n_rows + ytk::kPartTile - 1) / ytk::kPartTile);
    // The array still holds whatever the previous call left.  The flags, tile counts and totals are
    // rewritten below before they are read; the tile -> segment table must read all zero (there is
    // one segment), so it is cleared here rather than trusted.
    e = cudaMemsetAsync(tile_seg, 0, sizeof(int) * (size_t)(n_tiles + 1), st);
    if (e != cudaSuccess) return fail(YTK_ERR_CUDA, "partition_rows_static: memset: %s", cudaGetErrorString(e));
    ytk::SplitSeg v;
    v.begin = 0; v.end = n_rows; v.buf = 0; v.fid = fid; v.cond = cond; v.tile0 = 0; v.ntiles = n_tiles; v.wide = 0;
    static_begin_k<<<1, 1, 0, st>>>(seg, v);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(YTK_ERR_CUDA, "partition_rows_static: launch: %s", cudaGetErrorString(e));
    // buffer 0 is the caller's input list (only read when the segment's source is buffer 0),
    // buffer 1 its output list
    e = ytk::launch_partition(d_bins, stride, /*row-major*/ 0, ytk::WideCols{nullptr, 0, nullptr}, const_cast<int32_t*>(d_rows_in), d_rows_out, seg, 1, tile_seg, n_tiles,
                              flags, block_cnt, counts, st);
    if (e != cudaSuccess) return fail(YTK_ERR_CUDA, "partition_rows_static: launch: %s", cudaGetErrorString(e));
    static_end_k<<<1, 1, 0, st>>>(counts, reinterpret_cast<long long*>(d_left_cnt));
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(YTK_ERR_CUDA, "partition_rows_static: launch: %s", cudaGetErrorString(e));
    return YTK_OK;
}
