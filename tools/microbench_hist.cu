// Inner loop of hist_build_k in isolation (no global row loads): what does the shared-memory
// atomic path cost per 32-lane (g,h) update, and which variant of the carry bookkeeping is cheapest?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench_hist tools/microbench_hist.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ROWS_PER_WARP = 32 * 64;   // 64 iterations of 32 rows
constexpr int SMEM = 131072 + 32768;

__device__ __forceinline__ uint32_t hash32(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(uint32_t* out, long long* cyc, int nwarps_active, const uint4* __restrict__ gbuf) {
    extern __shared__ __align__(128) uint32_t sh[];
    for (int i = threadIdx.x; i < 131072 / 4; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int bsel = lane & 3;
    uint32_t selk[4];
#pragma unroll
    for (int kk = 0; kk < 4; kk++) selk[kk] = 0x4404u | ((uint32_t)(kk ^ bsel) << 4);
    uint32_t lane4 = (uint32_t)lane << 2;
    asm volatile("" : "+r"(lane4));
    char* const shb = reinterpret_cast<char*>(sh);
    constexpr uint32_t kHiOff = 65536;
    uint32_t acc = 0;
    long long t0 = clock64();
    if (warp < nwarps_active) {
        for (int it = 0; it < ROWS_PER_WARP / 32; it++) {
            uint32_t w[8];
            uint32_t seed = hash32((blockIdx.x * 1024 + threadIdx.x) * 977u + it);
            if (MODE == 7 || MODE == 8) {
                uint4 a, b;
                if (MODE == 8) {       // coalesced 128-bit global loads of a 1 KB tile per warp-iteration
                    const uint4* src = gbuf + ((size_t)(blockIdx.x * 32 + warp) * (ROWS_PER_WARP / 32) + it) * 64;
                    a = src[lane]; b = src[32 + lane];
                } else {
                    a = make_uint4(hash32(seed), hash32(seed + 1), hash32(seed + 2), hash32(seed + 3));
                    b = make_uint4(hash32(seed + 4), hash32(seed + 5), hash32(seed + 6), hash32(seed + 7));
                }
                char* tile = reinterpret_cast<char*>(sh) + 131072 + warp * 1024;
                __syncwarp();
                *reinterpret_cast<uint4*>(tile + lane * 16) = a;
                *reinterpret_cast<uint4*>(tile + 512 + lane * 16) = b;
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 8; j++) w[j] = *reinterpret_cast<const uint32_t*>(tile + lane * 32 + ((j ^ (lane >> 2)) << 2));
            } else {
#pragma unroll
            for (int j = 0; j < 8; j++) w[j] = hash32(seed + j * 0x9e3779b9u);
            }
            const int32_t gi = (int32_t)(hash32(seed ^ 0x1234567u) >> 8) - (1 << 23);   // signed 24-bit
            const int32_t hi_ = (int32_t)(hash32(seed ^ 0x7654321u) >> 9);              // 23-bit positive
            const uint32_t g = (uint32_t)gi, h = (uint32_t)hi_;
            const unsigned long long g64 = (unsigned long long)(long long)gi, h64 = (unsigned long long)(long long)hi_;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const uint32_t word = w[j];
                uint32_t off[4]; int32_t hg[4], hh[4];
#pragma unroll
                for (int kk = 0; kk < 4; kk++) {
                    off[kk] = (__byte_perm(word, 0u, selk[kk]) | lane4) ^ (uint32_t)((4 * j + kk) << 2);
                    uint32_t* p = reinterpret_cast<uint32_t*>(shb + off[kk]);
                    if (MODE == 2) {                 // no-return atomics only
                        atomicAdd(p, g); atomicAdd(p + 32, h); hg[kk] = hh[kk] = 0;
                    } else if (MODE == 4 || MODE == 7 || MODE == 8) {          // carry-free limbs: 4 REDs
                        atomicAdd(p, g & 0xffffu); atomicAdd((int32_t*)(p + 32), gi >> 16);
                        atomicAdd(p + 16384, h & 0xffffu); atomicAdd((int32_t*)(p + 16384 + 32), hi_ >> 16);
                        hg[kk] = hh[kk] = 0;
                    } else if (MODE == 5) {          // 3 REDs
                        atomicAdd(p, g & 0xffffu); atomicAdd((int32_t*)(p + 32), gi >> 16);
                        atomicAdd(p + 16384, h & 0xffffu);
                        hg[kk] = hh[kk] = 0;
                    } else if (MODE == 6) {          // 1 RED
                        atomicAdd(p, g & 0xffffu);
                        hg[kk] = hh[kk] = 0;
                    } else {
                        const uint32_t og = atomicAdd(p, g);
                        const uint32_t oh = atomicAdd(p + 32, h);
                        if (MODE == 1) { acc |= og | oh; hg[kk] = hh[kk] = 0; }
                        else {
                            hg[kk] = (int32_t)(((unsigned long long)og + g64) >> 32);
                            hh[kk] = (int32_t)(((unsigned long long)oh + h64) >> 32);
                        }
                    }
                }
                if (MODE == 0 || MODE == 3) {
                    if (((hg[0] | hg[1] | hg[2]) | (hg[3] | hh[0] | hh[1]) | (hh[2] | hh[3])) != 0) {
#pragma unroll
                        for (int kk = 0; kk < 4; kk++) {
                            if ((hg[kk] | hh[kk]) != 0) {
                                int32_t* p = reinterpret_cast<int32_t*>(shb + kHiOff + off[kk]);
                                if (hg[kk] != 0) atomicAdd(p, hg[kk]);
                                if (hh[kk] != 0) atomicAdd(p + 32, hh[kk]);
                            }
                        }
                    }
                }
            }
        }
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    out[blockIdx.x * 1024 + threadIdx.x] = acc + sh[threadIdx.x];
}

template <int MODE>
void run(const char* name, uint32_t* out, long long* cyc, int threads, int nwarps, const uint4* gbuf = nullptr) {
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    k<MODE><<<148, threads, SMEM>>>(out, cyc, nwarps, gbuf);
    cudaDeviceSynchronize();
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<MODE><<<148, threads, SMEM>>>(out, cyc, nwarps, gbuf);
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, a, b);
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; i++) avg += h[i]; avg /= 148;
    double rows = (double)nwarps * ROWS_PER_WARP;
    printf("%-58s %2d warps  %7.3f ms  %6.2f SM-cycles per row of 32 features\n", name, nwarps, ms, avg / rows);
}

int main() {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    run<0>("kernel step: 2 returning ATOMS + carry + rare branch", out, cyc, 1024, 32);
    run<1>("2 returning ATOMS, no carry bookkeeping", out, cyc, 1024, 32);
    run<2>("2 no-return ATOMS (RED), no carry bookkeeping", out, cyc, 1024, 32);
    run<0>("kernel step, 16 warps", out, cyc, 1024, 16);
    run<0>("kernel step, 24 warps", out, cyc, 1024, 24);
    run<0>("kernel step, 8 warps", out, cyc, 1024, 8);
    run<2>("RED only, 16 warps", out, cyc, 1024, 16);
    run<4>("4 REDs (16-bit limbs of g and h)", out, cyc, 1024, 32);
    run<5>("3 REDs", out, cyc, 1024, 32);
    run<6>("1 RED", out, cyc, 1024, 32);
    run<4>("4 REDs, 16 warps", out, cyc, 1024, 16);
    uint4* gbuf; size_t gb = (size_t)148 * 32 * (ROWS_PER_WARP / 32) * 64 * sizeof(uint4);
    cudaMalloc(&gbuf, gb); cudaMemset(gbuf, 0x5a, gb);
    run<7>("4 REDs + STS/LDS staging of the row tile", out, cyc, 1024, 32, gbuf);
    run<8>("4 REDs + staging + coalesced 128-bit global loads", out, cyc, 1024, 32, gbuf);
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
