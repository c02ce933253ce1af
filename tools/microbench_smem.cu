// Shared-memory update throughput on B200, the quantity that bounds the histogram kernel.
// Each test: 148 CTAs x 1024 threads, every warp issues ITER update instructions on a
// 64 KB shared array.  Prints SM-cycles per warp-instruction (per SM).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench_smem tools/microbench_smem.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ITER = 4096;
constexpr int WORDS = 16384;  // 64 KB

__device__ __forceinline__ uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(uint32_t* out, long long* cyc) {
    extern __shared__ uint32_t sm[];
    for (int i = threadIdx.x; i < WORDS; i += 1024) sm[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    uint32_t s = threadIdx.x * 2654435761u + blockIdx.x;
    uint32_t acc = 0;
    long long t0 = clock64();
#pragma unroll 8
    for (int it = 0; it < ITER; it++) {
        uint32_t r = lcg(s);
        uint32_t row = r & 511;                      // "bin"
        if (MODE == 0) {            // ATOMS.ADD no return, lane -> own bank
            atomicAdd(&sm[row * 32 + lane], r);
        } else if (MODE == 1) {     // ATOMS.ADD with return, lane -> own bank
            acc += atomicAdd(&sm[row * 32 + lane], r);
        } else if (MODE == 2) {     // ATOMS.ADD no return, random bank (lanes = rows)
            atomicAdd(&sm[(r >> 3) & (WORDS - 1)], r);
        } else if (MODE == 3) {     // float atomicAdd (CAS loop), own bank
            atomicAdd(reinterpret_cast<float*>(&sm[row * 32 + lane]), 1.0f);
        } else if (MODE == 4) {     // plain 64-bit RMW, own bank pair (no atomicity)
            uint2* p = reinterpret_cast<uint2*>(sm) + ((row & 255) * 32 + lane);
            uint2 v = *p; v.x += r; v.y += 1; *p = v;
        } else if (MODE == 5) {     // 2 returning atomics + carry logic (the kernel's update)
            uint32_t* p = &sm[((row & 127) * 4) * 32 + lane];
            int32_t g = (int32_t)r - (1 << 23);
            uint32_t old = atomicAdd(p, (uint32_t)g);
            uint32_t sum = old + (uint32_t)g;
            int32_t hi = (g >> 31) + (sum < old ? 1 : 0);
            if (hi) atomicAdd((int32_t*)(p + 32), hi);
            uint32_t o2 = atomicAdd(p + 64, r & 0xffff);
            if (o2 + (r & 0xffff) < o2) atomicAdd(p + 96, 1u);
        } else if (MODE == 6) {     // shuffle
            acc += __shfl_sync(0xffffffffu, r, it & 31);
        } else if (MODE == 7) {     // LDS.32 own bank
            acc += sm[row * 32 + lane];
        } else if (MODE == 8) {     // 2 no-return atomics, own bank (32-bit accumulators)
            uint32_t* p = &sm[((row & 255) * 2) * 32 + lane];
            atomicAdd(p, r); atomicAdd(p + 32, r >> 4);
        } else if (MODE == 9) {     // ATOMS.ADD no return, 2-way conflict (lane/2 bank)
            atomicAdd(&sm[row * 32 + (lane >> 1) * 2], r);
        }
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    out[blockIdx.x * 1024 + threadIdx.x] = acc + sm[threadIdx.x];
}

template <int MODE>
void run(const char* name, uint32_t* out, long long* cyc) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, WORDS * 4);
    k<MODE><<<148, 1024, WORDS * 4>>>(out, cyc);
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    k<MODE><<<148, 1024, WORDS * 4>>>(out, cyc);
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, a, b);
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; i++) avg += h[i]; avg /= 148;
    double per = avg / (32.0 * ITER);   // 32 warps per SM
    printf("%-52s %8.3f ms  %7.2f cyc/warp-instr/SM  (%.2f G lane-updates/s chip)\n", name, ms, per,
           148.0 * 1024 * ITER / (ms * 1e-3) / 1e9);
}

int main() {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    run<0>("ATOMS.ADD u32 no-return, conflict-free", out, cyc);
    run<1>("ATOMS.ADD u32 returning, conflict-free", out, cyc);
    run<2>("ATOMS.ADD u32 no-return, random banks", out, cyc);
    run<9>("ATOMS.ADD u32 no-return, 2-way conflict", out, cyc);
    run<3>("atomicAdd float (CAS loop), conflict-free", out, cyc);
    run<4>("plain LDS.64+STS.64 RMW, conflict-free", out, cyc);
    run<8>("2x ATOMS.ADD no-return (32-bit g,h), conflict-free", out, cyc);
    run<5>("2x returning ATOMS + carry (64-bit g,h), conflict-free", out, cyc);
    run<6>("SHFL", out, cyc);
    run<7>("LDS.32 conflict-free", out, cyc);
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
