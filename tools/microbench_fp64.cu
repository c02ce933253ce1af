// FP64 pipe on this B200: latency of a dependent DFMA chain and throughput with 32 warps/SM, next to FP32.
#include <cstdio>
#include <cuda_runtime.h>
template <typename T>
__global__ void chain(T* out, int iters, long long* cyc) {
    T a = (T)threadIdx.x * (T)1e-3 + (T)1.0001, b = (T)0.9999, c = (T)1e-7;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) a = a * b + c;
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void divchain(double* out, int iters, long long* cyc) {
    double a = threadIdx.x * 1e-3 + 1.0001, b = 1.0000001;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) a = __ddiv_rn(a, b);
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
    double* o; long long* c; cudaMalloc(&o, 148 * 1024 * 8); cudaMalloc(&c, 148 * 8);
    long long h[148]; const int it = 4096;
    auto rep = [&](const char* n, int threads) { cudaDeviceSynchronize(); cudaMemcpy(h, c, sizeof(h), cudaMemcpyDeviceToHost);
        printf("%-34s %4d thr/SM: %7.1f cycles per op (per warp)  -> %6.2f warp-ops/cycle/SM\n", n, threads, (double)h[0] / it, (threads / 32.0) * it / h[0]); };
    chain<double><<<148, 32>>>(o, it, c); rep("DFMA dependent chain", 32);
    chain<double><<<148, 1024>>>(o, it, c); rep("DFMA 32 warps", 1024);
    chain<float><<<148, 32>>>((float*)o, it, c); rep("FFMA dependent chain", 32);
    chain<float><<<148, 1024>>>((float*)o, it, c); rep("FFMA 32 warps", 1024);
    divchain<<<148, 32>>>(o, it, c); rep("DDIV dependent chain", 32);
    divchain<<<148, 1024>>>(o, it, c); rep("DDIV 32 warps", 1024);
    printf("status %s\n", cudaGetErrorString(cudaGetLastError()));
}
