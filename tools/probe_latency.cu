// probe_latency.cu — dependent-issue latencies that bound one Picard iteration of one trajectory (design input for the
// multi-warp FOM / PROM instances): DFMA chain, 64-bit shuffle chain, shared-memory round trip, named barrier of 2 / 4
// warps, MUFU.RCP64H + two Newton steps.  One CTA per SM, timing by clock64 inside the kernel.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/probe_latency tools/probe_latency.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double rcp2(double x) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    return fma(y, e, y);
}

template <int MODE>
__global__ void lat(long long *out, double seed, int iters) {
    __shared__ double sm[256];
    double v = seed + threadIdx.x * 1e-9;
    sm[threadIdx.x] = v;
    __syncthreads();
    const int nth = blockDim.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        if (MODE == 0) v = fma(v, 1.0000001, 1e-9);
        if (MODE == 1) v = __shfl_xor_sync(0xffffffffu, v, 1) + 1e-9;                 // shuffle + dependent DADD
        if (MODE == 2) { sm[threadIdx.x] = v; __syncwarp(); v = sm[threadIdx.x ^ 1] + 1e-9; __syncwarp(); }
        if (MODE == 3) { asm volatile("bar.sync 1, %0;" ::"r"(nth) : "memory"); }
        if (MODE == 4) v = rcp2(v) + 1.5;
        if (MODE == 5) { sm[threadIdx.x] = v; asm volatile("bar.sync 1, %0;" ::"r"(nth) : "memory"); v = sm[(threadIdx.x + 32) % nth] + 1e-9; asm volatile("bar.sync 1, %0;" ::"r"(nth) : "memory"); }
        if (MODE == 6) v = __shfl_xor_sync(0xffffffffu, v, 1);                          // bare 64-bit shuffle chain
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
    if (v == 123.456) out[1] = 1;
}

template <int MODE> static double run(long long *d, int threads, int iters) {
    lat<MODE><<<1, threads>>>(d, 1.0, iters);
    cudaDeviceSynchronize();
    lat<MODE><<<1, threads>>>(d, 1.0, iters);
    long long h = 0;
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    return (double)h / iters;
}

int main() {
    long long *d;
    cudaMalloc(&d, 64);
    const int it = 4096;
    printf("DFMA dependent chain            : %.1f cycles\n", run<0>(d, 32, it));
    printf("SHFL.64 + DADD dependent chain  : %.1f cycles\n", run<1>(d, 32, it));
    printf("bare SHFL.64 dependent chain    : %.1f cycles\n", run<6>(d, 32, it));
    printf("STS + syncwarp + LDS + DADD     : %.1f cycles\n", run<2>(d, 32, it));
    printf("rcp (MUFU + 2 Newton) + DADD    : %.1f cycles\n", run<4>(d, 32, it));
    for (int w : {1, 2, 4, 8}) printf("bar.sync of %d warp(s)           : %.1f cycles\n", w, run<3>(d, 32 * w, it));
    for (int w : {2, 4}) printf("STS + bar + LDS(other warp) + bar, %d warps : %.1f cycles\n", w, run<5>(d, 32 * w, it));
    return cudaGetLastError() != cudaSuccess;
}
