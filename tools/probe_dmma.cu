// probe_dmma.cu — measures on the device the issue rate of FP64 FMA (DFMA), of the FP64 tensor
// instruction mma.sync.m8n8k4.f64 (DMMA) and of the two mixed, plus the L1-hit bandwidth of
// read-only global loads against shared-memory loads.  Design input for csrc/bx_prom.cu.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/probe_dmma tools/probe_dmma.cu && /tmp/probe_dmma
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int MODE, int TILES>
__global__ void __launch_bounds__(1024) rate(double *out, int iters, double seed) {
    double c[TILES][2], f[8];
    const double a = seed + threadIdx.x * 1e-9, b = seed * 0.5;
#pragma unroll
    for (int t = 0; t < TILES; ++t) { c[t][0] = t; c[t][1] = -t; }
#pragma unroll
    for (int t = 0; t < 8; ++t) f[t] = t * seed;
    for (int i = 0; i < iters; ++i) {
        if (MODE == 0 || MODE == 2) {
#pragma unroll
            for (int t = 0; t < TILES; ++t) dmma884(c[t][0], c[t][1], a, b);
        }
        if (MODE == 1 || MODE == 2) {
#pragma unroll
            for (int t = 0; t < 8; ++t) f[t] = fma(f[t], a, b);
        }
    }
    double s = 0;
#pragma unroll
    for (int t = 0; t < TILES; ++t) s += c[t][0] + c[t][1];
#pragma unroll
    for (int t = 0; t < 8; ++t) s += f[t];
    if (s == 123.456) out[0] = s;
}

// bandwidth: every warp streams the same 112 KB table (L1-resident) with 8-byte loads
template <bool SHARED>
__global__ void __launch_bounds__(512) bw(const double *__restrict__ tab, int n, int reps, double *out) {
    extern __shared__ double sm[];
    if (SHARED) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) sm[i] = tab[i];
        __syncthreads();
    }
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    for (int r = 0; r < reps; ++r) {
        for (int i = threadIdx.x; i + 3 * blockDim.x < n; i += 4 * blockDim.x) {
            if (SHARED) {
                s0 += sm[i]; s1 += sm[i + blockDim.x]; s2 += sm[i + 2 * blockDim.x]; s3 += sm[i + 3 * blockDim.x];
            } else {
                s0 += __ldg(tab + i); s1 += __ldg(tab + i + blockDim.x);
                s2 += __ldg(tab + i + 2 * blockDim.x); s3 += __ldg(tab + i + 3 * blockDim.x);
            }
        }
    }
    if (s0 + s1 + s2 + s3 == 123.456) out[0] = s0;
}

template <typename F> static float timeit(F f) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

int main() {
    double *d; cudaMalloc(&d, 1 << 20);
    cudaMemset(d, 0, 1 << 20);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    printf("SMs %d, clock %.0f MHz\n", sms, khz / 1e3);
    const int iters = 20000;
    for (int warps : {4, 8, 16, 32}) {
        const int th = warps * 32;
        float t0 = timeit([&] { rate<0, 10><<<sms, th>>>(d, iters, 1.0); });
        float t1 = timeit([&] { rate<1, 10><<<sms, th>>>(d, iters, 1.0); });
        float t2 = timeit([&] { rate<2, 10><<<sms, th>>>(d, iters, 1.0); });
        const double dmma_fma = (double)sms * warps * iters * 10 * 256;  // FMAs in the DMMA stream
        const double dfma_fma = (double)sms * warps * iters * 8 * 32;
        printf("warps/SM %2d: DMMA only %.2f TFLOP/s (%.1f cyc/inst/SMSP) | DFMA only %.2f TFLOP/s | mixed: %.3f ms vs %.3f + %.3f -> DMMA %.2f + DFMA %.2f TFLOP/s\n",
               warps, 2 * dmma_fma / t0 * 1e-9, t0 * 1e-3 * khz * 1e3 / (iters * 10.0 * warps / 4.0),
               2 * dfma_fma / t1 * 1e-9, t2, t0, t1, 2 * dmma_fma / t2 * 1e-9, 2 * dfma_fma / t2 * 1e-9);
    }
    {
        float t0 = timeit([&] { rate<0, 2><<<sms, 128>>>(d, iters, 1.0); });
        printf("DMMA dependent-chain latency estimate (2 tiles, 1 warp/SMSP): %.1f cycles per DMMA pair\n",
               t0 * 1e-3 * khz * 1e3 / iters);
    }
    const int n = 14336;  // 512 x 28 doubles
    const int reps = 2000;
    cudaFuncSetAttribute(bw<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, n * 8);
    for (int th : {128, 256, 512}) {
        float ts = timeit([&] { bw<true><<<sms, th, n * 8>>>(d, n, reps, d + 20000); });
        float tg = timeit([&] { bw<false><<<sms, th>>>(d, n, reps, d + 20000); });
        const double bytes = (double)sms * reps * n * 8;
        printf("threads %3d: smem %.1f B/clk/SM, L1-hit ldg %.1f B/clk/SM\n", th,
               bytes / (ts * 1e-3) / sms / (khz * 1e3), bytes / (tg * 1e-3) / sms / (khz * 1e3));
    }
    return cudaGetLastError() != cudaSuccess;
}
