// probe_rcp.cu — measures on the device the accuracy of the MUFU.RCP64H seed (rcp.approx.ftz.f64)
// and of the Newton variants built on it, against IEEE 1.0/x.  Build & run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/probe_rcp tools/probe_rcp.cu && /tmp/probe_rcp
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>

__device__ double seed(double x) { double y; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x)); return y; }

__global__ void probe(unsigned long long n, double *out) {
    double m0 = 0, m3 = 0, m4 = 0, m5 = 0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        // mantissas uniformly over [1,2), exponents -60..60
        unsigned long long h = i * 0x9E3779B97F4A7C15ull; h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
        double x = 1.0 + (double)(h >> 11) * (1.0 / 9007199254740992.0);
        x = ldexp(x, (int)(i % 121) - 60);
        if (i & 1) x = -x;
        const double ex = 1.0 / x;
        double y = seed(x);
        m0 = fmax(m0, fabs((y - ex) / ex));
        double e = fma(-x, y, 1.0);
        double e2 = fma(e, e, e);
        double y3 = fma(y, e2, y);                       // 3 DFMA (cubic)
        double y4 = fma(y, e, y); double f = fma(-x, y4, 1.0); y4 = fma(y4, f, y4);  // 4 DFMA (2 Newton)
        double g = fma(-x, y3, 1.0); double y5 = fma(y3, g, y3);                      // 5 DFMA
        const double ulp = ldexp(1.0, -52);
        m3 = fmax(m3, fabs((y3 - ex) / ex) / ulp);
        m4 = fmax(m4, fabs((y4 - ex) / ex) / ulp);
        m5 = fmax(m5, fabs((y5 - ex) / ex) / ulp);
    }
    for (int s = 16; s > 0; s >>= 1) {
        m0 = fmax(m0, __shfl_xor_sync(~0u, m0, s)); m3 = fmax(m3, __shfl_xor_sync(~0u, m3, s));
        m4 = fmax(m4, __shfl_xor_sync(~0u, m4, s)); m5 = fmax(m5, __shfl_xor_sync(~0u, m5, s));
    }
    if ((threadIdx.x & 31) == 0) {
        // max of non-negative doubles = max of their bit patterns
        unsigned long long *o = (unsigned long long *)out;
        atomicMax(o + 0, (unsigned long long)__double_as_longlong(m0));
        atomicMax(o + 1, (unsigned long long)__double_as_longlong(m3));
        atomicMax(o + 2, (unsigned long long)__double_as_longlong(m4));
        atomicMax(o + 3, (unsigned long long)__double_as_longlong(m5));
    }
}

int main() {
    double *d; cudaMalloc(&d, 4 * sizeof(double)); cudaMemset(d, 0, 4 * sizeof(double));
    probe<<<148 * 8, 256>>>(1ull << 30, d);
    double h[4]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("rcp.approx.ftz.f64 seed: max rel err %.3e (2^%.1f)\n", h[0], log2(h[0]));
    printf("3-DFMA cubic: %.3f ulp   4-DFMA (2 Newton): %.3f ulp   5-DFMA: %.3f ulp\n", h[1], h[2], h[3]);
    return cudaGetLastError() != cudaSuccess;
}
