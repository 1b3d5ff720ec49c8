// probe_dmma_mix.cu — how well does the FP64 pipe stay busy when DFMA results feed DMMA operands, as in the
// PROM contraction loop (12 DFMA -> 10 DMMA per node step)?  Modes:
//   0: DMMA operands constant (no dependency on the DFMAs), DFMAs in independent chains
//   1: DMMA operands = results of this step's DFMAs (the PROM pattern)
//   2: as 1, but software-pipelined by hand: DMMAs of step s use the DFMA results of step s-1
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/probe_mix tools/probe_dmma_mix.cu && /tmp/probe_mix
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int MODE>
__global__ void __launch_bounds__(1024) mix(double *out, const double *in, int iters) {
    double acc[10][2];
#pragma unroll
    for (int t = 0; t < 10; ++t) { acc[t][0] = t; acc[t][1] = -t; }
    double pm[4], p0[4], pp[4], jq[4] = {1, 2, 3, 4};
#pragma unroll
    for (int b = 0; b < 4; ++b) { pm[b] = in[threadIdx.x + b]; p0[b] = in[threadIdx.x + 4 + b]; }
    const double ra = in[0], rb = in[1], rc = in[2];
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int b = 0; b < 4; ++b) pp[b] = pm[b] * 1.0000001 + rc;   // stands for the Phi row load
        double jp[4];
#pragma unroll
        for (int b = 0; b < 4; ++b) jp[b] = fma(ra, pm[b], fma(rb, p0[b], rc * pp[b]));
        double op[4];
#pragma unroll
        for (int b = 0; b < 4; ++b) op[b] = MODE == 0 ? ra + b : (MODE == 1 ? jp[b] : jq[b]);
        int ti = 0;
#pragma unroll
        for (int I = 0; I < 4; ++I)
#pragma unroll
            for (int J = I; J < 4; ++J) { dmma884(acc[ti][0], acc[ti][1], op[I], op[J]); ++ti; }
#pragma unroll
        for (int b = 0; b < 4; ++b) { pm[b] = p0[b]; p0[b] = pp[b]; jq[b] = jp[b]; }
    }
    double s = 0;
#pragma unroll
    for (int t = 0; t < 10; ++t) s += acc[t][0] + acc[t][1];
#pragma unroll
    for (int b = 0; b < 4; ++b) s += pm[b] + p0[b] + jq[b];
    if (s == 123.456) out[0] = s;
}

__device__ __forceinline__ double vfma(double a, double b, double c) {
    double d; asm volatile("fma.rn.f64 %0, %1, %2, %3;" : "=d"(d) : "d"(a), "d"(b), "d"(c)); return d;
}
__device__ __forceinline__ double vmul(double a, double b) {
    double d; asm volatile("mul.rn.f64 %0, %1, %2;" : "=d"(d) : "d"(a), "d"(b)); return d;
}

// mode 3: DMMAs of step s use the DFMA results of step s-1, and the three dependent DFMA stages of step s are
// placed BETWEEN the DMMAs in program order (all asm volatile, so the order below is the issue order)
__global__ void __launch_bounds__(1024) mix3(double *out, const double *in, int iters) {
    double acc[10][2];
#pragma unroll
    for (int t = 0; t < 10; ++t) { acc[t][0] = t; acc[t][1] = -t; }
    double pm[4], p0[4], pp[4], jq[4] = {1, 2, 3, 4};
#pragma unroll
    for (int b = 0; b < 4; ++b) { pm[b] = in[threadIdx.x + b]; p0[b] = in[threadIdx.x + 4 + b]; pp[b] = in[threadIdx.x + 8 + b]; }
    const double ra = in[0], rb = in[1], rc = in[2];
#pragma unroll 2
    for (int i = 0; i < iters; ++i) {
        double t1[4], t2[4], jp[4], pn[4];
#pragma unroll
        for (int b = 0; b < 4; ++b) t1[b] = vmul(rc, pp[b]);
        dmma884(acc[0][0], acc[0][1], jq[0], jq[0]);
        dmma884(acc[1][0], acc[1][1], jq[0], jq[1]);
        dmma884(acc[2][0], acc[2][1], jq[0], jq[2]);
#pragma unroll
        for (int b = 0; b < 4; ++b) t2[b] = vfma(rb, p0[b], t1[b]);
        dmma884(acc[3][0], acc[3][1], jq[0], jq[3]);
        dmma884(acc[4][0], acc[4][1], jq[1], jq[1]);
        dmma884(acc[5][0], acc[5][1], jq[1], jq[2]);
#pragma unroll
        for (int b = 0; b < 4; ++b) jp[b] = vfma(ra, pm[b], t2[b]);
        dmma884(acc[6][0], acc[6][1], jq[1], jq[3]);
        dmma884(acc[7][0], acc[7][1], jq[2], jq[2]);
#pragma unroll
        for (int b = 0; b < 4; ++b) pn[b] = vfma(pm[b], 1.0000001, rc);   // stands for the next Phi row load
        dmma884(acc[8][0], acc[8][1], jq[2], jq[3]);
        dmma884(acc[9][0], acc[9][1], jq[3], jq[3]);
#pragma unroll
        for (int b = 0; b < 4; ++b) { pm[b] = p0[b]; p0[b] = pp[b]; pp[b] = pn[b]; jq[b] = jp[b]; }
    }
    double s = 0;
#pragma unroll
    for (int t = 0; t < 10; ++t) s += acc[t][0] + acc[t][1];
#pragma unroll
    for (int b = 0; b < 4; ++b) s += pm[b] + p0[b] + jq[b];
    if (s == 123.456) out[0] = s;
}

// mode 4: groups of GS node steps; the DFMA burst of group g+1 (stage by stage: GS*4 independent instructions
// per stage) is issued before the DMMA burst of group g, whose operands were computed one group earlier
template <int GS>
__global__ void __launch_bounds__(384) mix4(double *out, const double *in, int iters) {
    double acc[10][2];
#pragma unroll
    for (int t = 0; t < 10; ++t) { acc[t][0] = t; acc[t][1] = -t; }
    double ph[GS + 2][4];       // Phi rows of the next group (+ the two rows before it)
    double jc[GS][4];           // J fragments of the current group
#pragma unroll
    for (int s = 0; s < GS + 2; ++s)
#pragma unroll
        for (int b = 0; b < 4; ++b) ph[s][b] = in[threadIdx.x + 4 * s + b];
#pragma unroll
    for (int s = 0; s < GS; ++s)
#pragma unroll
        for (int b = 0; b < 4; ++b) jc[s][b] = 1.0 + s + b;
    const double ra = in[0], rb = in[1], rc = in[2];
    for (int i = 0; i < iters; i += GS) {
        double t1[GS][4], jn[GS][4];
#pragma unroll
        for (int s = 0; s < GS; ++s)
#pragma unroll
            for (int b = 0; b < 4; ++b) t1[s][b] = vmul(rc, ph[s + 2][b]);
#pragma unroll
        for (int s = 0; s < GS; ++s)
#pragma unroll
            for (int b = 0; b < 4; ++b) t1[s][b] = vfma(rb, ph[s + 1][b], t1[s][b]);
#pragma unroll
        for (int s = 0; s < GS; ++s)
#pragma unroll
            for (int b = 0; b < 4; ++b) jn[s][b] = vfma(ra, ph[s][b], t1[s][b]);
#pragma unroll
        for (int s = 0; s < GS; ++s) {
            int ti = 0;
#pragma unroll
            for (int I = 0; I < 4; ++I)
#pragma unroll
                for (int J = I; J < 4; ++J) { dmma884(acc[ti][0], acc[ti][1], jc[s][I], jc[s][J]); ++ti; }
        }
        // next group's Phi rows (stand-in for the loads) and rotate
#pragma unroll
        for (int b = 0; b < 4; ++b) { ph[0][b] = ph[GS][b]; ph[1][b] = ph[GS + 1][b]; }
#pragma unroll
        for (int s = 2; s < GS + 2; ++s)
#pragma unroll
            for (int b = 0; b < 4; ++b) ph[s][b] = in[(threadIdx.x + 4 * s + b + i) & 1023];
#pragma unroll
        for (int s = 0; s < GS; ++s)
#pragma unroll
            for (int b = 0; b < 4; ++b) jc[s][b] = jn[s][b];
    }
    double sum = 0;
#pragma unroll
    for (int t = 0; t < 10; ++t) sum += acc[t][0] + acc[t][1];
#pragma unroll
    for (int s = 0; s < GS; ++s)
#pragma unroll
        for (int b = 0; b < 4; ++b) sum += jc[s][b];
    if (sum == 123.456) out[0] = sum;
}

template <typename F> static float timeit(F f) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

int main() {
    double *d; cudaMalloc(&d, 1 << 20); cudaMemset(d, 0, 1 << 20);
    int sms, khz; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const int iters = 20000;
    for (int warps : {4, 8, 12}) {
        const int th = warps * 32;
        float t0 = timeit([&] { mix<0><<<sms, th>>>(d, d + 4096, iters); });
        float t1 = timeit([&] { mix<1><<<sms, th>>>(d, d + 4096, iters); });
        float t2 = timeit([&] { mix<2><<<sms, th>>>(d, d + 4096, iters); });
        float t3 = timeit([&] { mix3<<<sms, th>>>(d, d + 4096, iters); });
        float t4 = timeit([&] { mix4<4><<<sms, th>>>(d, d + 4096, iters); });
        float t8 = th <= 384 ? timeit([&] { mix4<8><<<sms, th>>>(d, d + 4096, iters); }) : 0.f;
        // ideal pipe time per step per SMSP-warp: 10 DMMA * 16.1 + 16 DFMA/DMUL * 2.2 cycles
        const double ideal = (10 * 16.1 + 16 * 2.2) * iters * (warps / 4.0) / (khz * 1e3) * 1e3;
        printf("warps/SM %2d: ideal %.3f ms | independent %.3f (%.0f%%) | dependent %.3f (%.0f%%) | pipelined %.3f (%.0f%%) | interleaved %.3f (%.0f%%) | grouped x4 %.3f (%.0f%%) x8 %.3f (%.0f%%)\n",
               warps, ideal, t0, 100 * ideal / t0, t1, 100 * ideal / t1, t2, 100 * ideal / t2, t3, 100 * ideal / t3,
               t4, 100 * ideal / t4, t8, t8 > 0 ? 100 * ideal / t8 : 0.0);
    }
    return cudaGetLastError() != cudaSuccess;
}
