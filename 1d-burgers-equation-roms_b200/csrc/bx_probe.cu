// bx_probe.cu — FP64 peak probes for the roofline denominators of bench.py.
// MEASURED_PEAKS.json carries HBM bandwidth and bf16 tensor throughput only; every hot kernel of this path is
// bound by the FP64 pipe (DFMA) or the FP64 tensor instruction (DMMA, mma.sync.m8n8k4.f64 — tcgen05 has no FP64
// kind), so the denominators are measured in the same run: the caller times this launch with events on `stream`.
#include "bx_common.cuh"

namespace bx {

__device__ __forceinline__ void probe_dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// 16 warps per SM; every thread runs `iters` rounds of 10 independent DMMA tiles (kind 1) or 8 independent
// DFMA chains (kind 0): enough instruction-level parallelism to saturate the pipe (tools/probe_dmma.cu).
template <int KIND>
__global__ void __launch_bounds__(512) fp64_peak_kernel(double *sink, int iters, double seed) {
    double c[10][2], f[8];
    const double a = seed + threadIdx.x * 1e-9, b = seed * 0.5;
#pragma unroll
    for (int t = 0; t < 10; ++t) { c[t][0] = t; c[t][1] = -t; }
#pragma unroll
    for (int t = 0; t < 8; ++t) f[t] = t * seed;
    for (int i = 0; i < iters; ++i) {
        if (KIND == 1) {
#pragma unroll
            for (int t = 0; t < 10; ++t) probe_dmma(c[t][0], c[t][1], a, b);
        } else {
#pragma unroll
            for (int t = 0; t < 8; ++t) f[t] = fma(f[t], a, b);
        }
    }
    double s = 0;
#pragma unroll
    for (int t = 0; t < 10; ++t) s += c[t][0] + c[t][1];
#pragma unroll
    for (int t = 0; t < 8; ++t) s += f[t];
    if (s == 123.456) sink[0] = s;
}

}  // namespace bx

extern "C" long long bx_probe_fp64_peak(int kind, int iters, double *sink, void *stream) {
    using namespace bx;
    if (kind != BX_PROBE_DFMA && kind != BX_PROBE_DMMA) { set_error("bx_probe_fp64_peak: kind must be BX_PROBE_DFMA or BX_PROBE_DMMA"); return BX_ERR_INVALID; }
    if (iters < 1 || !sink) { set_error("bx_probe_fp64_peak: need iters >= 1 and a device sink"); return BX_ERR_INVALID; }
    DeviceInfo di;
    const int rc = device_info(di);
    if (rc != BX_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (kind == BX_PROBE_DMMA) fp64_peak_kernel<1><<<di.sms, 512, 0, st>>>(sink, iters, 1.0);
    else fp64_peak_kernel<0><<<di.sms, 512, 0, st>>>(sink, iters, 1.0);
    if (cudaGetLastError() != cudaSuccess) { set_error("bx_probe_fp64_peak: launch failed"); return BX_ERR_CUDA; }
    // flops of the launch: DMMA m8n8k4 = 256 FMA per warp instruction, DFMA = 32 FMA per warp instruction
    const long long warps = (long long)di.sms * 16;
    return kind == BX_PROBE_DMMA ? 2LL * warps * iters * 10 * 256 : 2LL * warps * iters * 8 * 32;
}
