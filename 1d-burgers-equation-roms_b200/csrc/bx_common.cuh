// bx_common.cuh — shared device helpers and host-side error plumbing for libburgers_b200.so
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdarg>

#include "../../include/burgers_b200.h"

namespace bx {

// ---- host: last-error text ------------------------------------------------------------
void set_error(const char *fmt, ...);

#define BX_REQUIRE(cond, ...)                                   \
    do {                                                        \
        if (!(cond)) {                                          \
            ::bx::set_error(__VA_ARGS__);                       \
            return BX_ERR_INVALID;                              \
        }                                                       \
    } while (0)

#define BX_CUDA(call)                                                                  \
    do {                                                                               \
        cudaError_t _e = (call);                                                       \
        if (_e != cudaSuccess) {                                                       \
            ::bx::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e),    \
                            __FILE__, __LINE__);                                       \
            return BX_ERR_CUDA;                                                        \
        }                                                                              \
    } while (0)

#define BX_LAUNCH_CHECK()                                                              \
    do {                                                                               \
        cudaError_t _e = cudaGetLastError();                                           \
        if (_e != cudaSuccess) {                                                       \
            ::bx::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e),\
                            __FILE__, __LINE__);                                       \
            return BX_ERR_CUDA;                                                        \
        }                                                                              \
    } while (0)

// ---- device -----------------------------------------------------------------------------
constexpr unsigned FULL = 0xffffffffu;

// Shape functions at the Gauss points xi = -+1/sqrt(3) (FEM/fem_burgers.py:319-321)
#define BX_GA 0.78867513459481288225   /* (1 + 1/sqrt 3)/2 */
#define BX_GB 0.21132486540518711775   /* (1 - 1/sqrt 3)/2 */

__device__ __forceinline__ double shfl_up(double v, int d) { return __shfl_up_sync(FULL, v, d); }
__device__ __forceinline__ double shfl_down(double v, int d) { return __shfl_down_sync(FULL, v, d); }
__device__ __forceinline__ double shfl_xor(double v, int d) { return __shfl_xor_sync(FULL, v, d); }
__device__ __forceinline__ double shfl_idx(double v, int l) { return __shfl_sync(FULL, v, l); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) v += shfl_xor(v, s);
    return v;
}

// Reciprocal, correctly rounded (MUFU.RCP64H + Newton + fix-up).
__device__ __forceinline__ double rcp(double x) { return __drcp_rn(x); }

// PCR over the 32 lanes of a warp for a tridiagonal system whose rows are already scaled to
// unit diagonal:   al * x[lane-1] + x[lane] + ga * x[lane+1] = rho.   Returns x[lane].
// Rows outside [0, 31] are treated as identity rows (al = ga = rho = 0).
__device__ __forceinline__ double warp_pcr_unit_diag(double al, double ga, double rho, int lane) {
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
        double al_m = shfl_up(al, s), ga_m = shfl_up(ga, s), rho_m = shfl_up(rho, s);
        double al_p = shfl_down(al, s), ga_p = shfl_down(ga, s), rho_p = shfl_down(rho, s);
        if (lane < s) { al_m = 0.0; ga_m = 0.0; rho_m = 0.0; }
        if (lane + s > 31) { al_p = 0.0; ga_p = 0.0; rho_p = 0.0; }
        double be = 1.0 - al * ga_m - ga * al_p;
        double rn = rho - al * rho_m - ga * rho_p;
        double an = -al * al_m;
        double gn = -ga * ga_p;
        double inv = rcp(be);
        al = an * inv; ga = gn * inv; rho = rn * inv;
    }
    return rho;
}

}  // namespace bx
