// bx_common.cuh — shared device helpers and host-side error plumbing for libburgers_b200.so
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdarg>

#include "../../include/burgers_b200.h"

namespace bx {

// ---- host: last-error text ------------------------------------------------------------
void set_error(const char *fmt, ...);

// ---- host: explicit tuning knobs (bx_set_tuning), cached device / kernel properties --------------
int tuning(int knob);   // value of a BX_TUNE_* knob, -1 = automatic
struct DeviceInfo { int device, sms, max_smem_optin, cooperative; };
int device_info(DeviceInfo &out);   // properties of the current device (queried once per device)
// opt-in shared memory + resident CTAs per SM of `kern` at (threads, smem): the runtime is asked once per
// (kernel, device, threads, smem), later launches hit a small cache
int kernel_occupancy(const void *kern, int threads, size_t smem, int *occ_out);
// one of the 64 work-queue counters in the 256-byte head of a caller workspace (rotates per launch)
unsigned int *queue_slot(void *workspace);

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

// Reciprocal without the special-case branch of __drcp_rn: MUFU.RCP64H seed (measured on B200:
// max relative error 2^-19.9, tools/probe_rcp.cu) + two Newton steps (4 DFMA).  Over 2^30 probe
// inputs with exponents in [-60, 60] the result is bit-identical to IEEE 1.0/x.  Every
// denominator on the hot path is a Thomas pivot, a continuant or max(|u|, 1e-10): normal, finite.
__device__ __forceinline__ double rcp(double x) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    return fma(y, e, y);
}

// Reciprocal to 1 ulp in 3 DFMA: the same seed, one cubic step  y (1 + e + e^2),  e = 1 - x y  (|e| <= 2^-19.9, so the
// truncation error is 2^-59.7).  Measured over the same 2^30 probe inputs: max error 1.000 ulp (profiles/r2_probe_rcp.txt).
// For the FOM kernels, whose tolerance against the reference (1e-10 per step) is six orders above it.
__device__ __forceinline__ double rcp1(double x) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double e = fma(-x, y, 1.0);
    return fma(y, fma(e, e, e), y);
}

// One division-free PCR step (stride s) on rows  al x[l-s] + be x[l] + ga x[l+s] = rho :
// row' = be_m be_p row - al be_p row_m - ga be_m row_p.  Rows whose neighbour falls outside the
// warp need no masking: by induction al = 0 on lanes < s and ga = 0 on lanes > 31 - s, so the
// (finite) values a shuffle returns for an out-of-range source only rescale the row.
__device__ __forceinline__ void pcr_step(double &al, double &be, double &ga, double &rho, int s) {
    const double al_m = shfl_up(al, s), be_m = shfl_up(be, s), ga_m = shfl_up(ga, s), rho_m = shfl_up(rho, s);
    const double al_p = shfl_down(al, s), be_p = shfl_down(be, s), ga_p = shfl_down(ga, s), rho_p = shfl_down(rho, s);
    const double bb = be_m * be_p, ab = al * be_p, gb = ga * be_m;
    be = fma(be, bb, -fma(ab, ga_m, gb * al_p));
    rho = fma(rho, bb, -fma(ab, rho_m, gb * rho_p));
    al = -ab * al_m;
    ga = -gb * ga_p;
}

// Solve the 32-row tridiagonal system (one row per lane) by PCR; returns x[lane].
// Division-free steps with one renormalisation half-way (bounds the growth of be to be^27).
__device__ __forceinline__ double warp_pcr(double al, double be, double ga, double rho) {
    pcr_step(al, be, ga, rho, 1);
    pcr_step(al, be, ga, rho, 2);
    {
        const double inv = rcp(be);
        al *= inv; ga *= inv; rho *= inv; be = 1.0;
    }
    pcr_step(al, be, ga, rho, 4);
    pcr_step(al, be, ga, rho, 8);
    pcr_step(al, be, ga, rho, 16);
    return rho * rcp(be);
}

// Partition solve, condensation part.  Input: the lane's M rows after the local normalised
// Thomas sweep,  fs[j] P + x[j] + cs[j] x[j+1] = rs[j]  (P = previous lane's last unknown,
// x[M] = next lane's first unknown).  Output: x[M-1] of this lane (xl) and P.
// upward sweep of a lane's M swept rows: row 0 as  F P + x[0] + G x[M-1] = R
template <int M>
__device__ __forceinline__ void partition_upward(const double (&cs)[M], const double (&rs)[M],
                                                 const double (&fs)[M], double &F, double &G, double &R) {
    if (M == 1) { F = fs[0]; G = 0.0; R = rs[0]; return; }   // x[0] is x[M-1]: F P + x[0] = R - cs[0] x_next, unused coupling
    F = fs[M > 1 ? M - 2 : 0]; G = cs[M > 1 ? M - 2 : 0]; R = rs[M > 1 ? M - 2 : 0];
#pragma unroll
    for (int j = M - 3; j >= 0; --j) {
        F = fma(-cs[j], F, fs[j]);
        G = -cs[j] * G;
        R = fma(-cs[j], R, rs[j]);
    }
}

template <int M>
__device__ __forceinline__ void partition_condense(const double (&cs)[M], const double (&rs)[M],
                                                   const double (&fs)[M], int lane, double &xl, double &P) {
    static_assert(M >= 2, "a lane needs at least two rows");
    double F, G, R;
    partition_upward<M>(cs, rs, fs, F, G, R);
    // eliminate the next lane's x[0] from this lane's last row
    const double Fn = shfl_down(F, 1), Gn = shfl_down(G, 1), Rn = shfl_down(R, 1);
    const double cl = (lane == 31) ? 0.0 : cs[M - 1];
    const double al = (lane == 0) ? 0.0 : fs[M - 1];
    xl = warp_pcr(al, fma(-cl, Fn, 1.0), -cl * Gn, fma(-cl, Rn, rs[M - 1]));
    P = shfl_up(xl, 1);
    if (lane == 0) P = 0.0;
}

// Continuant form of the local Thomas sweep: every recurrence in the chain is ONE fma/mul per
// row (no reciprocal in the dependent chain); the M reciprocals are independent.
//   th[j] = b[j] th[j-1] - a[j] c[j-1] th[j-2]      (leading-minor determinant; th[-1] = 1)
//   ro[j] = r[j] th[j-1] - a[j] ro[j-1]             (= rs[j] th[j])
//   ph[j] = -a[j] ph[j-1], ph[0] = a[0]             (= fs[j] th[j])
//   cs[j] = c[j] th[j-1] / th[j],  rs[j] = ro[j] / th[j],  fs[j] = ph[j] / th[j]
template <int M, bool ULP1 = false> struct Continuant {
    double thm1, thm2;  // th[j-1], th[j-2]
    double rom1, phm1, cm1;
    __device__ __forceinline__ void row(int j, double a, double b, double c, double r, double &cs, double &rs,
                                        double &fs) {
        double th, ro, ph, ct;
        if (j == 0) {
            th = b; ro = r; ph = a; ct = c;
            thm1 = 1.0;
        } else {
            const double t = -(a * cm1) * thm2;
            th = fma(b, thm1, t);
            ro = fma(-a, rom1, r * thm1);
            ph = -a * phm1;
            ct = c * thm1;
        }
        const double inv = ULP1 ? rcp1(th) : rcp(th);
        cs = ct * inv; rs = ro * inv; fs = ph * inv;
        thm2 = thm1; thm1 = th; rom1 = ro; phm1 = ph; cm1 = c;
    }
};

}  // namespace bx
