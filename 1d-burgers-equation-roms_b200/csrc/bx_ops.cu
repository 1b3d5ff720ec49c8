// bx_ops.cu — the L1 FE operators of FEMBurgers (FEM/fem_burgers.py:324-581) and one Picard
// linearisation with the bands materialised in HBM (the drop-in for callers that want A, b, R
// as data: compute_* methods, PROM variants outside the fused kernels).
//
// Layout: batched arrays are sample-major (b, i) -> [b*N + i]; a warp covers 31 consecutive
// nodes of one sample: lane L >= 1 owns node i = 31*tile + L - 1 and evaluates the element to
// its right (e = i); lane 0 evaluates the halo element e = 31*tile - 1, so every element
// integral (2 exp per element for the forcing) is computed once per warp and handed to the
// node on its right by one shuffle.  All global accesses are unit-stride across the warp.
#include "bx_common.cuh"

namespace bx {

struct Elem {       // closed forms of the 2-point Gauss integrals of one linear element
    double h6;      // h/6                      mass off-diagonal          (:346)
    double rh;      // 1/h                      diffusion                  (:380)
    double p, q;    // (2u1+u2)/6, (u1+2u2)/6   convection rows            (:418)
    double c0, c1;  // forcing contributions to local node 1 / 2           (:455)
    double g;       // SUPG: S_e = [-g, +g]                                 (:575)
};

template <bool CONV, bool FORCE, bool SUPG>
__device__ __forceinline__ Elem eval_elem(int e, int N, const double *__restrict__ X,
                                          const double *__restrict__ U, double mu2) {
    Elem r = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    if (e < 0 || e > N - 2) return r;
    const double x1 = X[e], x2 = X[e + 1];
    const double h = x2 - x1;
    r.h6 = h / 6.0;
    r.rh = 1.0 / h;
    double u1 = 0.0, u2 = 0.0;
    if (CONV || SUPG) { u1 = U[e]; u2 = U[e + 1]; }
    if (CONV) {
        r.p = (2.0 * u1 + u2) / 6.0;
        r.q = (u1 + 2.0 * u2) / 6.0;
    }
    if (FORCE || SUPG) {
        const double f1 = 0.02 * exp(mu2 * (BX_GA * x1 + BX_GB * x2));
        const double f2 = 0.02 * exp(mu2 * (BX_GB * x1 + BX_GA * x2));
        if (FORCE) {
            const double dV = 0.5 * h;
            r.c0 = (f1 * BX_GA + f2 * BX_GB) * dV;
            r.c1 = (f1 * BX_GB + f2 * BX_GA) * dV;
        }
        if (SUPG) {
            const double sm = u1 + u2;
            const double vel = fmax(fabs(0.5 * sm), 1.0e-10);
            const double tau = 0.5 * h / (2.0 * vel);
            const double dudx = (u2 - u1) / h;
            r.g = 0.5 * tau * (sm * dudx - (f1 + f2));
        }
    }
    return r;
}

__device__ __forceinline__ Elem shfl_up_elem(const Elem &e) {
    Elem r;
    r.h6 = shfl_up(e.h6, 1); r.rh = shfl_up(e.rh, 1); r.p = shfl_up(e.p, 1); r.q = shfl_up(e.q, 1);
    r.c0 = shfl_up(e.c0, 1); r.c1 = shfl_up(e.c1, 1); r.g = shfl_up(e.g, 1);
    return r;
}

enum OpKind { OP_MASS = 0, OP_DIFF = 1, OP_CONV = 2, OP_DCONV = 3 };

// bands of M, K, C(U) or dC/dU for B states
template <int KIND>
__global__ void __launch_bounds__(128) bands_kernel(int B, int N, const double *__restrict__ X,
                                                    const double *__restrict__ U, double *__restrict__ dl,
                                                    double *__restrict__ d, double *__restrict__ du) {
    const int tiles = (N + 30) / 31;
    const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= (long long)B * tiles) return;
    const int b = (int)(w / tiles), tile = (int)(w % tiles);
    const int e = 31 * tile + lane - 1;
    const double *Ub = (KIND == OP_CONV) ? U + (long long)b * N : nullptr;
    const Elem R = eval_elem<KIND == OP_CONV, false, false>(e, N, X, Ub, 0.0);
    const Elem Lf = shfl_up_elem(R);
    const int i = e;
    if (lane == 0 || i >= N) return;
    const bool hasL = i >= 1, hasR = i <= N - 2;
    double a = 0.0, dd = 0.0, c = 0.0;
    if (KIND == OP_MASS) {
        a = Lf.h6; c = R.h6; dd = 2.0 * Lf.h6 + 2.0 * R.h6;
    } else if (KIND == OP_DIFF) {
        a = -Lf.rh; c = -R.rh; dd = Lf.rh + R.rh;
    } else if (KIND == OP_CONV) {
        a = -Lf.q; c = R.p; dd = Lf.q - R.p;
    } else {
        a = hasL ? -0.5 : 0.0; c = hasR ? 0.5 : 0.0; dd = (hasL ? 0.5 : 0.0) - (hasR ? 0.5 : 0.0);
    }
    const long long o = (long long)b * N + i;
    dl[o] = a; d[o] = dd; du[o] = c;
}

// F(mu2) (SUPGV = false) or S(U, mu2) (SUPGV = true)
template <bool SUPGV>
__global__ void __launch_bounds__(128) vector_kernel(int B, int N, const double *__restrict__ X,
                                                     const double *__restrict__ U,
                                                     const double *__restrict__ mu2, double *__restrict__ out) {
    const int tiles = (N + 30) / 31;
    const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= (long long)B * tiles) return;
    const int b = (int)(w / tiles), tile = (int)(w % tiles);
    const int e = 31 * tile + lane - 1;
    const double *Ub = SUPGV ? U + (long long)b * N : nullptr;
    const Elem R = eval_elem<false, !SUPGV, SUPGV>(e, N, X, Ub, mu2[b]);
    const Elem Lf = shfl_up_elem(R);
    if (lane == 0 || e >= N) return;
    out[(long long)b * N + e] = SUPGV ? (Lf.g - R.g) : (Lf.c1 + R.c0);
}

// K1: A bands, b, R of one Picard linearisation (FEM/fem_burgers.py:667-689)
template <bool SUPGV>
__global__ void __launch_bounds__(128)
assemble_picard_kernel(int B, int N, const double *__restrict__ X, const double *__restrict__ Uk,
                       const double *__restrict__ Un, const double *__restrict__ mu1,
                       const double *__restrict__ mu2, double At, double E, double *__restrict__ dl,
                       double *__restrict__ d, double *__restrict__ du, double *__restrict__ bout,
                       double *__restrict__ Rout) {
    const int tiles = (N + 30) / 31;
    const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= (long long)B * tiles) return;
    const int b = (int)(w / tiles), tile = (int)(w % tiles);
    const int e = 31 * tile + lane - 1;
    const double *Ukb = Uk + (long long)b * N;
    const double *Unb = Un + (long long)b * N;
    const Elem R = eval_elem<true, true, SUPGV>(e, N, X, Ukb, mu2[b]);
    const Elem Lf = shfl_up_elem(R);
    const int i = e;
    // neighbours of U^n and U_k come from the lanes next door (halo lanes load them directly)
    const double uk = (i >= 0 && i < N) ? Ukb[i] : 0.0;
    const double un = (i >= 0 && i < N) ? Unb[i] : 0.0;
    double ukL = shfl_up(uk, 1), unL = shfl_up(un, 1);
    double ukR = shfl_down(uk, 1), unR = shfl_down(un, 1);
    if (lane == 31) {
        ukR = (i + 1 < N) ? Ukb[i + 1] : 0.0;
        unR = (i + 1 < N) ? Unb[i + 1] : 0.0;
    }
    if (lane == 0 || i >= N) return;
    double a, dd, c, rhs, res;
    if (i == 0) {  // Dirichlet row (:679-680, :686)
        a = 0.0; dd = 1.0; c = 0.0; rhs = mu1[b];
        res = uk - rhs;
    } else {
        const double ekL = At * E * Lf.rh, ekR = At * E * R.rh;
        a = Lf.h6 - At * Lf.q - ekL;
        c = R.h6 + At * R.p - ekR;
        dd = (2.0 * Lf.h6 + 2.0 * R.h6) + At * (Lf.q - R.p) + (ekL + ekR);
        rhs = Lf.h6 * unL + (2.0 * Lf.h6 + 2.0 * R.h6) * un + R.h6 * unR + At * (Lf.c1 + R.c0);
        if (SUPGV) rhs -= At * (Lf.g - R.g);
        res = a * ukL + dd * uk + c * ukR - rhs;
    }
    const long long o = (long long)b * N + i;
    dl[o] = a; d[o] = dd; du[o] = c; Rout[o] = res;
    if (bout) bout[o] = rhs;
}

static inline unsigned grid_for(int B, int N) {
    const long long warps = (long long)B * ((N + 30) / 31);
    return (unsigned)((warps * 32 + 127) / 128);
}

}  // namespace bx

using namespace bx;

#define BX_OPS_CHECK(B, N)                                       \
    BX_REQUIRE((B) >= 0, "B must be non-negative");              \
    BX_REQUIRE((N) >= 2, "need at least 2 nodes (got %d)", (N)); \
    if ((B) == 0) return BX_OK;

extern "C" int bx_mass_bands(int N, const double *X, double *dl, double *d, double *du, void *stream) {
    BX_OPS_CHECK(1, N); BX_REQUIRE(X && dl && d && du, "null device pointer");
    bands_kernel<OP_MASS><<<grid_for(1, N), 128, 0, (cudaStream_t)stream>>>(1, N, X, nullptr, dl, d, du);
    BX_LAUNCH_CHECK(); return BX_OK;
}
extern "C" int bx_diffusion_bands(int N, const double *X, double *dl, double *d, double *du, void *stream) {
    BX_OPS_CHECK(1, N); BX_REQUIRE(X && dl && d && du, "null device pointer");
    bands_kernel<OP_DIFF><<<grid_for(1, N), 128, 0, (cudaStream_t)stream>>>(1, N, X, nullptr, dl, d, du);
    BX_LAUNCH_CHECK(); return BX_OK;
}
extern "C" int bx_convection_bands(int B, int N, const double *X, const double *U, double *dl, double *d,
                                   double *du, void *stream) {
    BX_OPS_CHECK(B, N); BX_REQUIRE(X && U && dl && d && du, "null device pointer");
    bands_kernel<OP_CONV><<<grid_for(B, N), 128, 0, (cudaStream_t)stream>>>(B, N, X, U, dl, d, du);
    BX_LAUNCH_CHECK(); return BX_OK;
}
extern "C" int bx_convection_derivative_bands(int N, const double *X, double *dl, double *d, double *du,
                                              void *stream) {
    BX_OPS_CHECK(1, N); BX_REQUIRE(X && dl && d && du, "null device pointer");
    bands_kernel<OP_DCONV><<<grid_for(1, N), 128, 0, (cudaStream_t)stream>>>(1, N, X, nullptr, dl, d, du);
    BX_LAUNCH_CHECK(); return BX_OK;
}
extern "C" int bx_forcing_vector(int B, int N, const double *X, const double *mu2, double *F, void *stream) {
    BX_OPS_CHECK(B, N); BX_REQUIRE(X && mu2 && F, "null device pointer");
    vector_kernel<false><<<grid_for(B, N), 128, 0, (cudaStream_t)stream>>>(B, N, X, nullptr, mu2, F);
    BX_LAUNCH_CHECK(); return BX_OK;
}
extern "C" int bx_supg_term(int B, int N, const double *X, const double *U, const double *mu2, double *S,
                            void *stream) {
    BX_OPS_CHECK(B, N); BX_REQUIRE(X && U && mu2 && S, "null device pointer");
    vector_kernel<true><<<grid_for(B, N), 128, 0, (cudaStream_t)stream>>>(B, N, X, U, mu2, S);
    BX_LAUNCH_CHECK(); return BX_OK;
}
extern "C" int bx_assemble_picard(int B, int N, const double *X, const double *Uk, const double *Un,
                                  const double *mu1, const double *mu2, double At, double E, int supg,
                                  double *dl, double *d, double *du, double *b, double *R, void *stream) {
    BX_OPS_CHECK(B, N);
    BX_REQUIRE(X && Uk && Un && mu1 && mu2 && dl && d && du && R, "null device pointer");
    if (supg)
        assemble_picard_kernel<true><<<grid_for(B, N), 128, 0, (cudaStream_t)stream>>>(B, N, X, Uk, Un, mu1, mu2, At, E, dl, d, du, b, R);
    else
        assemble_picard_kernel<false><<<grid_for(B, N), 128, 0, (cudaStream_t)stream>>>(B, N, X, Uk, Un, mu1, mu2, At, E, dl, d, du, b, R);
    BX_LAUNCH_CHECK(); return BX_OK;
}
