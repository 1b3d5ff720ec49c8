// bx_reduced.cu — the reduced system of ONE nonlinear iteration for a caller-supplied tangent:
//     A = M + At C(U_k) + At E K  (row 0 <- e0^T),   b = M U^n + At F - At S(U_k),  b[0] = mu1,   R = A U_k - b
//     Galerkin  Ar = T^T A T,  br = T^T R          LSPG  Ar = (A T)^T (A T),  br = (A T)^T R
// i.e. lines :730-762 of FEMBurgers.pod_prom_burgers, which local_prom_burgers (:1022-1058), pod_ann_prom
// (:1225-1262), pod_rbf_prom (:1330-1375), lie_prom and ae_prom repeat verbatim with their own decoder tangent in
// the place of Phi.  The reference's native precedent has exactly this shape: compute_Ar_br(A, Phi, R, Ar, br)
// (Stanford_2D/Burgers_2D/cpp_files/mkl_sparse_dense_operations.cpp:13-20), in-place outputs.
//
// One CTA (8 warps) per sample.  The tridiagonal rows and R are assembled once into shared memory from the closed
// forms of the element integrals (same arithmetic as bx_prom_common.cuh:prom_row); the contraction runs on the FP64
// tensor pipe (DMMA m8n8k4, k index = node) over chunks of 32 nodes: the chunk of J = A T (column n carries R, so
// br is column n of the same accumulators) and, for Galerkin, of T is staged in shared memory, every warp owns
// up to 16 of the 8 x 8 output tiles per pass (accumulators in registers), tiles beyond 8 x 16 take further passes
// that re-stage the chunks (T is L2-resident).  n <= 240, N <= 2048.  T: (N, n) row-major, shared by all samples
// (T_stride = 0) or one per sample (T_stride doubles apart, e.g. N * n).
#include "bx_prom_common.cuh"

namespace bx {

struct ReducedParams {
    int B, N, n, NB, supg, lspg;
    const double *X, *Uk, *Un, *mu1, *mu2;
    double At, E;
    const double *T;
    long long T_stride;
    double *Ar, *br;
};

constexpr int RS_CH = 32;          // nodes per staged chunk
constexpr int RS_TPW = 16;         // output tiles per warp and pass
constexpr int RS_WARPS = 8;

__device__ __forceinline__ double rs_force(double mu2, double x1, double x2, double &f1, double &f2) {
    f1 = 0.02 * exp(mu2 * (BX_GA * x1 + BX_GB * x2));
    f2 = 0.02 * exp(mu2 * (BX_GB * x1 + BX_GA * x2));
    return 0.5 * (x2 - x1);
}

template <bool LSPG>
__global__ void __launch_bounds__(RS_WARPS * 32) reduced_system_kernel(const ReducedParams p) {
    extern __shared__ __align__(16) double smem[];
    const int N = p.N, n = p.n, NB = p.NB, NC = 8 * NB, LD = NC + 4;
    double *s_a = smem, *s_b = s_a + N, *s_c = s_b + N, *s_R = s_c + N;
    double *s_J = s_R + N;                 // RS_CH x LD
    double *s_T = s_J + RS_CH * LD;        // RS_CH x LD (Galerkin); before the contraction: U_k, U^n with halos
    const int smp = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t = lane & 3, g = lane >> 2;
    const double At = p.At;

    // ---- rows of A and R (FEM/fem_burgers.py:730-752) --------------------------------------------
    double *s_uk = s_J + 1, *s_un = s_J + (N + 2) + 1;      // [-1 .. N]; the J / T chunk area is free until the contraction
    for (int i = threadIdx.x - 1; i <= N; i += blockDim.x) {
        const bool in = i >= 0 && i < N;
        s_uk[i] = in ? p.Uk[(size_t)smp * N + i] : 0.0;
        s_un[i] = in ? p.Un[(size_t)smp * N + i] : 0.0;
    }
    __syncthreads();
    const double mu1 = p.mu1[smp], mu2 = p.mu2[smp];
    const double a6 = At / 6.0, c1 = 1.5 * At, ate = At * p.E;
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
        const double um = s_uk[i - 1], uc = s_uk[i], up = s_uk[i + 1];
        double h6L = 0.0, h6R = 0.0, keL = 0.0, keR = 0.0, F = 0.0, hfL = 0.0, hfR = 0.0;
        if (i >= 1) {
            const double x1 = p.X[i - 1], x2 = p.X[i];
            double f1, f2;
            const double dV = rs_force(mu2, x1, x2, f1, f2);
            h6L = (x2 - x1) / 6.0;
            F += (f1 * BX_GB + f2 * BX_GA) * dV;
            hfL = h6L * (f1 + f2);
            if (p.E != 0.0) keL = ate / (x2 - x1);
        }
        if (i <= N - 2) {
            const double x1 = p.X[i], x2 = p.X[i + 1];
            double f1, f2;
            const double dV = rs_force(mu2, x1, x2, f1, f2);
            h6R = (x2 - x1) / 6.0;
            F += (f1 * BX_GA + f2 * BX_GB) * dV;
            hfR = h6R * (f1 + f2);
            if (p.E != 0.0) keR = ate / (x2 - x1);
        }
        double pe = 0.0, Ge = 0.0, qL = 0.0, GL = 0.0;
        if (i <= N - 2) {
            const double sm = uc + up;
            pe = fma(a6, uc, a6 * sm);
            if (p.supg) Ge = (c1 * rcp(fmax(fabs(sm), 2.0e-10))) * fma(sm * (up - uc), 1.0 / 6.0, -hfR);
        }
        if (i >= 1) {
            const double sm = um + uc;
            qL = fma(a6, uc, a6 * sm);
            if (p.supg) GL = (c1 * rcp(fmax(fabs(sm), 2.0e-10))) * fma(sm * (uc - um), 1.0 / 6.0, -hfL);
        }
        const double bn = h6L * s_un[i - 1] + (2.0 * (h6L + h6R)) * s_un[i] + h6R * s_un[i + 1] + At * F;
        double a = (h6L - keL) - qL;
        double c = (h6R - keR) + pe;
        double b = (2.0 * (h6L + h6R) + (keL + keR)) + (qL - pe);
        double R = fma(a, um, fma(b, uc, c * up)) - (bn - GL + Ge);
        if (i == 0) { a = 0.0; b = 1.0; c = 0.0; R = uc - mu1; }
        s_a[i] = a; s_b[i] = b; s_c[i] = c; s_R[i] = R;
    }
    __syncthreads();

    // ---- contraction, tiles in passes ---------------------------------------------------------------
    const double *T = p.T + (size_t)smp * p.T_stride;
    const int ntiles = LSPG ? NB * (NB + 1) / 2 : NB * NB;
    const int per_pass = RS_WARPS * RS_TPW;
    for (int pass0 = 0; pass0 < ntiles; pass0 += per_pass) {
        // this warp's tiles of the pass: tile index -> (I, J)
        int tI[RS_TPW], tJ[RS_TPW];
        int mine = 0;
#pragma unroll
        for (int q = 0; q < RS_TPW; ++q) {
            const int ti = pass0 + q * RS_WARPS + warp;
            tI[q] = 0; tJ[q] = 0;
            if (ti < ntiles) {
                mine = q + 1;
                if (LSPG) {   // row-major upper triangle
                    int I = 0, rem = ti;
                    while (rem >= NB - I) { rem -= NB - I; ++I; }
                    tI[q] = I; tJ[q] = I + rem;
                } else {
                    tI[q] = ti / NB; tJ[q] = ti - (ti / NB) * NB;
                }
            }
        }
        double acc[RS_TPW][2];
#pragma unroll
        for (int q = 0; q < RS_TPW; ++q) { acc[q][0] = 0.0; acc[q][1] = 0.0; }

        for (int i0 = 0; i0 < N; i0 += RS_CH) {
            __syncthreads();   // the previous chunk (or the assembly's use of this area) is done with
            for (int idx = threadIdx.x; idx < RS_CH * NC; idx += blockDim.x) {
                const int r = idx / NC, c = idx - r * NC;
                const int i = i0 + r;
                double J = 0.0, T0 = 0.0;
                if (i < N) {
                    if (c < n) {
                        T0 = T[(size_t)i * n + c];
                        const double Tm = i >= 1 ? T[(size_t)(i - 1) * n + c] : 0.0;
                        const double Tp = i <= N - 2 ? T[(size_t)(i + 1) * n + c] : 0.0;
                        J = fma(s_a[i], Tm, fma(s_b[i], T0, s_c[i] * Tp));
                    } else if (c == n) {
                        J = s_R[i];    // R rides as column n of the right operand
                    }
                }
                s_J[r * LD + c] = J;
                if (!LSPG) s_T[r * LD + c] = T0;
            }
            __syncthreads();
            const double *L = LSPG ? s_J : s_T;
#pragma unroll
            for (int ks = 0; ks < RS_CH / 4; ++ks) {
                const double *Lr = L + (4 * ks + t) * LD + g;
                const double *Jr = s_J + (4 * ks + t) * LD + g;
#pragma unroll
                for (int q = 0; q < RS_TPW; ++q)
                    if (q < mine) dmma884(acc[q][0], acc[q][1], Lr[8 * tI[q]], Jr[8 * tJ[q]]);
            }
        }
        // ---- accumulators -> Ar, br ---------------------------------------------------------------------
        double *Ar = p.Ar + (size_t)smp * n * n;
        double *br = p.br + (size_t)smp * n;
#pragma unroll
        for (int q = 0; q < RS_TPW; ++q) {
            if (q < mine) {
                const int r = 8 * tI[q] + g;
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int c = 8 * tJ[q] + 2 * t + e;
                    const double v = acc[q][e];
                    if (r < n && c < n) {
                        Ar[(size_t)r * n + c] = v;
                        if (LSPG && tJ[q] > tI[q]) Ar[(size_t)c * n + r] = v;
                    }
                    if (r < n && c == n) br[r] = v;
                }
            }
        }
    }
}

}  // namespace bx

extern "C" int bx_reduced_system(int B, int N, int n, const double *X, const double *Uk, const double *Un,
                                 const double *mu1, const double *mu2, double At, double E, int supg,
                                 const double *T, long long T_stride, int projection, double *Ar, double *br,
                                 void *stream) {
    using namespace bx;
    if (projection != BX_GALERKIN && projection != BX_LSPG) {
        set_error("Projection method '%d' is not available. Please use 'Galerkin' or 'LSPG'.", projection);
        return BX_ERR_PROJECTION;
    }
    BX_REQUIRE(B >= 0, "B must be non-negative");
    BX_REQUIRE(N >= 2 && N <= 2048, "bx_reduced_system: 2 <= N <= 2048 (got %d)", N);
    BX_REQUIRE(n >= 1 && n <= 240 && n <= N, "bx_reduced_system: 1 <= n <= min(N, 240) (got %d)", n);
    BX_REQUIRE(T_stride == 0 || T_stride >= (long long)N * n, "T_stride must be 0 (shared tangent) or >= N * n");
    if (B == 0) return BX_OK;
    BX_REQUIRE(X && Uk && Un && mu1 && mu2 && T && Ar && br, "null device pointer");
    ReducedParams p;
    p.B = B; p.N = N; p.n = n; p.NB = n / 8 + 1; p.supg = supg ? 1 : 0; p.lspg = projection == BX_LSPG;
    p.X = X; p.Uk = Uk; p.Un = Un; p.mu1 = mu1; p.mu2 = mu2; p.At = At; p.E = E;
    p.T = T; p.T_stride = T_stride; p.Ar = Ar; p.br = br;
    const int LD = 8 * p.NB + 4;
    size_t chunk = (size_t)2 * RS_CH * LD;
    if (chunk < (size_t)2 * (N + 2) + 2) chunk = (size_t)2 * (N + 2) + 2;   // U_k, U^n with halos live there first
    const size_t smem = sizeof(double) * ((size_t)4 * N + chunk);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc;
    if (p.lspg) {
        rc = kernel_occupancy(reinterpret_cast<const void *>(reduced_system_kernel<true>), RS_WARPS * 32, smem, nullptr);
        if (rc != BX_OK) return rc;
        reduced_system_kernel<true><<<(unsigned)B, RS_WARPS * 32, smem, st>>>(p);
    } else {
        rc = kernel_occupancy(reinterpret_cast<const void *>(reduced_system_kernel<false>), RS_WARPS * 32, smem, nullptr);
        if (rc != BX_OK) return rc;
        reduced_system_kernel<false><<<(unsigned)B, RS_WARPS * 32, smem, st>>>(p);
    }
    BX_LAUNCH_CHECK();
    return BX_OK;
}
