// bx_fom.cu — fused FOM trajectory kernel: FEMBurgers.fom_burgers (FEM/fem_burgers.py:646-707)
// for B (mu1, mu2) samples at once.
//
// One WARP integrates one trajectory from u0 to step nsteps without leaving the SM:
//   * lane l owns the M = ceil(N/32) consecutive nodes l*M .. l*M+M-1 in registers;
//   * every Picard iteration assembles the tridiagonal A = M + At C(U_k) + At E K, the
//     right-hand side b = M U^n + At F - At S(U_k) and the residual R = A U_k - b on the fly
//     from the closed forms of the 2-point-Gauss element integrals (no band ever reaches HBM);
//   * A delta = -R is solved with a partition (Wang / SPIKE-type) method: each lane runs a
//     normalised Thomas sweep over its M rows carrying the coupling to the previous lane's last
//     unknown, an upward sweep condenses the lane to one row of a 32 x 32 tridiagonal system
//     in the lanes' last unknowns, that system is solved by parallel cyclic reduction with
//     warp shuffles, and a local back-substitution finishes;
//   * the update, |delta|^2 and |U_{k+1}|^2 are fused into the back-substitution, the
//     convergence test (tol, kmax of :663) is evaluated per warp, so every trajectory runs its
//     own iteration counts with no divergence inside a warp;
//   * the only HBM traffic is the 8N-byte snapshot per step (+ 4 B iteration count).
// Warps pull samples from an atomic work queue (persistent CTAs).
#include "bx_common.cuh"
#include <cstdlib>

namespace bx {

struct FomParams {
    int B, N, nsteps, kmax;
    const double *X;
    const double *u0;
    long long u0_stride;  // 0: shared initial condition, N: per-sample
    const double *mu1;
    const double *mu2;
    double At, E, tol2;
    double *U;
    int *iters;
    unsigned int *counter;
};

// shared-memory carve-up (doubles): element arrays have (M+1)*32 entries, indexed
// (j+1)*32 + lane for the element between local nodes j and j+1 (j = -1 is the element shared
// with the previous lane); node arrays have M*32 entries indexed j*32 + lane.  Both are
// bank-conflict free for 64-bit accesses.
template <int M> struct FomSmem {
    static constexpr int EL = (M + 1) * 32;
    static constexpr int NL = M * 32;
    static constexpr int MESH = 2 * EL + NL;     // off, h6 | dg
    static constexpr int PER_WARP = EL + 2 * NL; // hfs | AtF, bn
    static constexpr size_t bytes(int warps) { return sizeof(double) * (MESH + (size_t)warps * PER_WARP); }
};

// Sum two values over the warp with 7 instead of 10 64-bit shuffles: the two half-warps first
// trade the value they do not keep, each half reduces one quantity, lanes 0 / 16 broadcast.
__device__ __forceinline__ void warp_sum2(double &a, double &b, int lane) {
    const bool lo = lane < 16;
    double mine = lo ? a : b;
    const double give = lo ? b : a;
    mine += shfl_xor(give, 16);
#pragma unroll
    for (int s = 8; s > 0; s >>= 1) mine += shfl_xor(mine, s);
    a = shfl_idx(mine, 0);
    b = shfl_idx(mine, 16);
}

// EXACT: N == 32 M, so the only lane with a missing right-hand element is lane 31 (local row M-1).
template <int M, int WARPS, int MINB, bool EXACT>
__global__ void __launch_bounds__(WARPS * 32, MINB) fom_picard_kernel(const FomParams p) {
    using L = FomSmem<M>;
    extern __shared__ __align__(16) double smem[];
    double *s_off = smem;             // h/6 - At E / h   (off-diagonal constant of A; 0 where no element)
    double *s_h6 = s_off + L::EL;     // h/6              (mass off-diagonal; read once per time step)
    double *s_dg = s_h6 + L::EL;      // diagonal constant of A (1 for padded rows)
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    double *w_hfs = s_dg + L::NL + warp * L::PER_WARP;  // (h/6) (f_gp1 + f_gp2) per element
    double *w_atf = w_hfs + L::EL;                      // At * F per node
    double *w_bn = w_atf + L::NL;                       // M U^n + At F per node (per time step)

    const int N = p.N;
    const double At = p.At;

    // ---- mesh constants, once per CTA (X is shared by all samples) ----------------------
    for (int idx = threadIdx.x; idx < L::EL; idx += WARPS * 32) {
        const int l = idx & 31, jj = (idx >> 5) - 1;
        const int e = l * M + jj;
        double off = 0.0, h6 = 0.0;
        if (e >= 0 && e <= N - 2) {
            const double h = p.X[e + 1] - p.X[e];
            h6 = h / 6.0;
            off = h6 - At * p.E / h;
        }
        s_off[idx] = off; s_h6[idx] = h6;
    }
    for (int idx = threadIdx.x; idx < L::NL; idx += WARPS * 32) {
        const int l = idx & 31, j = idx >> 5;
        const int i = l * M + j;
        double dg = 1.0;
        if (i < N) {
            dg = 0.0;
            if (i >= 1) { const double h = p.X[i] - p.X[i - 1]; dg += 2.0 * (h / 6.0) + At * p.E / h; }
            if (i <= N - 2) { const double h = p.X[i + 1] - p.X[i]; dg += 2.0 * (h / 6.0) + At * p.E / h; }
        }
        s_dg[idx] = dg;
    }
    __syncthreads();

    const int node0 = lane * M;
    const bool vec_ok = (node0 + M <= N) && ((N & 1) == 0) && ((M & 1) == 0);
    // Elements that do not exist must contribute nothing.  Padded nodes carry u = 0, so every
    // missing element has p = q = G = 0 by itself except the one to the right of the last real
    // node (u1 != 0): its p, q, G are zeroed explicitly on local row jl of the lane that owns
    // node N-1.  The element to the left of a lane (j = -1) is missing on lane 0 only (row 0 is
    // the Dirichlet row there) and on fully padded lanes.
    const int jl = N - 1 - node0;   // local index of the last real row (may be outside [0, M))
    const double a6 = At / 6.0;     // convection: p = a6 (2 u1 + u2), q = a6 (u1 + 2 u2)
    const double c1 = 1.5 * At;     // SUPG: At g = c1 (s du / 6 - (h/6) fsum) / max(|s|, 2e-10),  s = u1 + u2
    const bool leftv = lane > 0 && jl >= 0;
    const double a6L = leftv ? a6 : 0.0;
    const double c1L = leftv ? c1 : 0.0;

    for (;;) {
        int s = 0;
        if (lane == 0) s = (int)atomicAdd(p.counter, 1u);
        s = __shfl_sync(FULL, s, 0);
        if (s >= p.B) break;

        const double mu1 = p.mu1[s];
        const double mu2 = p.mu2[s];

        // ---- per-sample forcing constants (FEM/fem_burgers.py:449-455, 565) -------------
        {
            double left = 0.0;  // contribution of the element on the left to the current node
#pragma unroll
            for (int jj = -1; jj < M; ++jj) {
                const int e = node0 + jj;
                double hfs = 0.0, c0 = 0.0, cr = 0.0;
                if (e >= 0 && e <= N - 2) {
                    const double x1 = p.X[e], x2 = p.X[e + 1];
                    const double f1 = 0.02 * exp(mu2 * (BX_GA * x1 + BX_GB * x2));
                    const double f2 = 0.02 * exp(mu2 * (BX_GB * x1 + BX_GA * x2));
                    const double dV = 0.5 * (x2 - x1);
                    hfs = ((x2 - x1) / 6.0) * (f1 + f2);
                    c0 = (f1 * BX_GA + f2 * BX_GB) * dV;  // local node 1 of element e -> node e
                    cr = (f1 * BX_GB + f2 * BX_GA) * dV;  // local node 2 of element e -> node e+1
                }
                w_hfs[(jj + 1) * 32 + lane] = hfs;
                if (jj >= 0) w_atf[jj * 32 + lane] = At * (left + c0);
                left = cr;
            }
        }

        // ---- initial condition -> registers + snapshot 0 -------------------------------
        double u[M];
        const double *u0p = p.u0 + (long long)s * p.u0_stride + node0;
        double *Us = p.U + (long long)s * (long long)(p.nsteps + 1) * N + node0;
#pragma unroll
        for (int j = 0; j < M; ++j) u[j] = (node0 + j < N) ? u0p[j] : 0.0;
        if (vec_ok) {
#pragma unroll
            for (int j = 0; j < M; j += 2) *reinterpret_cast<double2 *>(Us + j) = make_double2(u[j], u[j + 1]);
        } else {
#pragma unroll
            for (int j = 0; j < M; ++j) if (node0 + j < N) Us[j] = u[j];
        }
        __syncwarp();

        for (int n = 0; n < p.nsteps; ++n) {
            // ---- bn = M U^n + At F (constant during the Picard loop, :683) ---------------
            {
                const double uL = shfl_up(u[M - 1], 1), uR = shfl_down(u[0], 1);
                double h6L = s_h6[lane];
#pragma unroll
                for (int j = 0; j < M; ++j) {
                    const double h6R = s_h6[(j + 1) * 32 + lane];
                    const double um = (j == 0) ? uL : u[j > 0 ? j - 1 : 0];
                    const double up = (j == M - 1) ? uR : u[j < M - 1 ? j + 1 : j];
                    w_bn[j * 32 + lane] = h6L * um + (2.0 * (h6L + h6R)) * u[j] + h6R * up + w_atf[j * 32 + lane];
                    h6L = h6R;
                }
            }
            int k = 0;
            double sd, su;
            do {
                // ================= one Picard iteration (:667-698) ======================
                const double uL = shfl_up(u[M - 1], 1), uR = shfl_down(u[0], 1);
                double cs[M], rs[M], fs[M];
                Continuant<M> ct;
                // element j = -1 (shared with the previous lane)
                double offL, qL, GL;
                {
                    const double u1 = uL, u2 = u[0];
                    const double sm = u1 + u2;
                    qL = fma(a6L, u2, a6L * sm);
                    const double z = fma(sm * (u2 - u1), 1.0 / 6.0, -w_hfs[lane]);
                    GL = (c1L * rcp(fmax(fabs(sm), 2.0e-10))) * z;
                    offL = s_off[lane];
                }
#pragma unroll
                for (int j = 0; j < M; ++j) {
                    const int ei = (j + 1) * 32 + lane;
                    const double um = (j == 0) ? uL : u[j > 0 ? j - 1 : 0];
                    const double up = (j == M - 1) ? uR : u[j < M - 1 ? j + 1 : j];
                    const double u1 = u[j], u2 = up;
                    // element j: convection p, q (:418) and SUPG G = At*g (:569-575)
                    const double sm = u1 + u2;
                    const double t = a6 * sm;
                    double pe = fma(a6, u1, t);
                    double qe = fma(a6, u2, t);
                    const double z = fma(sm * (u2 - u1), 1.0 / 6.0, -w_hfs[ei]);
                    double Ge = (c1 * rcp(fmax(fabs(sm), 2.0e-10))) * z;
                    if ((!EXACT || j == M - 1) && j == jl) { pe = 0.0; qe = 0.0; Ge = 0.0; }
                    const double off = s_off[ei];
                    // row j of A, b and r = -R = b - A U_k
                    double a = offL - qL;
                    double c = off + pe;
                    double b = s_dg[j * 32 + lane] + (qL - pe);
                    const double rhs = w_bn[j * 32 + lane] - GL + Ge;
                    double r = rhs - fma(a, um, fma(b, u1, c * u2));
                    if (j == 0 && lane == 0) {  // Dirichlet row (:679-680, :686)
                        a = 0.0; b = 1.0; c = 0.0; r = mu1 - u1;
                    }
                    // local Thomas sweep in continuant form (no reciprocal in the dependent chain)
                    ct.row(j, a, b, c, r, cs[j], rs[j], fs[j]);
                    offL = off; qL = qe; GL = Ge;
                }
                // condense each lane to one row, PCR across the lanes, hand P to the next lane
                double xl, P;
                partition_condense<M>(cs, rs, fs, lane, xl, P);
                // back-substitution fused with the update and the two norms (:695-698)
                double x = xl;
                sd = 0.0; su = 0.0;
#pragma unroll
                for (int j = M - 1; j >= 0; --j) {
                    if (j < M - 1) x = fma(-cs[j], x, fma(-fs[j], P, rs[j]));
                    sd = fma(x, x, sd);
                    u[j] += x;
                    su = fma(u[j], u[j], su);
                }
                warp_sum2(sd, su, lane);
                ++k;
            } while (sd > p.tol2 * su && k < p.kmax);

            // ---- snapshot n+1 (:705) and iteration count -------------------------------
            double *Un1 = Us + (long long)(n + 1) * N;
            if (vec_ok) {
#pragma unroll
                for (int j = 0; j < M; j += 2) *reinterpret_cast<double2 *>(Un1 + j) = make_double2(u[j], u[j + 1]);
            } else {
#pragma unroll
                for (int j = 0; j < M; ++j) if (node0 + j < N) Un1[j] = u[j];
            }
            if (lane == 0 && p.iters) p.iters[(long long)s * p.nsteps + n] = k;
        }
    }
}

template <int M, int WARPS, int MINB, bool EXACT>
static int launch_fom_t(const FomParams &p, cudaStream_t st) {
    using L = FomSmem<M>;
    const size_t smem = L::bytes(WARPS);
    auto kern = fom_picard_kernel<M, WARPS, MINB, EXACT>;
    BX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int dev = 0, sms = 0, occ = 0;
    BX_CUDA(cudaGetDevice(&dev));
    BX_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    BX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, WARPS * 32, smem));
    BX_REQUIRE(occ >= 1, "fom kernel does not fit on an SM (smem %zu B)", smem);
    long long ctas = (long long)sms * occ;
    const long long need = ((long long)p.B + WARPS - 1) / WARPS;
    if (ctas > need) ctas = need;
    BX_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st));
    kern<<<(unsigned)ctas, WARPS * 32, smem, st>>>(p);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

template <int M, int WARPS, int MINB>
static int launch_fom(const FomParams &p, cudaStream_t st) {
    return (p.N == 32 * M) ? launch_fom_t<M, WARPS, MINB, true>(p, st) : launch_fom_t<M, WARPS, MINB, false>(p, st);
}

}  // namespace bx

extern "C" long long bx_workspace_bytes(void) { return 256; }

extern "C" int bx_fom_picard_trajectories(int B, int N, int nsteps, const double *X, const double *u0,
                                          int u0_batched, const double *mu1, const double *mu2,
                                          double At, double E, double tol, int kmax, double *U,
                                          int *iters, void *workspace, void *stream) {
    using namespace bx;
    BX_REQUIRE(B >= 0 && nsteps >= 0, "B and nsteps must be non-negative");
    BX_REQUIRE(N >= 2, "need at least 2 nodes (got N = %d)", N);
    BX_REQUIRE(kmax >= 1, "kmax must be >= 1 (the reference loop body defines U1)");
    if (B == 0) return BX_OK;
    BX_REQUIRE(X && u0 && mu1 && mu2 && U && workspace, "null device pointer");
    FomParams p;
    p.B = B; p.N = N; p.nsteps = nsteps; p.kmax = kmax;
    p.X = X; p.u0 = u0; p.u0_stride = u0_batched ? N : 0;
    p.mu1 = mu1; p.mu2 = mu2; p.At = At; p.E = E; p.tol2 = tol * tol;
    p.U = U; p.iters = iters; p.counter = static_cast<unsigned int *>(workspace);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (N <= 64) return launch_fom<2, 4, 4>(p, st);
    if (N <= 128) return launch_fom<4, 4, 4>(p, st);
    if (N <= 256) return launch_fom<8, 4, 4>(p, st);
    if (N <= 512) {
        // tuning knob (undocumented, for A/B runs on the box): BX_FOM_VARIANT=1 -> 16 warps/CTA at 128 regs
        const char *v = getenv("BX_FOM_VARIANT");
        if (v && v[0] == '1') return launch_fom<16, 16, 1>(p, st);
        if (v && v[0] == '2') return launch_fom<16, 8, 1>(p, st);
        return launch_fom<16, 4, 3>(p, st);
    }
    if (N <= 1024) return launch_fom<32, 4, 1>(p, st);
    set_error("bx_fom_picard_trajectories: N = %d > 1024 needs the multi-CTA cyclic-reduction path", N);
    return BX_ERR_UNSUPPORTED;
}
