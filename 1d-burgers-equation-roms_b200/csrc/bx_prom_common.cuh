// bx_prom_common.cuh — device helpers shared by the POD-PROM (bx_prom.cu) and quadratic-manifold
// PROM (bx_qm.cu) trajectory kernels: DMMA wrapper, node-run geometry, the per-lane reduced solve,
// accumulator-tile indexing and the one-node row assembly.
#pragma once
#include "bx_common.cuh"

namespace bx {

// parameter block of the warp-per-trajectory POD-PROM kernels (bx_prom.cu: n <= 31, bx_prom_wide.cu: 32 <= n <= 47)
struct PromParams {
    int B, N, n, nsteps, kmax, ldp, warps;
    const double *X;
    const double *u0;
    long long u0_stride;
    const double *mu1;
    const double *mu2;
    double At, E, tol;
    const double *Phi;   // (N, n) row-major (C order), shared by all samples
    double *atf;         // workspace: At * F            (B, N)
    double *hfs;         // workspace: (h/6)(f_gp1+f_gp2) (B, N), element e at [e]
    double *U;
    int *iters;
    double *err_hist;    // optional (B, nsteps, kmax): |dq| / |q| after every iteration (:776)
    unsigned int *counter;
    int *progress;       // workspace: (B,) segments completed per trajectory (-1: not finite any more); nseg > 1 only
    int nseg, seg;       // the nsteps time steps are cut into nseg segments of seg steps, handed out round-robin
};

constexpr int MAT_LD = 33;   // row stride of the dense reduced system (odd: conflict-free row reads)

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// geometry shared by host and device: S = nodes per k-lane run (power of two >= 8), NP = 4 S
struct PromGeom {
    int S, lgS, NP;
    __host__ __device__ static PromGeom make(int N) {
        PromGeom g;
        g.lgS = 3;
        while ((4 << g.lgS) < N) ++g.lgS;
        g.S = 1 << g.lgS;
        g.NP = 4 << g.lgS;
        return g;
    }
};

// skewed index of node (or element) i in [-1, NP]
__device__ __forceinline__ int sk(int i, int S, int lgS) { return i + 1 + ((i + S) >> lgS); }
// permuted Phi row of natural node i = t S + s  ->  4 s + t
__device__ __forceinline__ int perm_row(int i, int S, int lgS) { return ((i & (S - 1)) << 2) + (i >> lgS); }

// n x n solve, one matrix row per lane (lane < n), columns in registers.
// a[j]: row `lane` of Ar (columns >= n and rows >= n may hold anything finite: they never act as
// pivots and only receive harmless updates); rhs: -br[lane].  Returns x[lane].
// PIVOT: partial pivoting in the elimination order of LAPACK dgetf2 (first maximal |a_ik| among the
// remaining rows, scaling by the reciprocal pivot), as np.linalg.solve -> gesv does (:767); rows
// never move (implicit pivoting by lane).  !PIVOT: the LSPG normal matrix J^T J is symmetric
// positive definite, for which elimination in natural order is backward stable (it is the
// Cholesky recurrence without the square roots).
// The pivot row of step k is published through a double-buffered shared-memory row
// (pb: 2 x (NC+2) doubles, 16-byte aligned), moved with 128-bit accesses.
template <int NC, bool PIVOT>
__device__ __forceinline__ double lu_solve_rows(double (&a)[NC], double rhs, int n, int lane, double *pb) {
    constexpr int PLD = NC + 2;
    bool avail = lane < n;   // row not yet used as a pivot row
    int mystep = -1;         // elimination step at which this lane's row became the pivot row
    int pls[NC];             // pivot lane of every step (warp-uniform)
    double pinv = 0.0;       // 1 / pivot of the step this lane's row served
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        pls[k] = k;
        if (k < n) {
            int pl = k;
            if (PIVOT) {   // argmax |a[k]| over the available rows (idamax of dgetf2)
                const unsigned long long key = avail ? (unsigned long long)__double_as_longlong(fabs(a[k])) + 1ull : 0ull;
                const unsigned khi = (unsigned)(key >> 32), klo = (unsigned)key;
                const unsigned mhi = __reduce_max_sync(FULL, khi);
                const bool c1 = khi == mhi;
                const unsigned mlo = __reduce_max_sync(FULL, c1 ? klo : 0u);
                const unsigned ball = __ballot_sync(FULL, c1 && klo == mlo && avail);
                pl = ball ? (__ffs(ball) - 1) : 0;
                pls[k] = pl;
            }
            const int kk = k & ~1;
            double2 *row2 = reinterpret_cast<double2 *>(pb + (k & 1) * PLD);
            if (lane == pl) {
#pragma unroll
                for (int j = kk; j < NC; j += 2) row2[j >> 1] = make_double2(a[j], a[j + 1]);
                pb[(k & 1) * PLD + NC] = rhs;
            }
            __syncwarp();
            double pr[NC];
#pragma unroll
            for (int j = kk; j < NC; j += 2) {
                const double2 v = row2[j >> 1];
                pr[j] = v.x; pr[j + 1] = v.y;
            }
            const double inv = rcp(pr[k]);
            const double l = (avail && lane != pl) ? a[k] * inv : 0.0;
#pragma unroll
            for (int j = k + 1; j < NC; ++j) a[j] = fma(-l, pr[j], a[j]);
            rhs = fma(-l, pb[(k & 1) * PLD + NC], rhs);
            if (lane == pl) { avail = false; mystep = k; pinv = inv; }
        }
    }
    // back substitution, column oriented: the pivot row of step k yields x[k]
    double x = 0.0;
#pragma unroll
    for (int k = NC - 1; k >= 0; --k) {
        if (k < n) {
            const double xk = shfl_idx(rhs * pinv, pls[k]);
            if (mystep < k) rhs = fma(-a[k], xk, rhs);
            if (lane == k) x = xk;
        }
    }
    return x;
}

template <int NB, bool LSPG> struct PromTiles {
    static constexpr int COUNT = LSPG ? NB * (NB + 1) / 2 : NB * NB;
    __host__ __device__ static constexpr int index(int I, int J) {   // LSPG: I <= J
        return LSPG ? (I * NB - I * (I - 1) / 2 + (J - I)) : I * NB + J;
    }
};

// One row of A and of R for node i (skewed index o; om / op: its neighbours), FEM/fem_burgers.py:730-752.
struct PromRow { double a, b, c, R; };
struct PromRowCtx {
    const double *u, *bn, *hfs, *h6;
    double a6, c1, ate, mu1;
    int N, S, lgS;
    bool has_e;
    bool supg;   // false: no SUPG term in b (pod_quadratic_manifold, FEM/fem_burgers.py:1141)
};
__device__ __forceinline__ PromRow prom_row(const PromRowCtx &x, int i, int o) {
    const int om = sk(i - 1, x.S, x.lgS), op = sk(i + 1, x.S, x.lgS);
    const double um = x.u[om], uc = x.u[o], up = x.u[op];
    double pe = 0.0, Ge = 0.0, qL = 0.0, GL = 0.0;
    if (i <= x.N - 2) {   // right element (i, i+1): convection (:418) and SUPG (:569-575)
        const double sm = uc + up;
        pe = fma(x.a6, uc, x.a6 * sm);
        if (x.supg) {
            const double z = fma(sm * (up - uc), 1.0 / 6.0, -x.hfs[o]);
            Ge = (x.c1 * rcp(fmax(fabs(sm), 2.0e-10))) * z;
        }
    }
    if (i >= 1 && i <= x.N - 1) {   // left element (i-1, i)
        const double sm = um + uc;
        qL = fma(x.a6, uc, x.a6 * sm);
        if (x.supg) {
            const double z = fma(sm * (uc - um), 1.0 / 6.0, -x.hfs[om]);
            GL = (x.c1 * rcp(fmax(fabs(sm), 2.0e-10))) * z;
        }
    }
    const double h6L = x.h6[om], h6R = x.h6[o];
    double keL = 0.0, keR = 0.0;
    if (x.has_e) {
        keL = h6L > 0.0 ? x.ate / (6.0 * h6L) : 0.0;
        keR = h6R > 0.0 ? x.ate / (6.0 * h6R) : 0.0;
    }
    PromRow r;
    r.a = (h6L - keL) - qL;
    r.c = (h6R - keR) + pe;
    r.b = (2.0 * (h6L + h6R) + (keL + keR)) + (qL - pe);
    r.R = fma(r.a, um, fma(r.b, uc, r.c * up)) - (x.bn[o] - GL + Ge);
    if (i == 0) { r.a = 0.0; r.b = 1.0; r.c = 0.0; r.R = uc - x.mu1; }   // Dirichlet row (:741-742, :748)
    if (i >= x.N) { r.a = 0.0; r.b = 0.0; r.c = 0.0; r.R = 0.0; }
    return r;
}

}  // namespace bx
