// bx_prom_wide.cu — POD-PROM trajectory kernel for 32 <= n <= 47 modes, one WARP per trajectory:
// FEMBurgers.pod_prom_burgers (FEM/fem_burgers.py:709-785), Galerkin and LSPG.  The reference's own driver
// runs 40 modes (POD/Results_thesis/prom_pod.py:35-39, U_modes_tol_1e-03.npy).
//
// Same design as bx_prom.cu (the n <= 31 kernel: contraction on the FP64 tensor pipe with the k index = node, every
// accumulator tile in registers, no CTA-wide barrier, rows of A and R assembled eight nodes ahead, one node per lane),
// widened to NB = 5 or 6 column blocks:
//   * the basis does not fit next to the per-warp state any more (512 x 44 doubles = 180 KB): Phi rows come from
//     global memory (L1 / L2 resident: 164 KB at n = 40, shared by every warp of the device);
//   * n = 40 exactly (the shipped size) keeps NB = 5: R does not ride as column n of the right operand (that would
//     cost a sixth block: 21 instead of 15 LSPG tiles, 36 instead of 25 Galerkin tiles) — br = J^T R (LSPG) or
//     Phi^T R (Galerkin) is accumulated beside the tiles with NB DFMA per k-step and reduced over the four k-lanes
//     once per iteration;
//   * the n x n reduced system no longer fits "one row per lane": the accumulator tiles go to a per-warp
//     shared-memory matrix (row stride NC + 1, odd: conflict-free row access), LU with partial pivoting in LAPACK
//     dgetf2's order (first maximal |a_ik|, physical row swap, scaling by the reciprocal pivot: np.linalg.solve ->
//     gesv, :767) for Galerkin, natural order for the SPD LSPG normal matrix; lanes own rows lane and lane + 32;
//   * every per-mode vector (q, dq, q0) has two entries per lane (modes lane and lane + 32).
// Work units are segments of trajectories handed out round-robin, as in bx_prom.cu / bx_fom.cu.
#include "bx_prom_common.cuh"
#include <utility>

namespace bx {

constexpr int PW_MAXW = 8;

// compile-time loop: f(integral_constant<int, 0>) ... f(integral_constant<int, N - 1>), in order (the elimination
// steps index register arrays, so the step number has to be a constant: `#pragma unroll` gives up on loops this size)
template <class F, int... Is>
__device__ __forceinline__ void static_for_impl(F &&f, std::integer_sequence<int, Is...>) {
    (f(std::integral_constant<int, Is>{}), ...);
}
template <int N, class F> __device__ __forceinline__ void static_for(F &&f) {
    static_for_impl(static_cast<F &&>(f), std::make_integer_sequence<int, N>{});
}

struct PromWideSmem {
    int h6, gm, per_cta;
    int u, bn, hfs, mat, q, per_warp, arr, ldm;
    __host__ __device__ static PromWideSmem make(int NP, int NC) {
        PromWideSmem m;
        m.arr = NP + 8;
        m.ldm = NC + 1;                    // column NC = right-hand side
        int o = 0;
        m.h6 = o; o += m.arr;
        o = (o + 1) & ~1;
        m.gm = o; o += NC * m.ldm;         // Phi^T Phi
        m.per_cta = (o + 1) & ~1;
        o = 0;
        m.u = o; o += (m.arr + 1) & ~1;
        m.bn = o; o += (m.arr + 1) & ~1;
        m.hfs = o; o += (m.arr + 1) & ~1;
        m.mat = o; o += NC * m.ldm + 1;
        o = (o + 1) & ~1;
        m.q = o; o += NC + 8;
        m.per_warp = (o + 1) & ~1;
        return m;
    }
};

// n x n solve, 32 < n <= 64, TWO matrix rows per lane (rows `lane` and `lane + 32`), columns in registers: the
// two-row version of lu_solve_rows (bx_prom_common.cuh).  a0 / a1: the lane's rows of Ar (a1 all zero when
// lane + 32 >= n), rhs0 / rhs1 = -br.  PIVOT: partial pivoting in the elimination order of LAPACK dgetf2 (maximal
// |a_ik| among the remaining rows, scaling by the reciprocal pivot; np.linalg.solve -> gesv, :767), rows never move
// (implicit pivoting).  !PIVOT: natural order (the SPD LSPG normal matrix).  The pivot row of step k goes through a
// double-buffered shared-memory row (pb: 2 x (NC + 2) doubles, 16-byte aligned).  Returns x[lane], x[lane + 32].
template <int NC, bool PIVOT>
__device__ __forceinline__ void lu_solve_rows2(double (&a0)[NC], double (&a1)[NC], double rhs0, double rhs1, int n,
                                               int lane, double *pb, double &x0, double &x1) {
    constexpr int PLD = NC + 2;
    bool av0 = lane < n, av1 = lane + 32 < n;   // row not yet used as a pivot row
    int ms0 = -1, ms1 = -1;                     // elimination step at which the row became the pivot row
    double pinv0 = 0.0, pinv1 = 0.0;
    int prs[NC];                                // pivot row of every step (warp-uniform)
    static_for<NC>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        prs[k] = k;
        if (k < n) {
            int pr = k;
            if (PIVOT) {
                const double v0 = av0 ? fabs(a0[k]) : -1.0, v1 = av1 ? fabs(a1[k]) : -1.0;
                const bool take1 = v1 > v0;
                const double v = take1 ? v1 : v0;
                const unsigned long long key = v >= 0.0 ? (unsigned long long)__double_as_longlong(v) + 1ull : 0ull;
                const unsigned khi = (unsigned)(key >> 32), klo = (unsigned)key;
                const unsigned mhi = __reduce_max_sync(FULL, khi);
                const bool c1 = khi == mhi;
                const unsigned mlo = __reduce_max_sync(FULL, c1 ? klo : 0u);
                const unsigned cand = (c1 && klo == mlo && key != 0ull) ? (unsigned)(lane + (take1 ? 32 : 0)) : 0xffffu;
                const unsigned best = __reduce_min_sync(FULL, cand);
                pr = best == 0xffffu ? k : (int)best;
                prs[k] = pr;
            }
            const int pl = pr & 31;
            const bool ps = pr >= 32;
            const int kk = k & ~1;
            double2 *row2 = reinterpret_cast<double2 *>(pb + (k & 1) * PLD);
            if (lane == pl) {
#pragma unroll
                for (int j = kk; j < NC; j += 2)
                    row2[j >> 1] = ps ? make_double2(a1[j], a1[j + 1]) : make_double2(a0[j], a0[j + 1]);
                pb[(k & 1) * PLD + NC] = ps ? rhs1 : rhs0;
            }
            __syncwarp();
            double prow[NC];
#pragma unroll
            for (int j = kk; j < NC; j += 2) {
                const double2 v = row2[j >> 1];
                prow[j] = v.x; prow[j + 1] = v.y;
            }
            const double prhs = pb[(k & 1) * PLD + NC];
            const double inv = rcp(prow[k]);
            const bool self0 = lane == pl && !ps, self1 = lane == pl && ps;
            const double l0 = (av0 && !self0) ? a0[k] * inv : 0.0;
            const double l1 = (av1 && !self1) ? a1[k] * inv : 0.0;
#pragma unroll
            for (int j = k + 1; j < NC; ++j) { a0[j] = fma(-l0, prow[j], a0[j]); a1[j] = fma(-l1, prow[j], a1[j]); }
            rhs0 = fma(-l0, prhs, rhs0);
            rhs1 = fma(-l1, prhs, rhs1);
            if (self0) { av0 = false; ms0 = k; pinv0 = inv; }
            if (self1) { av1 = false; ms1 = k; pinv1 = inv; }
        }
    });
    x0 = 0.0; x1 = 0.0;
    static_for<NC>([&](auto kc) {
        constexpr int k = NC - 1 - decltype(kc)::value;
        if (k < n) {
            const int pr = prs[k];
            const double mine = pr >= 32 ? rhs1 * pinv1 : rhs0 * pinv0;
            const double xk = shfl_idx(mine, pr & 31);
            if (ms0 < k) rhs0 = fma(-a0[k], xk, rhs0);
            if (ms1 < k) rhs1 = fma(-a1[k], xk, rhs1);
            if (lane == (k & 31)) { if (k >= 32) x1 = xk; else x0 = xk; }
        }
    });
}

// Variant: the same solve out of shared memory (rows r < n at M + r * ldm, right-hand side in column NC), lanes own rows
// `lane` and `lane + 32`; physical row swaps in dgetf2's order.  Cost grows with n^3 (the register version always pays
// for NC columns) and it needs no registers beside the contraction's.  Returns with the solution in column NC.
template <int NC, bool PIVOT>
__device__ __forceinline__ void lu_solve_smem(double *M, int ldm, int n, int lane) {
    const int r0 = lane, r1 = lane + 32;
    for (int k = 0; k < n; ++k) {
        if (PIVOT) {   // idamax over rows k .. n-1 of column k: first maximal |a_ik|
            double v = -1.0; int idx = 0x7fffffff;
            if (r0 >= k && r0 < n) { v = fabs(M[r0 * ldm + k]); idx = r0; }
            if (r1 >= k && r1 < n) { const double w = fabs(M[r1 * ldm + k]); if (w > v) { v = w; idx = r1; } }
            const unsigned long long key = v >= 0.0 ? (unsigned long long)__double_as_longlong(v) + 1ull : 0ull;
            const unsigned khi = (unsigned)(key >> 32), klo = (unsigned)key;
            const unsigned mhi = __reduce_max_sync(FULL, khi);
            const bool c1 = khi == mhi;
            const unsigned mlo = __reduce_max_sync(FULL, c1 ? klo : 0u);
            const unsigned piv = __reduce_min_sync(FULL, (c1 && klo == mlo && key != 0ull) ? (unsigned)idx : 0x7fffffffu);
            const int pr = piv == 0x7fffffffu ? k : (int)piv;   // (all candidates NaN: keep the diagonal)
            if (pr != k) {   // physical swap of rows k and pr, columns k .. n-1 and the right-hand side
                for (int j = k + lane; j <= NC; j += 32) {
                    if (j < n || j == NC) {
                        const double a = M[k * ldm + j], b = M[pr * ldm + j];
                        M[k * ldm + j] = b; M[pr * ldm + j] = a;
                    }
                }
            }
            __syncwarp();
        }
        const double inv = rcp(M[k * ldm + k]);
        const double *pk = M + k * ldm;
        if (r0 > k && r0 < n) {
            double *row = M + r0 * ldm;
            const double l = row[k] * inv;
            for (int j = k + 1; j < n; ++j) row[j] = fma(-l, pk[j], row[j]);
            row[NC] = fma(-l, pk[NC], row[NC]);
        }
        if (r1 > k && r1 < n) {
            double *row = M + r1 * ldm;
            const double l = row[k] * inv;
            for (int j = k + 1; j < n; ++j) row[j] = fma(-l, pk[j], row[j]);
            row[NC] = fma(-l, pk[NC], row[NC]);
        }
        __syncwarp();
    }
    for (int k = n - 1; k >= 0; --k) {   // back substitution, column oriented
        const double xk = M[k * ldm + NC] * rcp(M[k * ldm + k]);
        __syncwarp();
        if (lane == 0) M[k * ldm + NC] = xk;
        if (r0 < k) M[r0 * ldm + NC] = fma(-M[r0 * ldm + k], xk, M[r0 * ldm + NC]);
        if (r1 < k) M[r1 * ldm + NC] = fma(-M[r1 * ldm + k], xk, M[r1 * ldm + NC]);
        __syncwarp();
    }
}

// RSEP: n == 8 NB, R does not ride as a column of the right operand (br accumulated beside the tiles).
// SLU: reduced solve out of shared memory (lu_solve_smem) instead of registers (lu_solve_rows2).  Measured at 4096
// samples (profiles/r2u_prom_wide_variants.txt): the pivoted solve is faster out of shared memory at every n (n = 40
// Galerkin 7.7 against 6.0 M it/s), the unpivoted LSPG solve is faster in registers from n = 36 on (n = 40: 11.1 against
// 9.2 M it/s; n = 32: 12.0 against 13.2).  A deeper prefetch of the Phi rows (three k-steps ahead) changed nothing.
template <int NB, bool LSPG, bool RSEP, bool SLU>
__global__ void __launch_bounds__(PW_MAXW * 32, 1) prom_wide_kernel(const PromParams p) {
    constexpr int NC = NB * 8;
    using Tl = PromTiles<NB, LSPG>;
    constexpr int TILES = Tl::COUNT;
    extern __shared__ __align__(16) double smem[];

    const int N = p.N, n = p.n;
    const PromGeom geo = PromGeom::make(N);
    const int S = geo.S, lgS = geo.lgS, NP = geo.NP;
    const PromWideSmem lay = PromWideSmem::make(NP, NC);
    const int ldm = lay.ldm;
    double *s_h6 = smem + lay.h6;
    double *s_gm = smem + lay.gm;          // Phi^T Phi, NC x ldm

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t = lane & 3, g = lane >> 2;
    double *wbase = smem + lay.per_cta + warp * lay.per_warp;
    double *w_u = wbase + lay.u;
    double *w_bn = wbase + lay.bn;
    double *w_hfs = wbase + lay.hfs;
    double *w_mat = wbase + lay.mat;
    double *w_q = wbase + lay.q;
    const double At = p.At;

    // ---- CTA prologue: mesh constants and Phi^T Phi ------------------------------------------------
    for (int idx = threadIdx.x; idx < lay.arr; idx += blockDim.x) s_h6[idx] = 0.0;
    for (int idx = threadIdx.x; idx < NC * NC; idx += blockDim.x) {
        const int j = idx / NC, m = idx - j * NC;
        double sum = 0.0;
        if (j < n && m < n)
            for (int i = 0; i < N; ++i) sum = fma(p.Phi[(size_t)i * n + j], p.Phi[(size_t)i * n + m], sum);
        s_gm[j * ldm + m] = sum;
    }
    __syncthreads();
    for (int e = threadIdx.x; e <= N - 2; e += blockDim.x) s_h6[sk(e, S, lgS)] = (p.X[e + 1] - p.X[e]) / 6.0;
    __syncthreads();

    const int run0 = t << lgS;
    const int ubase = t * (S + 1) + 2;                       // sk(run0 + s) = ubase + s
    const int gR = n & 7;                                    // (!RSEP) lane-group of the R column (its block is NB-1)
    const bool col_last = ((NB - 1) * 8 + g) < n;            // the ragged last block: columns >= n read as zero
    const int nblk = S >> 3;
    const int m0 = lane, m1 = lane + 32;                     // the two modes a lane owns in per-mode vectors
    const bool has1 = m1 < n;

    PromRowCtx rc;
    rc.u = w_u; rc.bn = w_bn; rc.hfs = w_hfs; rc.h6 = s_h6;
    rc.a6 = At / 6.0; rc.c1 = 1.5 * At; rc.ate = At * p.E; rc.has_e = p.E != 0.0; rc.supg = true;
    rc.N = N; rc.S = S; rc.lgS = lgS;

    for (;;) {
        unsigned unit = 0;
        if (lane == 0) unit = atomicAdd(p.counter, 1u);
        unit = __shfl_sync(FULL, unit, 0);
        if (unit >= (unsigned)p.B * (unsigned)p.nseg) break;
        const int sg = (int)(unit / (unsigned)p.B);
        const int smp = (int)(unit - (unsigned)sg * (unsigned)p.B);
        const int n_begin = sg * p.seg;
        const int n_end = min(p.nsteps, n_begin + p.seg);
        rc.mu1 = p.mu1[smp];
        const double *atf = p.atf + (size_t)smp * N;
        double *Us = p.U + (long long)smp * (long long)(p.nsteps + 1) * N;
        {
            const double *hf = p.hfs + (size_t)smp * N;
            const double *src = p.u0 + (long long)smp * p.u0_stride;
            double poison = 0.0;
            if (sg > 0) {
                int done;
                const volatile int *flag = p.progress + smp;
                while ((done = *flag) < sg && done >= 0) __nanosleep(500);
                done = __shfl_sync(FULL, done, 0);
                __threadfence();
                if (done < 0) poison = __longlong_as_double(0x7ff8000000000000ll);
                src = Us + (long long)n_begin * N;
            }
            for (int i = lane - 1; i <= NP; i += 32) {
                const int o = sk(i, S, lgS);
                w_hfs[o] = (i >= 0 && i <= N - 2) ? hf[i] : 0.0;
                const double v = (i >= 0 && i < N) ? __ldcg(src + i) + poison : 0.0;
                w_u[o] = v;
                w_bn[o] = 0.0;
                if (sg == 0 && i >= 0 && i < N) Us[i] = v;
            }
        }
        __syncwarp();
        // q0 = Phi^T U_k of the first iterate of the segment (:716 / :770), two modes per lane
        double q00, q01;
        {
            double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
            const double *ph0 = p.Phi + m0, *ph1 = p.Phi + (has1 ? m1 : m0);
            int i = 0;
            for (; i + 2 <= N; i += 2) {
                const double u0v = w_u[sk(i, S, lgS)], u1v = w_u[sk(i + 1, S, lgS)];
                a0 = fma(ph0[(size_t)i * n], u0v, a0); a1 = fma(ph0[(size_t)(i + 1) * n], u1v, a1);
                b0 = fma(ph1[(size_t)i * n], u0v, b0); b1 = fma(ph1[(size_t)(i + 1) * n], u1v, b1);
            }
            if (i < N) {
                const double u0v = w_u[sk(i, S, lgS)];
                a0 = fma(ph0[(size_t)i * n], u0v, a0); b0 = fma(ph1[(size_t)i * n], u0v, b0);
            }
            q00 = a0 + a1; q01 = has1 ? b0 + b1 : 0.0;
        }

        bool bad = false;
        for (int step = n_begin; step < n_end; ++step) {
            for (int i = lane; i < N; i += 32) {   // bn = M U^n + At F (:745)
                const double h6L = s_h6[sk(i - 1, S, lgS)], h6R = s_h6[sk(i, S, lgS)];
                const double um = w_u[sk(i - 1, S, lgS)], uc = w_u[sk(i, S, lgS)], up = w_u[sk(i + 1, S, lgS)];
                w_bn[sk(i, S, lgS)] = h6L * um + (2.0 * (h6L + h6R)) * uc + h6R * up + atf[i];
            }
            __syncwarp();
            int k = 0;
            bool finite = true;
            for (;;) {
                // ================= phase B: assembly + contraction (:730-762) ============
                double acc[TILES][2];
#pragma unroll
                for (int i = 0; i < TILES; ++i) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
                double brp[NB];
#pragma unroll
                for (int b = 0; b < NB; ++b) brp[b] = 0.0;
                // Phi rows of the nodes run0 + s - 1, s, s + 1 (rolling three-row window; the rows come from L1 / L2)
                double pm[NB], p0[NB], pp[NB];
                auto phi_row = [&](double (&dst)[NB], int inx) {
                    const bool ok = inx >= 0 && inx < N;
                    const double *rp = p.Phi + (size_t)(ok ? inx : 0) * n;
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        const bool cok = (b < NB - 1) || col_last;
                        dst[b] = (cok && ok) ? rp[b * 8 + g] : 0.0;
                    }
                };
                phi_row(pm, t > 0 ? run0 - 1 : -1);
                phi_row(p0, run0);
                PromRow nxt = prom_row(rc, run0 + g, ubase + g);
                for (int blk = 0; blk < nblk; ++blk) {
                    const PromRow cur = nxt;
                    if (blk + 1 < nblk) nxt = prom_row(rc, run0 + ((blk + 1) << 3) + g, ubase + ((blk + 1) << 3) + g);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        phi_row(pp, run0 + (blk << 3) + j + 1);
                        const int src = 4 * j + t;
                        const double ra = shfl_idx(cur.a, src), rb = shfl_idx(cur.b, src);
                        const double rcc = shfl_idx(cur.c, src), rR = shfl_idx(cur.R, src);
                        double jp[NB];
#pragma unroll
                        for (int b = 0; b < NB; ++b) jp[b] = fma(ra, pm[b], fma(rb, p0[b], rcc * pp[b]));
                        if (RSEP) {
#pragma unroll
                            for (int b = 0; b < NB; ++b) brp[b] = fma(LSPG ? jp[b] : p0[b], rR, brp[b]);
                        } else {
                            if (g == gR) jp[NB - 1] = rR;   // R rides as column n (block NB-1)
                        }
#pragma unroll
                        for (int I = 0; I < NB; ++I) {
#pragma unroll
                            for (int J = LSPG ? I : 0; J < NB; ++J) {
                                const int ti = Tl::index(I, J);
                                dmma884(acc[ti][0], acc[ti][1], LSPG ? jp[I] : p0[I], jp[J]);
                            }
                        }
#pragma unroll
                        for (int b = 0; b < NB; ++b) { pm[b] = p0[b]; p0[b] = pp[b]; }
                    }
                }
                // ================= phase C: reduced solve in shared memory =================
                __syncwarp();
#pragma unroll
                for (int I = 0; I < NB; ++I) {
#pragma unroll
                    for (int J = LSPG ? I : 0; J < NB; ++J) {
                        const int ti = Tl::index(I, J);
                        const int r = I * 8 + g, c = J * 8 + 2 * t;
                        w_mat[r * ldm + c] = acc[ti][0];
                        w_mat[r * ldm + c + 1] = acc[ti][1];
                        if (LSPG && J > I) {
                            w_mat[c * ldm + r] = acc[ti][0];
                            w_mat[(c + 1) * ldm + r] = acc[ti][1];
                        }
                    }
                }
                if (RSEP) {   // br: reduce the four k-lanes, lane t == 0 holds mode 8 b + g
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        double v = brp[b];
                        v += shfl_xor(v, 1); v += shfl_xor(v, 2);
                        if (t == 0) w_mat[(8 * b + g) * ldm + NC] = -v;
                    }
                    __syncwarp();
                } else {
                    __syncwarp();
                    if (m0 < n) w_mat[m0 * ldm + NC] = -w_mat[m0 * ldm + n];
                    if (has1) w_mat[m1 * ldm + NC] = -w_mat[m1 * ldm + n];
                    __syncwarp();
                }
                double dq0, dq1;
                if constexpr (SLU) {
                    lu_solve_smem<NC, !LSPG>(w_mat, ldm, n, lane);
                    dq0 = w_mat[m0 * ldm + NC];
                    dq1 = has1 ? w_mat[m1 * ldm + NC] : 0.0;
                } else {
                    double a0[NC], a1[NC];
                    const double *mr0 = w_mat + m0 * ldm, *mr1 = w_mat + (has1 ? m1 : m0) * ldm;
#pragma unroll
                    for (int j = 0; j < NC; ++j) { a0[j] = mr0[j]; a1[j] = has1 ? mr1[j] : 0.0; }
                    const double rhs0 = mr0[NC], rhs1 = has1 ? mr1[NC] : 0.0;
                    __syncwarp();   // every row is in registers: the matrix area now carries the pivot rows
                    lu_solve_rows2<NC, !LSPG>(a0, a1, rhs0, rhs1, n, lane, w_mat, dq0, dq1);
                    if (!has1) dq1 = 0.0;
                }
                const double qa = m0 < n ? q00 + dq0 : 0.0;   // q = Phi^T U_k + dq (:770)
                const double qb = has1 ? q01 + dq1 : 0.0;
                const double sd = warp_sum(fma(dq0, dq0, dq1 * dq1));
                const double sq = warp_sum(fma(qa, qa, qb * qb));
                const double err = sqrt(sd) / sqrt(sq);
                __syncwarp();
                w_q[m0] = qa;
                if (m1 < NC) w_q[m1] = qb;
                __syncwarp();
                // Phi^T U_{k+1} for the next iteration: U_{k+1} = Phi q, so it is (Phi^T Phi) q
                {
                    const double *g0 = s_gm + (m0 < n ? m0 : 0) * ldm, *g1 = s_gm + (has1 ? m1 : 0) * ldm;
                    double s0 = 0.0, s1 = 0.0;
                    for (int m = 0; m < n; ++m) { const double qm = w_q[m]; s0 = fma(g0[m], qm, s0); s1 = fma(g1[m], qm, s1); }
                    q00 = m0 < n ? s0 : 0.0; q01 = has1 ? s1 : 0.0;
                }
                // ================= phase D: lift U_{k+1} = Phi q (:773) ===================
                {
                    double qv[NB];
#pragma unroll
                    for (int b = 0; b < NB; ++b) qv[b] = w_q[b * 8 + g];   // zero for modes >= n
                    for (int blk = 0; blk < nblk; ++blk) {
                        double v[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int i = run0 + (blk << 3) + j;
                            const double *r0p = p.Phi + (size_t)(i < N ? i : 0) * n;
                            double vv = 0.0;
#pragma unroll
                            for (int b = 0; b < NB; ++b) {
                                const bool cok = (b < NB - 1) || col_last;
                                vv = fma((cok && i < N) ? r0p[b * 8 + g] : 0.0, qv[b], vv);
                            }
                            v[j] = vv;
                        }
                        double w4[4], w2[2];
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const double send = (g & 4) ? v[j] : v[j + 4];
                            const double keep = (g & 4) ? v[j + 4] : v[j];
                            w4[j] = keep + shfl_xor(send, 16);
                        }
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            const double send = (g & 2) ? w4[j] : w4[j + 2];
                            const double keep = (g & 2) ? w4[j + 2] : w4[j];
                            w2[j] = keep + shfl_xor(send, 8);
                        }
                        const double send = (g & 1) ? w2[0] : w2[1];
                        const double keep = (g & 1) ? w2[1] : w2[0];
                        w_u[ubase + (blk << 3) + g] = keep + shfl_xor(send, 4);
                    }
                }
                __syncwarp();
                if (p.err_hist && lane == 0) p.err_hist[((long long)smp * p.nsteps + step) * p.kmax + k] = err;
                ++k;
                finite = (sd + sq) <= 1.7976931348623157e308;   // false for inf and NaN
                if (!(err > p.tol) || k >= p.kmax) break;
            }
            double *Un1 = Us + (long long)(step + 1) * N;
            for (int i = lane; i < N; i += 32) Un1[i] = w_u[sk(i, S, lgS)];
            if (lane == 0 && p.iters) p.iters[(long long)smp * p.nsteps + step] = finite ? k : -1;
            if (!finite) {
                const double qnan = __longlong_as_double(0x7ff8000000000000ll);
                for (++step; step < n_end; ++step) {
                    double *Un2 = Us + (long long)(step + 1) * N;
                    for (int i = lane; i < N; i += 32) Un2[i] = qnan;
                    if (lane == 0 && p.iters) p.iters[(long long)smp * p.nsteps + step] = -1;
                }
                bad = true;
            }
        }
        if (p.nseg > 1) {
            __threadfence();
            __syncwarp();
            *reinterpret_cast<volatile int *>(p.progress + smp) = bad ? -1 : sg + 1;
        }
        __syncwarp();
    }
}

// defined in bx_prom.cu: per-sample forcing data At F and (h/6)(f1 + f2) into the workspace
int launch_prom_forcing(int B, int N, const double *X, const double *mu2, double At, double *atf, double *hfs,
                        cudaStream_t st);

template <int NB, bool LSPG, bool RSEP, bool SLU>
static int launch_wide_t(PromParams p, cudaStream_t st) {
    const PromGeom geo = PromGeom::make(p.N);
    const PromWideSmem lay = PromWideSmem::make(geo.NP, NB * 8);
    DeviceInfo di;
    {
        const int rcd = device_info(di);
        if (rcd != BX_OK) return rcd;
    }
    const long long avail = (long long)di.max_smem_optin / (long long)sizeof(double) - lay.per_cta;
    long long warps = avail > 0 ? avail / lay.per_warp : 0;
    if (warps > PW_MAXW) warps = PW_MAXW;
    if (warps < 1) {
        set_error("bx_prom_trajectories: N = %d, n = %d does not fit in shared memory (register kernel)", p.N, p.n);
        return BX_ERR_UNSUPPORTED;
    }
    {   // a sweep smaller than the device spreads over all SMs first
        const long long per_sm = ((long long)p.B + di.sms - 1) / di.sms;
        if (warps > per_sm) warps = per_sm < 1 ? 1 : per_sm;
    }
    p.warps = (int)warps;
    const size_t smem = sizeof(double) * ((size_t)lay.per_cta + (size_t)warps * lay.per_warp);
    auto kern = prom_wide_kernel<NB, LSPG, RSEP, SLU>;
    {
        const int rco = kernel_occupancy(reinterpret_cast<const void *>(kern), (int)warps * 32, smem, nullptr);
        if (rco != BX_OK) return rco;
    }
    long long ctas = di.sms;
    const long long need = ((long long)p.B + warps - 1) / warps;
    p.nseg = 1; p.seg = p.nsteps > 0 ? p.nsteps : 1;
    if (ctas >= need) {
        ctas = need;
    } else if (!p.err_hist) {
        int seg = tuning(BX_TUNE_FOM_SEGMENT_STEPS);
        if (seg < 0) seg = 25;
        if (seg > 0 && seg < p.nsteps) { p.seg = seg; p.nseg = (p.nsteps + seg - 1) / seg; }
    }
    BX_REQUIRE((long long)p.B * p.nseg < 0xffffffffLL, "too many work units (B = %d, %d segments)", p.B, p.nseg);
    BX_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st));
    if (p.nseg > 1) BX_CUDA(cudaMemsetAsync(p.progress, 0, sizeof(int) * (size_t)p.B, st));
    const int rcf = launch_prom_forcing(p.B, p.N, p.X, p.mu2, p.At, p.atf, p.hfs, st);
    if (rcf != BX_OK) return rcf;
    kern<<<(unsigned)ctas, (unsigned)warps * 32, smem, st>>>(p);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

template <bool LSPG>
static int launch_wide(const PromParams &p, cudaStream_t st) {
    constexpr bool G = !LSPG;   // Galerkin: pivoted solve, always out of shared memory
    if (p.n == 40) return launch_wide_t<5, LSPG, true, G>(p, st);
    if (p.n < 40) {
        if (LSPG && p.n <= 33) return launch_wide_t<5, LSPG, false, true>(p, st);
        return launch_wide_t<5, LSPG, false, G>(p, st);
    }
    if (p.n == 48) return launch_wide_t<6, LSPG, true, G>(p, st);
    return launch_wide_t<6, LSPG, false, G>(p, st);
}

int launch_prom_wide(const PromParams &p, bool lspg, cudaStream_t st) {
    if (p.n < 32 || p.n > 48) {
        set_error("bx_prom_trajectories: internal dispatch error (register kernel, n = %d)", p.n);
        return BX_ERR_UNSUPPORTED;
    }
    return lspg ? launch_wide<true>(p, st) : launch_wide<false>(p, st);
}

}  // namespace bx
