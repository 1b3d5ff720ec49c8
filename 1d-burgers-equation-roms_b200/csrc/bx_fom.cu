// bx_fom.cu — fused FOM trajectory kernel: FEMBurgers.fom_burgers (FEM/fem_burgers.py:646-707)
// for B (mu1, mu2) samples at once.
//
// W WARPS (1, 2 or 4) integrate one trajectory from u0 to step nsteps without leaving the SM:
//   * lane l of the group (l = 0 .. 32 W - 1) owns the M consecutive nodes l*M .. l*M+M-1 in
//     registers (32 W M >= N);
//   * every Picard iteration assembles the tridiagonal A = M + At C(U_k) + At E K, the
//     right-hand side b = M U^n + At F - At S(U_k) and the residual R = A U_k - b on the fly
//     from the closed forms of the 2-point-Gauss element integrals (no band ever reaches HBM);
//   * A delta = -R is solved with a partition (Wang / SPIKE-type) method: each lane runs a
//     Thomas sweep in continuant form over its M rows carrying the coupling to the previous lane's
//     last unknown, an upward sweep condenses the lane to one row of a 32 W-row tridiagonal system
//     in the lanes' last unknowns;
//       W = 1: that system is solved by parallel cyclic reduction with warp shuffles;
//       W > 1: the lanes publish their condensed rows in shared memory (ONE named barrier), every
//              warp of the group then solves the whole 32 W-row system redundantly — W rows per
//              lane, the same continuant sweep + 32-row PCR one level up — so no second exchange
//              is needed before the local back-substitution;
//   * the update, |delta|^2 and |U_{k+1}|^2 are fused into the back-substitution, the convergence
//     test (tol, kmax of :663) is evaluated per trajectory (W > 1: partial norms and the two halo
//     values per warp cross in shared memory behind a second named barrier, summed in warp order by
//     every warp, so all warps of a group take the same decision);
//   * the only HBM traffic is the 8N-byte snapshot per step (+ 4 B iteration count).
// W = 1 is the throughput instance (a full sweep keeps every SM sub-partition busy with independent
// trajectories).  W = 2 / 4 shorten the serial chain of one Picard iteration for sweeps with fewer
// trajectories than the GPU has warp slots (strong scaling: 4096 samples over 8 GPUs = 512 per GPU).
// Groups pull samples from an atomic work queue (persistent CTAs).
#include "bx_common.cuh"
#include <cmath>

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
    double *err_hist;     // optional (B, nsteps, kmax): |delta| / |U_{k+1}| after every iteration (:698)
    unsigned int *counter;
};

// shared-memory carve-up (doubles), L = 32 W lanes per trajectory group: element arrays have
// (M+1)*L entries, indexed (j+1)*L + l for the element between local nodes j and j+1 of group lane l
// (j = -1 is the element shared with the previous lane); node arrays have M*L entries indexed
// j*L + l.  Both are bank-conflict free for 64-bit accesses.
template <int M, int W> struct FomSmem {
    static constexpr int L = 32 * W;
    static constexpr int EL = (M + 1) * L;
    static constexpr int NL = M * L;
    static constexpr int MESH = 2 * EL + NL;                 // off, h6 | dg
    static constexpr int XCH = W > 1 ? 6 * L + W * L + 4 * W + 4 : 0;   // condensed rows | per-warp solution | halo, norms | sample
    static constexpr int PER_GROUP = EL + 2 * NL + XCH;      // hfs | AtF, bn | exchange
    static constexpr size_t bytes(int groups) { return sizeof(double) * (MESH + (size_t)groups * PER_GROUP); }
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

__device__ __forceinline__ void group_bar(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// EXACT: N == 32 W M, so the only lane with a missing right-hand element is the last (local row M-1).
// TRACE: also record |delta| / |U_{k+1}| of every iteration (the verbose mode of the legacy wrapper).
template <int M, int W, int WARPS, int MINB, bool EXACT, bool TRACE>
__global__ void __launch_bounds__(WARPS * 32, MINB) fom_picard_kernel(const FomParams p) {
    using L = FomSmem<M, W>;
    constexpr int LN = L::L;
    static_assert(WARPS % W == 0, "whole groups per CTA");
    extern __shared__ __align__(16) double smem[];
    double *s_off = smem;             // h/6 - At E / h   (off-diagonal constant of A; 0 where no element)
    double *s_h6 = s_off + L::EL;     // h/6              (mass off-diagonal; read once per time step)
    double *s_dg = s_h6 + L::EL;      // diagonal constant of A (1 for padded rows)
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int grp = warp / W, wq = warp % W;     // group in the CTA, warp in the group
    const int gl = wq * 32 + lane;               // lane in the group
    double *w_hfs = s_dg + L::NL + grp * L::PER_GROUP;  // (h/6) (f_gp1 + f_gp2) per element
    double *w_atf = w_hfs + L::EL;                      // At * F per node
    double *w_bn = w_atf + L::NL;                       // M U^n + At F per node (per time step)
    // exchange area of the group (W > 1)
    double *x_F = w_bn + L::NL;       // F, G, R of row 0 of lane l+1, stored at [l]
    double *x_G = x_F + LN;
    double *x_R = x_G + LN;
    double *x_fl = x_R + LN;          // last local row of lane l: coupling to P, to x[0] of lane l+1, rhs
    double *x_cl = x_fl + LN;
    double *x_rl = x_cl + LN;
    double *x_sol = x_rl + LN + wq * LN;   // this warp's copy of the 32 W condensed unknowns
    double *x_halo = x_rl + LN + W * LN;   // [2 w] = first node of warp w, [2 w + 1] = its last node
    double *x_nrm = x_halo + 2 * W;        // [2 w] = |delta|^2 partial, [2 w + 1] = |U|^2 partial
    int *x_smp = reinterpret_cast<int *>(x_nrm + 2 * W);
    const int bar_id = 1 + grp;

    const int N = p.N;
    const double At = p.At;

    // ---- mesh constants, once per CTA (X is shared by all samples) ----------------------
    for (int idx = threadIdx.x; idx < L::EL; idx += WARPS * 32) {
        const int l = idx % LN, jj = idx / LN - 1;
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
        const int l = idx % LN, j = idx / LN;
        const int i = l * M + j;
        double dg = 1.0;
        if (i < N) {
            dg = 0.0;
            if (i >= 1) { const double h = p.X[i] - p.X[i - 1]; dg += 2.0 * (h / 6.0) + At * p.E / h; }
            if (i <= N - 2) { const double h = p.X[i + 1] - p.X[i]; dg += 2.0 * (h / 6.0) + At * p.E / h; }
        }
        s_dg[idx] = dg;
    }
    if (W > 1 && gl == LN - 1) { x_F[gl] = 0.0; x_G[gl] = 0.0; x_R[gl] = 0.0; }   // no lane to the right of the last
    __syncthreads();

    const int node0 = gl * M;
    const bool vec_ok = (node0 + M <= N) && ((N & 1) == 0) && ((M & 1) == 0);
    // Elements that do not exist must contribute nothing.  Padded nodes carry u = 0, so every
    // missing element has p = q = G = 0 by itself except the one to the right of the last real
    // node (u1 != 0): its p, q, G are zeroed explicitly on local row jl of the lane that owns
    // node N-1.  The element to the left of a lane (j = -1) is missing on lane 0 only (row 0 is
    // the Dirichlet row there) and on fully padded lanes.
    const int jl = N - 1 - node0;   // local index of the last real row (may be outside [0, M))
    const double a6 = At / 6.0;     // convection: p = a6 (2 u1 + u2), q = a6 (u1 + 2 u2)
    const double c1 = 1.5 * At;     // SUPG: At g = c1 (s du / 6 - (h/6) fsum) / max(|s|, 2e-10),  s = u1 + u2
    const bool leftv = gl > 0 && jl >= 0;
    const double a6L = leftv ? a6 : 0.0;
    const double c1L = leftv ? c1 : 0.0;

    for (;;) {
        int s = 0;
        if (W == 1) {
            if (lane == 0) s = (int)atomicAdd(p.counter, 1u);
            s = __shfl_sync(FULL, s, 0);
        } else {
            if (gl == 0) *x_smp = (int)atomicAdd(p.counter, 1u);
            group_bar(bar_id, LN);
            s = *x_smp;
        }
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
                w_hfs[(jj + 1) * LN + gl] = hfs;
                if (jj >= 0) w_atf[jj * LN + gl] = At * (left + c0);
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
        if (W > 1) {
            if (lane == 0) x_halo[2 * wq] = u[0];
            if (lane == 31) x_halo[2 * wq + 1] = u[M - 1];
            group_bar(bar_id, LN);   // also orders the sample slot against the next fetch
        } else {
            __syncwarp();
        }

        int n = 0;
        bool finite = true;
        for (; n < p.nsteps; ++n) {
            // ---- bn = M U^n + At F (constant during the Picard loop, :683) ---------------
            {
                double uL = shfl_up(u[M - 1], 1), uR = shfl_down(u[0], 1);
                if (W > 1) {   // the neighbouring warps' boundary nodes (published behind the last barrier)
                    if (lane == 0 && wq > 0) uL = x_halo[2 * (wq - 1) + 1];
                    if (lane == 31 && wq < W - 1) uR = x_halo[2 * (wq + 1)];
                }
                double h6L = s_h6[gl];
#pragma unroll
                for (int j = 0; j < M; ++j) {
                    const double h6R = s_h6[(j + 1) * LN + gl];
                    const double um = (j == 0) ? uL : u[j > 0 ? j - 1 : 0];
                    const double up = (j == M - 1) ? uR : u[j < M - 1 ? j + 1 : j];
                    w_bn[j * LN + gl] = h6L * um + (2.0 * (h6L + h6R)) * u[j] + h6R * up + w_atf[j * LN + gl];
                    h6L = h6R;
                }
            }
            int k = 0;
            double sd, su;
            do {
                // ================= one Picard iteration (:667-698) ======================
                double uL = shfl_up(u[M - 1], 1), uR = shfl_down(u[0], 1);
                if (W > 1) {
                    if (lane == 0 && wq > 0) uL = x_halo[2 * (wq - 1) + 1];
                    if (lane == 31 && wq < W - 1) uR = x_halo[2 * (wq + 1)];
                }
                double cs[M], rs[M], fs[M];
                Continuant<M> ct;
                // element j = -1 (shared with the previous lane)
                double offL, qL, GL;
                {
                    const double u1 = uL, u2 = u[0];
                    const double sm = u1 + u2;
                    qL = fma(a6L, u2, a6L * sm);
                    const double z = fma(sm * (u2 - u1), 1.0 / 6.0, -w_hfs[gl]);
                    GL = (c1L * rcp(fmax(fabs(sm), 2.0e-10))) * z;
                    offL = s_off[gl];
                }
#pragma unroll
                for (int j = 0; j < M; ++j) {
                    const int ei = (j + 1) * LN + gl;
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
                    double b = s_dg[j * LN + gl] + (qL - pe);
                    const double rhs = w_bn[j * LN + gl] - GL + Ge;
                    double r = rhs - fma(a, um, fma(b, u1, c * u2));
                    if (j == 0 && gl == 0) {  // Dirichlet row (:679-680, :686)
                        a = 0.0; b = 1.0; c = 0.0; r = mu1 - u1;
                    }
                    // local Thomas sweep in continuant form (no reciprocal in the dependent chain)
                    ct.row(j, a, b, c, r, cs[j], rs[j], fs[j]);
                    offL = off; qL = qe; GL = Ge;
                }
                // condense each lane to one row, solve for the lanes' last unknowns, hand P to the next lane
                double xl, P;
                if constexpr (W == 1) {
                    partition_condense<M>(cs, rs, fs, lane, xl, P);
                } else {
                    double F, G, R;
                    partition_upward<M>(cs, rs, fs, F, G, R);
                    if (gl > 0) { x_F[gl - 1] = F; x_G[gl - 1] = G; x_R[gl - 1] = R; }
                    x_fl[gl] = (gl == 0) ? 0.0 : fs[M - 1];
                    x_cl[gl] = (gl == LN - 1) ? 0.0 : cs[M - 1];
                    x_rl[gl] = rs[M - 1];
                    group_bar(bar_id, LN);
                    // every warp: the whole 32 W-row system, W rows per lane
                    double cs2[W], rs2[W], fs2[W];
                    Continuant<W> c2;
#pragma unroll
                    for (int j = 0; j < W; ++j) {
                        const int q = W * lane + j;
                        const double cl = x_cl[q];
                        c2.row(j, x_fl[q], fma(-cl, x_F[q], 1.0), -cl * x_G[q], fma(-cl, x_R[q], x_rl[q]),
                               cs2[j], rs2[j], fs2[j]);
                    }
                    double x2, P2;
                    partition_condense<W>(cs2, rs2, fs2, lane, x2, P2);
                    x_sol[W * lane + W - 1] = x2;
#pragma unroll
                    for (int j = W - 2; j >= 0; --j) {
                        x2 = fma(-cs2[j], x2, fma(-fs2[j], P2, rs2[j]));
                        x_sol[W * lane + j] = x2;
                    }
                    __syncwarp();
                    xl = x_sol[gl];
                    P = (gl > 0) ? x_sol[gl - 1] : 0.0;
                }
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
                if (W > 1) {
                    if (lane == 0) { x_halo[2 * wq] = u[0]; x_nrm[2 * wq] = sd; }
                    if (lane == 31) { x_halo[2 * wq + 1] = u[M - 1]; x_nrm[2 * wq + 1] = su; }
                    group_bar(bar_id, LN);
                    sd = x_nrm[0]; su = x_nrm[1];
#pragma unroll
                    for (int w = 1; w < W; ++w) { sd += x_nrm[2 * w]; su += x_nrm[2 * w + 1]; }
                }
                if (TRACE && gl == 0)
                    p.err_hist[((long long)s * p.nsteps + n) * p.kmax + k] = sqrt(sd) / sqrt(su);
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
            finite = su <= 1.7976931348623157e308;   // false for inf and NaN
            if (gl == 0 && p.iters) p.iters[(long long)s * p.nsteps + n] = finite ? k : -1;
            if (!finite) { ++n; break; }
        }
        // A state that stopped being finite (the reference would go on printing nan, :664): the steps from
        // there on carry the flag -1 and NaN snapshots, and the batch goes on with the next sample.
        if (!finite) {
            const double qnan = __longlong_as_double(0x7ff8000000000000ll);
            for (; n < p.nsteps; ++n) {
                double *Un1 = Us + (long long)(n + 1) * N;
#pragma unroll
                for (int j = 0; j < M; ++j) if (node0 + j < N) Un1[j] = qnan;
                if (gl == 0 && p.iters) p.iters[(long long)s * p.nsteps + n] = -1;
            }
        }
    }
}

template <int M, int W, int WARPS, int MINB, bool EXACT, bool TRACE>
static int launch_fom_t(const FomParams &p, cudaStream_t st) {
    using L = FomSmem<M, W>;
    constexpr int GROUPS = WARPS / W;
    const size_t smem = L::bytes(GROUPS);
    auto kern = fom_picard_kernel<M, W, WARPS, MINB, EXACT, TRACE>;
    DeviceInfo di;
    int rc = device_info(di);
    if (rc != BX_OK) return rc;
    int occ = 0;
    rc = kernel_occupancy(reinterpret_cast<const void *>(kern), WARPS * 32, smem, &occ);
    if (rc != BX_OK) return rc;
    BX_REQUIRE(occ >= 1, "fom kernel does not fit on an SM (smem %zu B)", smem);
    long long ctas = (long long)di.sms * occ;
    const long long need = ((long long)p.B + GROUPS - 1) / GROUPS;
    if (ctas > need) ctas = need;
    BX_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st));
    kern<<<(unsigned)ctas, WARPS * 32, smem, st>>>(p);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

template <int M, int W, int WARPS, int MINB>
static int launch_fom(const FomParams &p, cudaStream_t st) {
    if (p.err_hist) {   // tracing runs are single-sample verbose calls: one general instance per mesh size
        if constexpr (W == 1) return launch_fom_t<M, 1, WARPS, MINB, false, true>(p, st);
    }
    return (p.N == 32 * W * M) ? launch_fom_t<M, W, WARPS, MINB, true, false>(p, st)
                               : launch_fom_t<M, W, WARPS, MINB, false, false>(p, st);
}

// Warps per trajectory: one warp is the throughput shape; with fewer trajectories than the device has
// warp slots for it (12 warps per SM at N = 512) the serial chain of a trajectory is what the sweep
// waits for, and 2 or 4 warps per trajectory shorten it (measured: profiles/r2a_fom_warps_per_trajectory.txt).
static int fom_warps_per_trajectory(int B, int N, int sms) {
    const int forced = tuning(BX_TUNE_FOM_WARPS_PER_TRAJECTORY);
    if (forced == 1 || forced == 2 || forced == 4) return (N <= 128 && forced > 1) ? 1 : forced;
    if (N <= 256) return 1;
    if ((long long)B <= 2LL * sms) return 4;
    if ((long long)B <= 6LL * sms) return 2;
    return 1;
}

}  // namespace bx

extern "C" long long bx_workspace_bytes(void) { return 256; }

extern "C" int bx_fom_picard_trajectories(int B, int N, int nsteps, const double *X, const double *u0,
                                          int u0_batched, const double *mu1, const double *mu2,
                                          double At, double E, double tol, int kmax, double *U,
                                          int *iters, double *err_hist, void *workspace, void *stream) {
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
    p.U = U; p.iters = iters; p.err_hist = err_hist; p.counter = queue_slot(workspace);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (N > 1024) {
        set_error("bx_fom_picard_trajectories: N = %d > 1024 needs the multi-CTA path (bx_fom_picard_finemesh)", N);
        return BX_ERR_UNSUPPORTED;
    }
    DeviceInfo di;
    const int rc = device_info(di);
    if (rc != BX_OK) return rc;
    const int W = err_hist ? 1 : fom_warps_per_trajectory(B, N, di.sms);
    if (N <= 64) return launch_fom<2, 1, 4, 4>(p, st);
    if (N <= 128) return launch_fom<4, 1, 4, 4>(p, st);
    if (N <= 256) {
        if (W == 2) return launch_fom<4, 2, 4, 4>(p, st);
        if (W == 4) return launch_fom<2, 4, 4, 4>(p, st);
        return launch_fom<8, 1, 4, 4>(p, st);
    }
    if (N <= 512) {
        if (W == 2) return launch_fom<8, 2, 4, 4>(p, st);
        if (W == 4) return launch_fom<4, 4, 4, 4>(p, st);
        return launch_fom<16, 1, 4, 3>(p, st);
    }
    if (W == 4) return launch_fom<8, 4, 4, 4>(p, st);
    if (W == 2) return launch_fom<16, 2, 4, 3>(p, st);
    return launch_fom<32, 1, 4, 1>(p, st);
}
