// bx_fom.cu — fused FOM trajectory kernel, throughput instance: FEMBurgers.fom_burgers
// (FEM/fem_burgers.py:646-707) for B (mu1, mu2) samples at once.
//
// ONE WARP integrates one trajectory without leaving the SM:
//   * lane l owns the M consecutive nodes l*M .. l*M+M-1 in registers (32 M >= N);
//   * every Picard iteration assembles the tridiagonal A = M + At C(U_k) + At E K, the
//     right-hand side b = M U^n + At F - At S(U_k) and the residual R = A U_k - b on the fly
//     from the closed forms of the 2-point-Gauss element integrals (no band ever reaches HBM);
//   * A delta = -R is solved with a partition (Wang / SPIKE-type) method: each lane runs a
//     Thomas sweep in continuant form over its M rows carrying the coupling to the previous lane's
//     last unknown, an upward sweep condenses the lane to one row of a 32-row tridiagonal system
//     in the lanes' last unknowns, solved by parallel cyclic reduction with warp shuffles;
//   * the update, |delta|^2 and |U_{k+1}|^2 are fused into the back-substitution, the convergence
//     test (tol, kmax of :663) is evaluated per trajectory;
//   * the only HBM traffic is the 8N-byte snapshot per step (+ 4 B iteration count).
// SCHEDULING.  A sweep with more trajectories than resident warps (config 2: 4096 against 12 per SM)
// is cut into SEGMENTS of `seg` time steps; the work units (segment g of trajectory s) are handed out
// round-robin — unit t = g * B + s from one atomic counter — so all trajectories advance together and
// all warp slots stay busy until the last segment (whole trajectories from a queue leave a tail of
// half-empty SMs: measured 55.3 ms against 42 ms of asymptotic throughput for config 2).  A trajectory
// resumes from its own last snapshot in U (the state IS the output), gated by a per-trajectory
// progress word that the producer publishes with release semantics after its snapshot stores; units
// are handed out in dependency order to warps that are already running, so the wait cannot deadlock.
// Sweeps with few trajectories per SM use the latency kernel (bx_fom_strong.cu) instead.
#include "bx_fom.cuh"
#include <cmath>

namespace bx {

// EXACT: N == 32 M, so the only lane with a missing right-hand element is the last (local row M-1).
// TRACE: also record |delta| / |U_{k+1}| of every iteration (the verbose mode of the legacy wrapper).
template <int M, int WARPS, int MINB, bool EXACT, bool TRACE>
__global__ void __launch_bounds__(WARPS * 32, MINB) fom_picard_kernel(const FomParams p) {
    using L = FomSmem<M>;
    extern __shared__ __align__(16) double smem[];
    double *s_off = smem;             // h/6 - At E / h   (off-diagonal constant of A; 0 where no element)
    double *s_h6 = s_off + L::EL;     // h/6              (mass off-diagonal; read once per time step)
    double *s_dg = s_h6 + L::EL;      // diagonal constant of A (1 for padded rows)
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    double *w_hfs = s_dg + L::NL + warp * L::PER_WARP;  // 1.5 At (h/6) (f_gp1 + f_gp2) per element
    double *w_atf = w_hfs + L::EL;                      // At * F per node
    double *w_bn = w_atf + L::NL;                       // M U^n + At F per node (per time step)
    double *w_mu1 = w_bn + L::NL;                       // Dirichlet value of the current trajectory

    const int N = p.N;
    const double At = p.At;

    // ---- mesh constants, once per CTA (X is shared by all samples) ----------------------
    for (int idx = threadIdx.x; idx < L::EL; idx += WARPS * 32) {
        const int l = idx % 32, jj = idx / 32 - 1;
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
        const int l = idx % 32, j = idx / 32;
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
    // convection: p = a6 (2 u1 + u2), q = a6 (u1 + 2 u2), a6 = At / 6;
    // SUPG: At g = (c16 s du - c1 (h/6) fsum) / max(|s|, 2e-10),  s = u1 + u2, c1 = 1.5 At, c16 = c1 / 6 = At / 4
    // (c1 (h/6) fsum is a per-element constant of the trajectory: it is what w_hfs holds).
    // EXACT: the left element is missing on lane 0 only, whose row 0 is the Dirichlet row and overrides whatever
    // the left element contributed, so no lane needs its own copy of the two constants.
    const bool leftv = EXACT || (lane > 0 && jl >= 0);
#define BX_A6L (leftv ? p.a6 : 0.0)

    for (;;) {
        unsigned t = 0;
        if (lane == 0) t = atomicAdd(p.counter, 1u);
        t = __shfl_sync(FULL, t, 0);
        if (t >= (unsigned)p.B * (unsigned)p.nseg) break;   // B * nseg < 2^32 (checked at launch)
        const int g = (int)(t / (unsigned)p.B);             // segment
        const int s = (int)(t - (unsigned)g * (unsigned)p.B);   // trajectory
        const int n_begin = g * p.seg;
        const int n_end = min(p.nsteps, n_begin + p.seg);

        if (lane == 0) { *w_mu1 = p.mu1[s]; reinterpret_cast<int *>(w_mu1 + 1)[0] = g; }
        const double mu2 = p.mu2[s];
        double *Us = p.U + (long long)s * (long long)(p.nsteps + 1) * N + node0;

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
                w_hfs[(jj + 1) * 32 + lane] = p.c1 * hfs;
                if (jj >= 0) w_atf[jj * 32 + lane] = At * (left + c0);
                left = cr;
            }
        }

        // ---- state at the start of the segment -> registers ------------------------------
        double u[M];
        bool finite = true;
        if (g == 0) {   // initial condition, snapshot 0
            const double *u0p = p.u0 + (long long)s * p.u0_stride + node0;
#pragma unroll
            for (int j = 0; j < M; ++j) u[j] = (node0 + j < N) ? u0p[j] : 0.0;
            if (vec_ok) {
#pragma unroll
                for (int j = 0; j < M; j += 2) *reinterpret_cast<double2 *>(Us + j) = make_double2(u[j], u[j + 1]);
            } else {
#pragma unroll
                for (int j = 0; j < M; ++j) if (node0 + j < N) Us[j] = u[j];
            }
        } else {        // resume from the snapshot the previous segment's warp left in U
            // every lane polls the same word (one broadcast transaction per probe; see the note at the publish
            // store below for why this is not left to lane 0), then lane 0's value decides for the warp
            int done;
            {
                const volatile int *flag = p.progress + s;
                while ((done = *flag) < g && done >= 0) __nanosleep(200);
            }
            done = __shfl_sync(FULL, done, 0);
            __threadfence();        // acquire side: the snapshot loads below stay behind the flag read
            // A negative word: the trajectory stopped being finite in an earlier segment.  Its state is poisoned
            // here as DATA (a select, not a branch: a warp-level branch on the polled word costs the compiler its
            // convergence proof for every shuffle of the hot loop), so the first step of this segment comes out
            // non-finite, takes the flag -1 and the segment is filled with NaN like the ones before it.
            const double poison = done < 0 ? __longlong_as_double(0x7ff8000000000000ll) : 0.0;
            const double *Ub = Us + (long long)n_begin * N;
#pragma unroll
            for (int j = 0; j < M; ++j) u[j] = (node0 + j < N) ? __ldcg(Ub + j) + poison : 0.0;
        }
        __syncwarp();

        int n = n_begin;
        for (; finite && n < n_end; ++n) {
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
                Continuant<M, true> ct;
                // element j = -1 (shared with the previous lane)
                double offL, qL, GL;
                {
                    const double u1 = uL, u2 = u[0];
                    const double sm = u1 + u2;
                    qL = fma(BX_A6L, u2, BX_A6L * sm);
                    const double z = fma(sm * (u2 - u1), p.c16, -w_hfs[lane]);
                    GL = leftv ? rcp1(fmax(fabs(sm), 2.0e-10)) * z : 0.0;
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
                    const double t6 = p.a6 * sm;
                    double pe = fma(p.a6, u1, t6);
                    double qe = fma(p.a6, u2, t6);
                    const double z = fma(sm * (u2 - u1), p.c16, -w_hfs[ei]);
                    double Ge = rcp1(fmax(fabs(sm), 2.0e-10)) * z;
                    if ((!EXACT || j == M - 1) && j == jl) { pe = 0.0; qe = 0.0; Ge = 0.0; }
                    const double off = s_off[ei];
                    // row j of A, b and r = -R = b - A U_k
                    double a = offL - qL;
                    double c = off + pe;
                    double b = s_dg[j * 32 + lane] + (qL - pe);
                    // r = b - A U_k as three FMAs onto the right-hand side (measured: 48.4 -> 46.3 ms for config 2 with the
                    // folded SUPG constant; summing A U_k first, or keeping the SUPG terms out of the chain, is slower here)
                    const double rhs = w_bn[j * 32 + lane] - GL + Ge;
                    double r = fma(-a, um, fma(-b, u1, fma(-c, u2, rhs)));
                    if (j == 0 && lane == 0) {  // Dirichlet row (:679-680, :686)
                        a = 0.0; b = 1.0; c = 0.0; r = *w_mu1 - u1;
                    }
                    // local Thomas sweep in continuant form (no reciprocal in the dependent chain)
                    ct.row(j, a, b, c, r, cs[j], rs[j], fs[j]);
                    offL = off; qL = qe; GL = Ge;
                }
                // condense each lane to one row, solve for the lanes' last unknowns, hand P to the next lane
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
                if (TRACE && lane == 0)
                    p.err_hist[((long long)s * p.nsteps + n) * p.kmax + k] = sqrt(sd) / sqrt(su);
                ++k;
            } while (sd > p.tol2 * su && k < p.kmax);
#undef BX_A6L

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
            if (lane == 0 && p.iters) p.iters[(long long)s * p.nsteps + n] = finite ? k : -1;
        }
        // A state that stopped being finite (the reference would go on printing nan, :664): the steps from
        // there on carry the flag -1 and NaN snapshots, and the batch goes on with the next unit.
        if (!finite) {
            const double qnan = __longlong_as_double(0x7ff8000000000000ll);
            for (; n < n_end; ++n) {
                double *Un1 = Us + (long long)(n + 1) * N;
#pragma unroll
                for (int j = 0; j < M; ++j) if (node0 + j < N) Un1[j] = qnan;
                if (lane == 0 && p.iters) p.iters[(long long)s * p.nsteps + n] = -1;
            }
        }
        if (p.nseg > 1) {   // publish the segment: snapshot stores of all lanes, then the progress word
            __threadfence();
            __syncwarp();
            // every lane stores the same word (one transaction); a lane-0-only store here gets threaded into the
            // lane-0-only atomicAdd at the top of the loop and costs the compiler its proof that the warp is
            // converged at the shuffles (BRA.DIV before each of them, 100 bytes more spills: measured 5 % slower)
            *reinterpret_cast<volatile int *>(p.progress + s) = finite ? reinterpret_cast<const int *>(w_mu1 + 1)[0] + 1 : -1;
        }
    }
}

template <int M, int WARPS, int MINB, bool EXACT, bool TRACE>
static int launch_fom_t(FomParams p, cudaStream_t st) {
    using L = FomSmem<M>;
    const size_t smem = L::bytes(WARPS);
    auto kern = fom_picard_kernel<M, WARPS, MINB, EXACT, TRACE>;
    DeviceInfo di;
    int rc = device_info(di);
    if (rc != BX_OK) return rc;
    int occ = 0;
    rc = kernel_occupancy(reinterpret_cast<const void *>(kern), WARPS * 32, smem, &occ);
    if (rc != BX_OK) return rc;
    BX_REQUIRE(occ >= 1, "fom kernel does not fit on an SM (smem %zu B)", smem);
    long long ctas = (long long)di.sms * occ;
    const long long need = ((long long)p.B + WARPS - 1) / WARPS;
    // segments only when the sweep does not fit the resident warps (otherwise every trajectory has its own warp
    // from the first step to the last and there is no tail to balance)
    p.nseg = 1; p.seg = p.nsteps > 0 ? p.nsteps : 1;
    if (ctas >= need) {
        ctas = need;
    } else if (!TRACE) {
        int seg = tuning(BX_TUNE_FOM_SEGMENT_STEPS);
        if (seg < 0) seg = 25;
        if (seg > 0 && seg < p.nsteps) { p.seg = seg; p.nseg = (p.nsteps + seg - 1) / seg; }
    }
    BX_REQUIRE((long long)p.B * p.nseg < 0xffffffffLL, "too many work units (B = %d, %d segments)", p.B, p.nseg);
    BX_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st));
    if (p.nseg > 1) BX_CUDA(cudaMemsetAsync(p.progress, 0, sizeof(int) * (size_t)p.B, st));
    kern<<<(unsigned)ctas, WARPS * 32, smem, st>>>(p);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

template <int M, int WARPS, int MINB>
static int launch_fom(const FomParams &p, cudaStream_t st) {
    if (p.err_hist)   // tracing runs are single-sample verbose calls: one general instance per mesh size
        return launch_fom_t<M, WARPS, MINB, false, true>(p, st);
    return (p.N == 32 * M) ? launch_fom_t<M, WARPS, MINB, true, false>(p, st)
                           : launch_fom_t<M, WARPS, MINB, false, false>(p, st);
}

// Warps per trajectory of the latency kernel: two warps while the device has a sub-partition to spare for the
// second one (measured: profiles/r2_fom_kernels_by_batch.txt); BX_TUNE_FOM_WARPS_PER_TRAJECTORY forces 1 / 2 / 4.
static int fom_warps_per_trajectory(int B, int N, int sms) {
    const int forced = tuning(BX_TUNE_FOM_WARPS_PER_TRAJECTORY);
    if (forced == 1 || forced == 2 || forced == 4) return forced;
    if (N <= 256) return 1;
    return ((long long)B <= 2LL * sms) ? 2 : 1;
}

}  // namespace bx

// 256-byte head (64 rotating work-queue counters) + one progress word per trajectory
extern "C" long long bx_workspace_bytes(void) { return 256; }
extern "C" long long bx_fom_workspace_bytes(int B) { return B < 0 ? BX_ERR_INVALID : 256 + 4LL * B; }

// which kernel and how many warps per trajectory a sweep of B samples on N nodes gets on the current device:
// returns 16 * kernel + warps (kernel 0 = throughput, 1 = latency), negative on error
extern "C" int bx_fom_plan(int B, int N) {
    using namespace bx;
    BX_REQUIRE(B >= 0 && N >= 2 && N <= 1024, "bx_fom_plan: B >= 0 and 2 <= N <= 1024");
    DeviceInfo di;
    const int rc = device_info(di);
    if (rc != BX_OK) return rc;
    if (N > 128) {
        const int forced = tuning(BX_TUNE_FOM_KERNEL);
        if (forced >= 0 ? forced == 1 : (long long)B <= 8LL * di.sms) return 16 + fom_warps_per_trajectory(B, N, di.sms);
    }
    return 1;
}

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
    p.a6 = At / 6.0; p.c1 = 1.5 * At; p.c16 = 0.25 * At;
    p.U = U; p.iters = iters; p.err_hist = err_hist; p.counter = queue_slot(workspace);
    p.progress = reinterpret_cast<int *>(static_cast<char *>(workspace) + 256);
    p.nseg = 1; p.seg = nsteps > 0 ? nsteps : 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (N > 1024) {
        set_error("bx_fom_picard_trajectories: N = %d > 1024 needs the multi-CTA path (bx_fom_picard_finemesh)", N);
        return BX_ERR_UNSUPPORTED;
    }
    DeviceInfo di;
    const int rc = device_info(di);
    if (rc != BX_OK) return rc;
    // latency kernel (bx_fom_strong.cu) when the sweep cannot fill the warp slots of this one
    if (!err_hist) {
        const int plan = bx_fom_plan(B, N);
        if (plan < 0) return plan;
        if (plan >= 16) return launch_fom_strong(p, plan & 15, st);
    }
    if (N <= 64) return launch_fom<2, 4, 4>(p, st);
    if (N <= 128) return launch_fom<4, 4, 4>(p, st);
    if (N <= 256) return launch_fom<8, 4, 4>(p, st);
    if (N <= 512) return launch_fom<16, 4, 3>(p, st);
    return launch_fom<32, 4, 1>(p, st);
}
