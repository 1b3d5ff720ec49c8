// bx_prom_cta.cu — POD-PROM trajectory kernel for LARGE bases (32 <= n <= 240 modes):
// FEMBurgers.pod_prom_burgers (FEM/fem_burgers.py:709-785), Galerkin and LSPG.  The reference's own driver
// runs 40 modes (POD/Results_thesis/prom_pod.py:35, tol 1e-3) and its stored truncations go up to 227.
//
// One CTA (8 or 16 warps) integrates one trajectory; CTAs pull samples from an atomic queue.  The basis
// Phi (N x n, up to 0.9 MB) no longer fits next to the work tiles, so it streams from L2 in chunks of KC nodes
// (a zero-padded N x NC copy of Phi in the workspace, rows moved by 1-D bulk copies of the TMA engine into a
// two- (LSPG) or three-slot (Galerkin) mbarrier ring); the state, the bands and the n x n reduced system stay in shared memory for the
// whole trajectory (the reduced system moves to an L2-resident scratch when it exceeds what is left, n >~ 140).
//
// One nonlinear iteration (:726-781):
//   A  rows of A = M + At C(U_k) + At E K (row 0 <- e0^T) and of R = A U_k - b, one node per thread (prom_row).
//   B  per chunk: J = A Phi from the three bands (3 DFMA per entry, R as column n), q0 += Phi^T U_k partials,
//      then the contraction on the FP64 tensor pipe (mma.sync.m8n8k4.f64), k index = node, 16 x 16 output
//      units (2 x 2 DMMA tiles) dealt round-robin to the warps, at most UB units per warp per pass over the
//      nodes:  LSPG  Ar = J^T J (upper unit triangle, mirrored), br = J^T R;  Galerkin  Ar = Phi^T J, br = Phi^T R.
//      Staging of chunk c+2, J of chunk c+1 and the DMMAs of chunk c share one barrier interval.
//   C  blocked right-looking LU of [Ar | br], panel width 8, one matrix row per thread in the panel (partial
//      pivoting in gesv's elimination order for Galerkin, :767; natural order for the SPD LSPG normal matrix;
//      rows never move — implicit pivoting), rank-8 trailing update with the U12 column in registers and the
//      L21 rows broadcast from a packed shared-memory panel; blocked back substitution.
//   D  q = q0 + dq, err = |dq| / |q| (:776), lift U_{k+1} = Phi q (:773), one warp per node.
// HBM traffic: the 8N-byte snapshot per time step; Phi is re-read from L2 twice per iteration.
#include "bx_prom_common.cuh"
#include "bx_async.cuh"
#include <cstdlib>

namespace bx {

// defined in bx_prom.cu: per-sample forcing data At F and (h/6)(f1 + f2) into the workspace
int launch_prom_forcing(int B, int N, const double *X, const double *mu2, double At, double *atf, double *hfs,
                        cudaStream_t st);

constexpr int PC_UB = 4;        // 16 x 16 output units a warp accumulates per pass (32 FP64 accumulators)
constexpr int PC_MAXW = 16;

struct PromCtaParams {
    int B, N, n, nsteps, kmax;
    int NBu, NC, ldp, ldm, KC, nchunks, NPad, units, passes, lgS;
    int mat_global;
    const double *X;
    const double *u0;
    long long u0_stride;
    const double *mu1;
    const double *mu2;
    double At, E, tol;
    const double *Phi;
    double *Phip;        // workspace: zero-padded copy of Phi, (NPad + 2) x NC, row 0 = node -1
    int CG, RR, RPT;     // J pass: warps per row run, row runs, rows per thread
    double *atf, *hfs;
    double *mat_ws;      // per-CTA reduced system when it does not fit in shared memory
    double *U;
    int *iters;
    double *err_hist;   // optional (B, nsteps, kmax): |dq| / |q| after every iteration (:776)
    unsigned int *counter;
};

// shared-memory carve-up (doubles), identical on host and device
struct PromCtaSmem {
    int h6, u, bn, hfs, row, bar, q, q0, x, prow, cand, red2, ints, work, phi, J, L, red, mat, total;
    int arr, phisz, Jsz;
    __host__ __device__ static PromCtaSmem make(int N, int n, int NC, int ldp, int ldm, int KC, int NPad, int lgS,
                                                int W, int mat_in_smem, int NS) {
        PromCtaSmem m;
        const int S = 1 << lgS;
        m.arr = S + 4;
        int o = 0;
        m.h6 = o; o += m.arr;
        m.u = o; o += m.arr;
        m.bn = o; o += m.arr;
        m.hfs = o; o += m.arr;
        o = (o + 1) & ~1;
        m.row = o; o += 4 * NPad;           // {a, b, c, R} per node, 32-byte records
        m.bar = o; o += 4;                  // 3 mbarriers
        m.q = o; o += NC;
        m.q0 = o; o += NC;
        m.x = o; o += NC;
        m.prow = o; o += 16;
        m.cand = o; o += 4 * W;             // 2 buffers x W x (key, index)
        m.red2 = o; o += 2 * W + 2;
        m.ints = o; o += NC + NC / 4 + 8;   // pivrow[NC] (int), act[NC] (int) -> NC doubles; cnt[W] + misc
        o = (o + 1) & ~1;
        m.work = o;
        // chunk ring
        m.phisz = (KC + 2) * ldp;
        m.Jsz = KC * ldp;
        m.phi = o; m.J = o + NS * m.phisz;   // NS ring slots: 3 (Galerkin reads Phi rows in the DMMA), 2 (LSPG)
        o += NS * m.phisz + 2 * m.Jsz;
        o = (o + 1) & ~1;
        // The Phi slots are refilled for the next contraction while the LU runs, the J tiles are not: the packed
        // L21 panel (NC x 8, 16-byte aligned rows) and the partial q0 per row run (RR x NC, RR <= KC, dead before
        // the LU starts) live in J tile 0 (KC x ldp >= 8 NC), which is dead from the last DMMA of an iteration to
        // the first J pass of the next.
        m.L = m.J;
        m.red = m.J;
        m.mat = o;
        if (mat_in_smem) o += n * ldm;
        m.total = o;
        return m;
    }
};

// zero-padded copy of Phi: row ip = node ip - 1, NC columns
__global__ void __launch_bounds__(256) prom_cta_pad_kernel(int N, int n, int NC, int rows, const double *Phi,
                                                           double *Phip) {
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    if (idx >= (long long)rows * NC) return;
    const int ip = (int)(idx / NC), j = (int)(idx - (long long)ip * NC);
    const int i = ip - 1;
    Phip[idx] = (i >= 0 && i < N && j < n) ? Phi[(size_t)i * n + j] : 0.0;
}

// DMMAs of one chunk for the first NU output units of a warp (the valid units form a prefix: u grows with s)
template <int NU, int UB>
__device__ __forceinline__ void prom_cta_mma(double (&acc)[UB][4][2], const int (&offA)[UB], const int (&offB)[UB],
                                             const double *Ac, const double *Jc, int KC, int ldp, int t) {
    for (int ks = 0; ks < KC; ks += 4) {
        const int ro = (ks + t) * ldp;
#pragma unroll
        for (int s = 0; s < NU; ++s) {
            const double a0 = Ac[ro + offA[s]], a1 = Ac[ro + offA[s] + 8];
            const double b0 = Jc[ro + offB[s]], b1 = Jc[ro + offB[s] + 8];
            dmma884(acc[s][0][0], acc[s][0][1], a0, b0);
            dmma884(acc[s][1][0], acc[s][1][1], a0, b1);
            dmma884(acc[s][2][0], acc[s][2][1], a1, b0);
            dmma884(acc[s][3][0], acc[s][3][1], a1, b1);
        }
    }
}

// UB = units per warp and pass.  UB = 1 (at most W = 8 units: LSPG up to 47 modes) and UB = 2 (at most 16 units:
// Galerkin up to 47 modes) are the light instances: 8 warps, at most 85 registers per thread, three CTAs per SM;
// UB = PC_UB is the general one (128 registers, 8 or 16 warps).
template <bool LSPG, int UB>
__global__ void __launch_bounds__(UB <= 2 ? 256 : PC_MAXW * 32, UB <= 2 ? 3 : 1) prom_cta_kernel(const PromCtaParams p) {
    extern __shared__ __align__(16) double smem[];
    const int N = p.N, n = p.n, NC = p.NC, ldp = p.ldp, ldm = p.ldm, KC = p.KC, lgS = p.lgS;
    const int S = 1 << lgS;
    const int T = blockDim.x, W = T >> 5;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int t = lane & 3, g = lane >> 2;
    // LSPG contracts J with itself, so a Phi chunk is dead once its J rows exist: two ring slots are enough
    // (chunk c + 2 lands in the slot chunk c was read from one barrier earlier); Galerkin still reads chunk c
    // as the left operand while c + 1 feeds the J pass and c + 2 is in flight
    constexpr unsigned NS = LSPG ? 2u : 3u;
    const PromCtaSmem lay = PromCtaSmem::make(N, n, NC, ldp, ldm, KC, p.NPad, lgS, W, p.mat_global ? 0 : 1, (int)NS);
    double *s_h6 = smem + lay.h6, *s_u = smem + lay.u, *s_bn = smem + lay.bn, *s_hfs = smem + lay.hfs;
    double2 *s_row = reinterpret_cast<double2 *>(smem + lay.row);
    uint64_t *s_bar = reinterpret_cast<uint64_t *>(smem + lay.bar);
    double *s_q = smem + lay.q, *s_q0 = smem + lay.q0, *s_x = smem + lay.x, *s_prow = smem + lay.prow;
    double *s_cand = smem + lay.cand, *s_red2 = smem + lay.red2;
    int *s_pivrow = reinterpret_cast<int *>(smem + lay.ints);
    int *s_act = s_pivrow + NC;
    int *s_cnt = s_act + NC;            // W warp counts, then [W] = next sample
    double *s_phi = smem + lay.phi, *s_J = smem + lay.J, *s_L = smem + lay.L, *s_red = smem + lay.red;
    double *mat = p.mat_global ? p.mat_ws + (size_t)blockIdx.x * (size_t)n * ldm : smem + lay.mat;

    // ---- CTA prologue: mesh constants, halo words ---------------------------------------------
    for (int i = tid; i < lay.arr; i += T) { s_h6[i] = 0.0; s_u[i] = 0.0; s_bn[i] = 0.0; s_hfs[i] = 0.0; }
    if (tid == 0) {
        for (int b = 0; b < 3; ++b) mbar_init(&s_bar[b], 1);
        fence_mbar_init();
    }
    unsigned gchunk = 0;   // chunks staged so far (ring slot = gchunk % NS, parity = (gchunk / NS) & 1)
    __syncthreads();
    for (int e = tid; e <= N - 2; e += T) s_h6[sk(e, S, lgS)] = (p.X[e + 1] - p.X[e]) / 6.0;

    PromRowCtx rc;
    rc.u = s_u; rc.bn = s_bn; rc.hfs = s_hfs; rc.h6 = s_h6;
    rc.a6 = p.At / 6.0; rc.c1 = 1.5 * p.At; rc.ate = p.At * p.E; rc.has_e = p.E != 0.0; rc.supg = true;
    rc.N = N; rc.S = S; rc.lgS = lgS;

    // J pass: thread = (column jcol, row run jrr); RPT consecutive rows per thread with a rolling 3-row window
    const int jcol = (warp % p.CG) * 32 + lane, jrr = warp / p.CG;
    const bool jact = jcol < NC && jrr < p.RR;
    const int RPT = p.RPT;

    // stage the KC + 2 Phi rows of chunk c (global number gc): one bulk copy per row, issued by warp 0
    auto stage_phi = [&](int c, unsigned gc) {
        if (warp == 0) {
            const unsigned slot = gc % NS;
            fence_proxy_async();   // the ring aliases buffers the generic proxy wrote (L21 panel, partial q0)
            if (lane == 0) mbar_arrive_expect_tx(&s_bar[slot], (uint32_t)((KC + 2) * NC * 8));
            __syncwarp();
            double *dst = s_phi + slot * lay.phisz;
            const double *src = p.Phip + (size_t)c * KC * NC;
            for (int r = lane; r < KC + 2; r += 32)
                bulk_g2s(dst + r * ldp, src + (size_t)r * NC, (uint32_t)(NC * 8), &s_bar[slot]);
        }
    };
    // invariant: the first two chunks of the next contraction pass are always in flight (Phi does not depend on
    // the state), so a pass never starts with an exposed L2 round trip
    stage_phi(0, gchunk);
    if (p.nchunks > 1) stage_phi(1, gchunk + 1);

    for (;;) {
        __syncthreads();
        if (tid == 0) s_cnt[W] = (int)atomicAdd(p.counter, 1u);
        __syncthreads();
        const int smp = s_cnt[W];
        if (smp >= p.B) {   // drain the two chunks in flight before the CTA gives up its shared memory
            mbar_wait(&s_bar[gchunk % NS], (gchunk / NS) & 1u);
            if (p.nchunks > 1) mbar_wait(&s_bar[(gchunk + 1) % NS], ((gchunk + 1) / NS) & 1u);
            break;
        }
        rc.mu1 = p.mu1[smp];
        const double *atf = p.atf + (size_t)smp * N;
        double *Us = p.U + (long long)smp * (long long)(p.nsteps + 1) * N;
        {
            const double *hf = p.hfs + (size_t)smp * N;
            const double *u0p = p.u0 + (long long)smp * p.u0_stride;
            for (int i = tid; i < N; i += T) {
                const int o = sk(i, S, lgS);
                s_hfs[o] = i <= N - 2 ? hf[i] : 0.0;
                const double v = u0p[i];
                s_u[o] = v;
                Us[i] = v;
            }
        }
        __syncthreads();

        for (int step = 0; step < p.nsteps; ++step) {
            // ---- bn = M U^n + At F (:745) ------------------------------------------------------
            for (int i = tid; i < N; i += T) {
                const double h6L = s_h6[sk(i - 1, S, lgS)], h6R = s_h6[sk(i, S, lgS)];
                const double um = s_u[sk(i - 1, S, lgS)], uc = s_u[sk(i, S, lgS)], up = s_u[sk(i + 1, S, lgS)];
                s_bn[sk(i, S, lgS)] = h6L * um + (2.0 * (h6L + h6R)) * uc + h6R * up + atf[i];
            }
            __syncthreads();
            int k = 0;
            bool finite = true;
            for (;;) {
                // ================= phase A: rows of A and R (:730-752) =========================
                for (int i = tid; i < p.NPad; i += T) {
                    PromRow r;
                    if (i < N) r = prom_row(rc, i, sk(i, S, lgS));
                    else { r.a = 0.0; r.b = 0.0; r.c = 0.0; r.R = 0.0; }
                    s_row[2 * i] = make_double2(r.a, r.b);
                    s_row[2 * i + 1] = make_double2(r.c, r.R);
                }
                // (the barrier after the first staging below orders these writes before their readers)

                // ================= phase B: J = A Phi, contraction (:756-762) ====================
                double q0acc = 0.0;
                auto compute_J = [&](int c, unsigned gc, bool with_q0) {
                    mbar_wait(&s_bar[gc % NS], (gc / NS) & 1u);
                    if (jact) {
                        const int kk0 = jrr * RPT;
                        const double *ph = s_phi + (gc % NS) * lay.phisz + kk0 * ldp + jcol;
                        double *jr = s_J + (c & 1) * lay.Jsz + kk0 * ldp + jcol;
                        const double2 *rw = s_row + 2 * (c * KC + kk0);
                        const double *up = s_u + c * KC + kk0 + 2;   // sk(i) = i + 2 for 0 <= i < S
                        double pm = ph[0], p0 = ph[ldp];
                        const bool isR = jcol == n;
#pragma unroll 4
                        for (int q = 0; q < RPT; ++q) {
                            const double pp = ph[(q + 2) * ldp];
                            const double2 ab = rw[2 * q], cR = rw[2 * q + 1];
                            double v = fma(ab.x, pm, fma(ab.y, p0, cR.x * pp));
                            if (isR) v = cR.y;
                            jr[q * ldp] = v;
                            if (with_q0) q0acc = fma(p0, up[q], q0acc);
                            pm = p0; p0 = pp;
                        }
                    }
                };

                for (int pass = 0; pass < p.passes; ++pass) {
                    double acc[UB][4][2];
                    int offA[UB], offB[UB];
                    bool uval[UB], udiag[UB];
                    int nu = 0;
#pragma unroll
                    for (int s = 0; s < UB; ++s) {
#pragma unroll
                        for (int q4 = 0; q4 < 4; ++q4) { acc[s][q4][0] = 0.0; acc[s][q4][1] = 0.0; }
                        const int u = (pass * UB + s) * W + warp;
                        uval[s] = u < p.units;
                        int Iu = 0, Ju = 0;
                        if (uval[s]) {
                            if (LSPG) {
                                int rem = u;
                                while (rem >= p.NBu - Iu) { rem -= p.NBu - Iu; ++Iu; }
                                Ju = Iu + rem;
                            } else {
                                Iu = u / p.NBu; Ju = u - Iu * p.NBu;
                            }
                        }
                        udiag[s] = Iu == Ju;
                        offA[s] = Iu * 16 + g; offB[s] = Ju * 16 + g;
                        nu += uval[s] ? 1 : 0;
                    }
                    __syncthreads();   // rows of phase A written
                    compute_J(0, gchunk, pass == 0);
                    __syncthreads();
                    for (int c = 0; c < p.nchunks; ++c) {
                        if (c + 2 < p.nchunks) stage_phi(c + 2, gchunk + c + 2);
                        if (c + 1 < p.nchunks) compute_J(c + 1, gchunk + c + 1, pass == 0);
                        const double *Jc = s_J + (c & 1) * lay.Jsz;
                        const double *Ac = LSPG ? Jc : s_phi + ((gchunk + c) % NS) * lay.phisz + ldp;
                        if constexpr (UB == 1) {
                            if (nu == 1) prom_cta_mma<1, UB>(acc, offA, offB, Ac, Jc, KC, ldp, t);
                        } else if constexpr (UB == 2) {
                            if (nu == 2) prom_cta_mma<2, UB>(acc, offA, offB, Ac, Jc, KC, ldp, t);
                            else if (nu == 1) prom_cta_mma<1, UB>(acc, offA, offB, Ac, Jc, KC, ldp, t);
                        } else {
                            switch (nu) {
                                case 1: prom_cta_mma<1, UB>(acc, offA, offB, Ac, Jc, KC, ldp, t); break;
                                case 2: prom_cta_mma<2, UB>(acc, offA, offB, Ac, Jc, KC, ldp, t); break;
                                case 3: prom_cta_mma<3, UB>(acc, offA, offB, Ac, Jc, KC, ldp, t); break;
                                case 4: prom_cta_mma<4, UB>(acc, offA, offB, Ac, Jc, KC, ldp, t); break;
                                default: break;
                            }
                        }
                        __syncthreads();
                    }
                    gchunk += (unsigned)p.nchunks;
                    stage_phi(0, gchunk);   // next pass / iteration / sample: ring slots are free after the barrier above
                    if (p.nchunks > 1) stage_phi(1, gchunk + 1);
                    // accumulator tiles -> [Ar | br]; column n = br
#pragma unroll
                    for (int s = 0; s < UB; ++s) {
                        if (uval[s]) {
#pragma unroll
                            for (int q4 = 0; q4 < 4; ++q4) {
                                const int r = (offA[s] - g) + (q4 >> 1) * 8 + g;
                                const int c0 = (offB[s] - g) + (q4 & 1) * 8 + 2 * t;
#pragma unroll
                                for (int e = 0; e < 2; ++e) {
                                    const int c = c0 + e;
                                    const double v = acc[s][q4][e];
                                    if (r < n && c <= n) mat[r * ldm + c] = v;
                                    if (LSPG && !udiag[s] && c < n && r < n) mat[c * ldm + r] = v;
                                }
                            }
                        }
                    }
                }
                // q0 = Phi^T U_k (:770): per-row-run partials summed in run order
                if (jact) s_red[jrr * NC + jcol] = q0acc;
                __syncthreads();
                if (tid < NC) {
                    double sum = 0.0;
                    for (int w = 0; w < p.RR; ++w) sum += s_red[w * NC + tid];
                    s_q0[tid] = tid < n ? sum : 0.0;
                }
                __syncthreads();

                // ================= phase C: blocked LU of [Ar | br], one row per thread ==========
                // Only the PW = ceil(n / 32) warps that own matrix rows take part in the panel factorisation and the
                // back substitution; they synchronise on named barrier 1, the other warps wait at the CTA barrier.
                const int PW = (n + 31) >> 5;
                const bool roww = warp < PW;
                auto rowsync = [&]() { asm volatile("bar.sync 1, %0;" ::"r"(PW * 32) : "memory"); };
                bool done = tid >= n;
                for (int k0 = 0; k0 < n; k0 += 8) {
                    const int w = n - k0 < 8 ? n - k0 : 8;
                  if (roww) {
                    const bool doneprev = done;
                    double pv[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) pv[j] = (!done && j < w) ? mat[tid * ldm + k0 + j] : 0.0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (j < w) {
                            int win = k0 + j;
                            if (!LSPG) {   // idamax of dgetf2 over the rows not yet used as pivots
                                const unsigned long long key =
                                    done ? 0ull : (unsigned long long)__double_as_longlong(fabs(pv[j])) + 1ull;
                                const unsigned khi = (unsigned)(key >> 32), klo = (unsigned)key;
                                const unsigned mhi = __reduce_max_sync(FULL, khi);
                                const bool c1 = khi == mhi;
                                const unsigned mlo = __reduce_max_sync(FULL, c1 ? klo : 0u);
                                const unsigned ball = __ballot_sync(FULL, c1 && klo == mlo && !done);
                                if (lane == 0) {
                                    double *cd = s_cand + (j & 1) * 2 * W + 2 * warp;
                                    cd[0] = __longlong_as_double(ball ? (long long)(((unsigned long long)mhi << 32) | mlo) : 0ll);
                                    cd[1] = (double)(ball ? (warp << 5) + (__ffs(ball) - 1) : -1);
                                }
                                rowsync();
                                unsigned long long best = 0ull;
                                win = -1;
                                for (int ww = 0; ww < PW; ++ww) {
                                    const double *cd = s_cand + (j & 1) * 2 * W + 2 * ww;
                                    const unsigned long long kk = (unsigned long long)__double_as_longlong(cd[0]);
                                    if (kk > best) { best = kk; win = (int)cd[1]; }
                                }
                                if (win < 0) win = k0 + j;   // only non-finite systems get here
                            }
                            if (tid == win) {
#pragma unroll
                                for (int jj = 0; jj < 8; ++jj) s_prow[(j & 1) * 8 + jj] = pv[jj];
                                done = true;
                                s_pivrow[k0 + j] = tid;
                            }
                            rowsync();
                            if (!done) {
                                const double *pr = s_prow + (j & 1) * 8;
                                const double l = pv[j] * rcp(pr[j]);
                                pv[j] = l;
#pragma unroll
                                for (int jj = j + 1; jj < 8; ++jj) pv[jj] = fma(-l, pr[jj], pv[jj]);
                            }
                        }
                    }
                    // pivot rows of this block back to the matrix (L11 / U11); multipliers to the packed panel
                    if (done && !doneprev) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) if (j < w) mat[tid * ldm + k0 + j] = pv[j];
                    }
                    int pos = -1;
                    if (LSPG) {
                        if (!done) pos = tid - (k0 + w);
                    } else {
                        const unsigned act = __ballot_sync(FULL, !done);
                        if (lane == 0) s_cnt[warp] = __popc(act);
                        rowsync();
                        if (!done) {
                            pos = __popc(act & ((1u << lane) - 1u));
                            for (int ww = 0; ww < warp; ++ww) pos += s_cnt[ww];
                        }
                    }
                    if (pos >= 0) {
                        double2 *Lr = reinterpret_cast<double2 *>(s_L + pos * 8);
                        Lr[0] = make_double2(pv[0], pv[1]); Lr[1] = make_double2(pv[2], pv[3]);
                        Lr[2] = make_double2(pv[4], pv[5]); Lr[3] = make_double2(pv[6], pv[7]);
                        s_act[pos] = tid;
                    }
                  }
                    __syncthreads();
                    const int nact = n - (k0 + w);
                    // U12 column in registers, then the rank-w update of the rows still in play
                    const int ncols = n - (k0 + w) + 1;
                    const int CW = (ncols + 31) & ~31;
                    const int RG = T / CW;
                    const int cidx = tid % CW, rg = tid / CW;
                    const bool cok = cidx < ncols && rg < RG;
                    const int col = k0 + w + cidx;
                    double uu[8];
                    int prw[8];
                    if (cok) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            uu[j] = 0.0; prw[j] = 0;
                            if (j < w) {
                                prw[j] = s_pivrow[k0 + j];
                                double v = mat[prw[j] * ldm + col];
#pragma unroll
                                for (int jj = 0; jj < j; ++jj) v = fma(-mat[prw[j] * ldm + k0 + jj], uu[jj], v);
                                uu[j] = v;
                            }
                        }
                        for (int ps = rg; ps < nact; ps += RG) {
                            const double2 *Lr = reinterpret_cast<const double2 *>(s_L + ps * 8);
                            const double2 l01 = Lr[0], l23 = Lr[1], l45 = Lr[2], l67 = Lr[3];
                            double *ap = mat + s_act[ps] * ldm + col;
                            double v = *ap;
                            v = fma(-l01.x, uu[0], v); v = fma(-l01.y, uu[1], v);
                            v = fma(-l23.x, uu[2], v); v = fma(-l23.y, uu[3], v);
                            v = fma(-l45.x, uu[4], v); v = fma(-l45.y, uu[5], v);
                            v = fma(-l67.x, uu[6], v); v = fma(-l67.y, uu[7], v);
                            *ap = v;
                        }
                    }
                    __syncthreads();
                    if (cok && rg == 0) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) if (j < w) mat[prw[j] * ldm + col] = uu[j];
                    }
                }
                __syncthreads();
                // ---- blocked back substitution: U x = y, thread k owns unknown k -----------------
                {
                    int pr = 0;
                    double y = 0.0;
                    if (tid < n) { pr = s_pivrow[tid]; y = mat[pr * ldm + n]; }
                    for (int k0 = ((n - 1) >> 3) << 3; roww && k0 >= 0; k0 -= 8) {
                        const int w = n - k0 < 8 ? n - k0 : 8;
                        if (warp == (k0 >> 5)) {
                            const int l0 = k0 & 31;
                            const double dinv = (tid < n) ? rcp(mat[pr * ldm + tid]) : 0.0;
                            for (int j = w - 1; j >= 0; --j) {
                                const double xj = shfl_idx(y * dinv, l0 + j);
                                if (lane >= l0 && lane < l0 + j) y = fma(-mat[pr * ldm + k0 + j], xj, y);
                                if (lane == l0 + j) s_x[k0 + j] = xj;
                            }
                        }
                        rowsync();
                        if (tid < k0) {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                if (j < w) y = fma(-mat[pr * ldm + k0 + j], s_x[k0 + j], y);
                        }
                    }
                    __syncthreads();   // x complete for everybody
                }
                // ================= phase D: q = q0 + dq, error, lift (:770-776) ====================
                double err;
                {
                    double dq = 0.0, q = 0.0;
                    if (tid < n) { dq = -s_x[tid]; q = s_q0[tid] + dq; }
                    if (tid < NC) s_q[tid] = q;
                    const double sd = warp_sum(dq * dq), sq = warp_sum(q * q);
                    if (lane == 0) { s_red2[2 * warp] = sd; s_red2[2 * warp + 1] = sq; }
                    __syncthreads();
                    double a = 0.0, b = 0.0;
                    for (int ww = 0; ww < W; ++ww) { a += s_red2[2 * ww]; b += s_red2[2 * ww + 1]; }
                    err = sqrt(a) / sqrt(b);
                    finite = (a + b) <= 1.7976931348623157e308;   // false for inf and NaN
                }
                for (int i0 = warp * 8; i0 < N; i0 += W * 8) {
                    double v[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r) v[r] = 0.0;
                    for (int j = 2 * lane; j < NC; j += 64) {   // padded rows: columns >= n and nodes >= N are zero
                        const double q0j = s_q[j], q1j = s_q[j + 1];
                        const double2 *pj = reinterpret_cast<const double2 *>(p.Phip + (size_t)(i0 + 1) * NC + j);
#pragma unroll
                        for (int r = 0; r < 8; ++r) {
                            const double2 ph2 = pj[(size_t)r * (NC / 2)];
                            v[r] = fma(ph2.x, q0j, fma(ph2.y, q1j, v[r]));
                        }
                    }
                    // recursive halving over lane bits 4, 3, 2, then bits 1, 0: lane 4 g + t ends with node i0 + g
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
                    double w1 = ((g & 1) ? w2[1] : w2[0]) + shfl_xor((g & 1) ? w2[0] : w2[1], 4);
                    w1 += shfl_xor(w1, 2);
                    w1 += shfl_xor(w1, 1);
                    if (t == 0 && i0 + g < N) s_u[sk(i0 + g, S, lgS)] = w1;
                }
                __syncthreads();
                if (p.err_hist && tid == 0) p.err_hist[((long long)smp * p.nsteps + step) * p.kmax + k] = err;
                ++k;
                if (!(err > p.tol) || k >= p.kmax) break;
            }
            // ---- snapshot step+1 (:783) and iteration count ------------------------------------
            double *Un1 = Us + (long long)(step + 1) * N;
            for (int i = tid; i < N; i += T) Un1[i] = s_u[sk(i, S, lgS)];
            if (tid == 0 && p.iters) p.iters[(long long)smp * p.nsteps + step] = finite ? k : -1;
            if (!finite) {   // singular reduced system / overflow (np.linalg.LinAlgError at :767): flag -1 and NaN
                const double qnan = __longlong_as_double(0x7ff8000000000000ll);   // snapshots from here on, batch goes on
                for (++step; step < p.nsteps; ++step) {
                    double *Un2 = Us + (long long)(step + 1) * N;
                    for (int i = tid; i < N; i += T) Un2[i] = qnan;
                    if (tid == 0 && p.iters) p.iters[(long long)smp * p.nsteps + step] = -1;
                }
            }
        }
    }
}

// ---- host: geometry, shared-memory fit, launch ----------------------------------------------------
int launch_prom_cta(int B, int N, int n, int nsteps, const double *X, const double *u0, int u0_batched,
                    const double *mu1, const double *mu2, double At, double E, const double *Phi, int projection,
                    double tol, int kmax, double *U, int *iters, double *err_hist, void *workspace, long long mat_offset_bytes,
                    cudaStream_t st, long long *mat_bytes_out) {
    PromCtaParams p;
    p.B = B; p.N = N; p.n = n; p.nsteps = nsteps; p.kmax = kmax;
    p.NBu = (n + 1 + 15) / 16;
    p.NC = 16 * p.NBu;
    p.ldp = p.NC + 4;                 // 4 mod 8: conflict-free DMMA fragment loads
    p.ldm = (n + 1) | 1;              // odd, >= n + 1
    p.units = projection == BX_LSPG ? p.NBu * (p.NBu + 1) / 2 : p.NBu * p.NBu;
    // Launch shape: W warps per trajectory and KC nodes per chunk chosen to keep the most warps resident per SM
    // (shared memory per CTA; 128 registers per thread = 512 threads per SM); ties go to the smaller CTA (more
    // trajectories in different phases per SM), then to the larger chunk (fewer barriers).  Measured on B200
    // (profiles/r02d_prom_cta_tune.txt): n = 40 -> (8, 16), n = 64 -> (8, 8), n = 96 -> (16, 16).
    const bool light = p.units <= 16;  // one or two units per warp at W = 8: the light instances, three CTAs per SM
    int Wmin = 8;   // at least 8 warps: staging, the J pass and the lift are per-warp instruction bound
    while (Wmin < PC_MAXW && (Wmin * PC_UB < p.units || Wmin * 32 < p.NC)) Wmin *= 2;
    int force_w = 0, force_kc = 0;
    {   // explicit tuning knobs (bx_set_tuning); invalid values are ignored
        const int w = tuning(BX_TUNE_PROM_CTA_WARPS), kc = tuning(BX_TUNE_PROM_CTA_CHUNK);
        if ((w == 8 || w == 16) && w >= Wmin) force_w = w;
        if (kc == 8 || kc == 12 || kc == 16 || kc == 24 || kc == 32) force_kc = kc;
    }
    p.lgS = 3;
    while ((1 << p.lgS) < N) ++p.lgS;
    DeviceInfo di;
    {
        const int rcd = device_info(di);
        if (rcd != BX_OK) return rcd;
    }
    const int sms = di.sms, maxsm = di.max_smem_optin;
    static const int kcs[5] = {32, 24, 16, 12, 8};   // multiples of 4 (DMMA k-steps), <= 32 (workspace rows)
    const int NS = projection == BX_LSPG ? 2 : 3;
    PromCtaSmem lay;
    bool found = false;
    int W = Wmin;
    int best_o = 0;   // CTAs per SM the chosen shape was scored with
    for (int mg = 0; mg < 2 && !found; ++mg) {   // reduced system in shared memory if it fits at all
        int best_score = 0, best_kc = 0, best_w = 0;
        best_o = 0;
        for (int w = Wmin; w <= PC_MAXW; w *= 2) {
            if (force_w && w != force_w) continue;
            for (int ki = 0; ki < 5; ++ki) {
                const int kc = kcs[ki];
                if (force_kc && kc != force_kc) continue;
                const int nch = (N + kc - 1) / kc;
                const PromCtaSmem l = PromCtaSmem::make(N, n, p.NC, p.ldp, p.ldm, kc, nch * kc, p.lgS, w, mg ? 0 : 1, NS);
                const long long bytes = (long long)l.total * (long long)sizeof(double);
                if (bytes > (long long)maxsm) continue;
                int o = (int)(233472LL / (bytes + 1024));
                const int regcap = (light && w == 8) ? 3 : 512 / (w * 32);   // 80 / 128 registers per thread
                if (o > regcap) o = regcap;
                if (o < 1) o = 1;
                if (o * w > best_score) { best_score = o * w; best_kc = kc; best_w = w; best_o = o; }
            }
        }
        if (best_score > 0) {
            found = true;
            W = best_w;
            p.KC = best_kc;
            p.nchunks = (N + p.KC - 1) / p.KC;
            p.NPad = p.nchunks * p.KC;
            p.mat_global = mg;
            lay = PromCtaSmem::make(N, n, p.NC, p.ldp, p.ldm, p.KC, p.NPad, p.lgS, W, mg ? 0 : 1, NS);
        }
    }
    // the light instances pay only where a third CTA fits (measured: n = 64 LSPG at two CTAs per SM is 11 % slower
    // with the 80-register build than with the general one)
    const int UBsel = (light && W == 8 && best_o >= 3) ? (p.units <= 8 ? 1 : 2) : PC_UB;
    p.passes = (p.units + W * UBsel - 1) / (W * UBsel);
    if (found) {
        p.CG = 1;
        while (p.CG * 32 < p.NC) p.CG *= 2;
        p.RR = W / p.CG;                    // W >= NC / 32 and both are powers of two
        while (p.KC % p.RR) p.RR /= 2;
        p.RPT = p.KC / p.RR;
    }
    // resident CTAs per SM can never exceed what shared memory allows (1 KB reserved per CTA)
    int occ_bound = found ? (int)(233472LL / ((long long)lay.total * (long long)sizeof(double) + 1024)) : 1;
    if (occ_bound < 1) occ_bound = 1;
    if (occ_bound > 4) occ_bound = 4;
    if (mat_bytes_out) {   // workspace query: one reduced system per resident CTA when it lives in global memory
        *mat_bytes_out = (found && p.mat_global) ? (long long)sms * occ_bound * (long long)n * p.ldm * (long long)sizeof(double) : 0;
        *mat_bytes_out += (long long)(N + 34) * p.NC * (long long)sizeof(double);   // padded Phi
        return BX_OK;
    }
    if (!found) {
        set_error("bx_prom_trajectories: N = %d, n = %d does not fit in shared memory", N, n);
        return BX_ERR_UNSUPPORTED;
    }
    p.X = X; p.u0 = u0; p.u0_stride = u0_batched ? N : 0;
    p.mu1 = mu1; p.mu2 = mu2; p.At = At; p.E = E; p.tol = tol;
    p.Phi = Phi; p.U = U; p.iters = iters; p.err_hist = err_hist;
    p.counter = queue_slot(workspace);
    p.atf = reinterpret_cast<double *>(static_cast<char *>(workspace) + 256);
    p.hfs = p.atf + (size_t)B * N;
    p.Phip = reinterpret_cast<double *>(static_cast<char *>(workspace) + mat_offset_bytes);
    p.mat_ws = p.Phip + (size_t)(N + 34) * p.NC;
    const size_t smem = sizeof(double) * (size_t)lay.total;
    int occ = 0;
    void (*kern)(const PromCtaParams) =
        projection == BX_LSPG
            ? (UBsel == 1 ? prom_cta_kernel<true, 1> : UBsel == 2 ? prom_cta_kernel<true, 2> : prom_cta_kernel<true, PC_UB>)
            : (UBsel == 1 ? prom_cta_kernel<false, 1> : UBsel == 2 ? prom_cta_kernel<false, 2> : prom_cta_kernel<false, PC_UB>);
    {
        const int rco = kernel_occupancy(reinterpret_cast<const void *>(kern), W * 32, smem, &occ);
        if (rco != BX_OK) return rco;
    }
    if (occ < 1) {
        set_error("bx_prom_trajectories: kernel does not fit on an SM (N = %d, n = %d)", N, n);
        return BX_ERR_UNSUPPORTED;
    }
    if (occ > occ_bound) occ = occ_bound;
    long long ctas = (long long)sms * occ;
    if (ctas > B) ctas = B;
    BX_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st));
    {
        const int rcf = launch_prom_forcing(B, N, X, mu2, At, p.atf, p.hfs, st);
        if (rcf != BX_OK) return rcf;
    }
    {
        const int rows = p.NPad + 2;   // <= N + 33
        prom_cta_pad_kernel<<<(unsigned)(((long long)rows * p.NC + 255) / 256), 256, 0, st>>>(N, n, p.NC, rows, Phi, p.Phip);
        BX_LAUNCH_CHECK();
    }
    kern<<<(unsigned)ctas, W * 32, smem, st>>>(p);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

}  // namespace bx
