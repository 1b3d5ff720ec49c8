// bx_qm.cu — quadratic-manifold PROM trajectory kernel: FEMBurgers.pod_quadratic_manifold
// (FEM/fem_burgers.py:1081-1175; helpers get_sym/get_single_Q/get_dQ_dq :263-312), Galerkin and
// LSPG, for B (mu1, mu2) samples at once.
//
// Decoder u(q) = Phi q + H sym(q x q), tangent T(q) = Phi + H dQ/dq.  With Hs_i the symmetric n x n
// matrix of row i of H (Hs_i[j][m] = H[i][idx(j,m)] (1 + delta_jm); monomial order i outer, j >= i
// inner, :263-289) both are ONE symmetric matrix-vector product per node:
//     T_i = Phi_i + Hs_i q            (dQ/dq has <= 2 non-zeros per row, :292-312)
//     u_i = (Phi_i + T_i) . q / 2     (q^T Hs_i q / 2 = sum_{a<=b} H[i][ab] q_a q_b)
// so each Gauss-Newton iteration makes one pass over H (k = n(n+1)/2 doubles per node).
//
// A CTA integrates EIGHT trajectories in lock step (one warp owns one trajectory for the contraction and the
// reduced solve; the tangent pass is shared):
//   T  tangent/decoder pass on the FP64 tensor pipe: for node i the eight tangent rows are the 8 x n product
//      [q_0 .. q_7]^T Hs_i, i.e. DMMA m8n8k4 with A = Hs_i (rows j, k index m, streamed ONCE per CTA from an
//      L2-resident unpacked copy laid out in fragment order: 64 contiguous bytes per lane and 8-row block),
//      B = the eight q vectors, C initialised with Phi_i.  Warp w takes nodes w, w + 8, ...; the next node's
//      fragments are in flight while the current one is multiplied.  u_i = (Phi_i + T_i) . q / 2 comes from the
//      same accumulators (reduction over the 8 row lanes).  T rows go to per-trajectory scratch in global
//      memory (L2), u to the owning warp's shared memory.  Trajectories that have converged keep taking part
//      with an unchanged q (the pass reproduces their T and u bit for bit) until all eight are done with the
//      time step (__syncthreads_and).
//   B  contraction on the FP64 tensor pipe (DMMA m8n8k4, k index = node) exactly as in bx_prom.cu with
//      T in the place of Phi and no SUPG term in b (:1141):
//         LSPG  Ar = (A T)^T (A T), br = (A T)^T R (:1153-1156);  Galerkin  Ar = T^T A T, br = T^T R (:1150-1151)
//   C  n x n LU with partial pivoting (np.linalg.solve, :1161), q += dq (:1162), then pass T again at the
//      new q (it yields u for :1163 and the tangent of the next iteration), rel = |dq| / max(1e-14, |q|) (:1165).
// A pre-kernel doubles the diagonal monomials of H once (Hp) and evaluates At F per sample.
#include "bx_prom_common.cuh"
#include <cstdint>

namespace bx {

struct QmParams {
    int B, N, n, k, nsteps, itmax, warps;
    int ngroups, nseg, seg;   // B / group lock-step groups; the time steps are cut into nseg segments of seg steps (round-robin units)
    int *progress;       // workspace: (ngroups,) segments completed per group
    int group;           // trajectories a CTA integrates in lock step (<= QM_MAXW; fewer when the sweep is smaller than the device)
    const double *X;
    const double *u0;
    long long u0_stride;
    const double *mu1;   // uxa per sample (left Dirichlet value)
    const double *mu2;
    double At, E, tol;
    const double *Phi;   // (N, n) row-major
    const double *Hs;    // workspace: (N, NB, 32 lanes, 2 NB) unpacked symmetric H rows in DMMA A-fragment order
    double *atf;         // workspace: At * F (B, N)
    double *Tg;          // workspace: per warp slot (NP + 1) x 32 tangent rows (row 0 = zero row for node -1)
    double *U;
    int *iters;
    double *err_hist;    // optional (B, nsteps, itmax): |dq| / max(1e-14, |q|) after every iteration (:1165)
    unsigned int *counter;
};

constexpr int QM_MAXW = 8;
constexpr int QM_TS_MAX_NB = 3;   // v3 is dispatched up to this many 8-column blocks (n <= 23)
constexpr int QM_PF = 4;      // tangent rows in flight ahead of the contraction (divides 8)

struct QmSmem {
    int h6, per_cta;
    int u, misc, bn, per_warp, arr;
    __host__ __device__ static QmSmem make(int NP) {
        QmSmem m;
        m.arr = NP + 8;
        int o = 0;
        m.h6 = o; o += m.arr;
        m.per_cta = (o + 1) & ~1;
        o = 0;
        int usz = m.arr < 16 * MAT_LD - 8 ? 16 * MAT_LD - 8 : m.arr;
        usz = (usz + 1) & ~1;
        m.u = o; o += usz;
        m.misc = o; o += 128;
        m.bn = o; o += m.arr;
        m.per_warp = (o + 1) & ~1;
        return m;
    }
};

// Hs: for node i, 8-row block jb, lane (g = lane / 4, t = lane % 4), k-step ks the entry
//   Hs_i[j = 8 jb + g][m = 4 ks + t] = H[i][idx(min(j, m), max(j, m))] (1 + delta_jm)   (0 beyond n),
// stored at ((i NB + jb) 32 + lane) 2 NB + ks — the A fragments of one (node, block) are 64-byte runs per lane.
// atf = At F.
__global__ void __launch_bounds__(256) qm_prep_kernel(int B, int N, int NB, int n, int k, const double *X,
                                                      const double *mu2v, double At, const double *H, double *Hs,
                                                      double *atf) {
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    const int NKS = 2 * NB;
    const long long nH = (long long)N * NB * 32 * NKS;
    if (idx < nH) {
        const int ks = (int)(idx % NKS);
        long long r = idx / NKS;
        const int lane = (int)(r & 31); r >>= 5;
        const int jb = (int)(r % NB), i = (int)(r / NB);
        const int j = 8 * jb + (lane >> 2), m = 4 * ks + (lane & 3);
        double v = 0.0;
        if (j < n && m < n) {
            const int a = j < m ? j : m, b = j < m ? m : j;
            const int c = a * n - a * (a - 1) / 2 + (b - a);
            v = H[(size_t)i * k + c] * (a == b ? 2.0 : 1.0);
        }
        Hs[idx] = v;
        return;
    }
    const long long j = idx - nH;
    if (j >= (long long)B * N) return;
    const int smp = (int)(j / N), i = (int)(j - (long long)smp * N);
    const double mu2 = mu2v[smp];
    double F = 0.0;
    if (i >= 1) {
        const double x1 = X[i - 1], x2 = X[i];
        const double f1 = 0.02 * exp(mu2 * (BX_GA * x1 + BX_GB * x2));
        const double f2 = 0.02 * exp(mu2 * (BX_GB * x1 + BX_GA * x2));
        F += (f1 * BX_GB + f2 * BX_GA) * (0.5 * (x2 - x1));
    }
    if (i <= N - 2) {
        const double x1 = X[i], x2 = X[i + 1];
        const double f1 = 0.02 * exp(mu2 * (BX_GA * x1 + BX_GB * x2));
        const double f2 = 0.02 * exp(mu2 * (BX_GB * x1 + BX_GA * x2));
        F += (f1 * BX_GA + f2 * BX_GB) * (0.5 * (x2 - x1));
    }
    atf[j] = At * F;
}

// ---- tangent pass building block shared by qm_kernel (v2) and qm_kernel_ts (v3) ---------------------------
// A node is multiplied in two halves of HB = ceil(NB / 2) eight-row blocks: 16 instead of 32 fragment doubles
// live per lane, which is what lets the tangent pass coexist with the 32 accumulators of an NB = 4 contraction.
// qm_half: eight tangent half-rows T_i[8 jb + g], jb in [hh HB, hh HB + HB), of the trajectories 2t, 2t+1 (c0 / c1),
// C initialised with Phi_i; while the DMMAs of k-steps 2h, 2h+1 retire, the fragment registers they read are
// refilled with those of the next half (node inext, half hnext; inext < 0: none) — one buffer, a whole half
// of prefetch distance.  s0 / s1 accumulate the decoder partial sums (Phi_i + T_i) . q of the two trajectories.
template <int NB>
__device__ __forceinline__ void qm_load_half(double (&A)[(NB + 1) / 2][2 * NB], const double *Hs, int lane, int i, int hh) {
    constexpr int HB = (NB + 1) / 2, NKS = 2 * NB;
#pragma unroll
    for (int b = 0; b < HB; ++b) {
        const int jb = hh * HB + b;
        if (jb < NB) {
            const double2 *src = reinterpret_cast<const double2 *>(Hs + (((size_t)i * NB + jb) * 32 + lane) * NKS);
#pragma unroll
            for (int h = 0; h < NB; ++h) {
                const double2 v = src[h];
                A[b][2 * h] = v.x; A[b][2 * h + 1] = v.y;
            }
        }
    }
}

template <int NB>
__device__ __forceinline__ void qm_half(double (&A)[(NB + 1) / 2][2 * NB], const double (&Bq)[2 * NB],
                                        const double (&qC)[NB][2], const double *Hs, const double *Phi, int n, int lane,
                                        int i, int hh, int inext, int hnext, double (&c)[(NB + 1) / 2][2],
                                        double &s0, double &s1) {
    constexpr int HB = (NB + 1) / 2, NKS = 2 * NB;
    const int g = lane >> 2;
    double ph[HB];
#pragma unroll
    for (int b = 0; b < HB; ++b) {
        const int jb = hh * HB + b;
        ph[b] = (jb < NB && 8 * jb + g < n) ? Phi[(size_t)i * n + 8 * jb + g] : 0.0;
        c[b][0] = ph[b]; c[b][1] = ph[b];
    }
#pragma unroll
    for (int h = 0; h < NB; ++h) {
#pragma unroll
        for (int b = 0; b < HB; ++b)
            if (hh * HB + b < NB) dmma884(c[b][0], c[b][1], A[b][2 * h], Bq[2 * h]);
#pragma unroll
        for (int b = 0; b < HB; ++b)
            if (hh * HB + b < NB) dmma884(c[b][0], c[b][1], A[b][2 * h + 1], Bq[2 * h + 1]);
        if (inext >= 0) {
#pragma unroll
            for (int b = 0; b < HB; ++b) {
                const int jn = hnext * HB + b;
                if (jn < NB) {
                    const double2 v = reinterpret_cast<const double2 *>(Hs + (((size_t)inext * NB + jn) * 32 + lane) * NKS)[h];
                    A[b][2 * h] = v.x; A[b][2 * h + 1] = v.y;
                }
            }
        }
    }
#pragma unroll
    for (int b = 0; b < HB; ++b) {
        const int jb = hh * HB + b;
        if (jb < NB) {
            s0 = fma(ph[b] + c[b][0], (hh ? qC[HB + b < NB ? HB + b : 0][0] : qC[b][0]), s0);
            s1 = fma(ph[b] + c[b][1], (hh ? qC[HB + b < NB ? HB + b : 0][1] : qC[b][1]), s1);
        }
    }
}

template <int NB, bool LSPG>
__global__ void __launch_bounds__(QM_MAXW * 32, 1) qm_kernel(const QmParams p) {
    constexpr int NC = NB * 8;
    using Tl = PromTiles<NB, LSPG>;
    constexpr int TILES = Tl::COUNT;
    extern __shared__ __align__(16) double smem[];

    const int N = p.N, n = p.n;
    const PromGeom geo = PromGeom::make(N);
    const int S = geo.S, lgS = geo.lgS, NP = geo.NP;
    const QmSmem lay = QmSmem::make(NP);
    double *s_h6 = smem + lay.h6;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t = lane & 3, g = lane >> 2;
    double *wbase = smem + lay.per_cta + warp * lay.per_warp;
    double *w_u = wbase + lay.u;
    double *w_misc = wbase + lay.misc;
    double *w_bn = wbase + lay.bn;
    double *w_mat = w_u;                 // 16 x MAT_LD, over the dead u (+ the first words of misc)
    double *w_piv = w_misc + 16;         // pivot rows, 2 x (NC+2) (<= 68)
    double *w_q = w_misc + 96;           // q (32), zero beyond n
    // tangent rows of this warp: row r = node r - 1, r = 0 .. NP + 1 (rows 0 and > N stay zero)
    const size_t tstride = (size_t)(NP + 2 + QM_PF) * 32;
    double *Tcta = p.Tg + (size_t)blockIdx.x * QM_MAXW * tstride;
    double *Tw = Tcta + warp * tstride;

    const double At = p.At;

    for (int idx = threadIdx.x; idx < lay.arr; idx += blockDim.x) s_h6[idx] = 0.0;
    __syncthreads();
    for (int e = threadIdx.x; e <= N - 2; e += blockDim.x) s_h6[sk(e, S, lgS)] = (p.X[e + 1] - p.X[e]) / 6.0;
    for (int r = lane; r < (NP + 2 + QM_PF) * 32; r += 32)
        if (r < 32 || r >= (N + 1) * 32) Tw[r] = 0.0;
    __syncthreads();

    const int run0 = t << lgS;
    const int ubase = t * (S + 1) + 2;                       // sk(run0 + s) = ubase + s
    const int gR = n & 7;                                    // lane-group of the R column (its block is NB-1)
    const int nblk = S >> 3;
    const double *t_run = Tw + (size_t)(run0 + 1) * 32;      // tangent row of node run0

    PromRowCtx rc;
    rc.u = w_u; rc.bn = w_bn; rc.hfs = nullptr; rc.h6 = s_h6;
    rc.a6 = At / 6.0; rc.c1 = 0.0; rc.ate = At * p.E; rc.has_e = p.E != 0.0; rc.supg = false;
    rc.N = N; rc.S = S; rc.lgS = lgS;

    const int jc = lane < n ? lane : n - 1;   // column of Phi^T U this lane owns (lanes >= n shadow n - 1 and discard)

    // ---- tangent / decoder pass for the eight trajectories of the CTA at the q's published in the w_q's ----
    constexpr int NKS = 2 * NB;
    constexpr int HB = (NB + 1) / 2;
    auto tangent_pass = [&]() {
        __syncthreads();   // every q published; nobody still reads T / u of the previous pass
        double Bq[NKS], qC[NB][2];
        {
            const double *qg = smem + lay.per_cta + g * lay.per_warp + lay.misc + 96;         // q of trajectory g
            const double *q0 = smem + lay.per_cta + (2 * t) * lay.per_warp + lay.misc + 96;   // trajectories 2t, 2t+1
            const double *q1 = q0 + lay.per_warp;
#pragma unroll
            for (int ks = 0; ks < NKS; ++ks) Bq[ks] = qg[4 * ks + t];
#pragma unroll
            for (int jb = 0; jb < NB; ++jb) { qC[jb][0] = q0[8 * jb + g]; qC[jb][1] = q1[8 * jb + g]; }
        }
        double *u0s = smem + lay.per_cta + (2 * t) * lay.per_warp + lay.u;   // u of trajectory 2t (2t+1: + per_warp)
        double A[HB][NKS];
        if (warp < N) qm_load_half<NB>(A, p.Hs, lane, warp, 0);
        for (int i = warp; i < N; i += QM_MAXW) {
            double s0 = 0.0, s1 = 0.0;
            double *T0 = Tcta + (size_t)(2 * t) * tstride + (size_t)(i + 1) * 32 + g;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                double c[HB][2];
                const int inext = hh == 0 ? i : (i + QM_MAXW < N ? i + QM_MAXW : -1);
                qm_half<NB>(A, Bq, qC, p.Hs, p.Phi, n, lane, i, hh, inext, 1 - hh, c, s0, s1);
#pragma unroll
                for (int b2 = 0; b2 < HB; ++b2) {
                    const int jb = hh * HB + b2;
                    if (jb < NB) { T0[8 * jb] = c[b2][0]; T0[tstride + 8 * jb] = c[b2][1]; }
                }
            }
            s0 += shfl_xor(s0, 4); s1 += shfl_xor(s1, 4);
            s0 += shfl_xor(s0, 8); s1 += shfl_xor(s1, 8);
            s0 += shfl_xor(s0, 16); s1 += shfl_xor(s1, 16);
            if (g == 0) {
                const int o = sk(i, S, lgS);
                u0s[o] = 0.5 * s0;
                u0s[lay.per_warp + o] = 0.5 * s1;
            }
        }
        __syncthreads();   // T rows (global) and u (shared) of all trajectories written
    };

    __shared__ int s_base, s_seg;
    for (;;) {
        __syncthreads();
        // work unit = (segment of time steps, lock-step group), handed out round-robin so that all groups advance
        // together (same scheme as the FOM sweep, csrc/bx_fom.cu); a group resumes from its own last snapshots
        if (threadIdx.x == 0) {
            const unsigned unit = atomicAdd(p.counter, 1u);
            int sgv = -1;
            if (unit < (unsigned)p.ngroups * (unsigned)p.nseg) {
                sgv = (int)(unit / (unsigned)p.ngroups);
                const int gi = (int)(unit - (unsigned)sgv * (unsigned)p.ngroups);
                if (sgv > 0) {
                    const volatile int *flag = p.progress + gi;
                    while (*flag < sgv) __nanosleep(500);
                    __threadfence();
                }
                s_base = gi * p.group;
            }
            s_seg = sgv;
        }
        __syncthreads();
        if (s_seg < 0) break;
        const int sg = s_seg;
        const int n_begin = sg * p.seg;
        const int n_end = min(p.nsteps, n_begin + p.seg);
        const bool active = warp < p.group && s_base + warp < p.B;   // idle warps (ragged last group, small groups) carry q = 0
        const int smp = active ? s_base + warp : p.B - 1;
        rc.mu1 = p.mu1[smp];
        const double *atf = p.atf + (size_t)smp * N;

        double *Us = p.U + (long long)smp * (long long)(p.nsteps + 1) * N;
        bool bad = false;   // this warp's trajectory stopped being finite
        {
            const double *u0p = sg == 0 ? p.u0 + (long long)smp * p.u0_stride : Us + (long long)n_begin * N;
            bool nan_seen = false;
            for (int i = lane - 1; i <= NP; i += 32) {
                const int o = sk(i, S, lgS);
                const double v = (active && i >= 0 && i < N) ? __ldcg(u0p + i) : 0.0;
                nan_seen |= v != v;
                w_u[o] = v;
                w_bn[o] = 0.0;
                if (sg == 0 && active && i >= 0 && i < N) Us[i] = v;
            }
            bad = sg > 0 && __any_sync(FULL, nan_seen);   // a trajectory flagged in an earlier segment left NaN snapshots
            if (bad)
                for (int i = lane - 1; i <= NP; i += 32) w_u[sk(i, S, lgS)] = 0.0;
        }
        __syncwarp();

        for (int step = n_begin; step < n_end; ++step) {
            // ---- bn = M U^m + At F (:1140) and the initial guess q = Phi^T U^m (:1129) --------
            for (int i = lane; i < N; i += 32) {
                const double h6L = s_h6[sk(i - 1, S, lgS)], h6R = s_h6[sk(i, S, lgS)];
                const double um = w_u[sk(i - 1, S, lgS)], uc = w_u[sk(i, S, lgS)], up = w_u[sk(i + 1, S, lgS)];
                w_bn[sk(i, S, lgS)] = h6L * um + (2.0 * (h6L + h6R)) * uc + h6R * up + atf[i];
            }
            double q;
            {
                double q4[4] = {0.0, 0.0, 0.0, 0.0};
                const double *ph = p.Phi + jc;
                int i = 0;
                for (; i + 4 <= N; i += 4) {
#pragma unroll
                    for (int r = 0; r < 4; ++r) q4[r] = fma(ph[(size_t)(i + r) * n], w_u[sk(i + r, S, lgS)], q4[r]);
                }
                for (; i < N; ++i) q4[0] = fma(ph[(size_t)i * n], w_u[sk(i, S, lgS)], q4[0]);
                q = (active && !bad && lane < n) ? (q4[0] + q4[1]) + (q4[2] + q4[3]) : 0.0;
            }
            __syncwarp();
            w_q[lane] = q;
            tangent_pass();   // u = decoder(q) (:1130), T(q)

            int it_done = 0;
            bool done = !active || bad;
            for (int it = 0; it < p.itmax; ++it) {
              if (!done) {
                // ================= contraction (:1134-1156) ==================================
                double acc[TILES][2];
#pragma unroll
                for (int i = 0; i < TILES; ++i) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
                double pm[NB], p0[NB], pp[NB];
                // tangent rows come from L2 / HBM (written by the tangent pass of all eight warps): a four-deep
                // register ring keeps the loads QM_PF node steps ahead of the DMMAs that consume them
                double ring[QM_PF][NB];
#pragma unroll
                for (int b = 0; b < NB; ++b) {
                    pm[b] = t_run[-32 + b * 8 + g];
                    p0[b] = t_run[b * 8 + g];
#pragma unroll
                    for (int r = 0; r < QM_PF; ++r) ring[r][b] = t_run[(r + 1) * 32 + b * 8 + g];
                }
                PromRow nxt = prom_row(rc, run0 + g, ubase + g);
                for (int blk = 0; blk < nblk; ++blk) {
                    const PromRow cur = nxt;
                    if (blk + 1 < nblk) nxt = prom_row(rc, run0 + ((blk + 1) << 3) + g, ubase + ((blk + 1) << 3) + g);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int s = (blk << 3) + j;
                        const double *rp = t_run + (size_t)(s + 1 + QM_PF) * 32;   // (rows past the run end: padding)
#pragma unroll
                        for (int b = 0; b < NB; ++b) {
                            pp[b] = ring[j % QM_PF][b];
                            ring[j % QM_PF][b] = rp[b * 8 + g];
                        }
                        const int src = 4 * j + t;
                        const double ra = shfl_idx(cur.a, src), rb = shfl_idx(cur.b, src);
                        const double rcc = shfl_idx(cur.c, src), rR = shfl_idx(cur.R, src);
                        double jp[NB];
#pragma unroll
                        for (int b = 0; b < NB; ++b) jp[b] = fma(ra, pm[b], fma(rb, p0[b], rcc * pp[b]));
                        if (g == gR) jp[NB - 1] = rR;   // R rides as column n (block NB-1)
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

                // ================= reduced solve (:1161-1165) ===============================
                double a[NC];
                double rhs = 0.0;
#pragma unroll
                for (int pass = 0; pass < (NB + 1) / 2; ++pass) {
                    __syncwarp();
#pragma unroll
                    for (int I = 0; I < NB; ++I) {
#pragma unroll
                        for (int J = LSPG ? I : 0; J < NB; ++J) {
                            const int ti = Tl::index(I, J);
                            const int r = I * 8 + g, c = J * 8 + 2 * t;
                            if ((I >> 1) == pass) {
                                w_mat[(r - 16 * pass) * MAT_LD + c] = acc[ti][0];
                                w_mat[(r - 16 * pass) * MAT_LD + c + 1] = acc[ti][1];
                            }
                            if (LSPG && J > I && (J >> 1) == pass) {
                                w_mat[(c - 16 * pass) * MAT_LD + r] = acc[ti][0];
                                w_mat[(c + 1 - 16 * pass) * MAT_LD + r] = acc[ti][1];
                            }
                        }
                    }
                    __syncwarp();
                    if (pass == 0 || lane >= 16) {
                        const double *mr = w_mat + (lane & 15) * MAT_LD;
#pragma unroll
                        for (int j = 0; j < NC; ++j) a[j] = mr[j];
                        rhs = -mr[n];
                    }
                }
                __syncwarp();
                const double dq = lu_solve_rows<NC, true>(a, rhs, n, lane, w_piv);
                q = (lane < n) ? q + dq : 0.0;
                const double sd = warp_sum((lane < n) ? dq * dq : 0.0);
                const double sq = warp_sum(q * q);
                const double rel = sqrt(sd) / fmax(1e-14, sqrt(sq));
                if (p.err_hist && lane == 0) p.err_hist[((long long)smp * p.nsteps + step) * p.itmax + it] = rel;
                if (!((sd + sq) <= 1.7976931348623157e308)) { bad = true; done = true; q = 0.0; }   // inf / NaN
                __syncwarp();
                w_q[lane] = q;
                if (lane == 0) { w_u[sk(-1, S, lgS)] = 0.0; w_u[sk(NP, S, lgS)] = 0.0; }   // halo words the matrix overwrote
                for (int i = N + lane; i < NP; i += 32) w_u[sk(i, S, lgS)] = 0.0;           // padded nodes likewise
                it_done = it + 1;
                if (rel < p.tol) done = true;
              }
                tangent_pass();   // u = decoder(q) (:1163) and the tangent of the next iteration
                if (__syncthreads_and(done ? 1 : 0)) break;
            }
            if (active) {
                double *Un1 = Us + (long long)(step + 1) * N;
                // a trajectory whose reduced system went singular / overflowed (np.linalg.LinAlgError at :1161) carries
                // the flag -1 and NaN snapshots from that step on; the other seven of the group are unaffected
                const double qnan = __longlong_as_double(0x7ff8000000000000ll);
                for (int i = lane; i < N; i += 32) Un1[i] = bad ? qnan : w_u[sk(i, S, lgS)];
                if (lane == 0 && p.iters) p.iters[(long long)smp * p.nsteps + step] = bad ? -1 : it_done;
            }
        }
        if (p.nseg > 1) {   // publish the segment: every thread's snapshot stores, then the group's progress word
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) *reinterpret_cast<volatile int *>(p.progress + s_base / p.group) = sg + 1;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// v3: same lock-step design, but the tangent rows never leave the SM.  The nodes are cut into chunks that follow
// the contraction's own order (its four k-lanes walk four runs of S nodes in parallel): chunk blk holds, for
// run t and j = 0..7, the tangent row of node t S + 8 blk + j + 1 — exactly the rows block blk of the contraction
// consumes as its "next node" operand.  Per barrier interval every warp first multiplies its four nodes of chunk
// blk + 1 (DMMA, Hs fragments from L2) into one half of a double-buffered shared tile [8 trajectories][32 nodes]
// [32 columns] (XOR-swizzled: conflict-free for the fragment stores and for the contraction's loads), then runs
// block blk of its own trajectory's contraction out of the other half.  The seven run-boundary rows a contraction
// starts from go through the small global scratch (prologue).  A final tangent-only pass per time step leaves
// u = decoder(q) of every trajectory in shared memory (:1163).
struct QmTsSmem {
    int T, total;
    __host__ __device__ static int trajstride() { return 32 * 32 + 2; }   // 2 mod 16: fragment stores conflict-free
    __host__ __device__ static QmTsSmem make(const QmSmem &lay) {
        QmTsSmem m;
        int o = lay.per_cta + QM_MAXW * lay.per_warp;
        o = (o + 1) & ~1;
        m.T = o; o += 2 * QM_MAXW * trajstride();
        m.total = o;
        return m;
    }
};

// CB = 0: tangent rows through the shared tile, one 8-step block per chunk (above).  CB > 0: through the global
// scratch in chunks of CB blocks, produced TWO chunks ahead of the contraction (so its register ring may read into
// the next chunk): only 3 chunks of rows are live at a time and they stay in L2 — for the mode counts whose
// accumulators leave no room for the shared tile's extra pressure (NB = 4).
template <int NB, bool LSPG, int CB>
__global__ void __launch_bounds__(QM_MAXW * 32, 1) qm_kernel_ts(const QmParams p) {
    constexpr int NC = NB * 8;
    constexpr int NKS = 2 * NB;
    using Tl = PromTiles<NB, LSPG>;
    constexpr int TILES = Tl::COUNT;
    extern __shared__ __align__(16) double smem[];

    const int N = p.N, n = p.n;
    const PromGeom geo = PromGeom::make(N);
    const int S = geo.S, lgS = geo.lgS, NP = geo.NP;
    const QmSmem lay = QmSmem::make(NP);
    const QmTsSmem tl = QmTsSmem::make(lay);
    constexpr int TSTR = 32 * 32 + 2;
    constexpr int CBB = CB ? CB : 1;       // blocks per chunk
    constexpr int LA = CB ? 2 : 1;         // chunks the tangent runs ahead of the contraction
    double *s_h6 = smem + lay.h6;
    double *s_T = smem + tl.T;             // (CB = 0 only)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t = lane & 3, g = lane >> 2;
    double *wbase = smem + lay.per_cta + warp * lay.per_warp;
    double *w_u = wbase + lay.u;
    double *w_misc = wbase + lay.misc;
    double *w_bn = wbase + lay.bn;
    double *w_mat = w_u;
    double *w_piv = w_misc + 16;
    double *w_q = w_misc + 96;
    // global scratch: only the run-boundary rows are used (row r = node r - 1; rows 0 and > N stay zero)
    const size_t tstride = (size_t)(NP + 2 + QM_PF) * 32;
    double *Tcta = p.Tg + (size_t)blockIdx.x * QM_MAXW * tstride;
    double *Tw = Tcta + warp * tstride;

    const double At = p.At;
    for (int idx = threadIdx.x; idx < lay.arr; idx += blockDim.x) s_h6[idx] = 0.0;
    __syncthreads();
    for (int e = threadIdx.x; e <= N - 2; e += blockDim.x) s_h6[sk(e, S, lgS)] = (p.X[e + 1] - p.X[e]) / 6.0;
    for (int r = lane; r < (NP + 2 + QM_PF) * 32; r += 32)
        if (r < 32 || r >= (N + 1) * 32) Tw[r] = 0.0;
    __syncthreads();

    const int run0 = t << lgS;
    const int ubase = t * (S + 1) + 2;
    const int gR = n & 7;
    const int nblk = S >> 3;
    const double *t_run = Tw + (size_t)(run0 + 1) * 32;
    // this warp's trajectory inside a shared T buffer: row of (j, t) at (4 j + t) 32, column c at c ^ 4 t
    const int myT = warp * TSTR + t * 32;

    PromRowCtx rc;
    rc.u = w_u; rc.bn = w_bn; rc.hfs = nullptr; rc.h6 = s_h6;
    rc.a6 = At / 6.0; rc.c1 = 0.0; rc.ate = At * p.E; rc.has_e = p.E != 0.0; rc.supg = false;
    rc.N = N; rc.S = S; rc.lgS = lgS;
    const int jc = lane < n ? lane : n - 1;

    // q fragments of the eight trajectories (owners' shared memory), loaded once per chunk
    const double *qB = smem + lay.per_cta + g * lay.per_warp + lay.misc + 96 + t;            // B operand: q of trajectory g
    const double *qC0 = smem + lay.per_cta + (2 * t) * lay.per_warp + lay.misc + 96 + g;     // trajectories 2t, 2t+1
    auto load_q = [&](double (&Bq)[NKS], double (&qC)[NB][2]) {
#pragma unroll
        for (int ks = 0; ks < NKS; ++ks) Bq[ks] = qB[4 * ks];
#pragma unroll
        for (int jb = 0; jb < NB; ++jb) { qC[jb][0] = qC0[8 * jb]; qC[jb][1] = qC0[lay.per_warp + 8 * jb]; }
    };
    constexpr int HB = (NB + 1) / 2;
    double *u0s = smem + lay.per_cta + (2 * t) * lay.per_warp + lay.u;
    // one node for the eight trajectories (two halves); dstS != nullptr: rows to the shared tile (row base dstS of
    // trajectory 0, swizzle sw), else to the global scratch.  A holds half 0 of node i on entry and half 0 of node
    // inext (if >= 0) on exit.
    auto node = [&](double (&A)[HB][NKS], const double (&Bq)[NKS], const double (&qC)[NB][2], int i, int inext,
                    double *dstS, int sw) {
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            double c[HB][2];
            qm_half<NB>(A, Bq, qC, p.Hs, p.Phi, n, lane, i, hh, hh == 0 ? i : inext, 1 - hh, c, s0, s1);
            if (dstS) {
                double *T0 = dstS + (2 * t) * TSTR;
#pragma unroll
                for (int b2 = 0; b2 < HB; ++b2) {
                    const int jb = hh * HB + b2;
                    if (jb < NB) { T0[(8 * jb + g) ^ sw] = c[b2][0]; T0[TSTR + ((8 * jb + g) ^ sw)] = c[b2][1]; }
                }
            } else {
                double *T0 = Tcta + (size_t)(2 * t) * tstride + (size_t)(i + 1) * 32 + g;
#pragma unroll
                for (int b2 = 0; b2 < HB; ++b2) {
                    const int jb = hh * HB + b2;
                    if (jb < NB) { T0[8 * jb] = c[b2][0]; T0[tstride + 8 * jb] = c[b2][1]; }
                }
            }
        }
        s0 += shfl_xor(s0, 4); s1 += shfl_xor(s1, 4);
        s0 += shfl_xor(s0, 8); s1 += shfl_xor(s1, 8);
        s0 += shfl_xor(s0, 16); s1 += shfl_xor(s1, 16);
        if (g == 0) {
            const int o = sk(i, S, lgS);
            u0s[o] = 0.5 * s0;
            u0s[lay.per_warp + o] = 0.5 * s1;
        }
    };
    // tangent rows of chunk blk into shared buffer `buf`: this warp takes rows nidx = warp, warp + 8, +16, +24
    // (node index grows with r, so the rows past N form a tail)
    const int nch = (nblk + CBB - 1) / CBB;
    auto chunk_pass = [&](int c, int buf) {
        double *base = CB ? nullptr : s_T + buf * (QM_MAXW * TSTR);
        double A[HB][NKS], Bq[NKS], qC[NB][2];
        load_q(Bq, qC);
        auto nd = [&](int r) { const int nidx = warp + 8 * r; return ((nidx & 3) << lgS) + c * (8 * CBB) + (nidx >> 2) + 1; };
        bool have = false;
#pragma unroll
        for (int r = 0; r < 4 * CBB; ++r) {
            const int nidx = warp + 8 * r, i = nd(r);
            if (i < N) {
                if (!have) qm_load_half<NB>(A, p.Hs, lane, i, 0);
                const int inx = r < 4 * CBB - 1 ? nd(r + 1) : N;
                have = inx < N;
                node(A, Bq, qC, i, have ? inx : -1, CB ? nullptr : base + nidx * 32, 4 * (nidx & 3));
            } else {
                have = false;
                if (!CB) {
                    double *d = base + nidx * 32 + lane;   // all 32 columns (the swizzle permutes within a row)
#pragma unroll
                    for (int tr = 0; tr < QM_MAXW; ++tr) d[tr * TSTR] = 0.0;
                }   // (global scratch: rows past N were zeroed once and are never written)
            }
        }
    };

    __shared__ int s_base, s_seg;
    for (;;) {
        __syncthreads();
        // work unit = (segment of time steps, lock-step group), handed out round-robin so that all groups advance
        // together (same scheme as the FOM sweep, csrc/bx_fom.cu); a group resumes from its own last snapshots
        if (threadIdx.x == 0) {
            const unsigned unit = atomicAdd(p.counter, 1u);
            int sgv = -1;
            if (unit < (unsigned)p.ngroups * (unsigned)p.nseg) {
                sgv = (int)(unit / (unsigned)p.ngroups);
                const int gi = (int)(unit - (unsigned)sgv * (unsigned)p.ngroups);
                if (sgv > 0) {
                    const volatile int *flag = p.progress + gi;
                    while (*flag < sgv) __nanosleep(500);
                    __threadfence();
                }
                s_base = gi * p.group;
            }
            s_seg = sgv;
        }
        __syncthreads();
        if (s_seg < 0) break;
        const int sg = s_seg;
        const int n_begin = sg * p.seg;
        const int n_end = min(p.nsteps, n_begin + p.seg);
        const bool active = warp < p.group && s_base + warp < p.B;   // idle warps (ragged last group, small groups) carry q = 0
        const int smp = active ? s_base + warp : p.B - 1;
        rc.mu1 = p.mu1[smp];
        const double *atf = p.atf + (size_t)smp * N;
        double *Us = p.U + (long long)smp * (long long)(p.nsteps + 1) * N;
        bool bad = false;   // this warp's trajectory stopped being finite
        {
            const double *u0p = sg == 0 ? p.u0 + (long long)smp * p.u0_stride : Us + (long long)n_begin * N;
            bool nan_seen = false;
            for (int i = lane - 1; i <= NP; i += 32) {
                const int o = sk(i, S, lgS);
                const double v = (active && i >= 0 && i < N) ? __ldcg(u0p + i) : 0.0;
                nan_seen |= v != v;
                w_u[o] = v;
                w_bn[o] = 0.0;
                if (sg == 0 && active && i >= 0 && i < N) Us[i] = v;
            }
            bad = sg > 0 && __any_sync(FULL, nan_seen);   // a trajectory flagged in an earlier segment left NaN snapshots
            if (bad)
                for (int i = lane - 1; i <= NP; i += 32) w_u[sk(i, S, lgS)] = 0.0;
        }
        __syncwarp();

        for (int step = n_begin; step < n_end; ++step) {
            for (int i = lane; i < N; i += 32) {
                const double h6L = s_h6[sk(i - 1, S, lgS)], h6R = s_h6[sk(i, S, lgS)];
                const double um = w_u[sk(i - 1, S, lgS)], uc = w_u[sk(i, S, lgS)], up = w_u[sk(i + 1, S, lgS)];
                w_bn[sk(i, S, lgS)] = h6L * um + (2.0 * (h6L + h6R)) * uc + h6R * up + atf[i];
            }
            double q;
            {
                double q4[4] = {0.0, 0.0, 0.0, 0.0};
                const double *ph = p.Phi + jc;
                int i = 0;
                for (; i + 4 <= N; i += 4) {
#pragma unroll
                    for (int r = 0; r < 4; ++r) q4[r] = fma(ph[(size_t)(i + r) * n], w_u[sk(i + r, S, lgS)], q4[r]);
                }
                for (; i < N; ++i) q4[0] = fma(ph[(size_t)i * n], w_u[sk(i, S, lgS)], q4[0]);
                q = (active && !bad && lane < n) ? (q4[0] + q4[1]) + (q4[2] + q4[3]) : 0.0;
            }
            __syncwarp();
            w_q[lane] = q;

            int it_done = 0;
            bool done = !active || bad;
            for (int it = 0; it < p.itmax; ++it) {
                // ---- prologue: run-boundary rows (global scratch) and chunk 0 ----------------------------
                __syncthreads();   // every q published; previous contraction / solve finished
                if (warp < 7) {
                    const int tb = (warp + 1) >> 1;                       // 0, 1, 1, 2, 2, 3, 3
                    const int i = warp == 0 ? 0 : (tb << lgS) - (warp & 1);   // 0, S-1, S, 2S-1, 2S, 3S-1, 3S
                    if (i < N) {
                        double Ab[HB][NKS], Bq[NKS], qC[NB][2];
                        load_q(Bq, qC);
                        qm_load_half<NB>(Ab, p.Hs, lane, i, 0);
                        node(Ab, Bq, qC, i, -1, nullptr, 0);
                    }
                }
                chunk_pass(0, 0);
                if (LA == 2 && nch > 1) chunk_pass(1, 1);
                __syncthreads();
                double acc[TILES][2];
#pragma unroll
                for (int i = 0; i < TILES; ++i) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
                double pm[NB], p0[NB], pp[NB];
#pragma unroll
                for (int b = 0; b < NB; ++b) {
                    pm[b] = t_run[-32 + b * 8 + g];
                    p0[b] = t_run[b * 8 + g];
                }
                for (int c = 0; c < nch; ++c) {
                    if (c + LA < nch) chunk_pass(c + LA, (c + LA) & 1);
                    if (!done) {
                        const int bend = (c + 1) * CBB < nblk ? (c + 1) * CBB : nblk;
                        // the ring is refilled per chunk so that it is not live across the tangent pass above
                        double ring[CB ? QM_PF : 1][NB];
                        if (CB) {
#pragma unroll
                            for (int b = 0; b < NB; ++b) {
#pragma unroll
                                for (int r = 0; r < QM_PF; ++r)
                                    ring[r][b] = t_run[(size_t)(c * CBB * 8 + r + 1) * 32 + b * 8 + g];
                            }
                        }
                        for (int blk = c * CBB; blk < bend; ++blk) {
                            const PromRow cur = prom_row(rc, run0 + (blk << 3) + g, ubase + (blk << 3) + g);
                            const double *Tb = s_T + (blk & 1) * (QM_MAXW * TSTR) + myT;
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                if (CB) {   // register ring QM_PF node steps ahead (rows of the next chunk are complete)
                                    const double *rp = t_run + (size_t)((blk << 3) + j + 1 + QM_PF) * 32;
#pragma unroll
                                    for (int b = 0; b < NB; ++b) {
                                        pp[b] = ring[j % QM_PF][b];
                                        ring[j % QM_PF][b] = rp[b * 8 + g];
                                    }
                                } else {
#pragma unroll
                                    for (int b = 0; b < NB; ++b) pp[b] = Tb[j * 128 + ((b * 8 + g) ^ (4 * t))];
                                }
                                const int src = 4 * j + t;
                                const double ra = shfl_idx(cur.a, src), rb = shfl_idx(cur.b, src);
                                const double rcc = shfl_idx(cur.c, src), rR = shfl_idx(cur.R, src);
                                double jp[NB];
#pragma unroll
                                for (int b = 0; b < NB; ++b) jp[b] = fma(ra, pm[b], fma(rb, p0[b], rcc * pp[b]));
                                if (g == gR) jp[NB - 1] = rR;
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
                    }
                    __syncthreads();
                }
                if (!done) {
                    // ================= reduced solve (:1161-1165) ===============================
                    double a[NC];
                    double rhs = 0.0;
#pragma unroll
                    for (int pass = 0; pass < (NB + 1) / 2; ++pass) {
                        __syncwarp();
#pragma unroll
                        for (int I = 0; I < NB; ++I) {
#pragma unroll
                            for (int J = LSPG ? I : 0; J < NB; ++J) {
                                const int ti = Tl::index(I, J);
                                const int r = I * 8 + g, c = J * 8 + 2 * t;
                                if ((I >> 1) == pass) {
                                    w_mat[(r - 16 * pass) * MAT_LD + c] = acc[ti][0];
                                    w_mat[(r - 16 * pass) * MAT_LD + c + 1] = acc[ti][1];
                                }
                                if (LSPG && J > I && (J >> 1) == pass) {
                                    w_mat[(c - 16 * pass) * MAT_LD + r] = acc[ti][0];
                                    w_mat[(c + 1 - 16 * pass) * MAT_LD + r] = acc[ti][1];
                                }
                            }
                        }
                        __syncwarp();
                        if (pass == 0 || lane >= 16) {
                            const double *mr = w_mat + (lane & 15) * MAT_LD;
#pragma unroll
                            for (int j = 0; j < NC; ++j) a[j] = mr[j];
                            rhs = -mr[n];
                        }
                    }
                    __syncwarp();
                    const double dq = lu_solve_rows<NC, true>(a, rhs, n, lane, w_piv);
                    q = (lane < n) ? q + dq : 0.0;
                    const double sd = warp_sum((lane < n) ? dq * dq : 0.0);
                    const double sq = warp_sum(q * q);
                    const double rel = sqrt(sd) / fmax(1e-14, sqrt(sq));
                    if (p.err_hist && lane == 0) p.err_hist[((long long)smp * p.nsteps + step) * p.itmax + it] = rel;
                    if (!((sd + sq) <= 1.7976931348623157e308)) { bad = true; done = true; q = 0.0; }   // inf / NaN
                    __syncwarp();
                    w_q[lane] = q;
                    if (lane == 0) { w_u[sk(-1, S, lgS)] = 0.0; w_u[sk(NP, S, lgS)] = 0.0; }
                    for (int i = N + lane; i < NP; i += 32) w_u[sk(i, S, lgS)] = 0.0;
                    it_done = it + 1;
                    if (rel < p.tol) done = true;
                }
                if (__syncthreads_and(done ? 1 : 0)) break;
            }
            // ---- u = decoder(q) at the final q of every trajectory (:1163 / :1130): tangent-only pass -----
            __syncthreads();
            for (int c = 0; c < nch; ++c) chunk_pass(c, 0);
            if (warp == 0) {   // node 0 is in no chunk (chunks hold rows 8 blk + j + 1)
                double Ab[HB][NKS], Bq[NKS], qC[NB][2];
                load_q(Bq, qC);
                qm_load_half<NB>(Ab, p.Hs, lane, 0, 0);
                node(Ab, Bq, qC, 0, -1, nullptr, 0);
            }
            __syncthreads();
            if (active) {
                double *Un1 = Us + (long long)(step + 1) * N;
                // a trajectory whose reduced system went singular / overflowed (np.linalg.LinAlgError at :1161) carries
                // the flag -1 and NaN snapshots from that step on; the other seven of the group are unaffected
                const double qnan = __longlong_as_double(0x7ff8000000000000ll);
                for (int i = lane; i < N; i += 32) Un1[i] = bad ? qnan : w_u[sk(i, S, lgS)];
                if (lane == 0 && p.iters) p.iters[(long long)smp * p.nsteps + step] = bad ? -1 : it_done;
            }
        }
        if (p.nseg > 1) {   // publish the segment: every thread's snapshot stores, then the group's progress word
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) *reinterpret_cast<volatile int *>(p.progress + s_base / p.group) = sg + 1;
        }
    }
}

struct QmPlan {
    int sms, maxsm, NP, NB;
    size_t off_hs, off_atf, off_tg, off_prog, total;
};

static int qm_plan(int B, int N, int n, QmPlan &pl) {
    DeviceInfo di;
    {
        const int rcd = device_info(di);
        if (rcd != BX_OK) return rcd;
    }
    pl.sms = di.sms;
    pl.maxsm = di.max_smem_optin;
    const PromGeom geo = PromGeom::make(N);
    pl.NP = geo.NP;
    pl.NB = n / 8 + 1;   // 8 NB > n: column n of the contraction carries R
    size_t o = 256;
    pl.off_hs = o; o += sizeof(double) * (size_t)N * pl.NB * 32 * 2 * pl.NB; o = (o + 255) & ~(size_t)255;
    pl.off_atf = o; o += sizeof(double) * (size_t)B * N; o = (o + 255) & ~(size_t)255;
    pl.off_tg = o; o += sizeof(double) * (size_t)pl.sms * QM_MAXW * (size_t)(pl.NP + 2 + QM_PF) * 32;
    o = (o + 255) & ~(size_t)255;
    pl.off_prog = o; o += sizeof(int) * (size_t)(B > 0 ? B : 1);
    pl.total = o;
    return BX_OK;
}

template <int NB, bool LSPG>
static int launch_qm_t(QmParams p, const QmPlan &pl, const double *H, cudaStream_t st) {
    const QmSmem lay = QmSmem::make(pl.NP);
    const int maxsm = pl.maxsm;
    const long long warps = QM_MAXW;   // eight trajectories per CTA in lock step (the M of the tangent DMMA)
    if ((long long)sizeof(double) * (lay.per_cta + warps * lay.per_warp) > (long long)maxsm) {
        set_error("bx_qm_prom_trajectories: N = %d, n = %d does not fit in shared memory", p.N, p.n);
        return BX_ERR_UNSUPPORTED;
    }
    p.warps = (int)warps;
    size_t smem = sizeof(double) * ((size_t)lay.per_cta + (size_t)warps * lay.per_warp);
    auto kern = qm_kernel<NB, LSPG>;
    const QmTsSmem tl = QmTsSmem::make(lay);
    // v3 with the shared tile for NB <= QM_TS_MAX_NB (when it fits), else v2.  The chunked-global variant of v3 is kept
    // for A/B runs only: measured slower than v2 at n = 28 (4.8 / 5.1 vs 5.3 / 5.7 M it/s, profiles/r02n_qm_modes.txt).
    // bx_set_tuning(BX_TUNE_QM_SCHEDULE, 0 (v2) / 1 (shared tile) / 2 (chunked global)) forces a choice.
    int mode = (sizeof(double) * (size_t)tl.total <= (size_t)maxsm && NB <= QM_TS_MAX_NB) ? 1 : 0;
    {
        const int f = tuning(BX_TUNE_QM_SCHEDULE);
        if (f == 0) mode = 0;
        if (f == 1 && sizeof(double) * (size_t)tl.total <= (size_t)maxsm) mode = 1;
        if (f == 2) mode = 2;
    }
    if (mode == 1) {
        kern = qm_kernel_ts<NB, LSPG, 0>;
        smem = sizeof(double) * (size_t)tl.total;
    } else if (mode == 2) {
        kern = qm_kernel_ts<NB, LSPG, 2>;
    }
    {
        const int rco = kernel_occupancy(reinterpret_cast<const void *>(kern), (int)warps * 32, smem, nullptr);
        if (rco != BX_OK) return rco;
    }
    // a sweep smaller than the device (strong scaling: 512 samples per GPU) spreads over ALL SMs: lock-step groups of
    // ceil(B / SMs) trajectories instead of full groups of eight on a fraction of the SMs (the idle rows of the
    // tangent DMMA cost nothing there: the SM would otherwise be empty)
    {
        const long long per_sm = ((long long)p.B + pl.sms - 1) / pl.sms;
        p.group = (int)(per_sm < 1 ? 1 : per_sm > QM_MAXW ? QM_MAXW : per_sm);
        const int f = tuning(BX_TUNE_QM_GROUP);
        if (f >= 1 && f <= QM_MAXW) p.group = f;
    }
    long long ctas = pl.sms;
    const long long need = ((long long)p.B + p.group - 1) / p.group;
    p.ngroups = (int)need;
    p.nseg = 1; p.seg = p.nsteps > 0 ? p.nsteps : 1;
    if (ctas >= need) {
        ctas = need;
    } else if (!p.err_hist) {   // more groups than SMs: round-robin segments of time steps
        int seg = tuning(BX_TUNE_FOM_SEGMENT_STEPS);
        if (seg < 0) seg = 25;
        if (seg > 0 && seg < p.nsteps) { p.seg = seg; p.nseg = (p.nsteps + seg - 1) / seg; }
    }
    BX_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st));
    if (p.nseg > 1) BX_CUDA(cudaMemsetAsync(p.progress, 0, sizeof(int) * (size_t)p.ngroups, st));
    const long long work = (long long)p.N * NB * 32 * 2 * NB + (long long)p.B * p.N;
    qm_prep_kernel<<<(unsigned)((work + 255) / 256), 256, 0, st>>>(p.B, p.N, NB, p.n, p.k, p.X, p.mu2, p.At, H,
                                                                 const_cast<double *>(p.Hs), p.atf);
    BX_LAUNCH_CHECK();
    kern<<<(unsigned)ctas, (unsigned)warps * 32, smem, st>>>(p);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

template <bool LSPG>
static int launch_qm(const QmParams &p, const QmPlan &pl, const double *H, cudaStream_t st) {
    if (p.n <= 7) return launch_qm_t<1, LSPG>(p, pl, H, st);
    if (p.n <= 15) return launch_qm_t<2, LSPG>(p, pl, H, st);
    if (p.n <= 23) return launch_qm_t<3, LSPG>(p, pl, H, st);
    if (p.n <= 31) return launch_qm_t<4, LSPG>(p, pl, H, st);
    set_error("bx_qm_prom_trajectories: n = %d modes > 31 is not supported by this kernel", p.n);
    return BX_ERR_UNSUPPORTED;
}

}  // namespace bx

extern "C" int bx_qm_schedule_available(int N, int n, int schedule) {
    using namespace bx;
    if (N < 2 || n < 1 || n > 31) return 0;
    if (schedule == 0 || schedule == 2) return 1;
    if (schedule != 1) return 0;
    QmPlan pl;
    if (qm_plan(1, N, n, pl) != BX_OK) return 0;
    const QmTsSmem tl = QmTsSmem::make(QmSmem::make(pl.NP));
    return sizeof(double) * (size_t)tl.total <= (size_t)pl.maxsm ? 1 : 0;
}

extern "C" long long bx_qm_workspace_bytes(int B, int N, int n) {
    using namespace bx;
    if (B < 0 || N < 2 || n < 1) return -1;
    QmPlan pl;
    if (qm_plan(B, N, n, pl) != BX_OK) return -1;
    return (long long)pl.total;
}

extern "C" int bx_qm_prom_trajectories(int B, int N, int n, int nsteps, const double *X, const double *u0,
                                       int u0_batched, const double *uxa, const double *mu2, double At, double E,
                                       const double *Phi, const double *H, int projection, double tol, int itmax,
                                       double *U, int *iters, double *err_hist, void *workspace, void *stream) {
    using namespace bx;
    if (projection != BX_GALERKIN && projection != BX_LSPG) {
        set_error("projection must be 'Galerkin' or 'LSPG'");
        return BX_ERR_PROJECTION;
    }
    BX_REQUIRE(B >= 0 && nsteps >= 0, "B and nsteps must be non-negative");
    BX_REQUIRE(N >= 2, "need at least 2 nodes (got N = %d)", N);
    BX_REQUIRE(n >= 1 && n <= N, "need 1 <= n <= N modes (got n = %d)", n);
    BX_REQUIRE(itmax >= 0, "newton_itmax must be >= 0");
    if (B == 0) return BX_OK;
    BX_REQUIRE(X && u0 && uxa && mu2 && Phi && H && U && workspace, "null device pointer");
    QmPlan pl;
    int rc = qm_plan(B, N, n, pl);
    if (rc != BX_OK) return rc;
    char *ws = static_cast<char *>(workspace);
    BX_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 15u) == 0, "workspace must be 16-byte aligned");
    QmParams p;
    p.B = B; p.N = N; p.n = n; p.k = n * (n + 1) / 2; p.nsteps = nsteps; p.itmax = itmax; p.warps = 0;
    p.X = X; p.u0 = u0; p.u0_stride = u0_batched ? N : 0;
    p.mu1 = uxa; p.mu2 = mu2; p.At = At; p.E = E; p.tol = tol;
    p.Phi = Phi;
    p.counter = queue_slot(ws);
    p.Hs = reinterpret_cast<double *>(ws + pl.off_hs);
    p.atf = reinterpret_cast<double *>(ws + pl.off_atf);
    p.Tg = reinterpret_cast<double *>(ws + pl.off_tg);
    p.progress = reinterpret_cast<int *>(ws + pl.off_prog);
    p.U = U; p.iters = iters; p.err_hist = err_hist;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return projection == BX_LSPG ? launch_qm<true>(p, pl, H, st) : launch_qm<false>(p, pl, H, st);
}
