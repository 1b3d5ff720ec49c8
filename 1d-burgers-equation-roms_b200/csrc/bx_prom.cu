// bx_prom.cu — fused POD-PROM trajectory kernel: FEMBurgers.pod_prom_burgers
// (FEM/fem_burgers.py:709-785), Galerkin and LSPG, for B (mu1, mu2) samples at once.
//
// One WARP integrates one trajectory from u0 to step nsteps without leaving the SM and without
// any CTA-wide barrier; up to 8 warps share one CTA and one shared-memory copy of the basis Phi
// (two warps per SM sub-partition, so the latency-bound reduced solve of one trajectory hides
// behind the FP64-pipe-bound contraction of the other).  Warps pull samples from an atomic queue.
//
// One nonlinear iteration (:726-781):
//   B  contraction on the FP64 tensor pipe (DMMA, mma.sync.m8n8k4.f64), k index = node.  Thread
//      (t = lane%4, g = lane/4) walks the S = NP/4 consecutive nodes t S .. t S + S - 1 and owns
//      columns 8 blk + g, so every Phi entry is loaded once per pass (rolling 3-row window) and no
//      operand needs a shuffle.  Rows of the tridiagonal A = M + At C(U_k) + At E K (row 0 <- e0^T)
//      and of R = A U_k - b, b = M U^n + At F - At S(U_k), b[0] = mu1 are assembled eight nodes
//      ahead, ONE node per lane (closed forms of the 2-point Gauss element integrals), and handed
//      to the four k-lanes of each MMA step through a 1 KB staging tile.  J = A Phi is built on
//      the fly from the three bands:
//         LSPG     Ar = J^T J  (upper block triangle),  br = J^T R        (:760-762)
//         Galerkin Ar = Phi^T J,                         br = Phi^T R      (:756-757)
//      R rides along as column n of the B operand, so br is column n of the same accumulators.
//      q0 = Phi^T U_k (:770): U_k = Phi q_k for every iterate but the raw u0, so it is (Phi^T Phi) q_k with
//      Phi^T Phi formed once per CTA; only the first iterate of a trajectory takes an explicit pass.
//   C  accumulator tiles -> one matrix row per lane (through the dead U_k buffer), n x n LU with
//      partial pivoting (LAPACK gesv, :767) in registers, pivot rows broadcast through shared
//      memory, q = q0 + dq, err = |dq| / |q| (:776).
//   D  lift U_{k+1} = Phi q (:773): same thread-to-Phi mapping as B, eight nodes per lane at a
//      time, recursive-halving reduction over the eight column lanes.
// Forcing data (At F and (h/6)(f1+f2), FEM/fem_burgers.py:449-455, 565) depend on mu2 only and are
// evaluated once per sample by a small pre-kernel into the caller's workspace.
// HBM traffic: the 8N-byte snapshot per time step (+ 4 B iteration count).
#include "bx_prom_common.cuh"

namespace bx {

constexpr int PROM_MAXW = 8;

// shared-memory carve-up, in doubles.  Node/element arrays are skewed by one word per run so
// that the four k-lanes of an MMA step (nodes S apart) hit different banks.
struct PromSmem {
    int phi, h6, gm, per_cta;             // CTA-wide: permuted Phi (NP x ldp), h/6 per element, Phi^T Phi
    int u, misc, bn, hfs, per_warp, arr;
    __host__ __device__ static PromSmem make(int NP, int ldp, int phi_smem, int NC) {
        PromSmem m;
        m.arr = NP + 8;
        int o = 0;
        m.phi = o; o += phi_smem ? (NP + 4) * ldp : 0;   // + 4 wrap rows o = (o + 1) & ~1;
        m.h6 = o; o += m.arr;
        o = (o + 1) & ~1;
        m.gm = o; o += NC * MAT_LD;
        m.per_cta = (o + 1) & ~1;
        o = 0;
        int usz = m.arr < 16 * MAT_LD - 8 ? 16 * MAT_LD - 8 : m.arr;   // u + first 8 of misc hold the 16 x 33 matrix
        usz = (usz + 1) & ~1;
        m.u = o; o += usz;
        m.misc = o; o += 128;             // q0 / pivot rows / q (and the tail of the matrix)
        m.bn = o; o += m.arr;             // M U^n + At F
        m.hfs = o; o += m.arr;            // (h/6)(f_gp1 + f_gp2) per element
        m.per_warp = (o + 1) & ~1;
        return m;
    }
};

// ---- per-sample forcing data -------------------------------------------------------------
__global__ void __launch_bounds__(256) prom_forcing_kernel(int B, int N, const double *X, const double *mu2v,
                                                           double At, double *atf, double *hfs) {
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    if (idx >= (long long)B * N) return;
    const int smp = (int)(idx / N), i = (int)(idx - (long long)smp * N);
    const double mu2 = mu2v[smp];
    double F = 0.0, h = 0.0;
    if (i >= 1) {   // local node 2 of element i-1
        const double x1 = X[i - 1], x2 = X[i];
        const double f1 = 0.02 * exp(mu2 * (BX_GA * x1 + BX_GB * x2));
        const double f2 = 0.02 * exp(mu2 * (BX_GB * x1 + BX_GA * x2));
        F += (f1 * BX_GB + f2 * BX_GA) * (0.5 * (x2 - x1));
    }
    if (i <= N - 2) {   // local node 1 of element i
        const double x1 = X[i], x2 = X[i + 1];
        const double f1 = 0.02 * exp(mu2 * (BX_GA * x1 + BX_GB * x2));
        const double f2 = 0.02 * exp(mu2 * (BX_GB * x1 + BX_GA * x2));
        F += (f1 * BX_GA + f2 * BX_GB) * (0.5 * (x2 - x1));
        h = ((x2 - x1) / 6.0) * (f1 + f2);
    }
    atf[idx] = At * F;
    hfs[idx] = h;
}


int launch_prom_forcing(int B, int N, const double *X, const double *mu2, double At, double *atf, double *hfs,
                        cudaStream_t st) {
    prom_forcing_kernel<<<(unsigned)(((long long)B * N + 255) / 256), 256, 0, st>>>(B, N, X, mu2, At, atf, hfs);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

// bx_prom_wide.cu: one warp per trajectory, 32 <= n <= 47 (register accumulators, reduced solve in shared memory)
int launch_prom_wide(const PromParams &p, bool lspg, cudaStream_t st);

// bx_prom_cta.cu: one CTA per trajectory, 32 <= n <= 240
int launch_prom_cta(int B, int N, int n, int nsteps, const double *X, const double *u0, int u0_batched,
                    const double *mu1, const double *mu2, double At, double E, const double *Phi, int projection,
                    double tol, int kmax, double *U, int *iters, double *err_hist, void *workspace, long long mat_offset_bytes,
                    cudaStream_t st, long long *mat_bytes_out);

template <int NB, bool LSPG, bool PHIS>
__global__ void __launch_bounds__(PROM_MAXW * 32, 1) prom_kernel(const PromParams p) {
    constexpr int NC = NB * 8;
    using Tl = PromTiles<NB, LSPG>;
    constexpr int TILES = Tl::COUNT;
    extern __shared__ __align__(16) double smem[];

    const int N = p.N, n = p.n, ldp = p.ldp;
    const PromGeom geo = PromGeom::make(N);
    const int S = geo.S, lgS = geo.lgS, NP = geo.NP;
    const PromSmem lay = PromSmem::make(NP, ldp, PHIS ? 1 : 0, NC);
    double *s_phi = smem + lay.phi;
    double *s_h6 = smem + lay.h6;
    double *s_gm = smem + lay.gm;          // Phi^T Phi, NC x MAT_LD

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t = lane & 3, g = lane >> 2;
    double *wbase = smem + lay.per_cta + warp * lay.per_warp;
    double *w_u = wbase + lay.u;
    double *w_misc = wbase + lay.misc;
    double *w_bn = wbase + lay.bn;
    double *w_hfs = wbase + lay.hfs;
    double *w_mat = w_u;                 // 16 x MAT_LD, over the dead U_k (+ the first words of misc)
    double *w_piv = w_misc + 16;         // pivot rows, 2 x (NC+2) (<= 68)
    double *w_q = w_misc + 16;           // q after the solve     (32), over the dead pivot rows

    const double At = p.At;

    // ---- CTA prologue: permuted Phi (+ 4 wrap rows) and mesh constants ---------------------
    if (PHIS) {
        for (int idx = threadIdx.x; idx < (NP + 4) * ldp; idx += blockDim.x) {
            const int r = idx / ldp, j = idx - r * ldp;
            // row 4 s + t holds node t S + s; wrap row NP + t holds node (t+1) S (the row after run t)
            const int i = r < NP ? ((r & 3) << lgS) + (r >> 2) : (r - NP < 3 ? (r - NP + 1) << lgS : N);
            s_phi[idx] = (i < N && j < n) ? p.Phi[(size_t)i * n + j] : 0.0;
        }
    }
    for (int idx = threadIdx.x; idx < lay.arr; idx += blockDim.x) s_h6[idx] = 0.0;
    for (int idx = threadIdx.x; idx < NC * NC; idx += blockDim.x) {
        const int j = idx / NC, m = idx - j * NC;
        double sum = 0.0;
        if (j < n && m < n)
            for (int i = 0; i < N; ++i) sum = fma(p.Phi[(size_t)i * n + j], p.Phi[(size_t)i * n + m], sum);
        s_gm[j * MAT_LD + m] = sum;
    }
    __syncthreads();
    for (int e = threadIdx.x; e <= N - 2; e += blockDim.x) s_h6[sk(e, S, lgS)] = (p.X[e + 1] - p.X[e]) / 6.0;
    __syncthreads();

    // this thread's run of nodes, its Phi rows, its columns
    const int run0 = t << lgS;
    const double *phi_run = PHIS ? s_phi + t * ldp : p.Phi + (size_t)run0 * n;
    const int row_stride = PHIS ? 4 * ldp : n;
    const int ubase = t * (S + 1) + 2;                       // sk(run0 + s) = ubase + s
    const int gR = n & 7;                                    // lane-group of the R column (its block is NB-1)
    // Shared-memory Phi is padded with zero columns up to ldp; a ragged last block may still read a few
    // words of the next row (finite values): they only reach accumulator rows / columns >= n, which
    // nothing reads.  The global-memory path predicates instead (it must not read past the array).
    const bool col_last = PHIS || ((NB - 1) * 8 + g) < n;
    const int nblk = S >> 3;

    PromRowCtx rc;
    rc.u = w_u; rc.bn = w_bn; rc.hfs = w_hfs; rc.h6 = s_h6;
    rc.a6 = At / 6.0; rc.c1 = 1.5 * At; rc.ate = At * p.E; rc.has_e = p.E != 0.0; rc.supg = true;
    rc.N = N; rc.S = S; rc.lgS = lgS;

    // Work units are SEGMENTS of trajectories, handed out round-robin (unit = segment * B + sample) so that all
    // trajectories advance together and every warp slot stays busy until the last segment; a trajectory resumes from
    // its own last snapshot, gated by a progress word (same scheme as the FOM sweep, csrc/bx_fom.cu).
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

        // ---- per-sample forcing constants and the state at the start of the segment ------
        double *Us = p.U + (long long)smp * (long long)(p.nsteps + 1) * N;
        {
            const double *hf = p.hfs + (size_t)smp * N;
            const double *src = p.u0 + (long long)smp * p.u0_stride;
            double poison = 0.0;
            if (sg > 0) {   // every lane polls the same word (see csrc/bx_fom.cu for why not lane 0 alone)
                int done;
                const volatile int *flag = p.progress + smp;
                while ((done = *flag) < sg && done >= 0) __nanosleep(500);
                done = __shfl_sync(FULL, done, 0);
                __threadfence();
                // a trajectory that stopped being finite earlier is poisoned as data: its first step here takes the flag
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
        // q0 = Phi^T U_k of the first iterate (the raw u0, which is not in the span of Phi, :716)
        double q0;
        {
            double q4[4] = {0.0, 0.0, 0.0, 0.0};
            const double *phc = p.Phi + (lane < n ? lane : 0);
            int i = 0;
            for (; i + 4 <= N; i += 4) {
#pragma unroll
                for (int r = 0; r < 4; ++r) q4[r] = fma(phc[(size_t)(i + r) * n], w_u[sk(i + r, S, lgS)], q4[r]);
            }
            for (; i < N; ++i) q4[0] = fma(phc[(size_t)i * n], w_u[sk(i, S, lgS)], q4[0]);
            q0 = lane < n ? (q4[0] + q4[1]) + (q4[2] + q4[3]) : 0.0;
        }

        bool bad = false;
        for (int step = n_begin; step < n_end; ++step) {
            // ---- bn = M U^n + At F (:745), constant during the nonlinear loop ------------
            for (int i = lane; i < N; i += 32) {
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
                double pm[NB], p0[NB], pp[NB];
                {
                    const double *rm = PHIS ? s_phi + (((S - 1) << 2) + (t > 0 ? t - 1 : 0)) * ldp
                                            : p.Phi + (size_t)(run0 > 0 ? run0 - 1 : 0) * n;
                    const bool okm = t > 0 && (PHIS || run0 - 1 < N);
                    const bool ok0 = PHIS || run0 < N;
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        const bool cok = (b < NB - 1) || col_last;
                        pm[b] = (cok && okm) ? rm[b * 8 + g] : 0.0;
                        p0[b] = (cok && ok0) ? phi_run[b * 8 + g] : 0.0;
                    }
                }
                // rows are assembled one block (8 nodes per k-lane) ahead, ONE node per lane
                // (node run0 + 8 blk + g), and handed to the k-lanes by shuffles
                PromRow nxt = prom_row(rc, run0 + g, ubase + g);
                for (int blk = 0; blk < nblk; ++blk) {
                    const PromRow cur = nxt;
                    if (blk + 1 < nblk) nxt = prom_row(rc, run0 + ((blk + 1) << 3) + g, ubase + ((blk + 1) << 3) + g);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int s = (blk << 3) + j;
                        // Phi row of node run0 + s + 1 (wrap rows make the end of a run uniform)
                        if (PHIS) {
                            const double *rp = phi_run + (size_t)(s + 1) * row_stride;
#pragma unroll
                            for (int b = 0; b < NB; ++b) pp[b] = rp[b * 8 + g];
                        } else {
                            const int inx = run0 + s + 1;
                            const double *rp = p.Phi + (size_t)(inx < N ? inx : 0) * n;
#pragma unroll
                            for (int b = 0; b < NB; ++b) {
                                const bool cok = (b < NB - 1) || col_last;
                                pp[b] = (cok && inx < N) ? rp[b * 8 + g] : 0.0;
                            }
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
                // ================= phase C: reduced solve =================================
                // accumulator tiles -> one matrix row per lane, 16 rows per pass; column n = br
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
                const double dq = lu_solve_rows<NC, !LSPG>(a, rhs, n, lane, w_piv);
                const double q = (lane < n) ? q0 + dq : 0.0;   // q0 = Phi^T U_k (:770)
                const double sd = warp_sum((lane < n) ? dq * dq : 0.0);
                const double sq = warp_sum(q * q);
                const double err = sqrt(sd) / sqrt(sq);
                __syncwarp();
                w_q[lane] = q;
                if (lane == 0) { w_u[sk(-1, S, lgS)] = 0.0; w_u[sk(NP, S, lgS)] = 0.0; }   // halo words the matrix overwrote
                __syncwarp();
                // Phi^T U_{k+1} for the next iteration (or time step): U_{k+1} = Phi q, so it is (Phi^T Phi) q
                {
                    const double *gr = s_gm + (lane < n ? lane : 0) * MAT_LD;
                    double s0 = 0.0, s1 = 0.0;
                    int m = 0;
                    for (; m + 2 <= n; m += 2) { s0 = fma(gr[m], w_q[m], s0); s1 = fma(gr[m + 1], w_q[m + 1], s1); }
                    if (m < n) s0 = fma(gr[m], w_q[m], s0);
                    q0 = (lane < n) ? s0 + s1 : 0.0;
                }

                // ================= phase D: lift U_{k+1} = Phi q (:773) ===================
                {
                    double qv[NB];
#pragma unroll
                    for (int b = 0; b < NB; ++b) qv[b] = w_q[b * 8 + g];   // zero for columns >= n
                    for (int blk = 0; blk < nblk; ++blk) {
                        double v[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int s = (blk << 3) + j;
                            const double *r0 = phi_run + (size_t)s * row_stride;
                            const bool rowok = PHIS || run0 + s < N;
                            double vv = 0.0;
#pragma unroll
                            for (int b = 0; b < NB; ++b) {
                                const bool cok = (b < NB - 1) || col_last;
                                vv = fma((cok && rowok) ? r0[b * 8 + g] : 0.0, qv[b], vv);
                            }
                            v[j] = vv;
                        }
                        // recursive halving over the eight column lanes: lane g ends with node 8 blk + g
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
            // ---- snapshot step+1 (:783) and iteration count --------------------------------
            double *Un1 = Us + (long long)(step + 1) * N;
            for (int i = lane; i < N; i += 32) Un1[i] = w_u[sk(i, S, lgS)];
            if (lane == 0 && p.iters) p.iters[(long long)smp * p.nsteps + step] = finite ? k : -1;
            if (!finite) {   // singular reduced system / overflow (np.linalg.LinAlgError at :767): flag -1 and NaN
                const double qnan = __longlong_as_double(0x7ff8000000000000ll);   // snapshots from here on, batch goes on
                for (++step; step < n_end; ++step) {
                    double *Un2 = Us + (long long)(step + 1) * N;
                    for (int i = lane; i < N; i += 32) Un2[i] = qnan;
                    if (lane == 0 && p.iters) p.iters[(long long)smp * p.nsteps + step] = -1;
                }
                bad = true;
            }
        }
        if (p.nseg > 1) {   // publish the segment: the snapshot stores of all lanes, then the progress word (all lanes)
            __threadfence();
            __syncwarp();
            *reinterpret_cast<volatile int *>(p.progress + smp) = bad ? -1 : sg + 1;
        }
        __syncwarp();
    }
}

template <int NB, bool LSPG, bool PHIS>
static int launch_prom_t(PromParams p, int min_warps, cudaStream_t st, bool *fits) {
    const PromGeom geo = PromGeom::make(p.N);
    const PromSmem lay = PromSmem::make(geo.NP, p.ldp, PHIS ? 1 : 0, NB * 8);
    DeviceInfo di;
    {
        const int rcd = device_info(di);
        if (rcd != BX_OK) return rcd;
    }
    const int sms = di.sms, maxsm = di.max_smem_optin;
    const long long avail = (long long)maxsm / (long long)sizeof(double) - lay.per_cta;
    long long warps = avail > 0 ? avail / lay.per_warp : 0;
    if (warps > PROM_MAXW) warps = PROM_MAXW;
    *fits = warps >= min_warps;
    if (!*fits) return BX_OK;
    // a sweep smaller than the device (strong scaling: 512 samples per GPU) spreads over ALL SMs first: as many
    // warps per CTA as it takes to seat B trajectories on `sms` CTAs, not full CTAs on a fraction of the SMs
    {
        const long long per_sm = ((long long)p.B + sms - 1) / sms;
        if (warps > per_sm) warps = per_sm < 1 ? 1 : per_sm;
    }
    p.warps = (int)warps;
    const size_t smem = sizeof(double) * ((size_t)lay.per_cta + (size_t)warps * lay.per_warp);
    auto kern = prom_kernel<NB, LSPG, PHIS>;
    {
        const int rco = kernel_occupancy(reinterpret_cast<const void *>(kern), (int)warps * 32, smem, nullptr);
        if (rco != BX_OK) return rco;
    }
    long long ctas = sms;
    const long long need = ((long long)p.B + warps - 1) / warps;
    p.nseg = 1; p.seg = p.nsteps > 0 ? p.nsteps : 1;
    if (ctas >= need) {
        ctas = need;
    } else if (!p.err_hist) {   // more trajectories than warp slots: round-robin segments
        int seg = tuning(BX_TUNE_FOM_SEGMENT_STEPS);
        if (seg < 0) seg = 25;
        if (seg > 0 && seg < p.nsteps) { p.seg = seg; p.nseg = (p.nsteps + seg - 1) / seg; }
    }
    BX_REQUIRE((long long)p.B * p.nseg < 0xffffffffLL, "too many work units (B = %d, %d segments)", p.B, p.nseg);
    BX_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st));
    if (p.nseg > 1) BX_CUDA(cudaMemsetAsync(p.progress, 0, sizeof(int) * (size_t)p.B, st));
    prom_forcing_kernel<<<(unsigned)(((long long)p.B * p.N + 255) / 256), 256, 0, st>>>(p.B, p.N, p.X, p.mu2, p.At,
                                                                                        p.atf, p.hfs);
    BX_LAUNCH_CHECK();
    kern<<<(unsigned)ctas, (unsigned)warps * 32, smem, st>>>(p);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

template <int NB, bool LSPG>
static int launch_prom_nb(const PromParams &p, cudaStream_t st) {
    bool fits = false;
    // Phi in shared memory if at least four trajectories still fit next to it; else Phi from global / L2
    int rc = launch_prom_t<NB, LSPG, true>(p, 4, st, &fits);
    if (rc != BX_OK || fits) return rc;
    rc = launch_prom_t<NB, LSPG, false>(p, 1, st, &fits);
    if (rc != BX_OK || fits) return rc;
    set_error("bx_prom_trajectories: N = %d, n = %d does not fit in shared memory", p.N, p.n);
    return BX_ERR_UNSUPPORTED;
}

template <bool LSPG>
static int launch_prom(const PromParams &p, cudaStream_t st) {
    if (p.n <= 7) return launch_prom_nb<1, LSPG>(p, st);
    if (p.n <= 15) return launch_prom_nb<2, LSPG>(p, st);
    if (p.n <= 23) return launch_prom_nb<3, LSPG>(p, st);
    if (p.n <= 31) return launch_prom_nb<4, LSPG>(p, st);
    set_error("bx_prom_trajectories: internal dispatch error (n = %d)", p.n);
    return BX_ERR_UNSUPPORTED;
}

}  // namespace bx

// 256-byte head | At F (B, N) | (h/6)(f1 + f2) (B, N) | progress words (B,)
static long long prom_forcing_bytes(int B, int N) {
    return (256 + 2LL * (long long)B * (long long)N * (long long)sizeof(double) + 4LL * B + 15) & ~15LL;
}

extern "C" long long bx_prom_workspace_bytes(int B, int N, int n) {
    if (B < 0 || N < 2 || n < 1) return -1;
    long long mat = 0;
    if (n >= 32 && n <= 240) {   // the two projections pick different launch shapes: size for the larger need
        static const int projs[2] = {BX_GALERKIN, BX_LSPG};
        for (int pi = 0; pi < 2; ++pi) {
            long long m = 0;
            if (bx::launch_prom_cta(B, N, n, 0, nullptr, nullptr, 0, nullptr, nullptr, 0.0, 0.0, nullptr, projs[pi], 0.0, 1,
                                    nullptr, nullptr, nullptr, nullptr, 0, nullptr, &m) != BX_OK)
                return -1;
            if (m > mat) mat = m;
        }
    }
    return prom_forcing_bytes(B, N) + mat;
}

extern "C" int bx_prom_trajectories(int B, int N, int n, int nsteps, const double *X, const double *u0,
                                    int u0_batched, const double *mu1, const double *mu2, double At,
                                    double E, const double *Phi, int projection, double tol, int kmax,
                                    double *U, int *iters, double *err_hist, void *workspace, void *stream) {
    using namespace bx;
    if (projection != BX_GALERKIN && projection != BX_LSPG) {
        set_error("Projection method '%d' is not available. Please use 'Galerkin' or 'LSPG'.", projection);
        return BX_ERR_PROJECTION;
    }
    BX_REQUIRE(B >= 0 && nsteps >= 0, "B and nsteps must be non-negative");
    BX_REQUIRE(N >= 2, "need at least 2 nodes (got N = %d)", N);
    BX_REQUIRE(n >= 1 && n <= N, "need 1 <= n <= N modes (got n = %d)", n);
    BX_REQUIRE(kmax >= 1, "kmax must be >= 1 (the reference loop body defines U1)");
    if (B == 0) return BX_OK;
    BX_REQUIRE(X && u0 && mu1 && mu2 && Phi && U && workspace, "null device pointer");
    // 32 <= n <= 48: the register kernel (bx_prom_wide.cu); BX_TUNE_PROM_WARPS_PER_TRAJECTORY = 8 forces the CTA kernel
    const bool wide = n >= 32 && n <= 48 && tuning(BX_TUNE_PROM_WARPS_PER_TRAJECTORY) != 8;
    if (n >= 32 && !wide) {
        if (n > 240) {
            set_error("bx_prom_trajectories: n = %d modes > 240 is not supported by this build", n);
            return BX_ERR_UNSUPPORTED;
        }
        return launch_prom_cta(B, N, n, nsteps, X, u0, u0_batched, mu1, mu2, At, E, Phi, projection, tol, kmax, U,
                               iters, err_hist, workspace, prom_forcing_bytes(B, N), static_cast<cudaStream_t>(stream), nullptr);
    }
    PromParams p;
    p.B = B; p.N = N; p.n = n; p.nsteps = nsteps; p.kmax = kmax;
    p.ldp = ((n + 3) / 8) * 8 + 4;   // smallest stride >= n that is 4 mod 8: conflict-free fragment loads
    p.warps = 0;
    p.X = X; p.u0 = u0; p.u0_stride = u0_batched ? N : 0;
    p.mu1 = mu1; p.mu2 = mu2; p.At = At; p.E = E; p.tol = tol;
    p.Phi = Phi; p.U = U; p.iters = iters; p.err_hist = err_hist;
    p.counter = queue_slot(workspace);
    p.atf = reinterpret_cast<double *>(static_cast<char *>(workspace) + 256);
    p.hfs = p.atf + (size_t)B * N;
    p.progress = reinterpret_cast<int *>(p.hfs + (size_t)B * N);
    p.nseg = 1; p.seg = nsteps > 0 ? nsteps : 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (wide) return launch_prom_wide(p, projection == BX_LSPG, st);
    return projection == BX_LSPG ? launch_prom<true>(p, st) : launch_prom<false>(p, st);
}
