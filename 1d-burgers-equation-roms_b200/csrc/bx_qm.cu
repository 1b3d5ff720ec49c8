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
// One WARP integrates one trajectory (up to 8 per CTA, no CTA-wide barrier in the time loop):
//   T  tangent/decoder pass: four nodes at a time, their packed H rows staged in shared memory,
//      lane j owns column j; T rows go to a per-warp scratch in global memory (L2 resident), u to
//      shared memory.
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
    const double *X;
    const double *u0;
    long long u0_stride;
    const double *mu1;   // uxa per sample (left Dirichlet value)
    const double *mu2;
    double At, E, tol;
    const double *Phi;   // (N, n) row-major
    const double *Hp;    // workspace: (NP4, k) H with doubled diagonal monomials, zero rows beyond N
    double *atf;         // workspace: At * F (B, N)
    double *Tg;          // workspace: per warp slot (NP + 1) x 32 tangent rows (row 0 = zero row for node -1)
    double *U;
    int *iters;
    unsigned int *counter;
};

constexpr int QM_MAXW = 8;
constexpr int QM_NODES = 4;   // nodes per tangent step

struct QmSmem {
    int h6, per_cta;
    int u, misc, bn, hrow, per_warp, arr;
    __host__ __device__ static QmSmem make(int NP, int k) {
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
        o = (o + 1) & ~1;
        m.hrow = o; o += (QM_NODES * k + 1) & ~1;
        m.per_warp = (o + 1) & ~1;
        return m;
    }
};

// Hp[i][c] = H[i][c] * (2 if monomial c is a square q_a^2 else 1); rows N .. NP4-1 are zero.  atf = At F.
__global__ void __launch_bounds__(256) qm_prep_kernel(int B, int N, int NP4, int n, int k, const double *X,
                                                      const double *mu2v, double At, const double *H, double *Hp,
                                                      double *atf) {
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    const long long nH = (long long)NP4 * k;
    if (idx < nH) {
        const int i = (int)(idx / k), c = (int)(idx - (long long)i * k);
        // c is a diagonal monomial iff c = a n - a (a - 1) / 2 for some a
        int a = 0, base = 0;
        while (base + (n - a) <= c) { base += n - a; ++a; }
        Hp[idx] = (i < N) ? H[(size_t)i * k + c] * (c == base ? 2.0 : 1.0) : 0.0;
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

template <int NB, bool LSPG>
__global__ void __launch_bounds__(QM_MAXW * 32, 1) qm_kernel(const QmParams p) {
    constexpr int NC = NB * 8;
    using Tl = PromTiles<NB, LSPG>;
    constexpr int TILES = Tl::COUNT;
    extern __shared__ __align__(16) double smem[];

    const int N = p.N, n = p.n, kq = p.k;
    const PromGeom geo = PromGeom::make(N);
    const int S = geo.S, lgS = geo.lgS, NP = geo.NP;
    const int NP4 = (N + QM_NODES - 1) / QM_NODES * QM_NODES;
    const QmSmem lay = QmSmem::make(NP, kq);
    double *s_h6 = smem + lay.h6;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t = lane & 3, g = lane >> 2;
    double *wbase = smem + lay.per_cta + warp * lay.per_warp;
    double *w_u = wbase + lay.u;
    double *w_misc = wbase + lay.misc;
    double *w_bn = wbase + lay.bn;
    double *w_h = wbase + lay.hrow;
    double *w_mat = w_u;                 // 16 x MAT_LD, over the dead u (+ the first words of misc)
    double *w_piv = w_misc + 16;         // pivot rows, 2 x (NC+2) (<= 68)
    double *w_q = w_misc + 96;           // q (32), zero beyond n
    // tangent rows of this warp: row r = node r - 1, r = 0 .. NP + 1 (rows 0 and > N stay zero)
    double *Tw = p.Tg + ((size_t)blockIdx.x * p.warps + warp) * (size_t)(NP + 2) * 32;

    const double At = p.At;

    for (int idx = threadIdx.x; idx < lay.arr; idx += blockDim.x) s_h6[idx] = 0.0;
    __syncthreads();
    for (int e = threadIdx.x; e <= N - 2; e += blockDim.x) s_h6[sk(e, S, lgS)] = (p.X[e + 1] - p.X[e]) / 6.0;
    for (int r = lane; r < (NP + 2) * 32; r += 32)
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

    // column owned by this lane in the tangent pass; lanes >= n shadow column n-1 and discard
    const int jc = lane < n ? lane : n - 1;

    // tangent/decoder pass at the q published in w_q: T rows -> Tw, u -> w_u
    auto tangent_pass = [&]() {
        const double qj = lane < n ? w_q[lane] : 0.0;
        for (int i0 = 0; i0 < NP4; i0 += QM_NODES) {
            {
                const double2 *src = reinterpret_cast<const double2 *>(p.Hp + (size_t)i0 * kq);
                double2 *dst = reinterpret_cast<double2 *>(w_h);
                for (int x = lane; x < (QM_NODES / 2) * kq; x += 32) dst[x] = src[x];
            }
            __syncwarp();
            double phi[QM_NODES], acc[QM_NODES];
#pragma unroll
            for (int r = 0; r < QM_NODES; ++r) {
                phi[r] = (lane < n && i0 + r < N) ? p.Phi[(size_t)(i0 + r) * n + lane] : 0.0;
                acc[r] = phi[r];
            }
            int idx = jc;
            for (int m = 0; m < n; ++m) {
                const double qm = w_q[m];
#pragma unroll
                for (int r = 0; r < QM_NODES; ++r) acc[r] = fma(w_h[r * kq + idx], qm, acc[r]);
                idx += (m < jc) ? (n - m - 1) : 1;
            }
            double part[QM_NODES];
#pragma unroll
            for (int r = 0; r < QM_NODES; ++r) {
                const double tv = lane < n ? acc[r] : 0.0;
                Tw[(size_t)(i0 + r + 1) * 32 + lane] = tv;
                part[r] = (phi[r] + tv) * qj;
            }
            // sum the four node values over the warp: recursive halving, then a butterfly
            double w2[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const double send = (lane & 16) ? part[j] : part[j + 2];
                const double keep = (lane & 16) ? part[j + 2] : part[j];
                w2[j] = keep + shfl_xor(send, 16);
            }
            double v;
            {
                const double send = (lane & 8) ? w2[0] : w2[1];
                const double keep = (lane & 8) ? w2[1] : w2[0];
                v = keep + shfl_xor(send, 8);
            }
            v += shfl_xor(v, 4);
            v += shfl_xor(v, 2);
            v += shfl_xor(v, 1);
            const int node = i0 + ((lane >> 3) & 3);
            if ((lane & 7) == 0 && node < N) w_u[sk(node, S, lgS)] = 0.5 * v;
            __syncwarp();
        }
    };

    for (;;) {
        int smp = 0;
        if (lane == 0) smp = (int)atomicAdd(p.counter, 1u);
        smp = __shfl_sync(FULL, smp, 0);
        if (smp >= p.B) break;
        rc.mu1 = p.mu1[smp];
        const double *atf = p.atf + (size_t)smp * N;

        double *Us = p.U + (long long)smp * (long long)(p.nsteps + 1) * N;
        {
            const double *u0p = p.u0 + (long long)smp * p.u0_stride;
            for (int i = lane - 1; i <= NP; i += 32) {
                const int o = sk(i, S, lgS);
                const double v = (i >= 0 && i < N) ? u0p[i] : 0.0;
                w_u[o] = v;
                w_bn[o] = 0.0;
                if (i >= 0 && i < N) Us[i] = v;
            }
        }
        __syncwarp();

        for (int step = 0; step < p.nsteps; ++step) {
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
                q = lane < n ? (q4[0] + q4[1]) + (q4[2] + q4[3]) : 0.0;
            }
            __syncwarp();
            w_q[lane] = q;
            __syncwarp();
            tangent_pass();   // u = decoder(q) (:1130), T(q)

            int it_done = 0;
            for (int it = 0; it < p.itmax; ++it) {
                // ================= contraction (:1134-1156) ==================================
                double acc[TILES][2];
#pragma unroll
                for (int i = 0; i < TILES; ++i) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
                double pm[NB], p0[NB], pp[NB];
#pragma unroll
                for (int b = 0; b < NB; ++b) {
                    pm[b] = t_run[-32 + b * 8 + g];
                    p0[b] = t_run[b * 8 + g];
                }
                PromRow nxt = prom_row(rc, run0 + g, ubase + g);
                for (int blk = 0; blk < nblk; ++blk) {
                    const PromRow cur = nxt;
                    if (blk + 1 < nblk) nxt = prom_row(rc, run0 + ((blk + 1) << 3) + g, ubase + ((blk + 1) << 3) + g);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int s = (blk << 3) + j;
                        const double *rp = t_run + (size_t)(s + 1) * 32;
#pragma unroll
                        for (int b = 0; b < NB; ++b) pp[b] = rp[b * 8 + g];
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
                __syncwarp();
                w_q[lane] = q;
                if (lane == 0) { w_u[sk(-1, S, lgS)] = 0.0; w_u[sk(NP, S, lgS)] = 0.0; }   // halo words the matrix overwrote
                for (int i = N + lane; i < NP; i += 32) w_u[sk(i, S, lgS)] = 0.0;           // padded nodes likewise
                __syncwarp();
                tangent_pass();   // u = decoder(q) (:1163) and the tangent of the next iteration
                it_done = it + 1;
                if (rel < p.tol) break;
            }
            double *Un1 = Us + (long long)(step + 1) * N;
            for (int i = lane; i < N; i += 32) Un1[i] = w_u[sk(i, S, lgS)];
            if (lane == 0 && p.iters) p.iters[(long long)smp * p.nsteps + step] = it_done;
        }
        __syncwarp();
    }
}

struct QmPlan {
    int sms, NP, NP4;
    size_t off_hp, off_atf, off_tg, total;
};

static int qm_plan(int B, int N, int n, QmPlan &pl) {
    int dev = 0;
    BX_CUDA(cudaGetDevice(&dev));
    BX_CUDA(cudaDeviceGetAttribute(&pl.sms, cudaDevAttrMultiProcessorCount, dev));
    const PromGeom geo = PromGeom::make(N);
    pl.NP = geo.NP;
    pl.NP4 = (N + QM_NODES - 1) / QM_NODES * QM_NODES;
    const size_t k = (size_t)n * (n + 1) / 2;
    size_t o = 256;
    pl.off_hp = o; o += sizeof(double) * (size_t)pl.NP4 * k; o = (o + 255) & ~(size_t)255;
    pl.off_atf = o; o += sizeof(double) * (size_t)B * N; o = (o + 255) & ~(size_t)255;
    pl.off_tg = o; o += sizeof(double) * (size_t)pl.sms * QM_MAXW * (size_t)(pl.NP + 2) * 32;
    pl.total = o;
    return BX_OK;
}

template <int NB, bool LSPG>
static int launch_qm_t(QmParams p, const QmPlan &pl, const double *H, cudaStream_t st) {
    const QmSmem lay = QmSmem::make(pl.NP, p.k);
    int dev = 0, maxsm = 0;
    BX_CUDA(cudaGetDevice(&dev));
    BX_CUDA(cudaDeviceGetAttribute(&maxsm, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    const long long avail = (long long)maxsm / (long long)sizeof(double) - lay.per_cta;
    long long warps = avail > 0 ? avail / lay.per_warp : 0;
    if (warps > QM_MAXW) warps = QM_MAXW;
    if (warps < 1) {
        set_error("bx_qm_prom_trajectories: N = %d, n = %d does not fit in shared memory", p.N, p.n);
        return BX_ERR_UNSUPPORTED;
    }
    if (warps > p.B) warps = p.B;
    p.warps = (int)warps;
    const size_t smem = sizeof(double) * ((size_t)lay.per_cta + (size_t)warps * lay.per_warp);
    auto kern = qm_kernel<NB, LSPG>;
    BX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long ctas = pl.sms;
    const long long need = ((long long)p.B + warps - 1) / warps;
    if (ctas > need) ctas = need;
    BX_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st));
    const long long work = (long long)pl.NP4 * p.k + (long long)p.B * p.N;
    qm_prep_kernel<<<(unsigned)((work + 255) / 256), 256, 0, st>>>(p.B, p.N, pl.NP4, p.n, p.k, p.X, p.mu2, p.At, H,
                                                                 const_cast<double *>(p.Hp), p.atf);
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
                                       double *U, int *iters, void *workspace, void *stream) {
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
    p.counter = reinterpret_cast<unsigned int *>(ws);
    p.Hp = reinterpret_cast<double *>(ws + pl.off_hp);
    p.atf = reinterpret_cast<double *>(ws + pl.off_atf);
    p.Tg = reinterpret_cast<double *>(ws + pl.off_tg);
    p.U = U; p.iters = iters;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return projection == BX_LSPG ? launch_qm<true>(p, pl, H, st) : launch_qm<false>(p, pl, H, st);
}
