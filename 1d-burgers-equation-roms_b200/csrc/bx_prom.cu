// bx_prom.cu — fused POD-PROM trajectory kernel: FEMBurgers.pod_prom_burgers
// (FEM/fem_burgers.py:709-785), Galerkin and LSPG, for B (mu1, mu2) samples at once.
//
// A GROUP of W = 4 warps integrates one trajectory from u0 to step nsteps without leaving the SM;
// G groups share one CTA (and one shared-memory copy of the basis Phi) and run independently of
// each other behind named barriers, so the latency-bound reduced solve of one group overlaps the
// FP64-pipe-bound contraction of the others.  Groups pull samples from an atomic work queue.
//
// One nonlinear iteration (:726-781) is four phases:
//   A  assembly: thread per node.  Tridiagonal A = M + At C(U_k) + At E K (row 0 <- e0^T) and the
//      residual R = A U_k - b, b = M U^n + At F - At S(U_k), b[0] = mu1, from the closed forms of
//      the 2-point Gauss element integrals; (dl, d, du, R) of every row go to shared memory.
//   B  contraction on the FP64 tensor pipe (DMMA, mma.sync.m8n8k4.f64).  With J = A Phi built on
//      the fly from the three bands and a rolling 3-row window of Phi,
//         LSPG     Ar = J^T J  (upper block triangle),  br = J^T R        (:760-762)
//         Galerkin Ar = Phi^T J,                         br = Phi^T R      (:756-757)
//      R rides along as column n of the B operand, so br is column n of the same accumulators.
//      The DMMA k index is a node; thread (t = lane%4, g = lane/4) of warp w walks the S
//      consecutive nodes (4w + t) S .. (4w + t) S + S - 1 and owns columns 8 blk + g, so each
//      Phi entry is loaded once per pass and no operand needs a shuffle.  Phi lives in shared
//      memory with its rows permuted to (w, s, t)-major order: conflict-free 64-bit loads.
//      q0 = Phi^T U_k (:770) is accumulated in the same pass.
//   C  warp 0: fixed-order reduction of the four warps' accumulators, n x n LU with partial
//      pivoting (LAPACK gesv, :767) one matrix row per lane in registers, q = q0 + dq,
//      err = |dq| / |q| (:776).
//   D  lift U_{k+1} = Phi q (:773), same thread-to-Phi mapping as B.
// HBM traffic: the 8N-byte snapshot per time step (+ 4 B iteration count); Phi is read once per CTA.
#include "bx_common.cuh"

namespace bx {

struct PromParams {
    int B, N, n, nsteps, kmax;
    const double *X;
    const double *u0;
    long long u0_stride;
    const double *mu1;
    const double *mu2;
    double At, E, tol;
    const double *Phi;  // (N, n) row-major (C order), shared by all samples
    int phi_smem;       // 1: the CTA keeps a permuted copy of Phi in shared memory; 0: read from global / L2
    double *U;
    int *iters;
    unsigned int *counter;
};

constexpr int PW = 4;          // warps per group
constexpr int PGT = PW * 32;   // threads per group

__device__ __forceinline__ void group_sync(int grp) {
    asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "r"(PGT) : "memory");
}

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// geometry shared by host and device: S = nodes per (warp, t) run (power of two >= 8),
// NP = 16 S padded node count
struct PromGeom {
    int S, lgS, NP;
    __host__ __device__ static PromGeom make(int N) {
        PromGeom g;
        g.lgS = 3;
        while ((16 << g.lgS) < N) ++g.lgS;
        g.S = 1 << g.lgS;
        g.NP = 16 << g.lgS;
        return g;
    }
};

// shared-memory carve-up, in doubles
struct PromSmem {
    int phi, h6, ke, per_cta;  // CTA-wide: permuted Phi (NP x n), h/6 and At E / h per element
    int u, rows, bn, hfs, atf, q, misc, per_group;
    __host__ __device__ static PromSmem make(int NP, int S, int n, int NC, int tiles, int phi_smem) {
        PromSmem m;
        int o = 0;
        m.phi = o; o += phi_smem ? NP * n : 0; o = (o + 1) & ~1;
        m.h6 = o; o += NP + 2;   // element e at index e + 1 (e = -1 .. NP-1)
        m.ke = o; o += NP + 2;
        m.per_cta = (o + 1) & ~1;
        o = 0;
        // (dl, d, du, R) per permuted row; reused as scratch for the cross-warp reduction
        // ([2][tiles][32][2]) and then for the dense reduced system (NC x (NC+1))
        int rows = 4 * NP;
        if (rows < 2 * tiles * 64) rows = 2 * tiles * 64;
        if (rows < NC * (NC + 1)) rows = NC * (NC + 1);
        m.rows = o; o += (rows + 1) & ~1;
        m.u = o; o += NP + NP / S + 4;   // U_k, skewed natural order, one halo entry on each side
        o = (o + 1) & ~1;
        m.bn = o; o += NP;               // M U^n + At F
        m.hfs = o; o += NP + 2;          // (h/6)(f_gp1 + f_gp2) per element, index e + 1
        m.atf = o; o += NP;              // At F
        m.q = o; o += 2 * NC + 4 * NC;   // q (padded to 2 NC) and the four warps' q0 partials
        m.misc = o; o += 4;              // sample id / loop flags (as ints)
        m.per_group = (o + 1) & ~1;
        return m;
    }
};

// permuted row of natural node i:  i = (4w + t) S + s  ->  r = (w S + s) 4 + t
__device__ __forceinline__ int perm_row(int i, int S, int lgS) {
    const int s = i & (S - 1), wt = i >> lgS;
    return (((wt >> 2) << lgS) + s) * 4 + (wt & 3);
}
// index of natural node i (may be -1 or NP) in the skewed U array
__device__ __forceinline__ int u_idx(int i, int lgS) { return i + 1 + ((i + (1 << lgS)) >> lgS); }

// n x n solve with partial pivoting, one matrix row per lane (lane < n), columns in registers.
// a[j] (j < n): row `lane` of Ar; rhs: -br[lane].  Returns x[lane] (solution component `lane`).
template <int NC>
__device__ __forceinline__ double lu_solve_rows(double (&a)[NC], double rhs, int n, int lane) {
    bool avail = lane < n;   // row not yet used as a pivot row
    int mystep = -1;         // elimination step at which this lane's row became the pivot row
    int piv_of_k = 0;        // lane k remembers the pivot lane of step k
    double pinv = 0.0;       // 1 / pivot of the step this lane's row served
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        if (k < n) {
            // argmax |a[k]| over the available rows (idamax of dgetf2)
            const unsigned long long key = avail ? (unsigned long long)__double_as_longlong(fabs(a[k])) + 1ull : 0ull;
            const unsigned khi = (unsigned)(key >> 32), klo = (unsigned)key;
            const unsigned mhi = __reduce_max_sync(FULL, khi);
            const bool c1 = khi == mhi;
            const unsigned mlo = __reduce_max_sync(FULL, c1 ? klo : 0u);
            const unsigned ball = __ballot_sync(FULL, c1 && klo == mlo && avail);
            const int pl = ball ? (__ffs(ball) - 1) : 0;
            const double pv = shfl_idx(a[k], pl);
            const double inv = 1.0 / pv;
            const double l = (avail && lane != pl) ? a[k] * inv : 0.0;
#pragma unroll
            for (int j = k + 1; j < NC; ++j) {
                if (j < n) {
                    const double pj = shfl_idx(a[j], pl);
                    a[j] = fma(-l, pj, a[j]);
                }
            }
            const double pr = shfl_idx(rhs, pl);
            rhs = fma(-l, pr, rhs);
            if (lane == pl) { avail = false; mystep = k; pinv = inv; }
            if (lane == k) piv_of_k = pl;
        }
    }
    // back substitution, column oriented: the pivot row of step k yields x[k]
    double x = 0.0;
#pragma unroll
    for (int k = NC - 1; k >= 0; --k) {
        if (k < n) {
            const int pl = __shfl_sync(FULL, piv_of_k, k);
            const double xk = shfl_idx(rhs * pinv, pl);
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

template <int NB, int G, bool LSPG>
__global__ void __launch_bounds__(G * PGT, 1) prom_kernel(const PromParams p) {
    constexpr int NC = NB * 8;
    using Tl = PromTiles<NB, LSPG>;
    constexpr int TILES = Tl::COUNT;
    extern __shared__ __align__(16) double smem[];

    const int N = p.N, n = p.n;
    const PromGeom geo = PromGeom::make(N);
    const int S = geo.S, lgS = geo.lgS, NP = geo.NP;
    const PromSmem lay = PromSmem::make(NP, S, n, NC, TILES, p.phi_smem);
    double *s_phi = smem + lay.phi;
    double *s_h6 = smem + lay.h6;
    double *s_ke = smem + lay.ke;

    const int grp = threadIdx.x / PGT;
    const int gt = threadIdx.x - grp * PGT;
    const int w = gt >> 5, lane = gt & 31;
    const int t = lane & 3, g = lane >> 2;
    double *gbase = smem + lay.per_cta + grp * lay.per_group;
    double *g_rows = gbase + lay.rows;
    double *g_u = gbase + lay.u;
    double *g_bn = gbase + lay.bn;
    double *g_hfs = gbase + lay.hfs;
    double *g_atf = gbase + lay.atf;
    double *g_q = gbase + lay.q;
    volatile int *g_misc = reinterpret_cast<volatile int *>(gbase + lay.misc);

    const double At = p.At;

    // ---- CTA prologue: permuted Phi and mesh constants ------------------------------------
    if (p.phi_smem) {
        for (int idx = threadIdx.x; idx < NP * n; idx += G * PGT) {
            const int i = idx / n, j = idx - i * n;
            s_phi[perm_row(i, S, lgS) * n + j] = (i < N) ? p.Phi[idx] : 0.0;
        }
    }
    for (int idx = threadIdx.x; idx < NP + 2; idx += G * PGT) {
        const int e = idx - 1;
        double h6 = 0.0, ke = 0.0;
        if (e >= 0 && e <= N - 2) {
            const double h = p.X[e + 1] - p.X[e];
            h6 = h / 6.0;
            ke = At * p.E / h;
        }
        s_h6[idx] = h6; s_ke[idx] = ke;
    }
    __syncthreads();

    const double a6 = At / 6.0;
    const double c1 = 1.5 * At;
    // phase-B/D constants of this thread: first node of its run, its Phi rows, its columns
    const int run0 = ((w << 2) + t) << lgS;                 // natural index of the run's first node
    // Phi row of natural node i: permuted shared-memory copy, or the caller's array (rows >= N are
    // never dereferenced there: see rowok below)
    const bool phis = p.phi_smem != 0;
    const double *phi_run = phis ? s_phi + (size_t)(((w << lgS) << 2) + t) * n   // permuted row of (w, s = 0, t)
                                 : p.Phi + (size_t)run0 * n;
    const int row_stride = phis ? 4 * n : n;                // s -> s + 1
    const int nreal = phis ? NP : N;                        // rows that may be dereferenced
    const int jR = n >> 3, gR = n & 7;                      // block / lane-group of the R column
    bool colok[NB];
#pragma unroll
    for (int b = 0; b < NB; ++b) colok[b] = (b * 8 + g) < n;

    for (;;) {
        if (gt == 0) g_misc[0] = (int)atomicAdd(p.counter, 1u);
        group_sync(grp);
        const int smp = g_misc[0];
        if (smp >= p.B) break;
        const double mu1 = p.mu1[smp];
        const double mu2 = p.mu2[smp];

        // ---- per-sample forcing constants (:449-455, :565) and the initial condition ----
        for (int idx = gt; idx < NP + 2; idx += PGT) {
            const int e = idx - 1;
            double hfs = 0.0;
            if (e >= 0 && e <= N - 2) {
                const double x1 = p.X[e], x2 = p.X[e + 1];
                const double f1 = 0.02 * exp(mu2 * (BX_GA * x1 + BX_GB * x2));
                const double f2 = 0.02 * exp(mu2 * (BX_GB * x1 + BX_GA * x2));
                hfs = ((x2 - x1) / 6.0) * (f1 + f2);
            }
            g_hfs[idx] = hfs;
        }
        for (int i = gt; i < NP; i += PGT) {
            double F = 0.0;
            if (i < N) {
                if (i >= 1) {   // local node 2 of element i-1
                    const double x1 = p.X[i - 1], x2 = p.X[i];
                    const double f1 = 0.02 * exp(mu2 * (BX_GA * x1 + BX_GB * x2));
                    const double f2 = 0.02 * exp(mu2 * (BX_GB * x1 + BX_GA * x2));
                    F += (f1 * BX_GB + f2 * BX_GA) * (0.5 * (x2 - x1));
                }
                if (i <= N - 2) {   // local node 1 of element i
                    const double x1 = p.X[i], x2 = p.X[i + 1];
                    const double f1 = 0.02 * exp(mu2 * (BX_GA * x1 + BX_GB * x2));
                    const double f2 = 0.02 * exp(mu2 * (BX_GB * x1 + BX_GA * x2));
                    F += (f1 * BX_GA + f2 * BX_GB) * (0.5 * (x2 - x1));
                }
            }
            g_atf[i] = At * F;
        }
        double *Us = p.U + (long long)smp * (long long)(p.nsteps + 1) * N;
        {
            const double *u0p = p.u0 + (long long)smp * p.u0_stride;
            for (int i = gt - 1; i <= NP; i += PGT) {   // i = -1 .. NP (halo entries are zero)
                const double v = (i >= 0 && i < N) ? u0p[i] : 0.0;
                g_u[u_idx(i, lgS)] = v;
                if (i >= 0 && i < N) Us[i] = v;
            }
            if (gt == 0) g_u[u_idx(NP, lgS)] = 0.0;
        }
        group_sync(grp);

        for (int step = 0; step < p.nsteps; ++step) {
            // ---- bn = M U^n + At F (:745), constant during the nonlinear loop ------------
            for (int i = gt; i < NP; i += PGT) {
                const double h6L = s_h6[i], h6R = s_h6[i + 1];
                const double um = g_u[u_idx(i - 1, lgS)], uc = g_u[u_idx(i, lgS)], up = g_u[u_idx(i + 1, lgS)];
                g_bn[i] = h6L * um + (2.0 * (h6L + h6R)) * uc + h6R * up + g_atf[i];
            }
            int k = 0;
            for (;;) {
                // ================= phase A: assembly (:730-752) =========================
                for (int i = gt; i < NP; i += PGT) {
                    const double um = g_u[u_idx(i - 1, lgS)], uc = g_u[u_idx(i, lgS)], up = g_u[u_idx(i + 1, lgS)];
                    // right element (i, i+1); the left one comes from the lane below
                    double pe = 0.0, qe = 0.0, Ge = 0.0;
                    if (i <= N - 2) {
                        const double sm = uc + up;
                        const double tt = a6 * sm;
                        pe = fma(a6, uc, tt);
                        qe = fma(a6, up, tt);
                        const double z = fma(sm * (up - uc), 1.0 / 6.0, -g_hfs[i + 1]);
                        Ge = (c1 * rcp(fmax(fabs(sm), 2.0e-10))) * z;
                    }
                    double qL = shfl_up(qe, 1), GL = shfl_up(Ge, 1);
                    if (lane == 0) {
                        qL = 0.0; GL = 0.0;
                        if (i >= 1 && i <= N - 1) {
                            const double sm = um + uc;
                            qL = fma(a6, uc, a6 * sm);
                            const double z = fma(sm * (uc - um), 1.0 / 6.0, -g_hfs[i]);
                            GL = (c1 * rcp(fmax(fabs(sm), 2.0e-10))) * z;
                        }
                    }
                    const double h6L = s_h6[i], h6R = s_h6[i + 1], keL = s_ke[i], keR = s_ke[i + 1];
                    double a = (h6L - keL) - qL;
                    double c = (h6R - keR) + pe;
                    double b = (2.0 * (h6L + h6R) + (keL + keR)) + (qL - pe);
                    double R = fma(a, um, fma(b, uc, c * up)) - (g_bn[i] - GL + Ge);
                    if (i == 0) { a = 0.0; b = 1.0; c = 0.0; R = uc - mu1; }   // Dirichlet row (:741-742, :748)
                    if (i >= N) { a = 0.0; b = 0.0; c = 0.0; R = 0.0; }
                    double2 *dst = reinterpret_cast<double2 *>(g_rows + 4 * perm_row(i, S, lgS));
                    dst[0] = make_double2(a, b);
                    dst[1] = make_double2(c, R);
                }
                group_sync(grp);

                // ================= phase B: contraction on the FP64 tensor pipe =========
                double acc[TILES][2];
#pragma unroll
                for (int i = 0; i < TILES; ++i) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
                double pq[NB];
#pragma unroll
                for (int b = 0; b < NB; ++b) pq[b] = 0.0;
                {
                    double pm[NB], p0[NB], pp[NB];
                    {
                        const int im = run0 > 0 ? run0 - 1 : 0;
                        const double *rm = phis ? s_phi + (size_t)perm_row(im, S, lgS) * n : p.Phi + (size_t)im * n;
#pragma unroll
                        for (int b = 0; b < NB; ++b) {
                            pm[b] = (colok[b] && run0 > 0 && im < nreal) ? rm[b * 8 + g] : 0.0;
                            p0[b] = (colok[b] && run0 < nreal) ? phi_run[b * 8 + g] : 0.0;
                        }
                    }
                    const double2 *rows = reinterpret_cast<const double2 *>(g_rows) + 2 * ((((w << lgS)) << 2) + t);
                    const double *up = g_u + u_idx(run0, lgS);
#pragma unroll 2
                    for (int s = 0; s < S; ++s) {
                        // Phi row of node run0 + s + 1 (next run's first row at the end of this one)
                        const int inx = run0 + s + 1;
                        const double *rp = (s + 1 < S) ? phi_run + (size_t)(s + 1) * row_stride
                                           : phis  ? s_phi + (size_t)perm_row(inx < NP ? inx : 0, S, lgS) * n
                                                   : p.Phi + (size_t)(inx < N ? inx : 0) * n;
                        const bool rowok = inx < nreal;
#pragma unroll
                        for (int b = 0; b < NB; ++b) pp[b] = (colok[b] && rowok) ? rp[b * 8 + g] : 0.0;
                        const double2 ab = rows[8 * s], cr = rows[8 * s + 1];
                        const double uk = up[s];
                        double jp[NB];
#pragma unroll
                        for (int b = 0; b < NB; ++b) {
                            jp[b] = fma(ab.x, pm[b], fma(ab.y, p0[b], cr.x * pp[b]));
                            pq[b] = fma(p0[b], uk, pq[b]);
                        }
#pragma unroll
                        for (int b = 0; b < NB; ++b)
                            if (b == jR && g == gR) jp[b] = cr.y;   // R rides as column n
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
                // q0 partials: reduce over t (lanes differing in bits 0, 1); lanes t == 0 hold the warp sum
#pragma unroll
                for (int b = 0; b < NB; ++b) {
                    pq[b] += shfl_xor(pq[b], 1);
                    pq[b] += shfl_xor(pq[b], 2);
                }
                group_sync(grp);   // everyone is done reading g_rows: reuse it as scratch

                // ---- cross-warp reduction, fixed order (w0 + w2) + (w1 + w3) ----------------
                double *scr = g_rows;                       // [2][TILES][32][2]
                double *scq = g_q + 2 * NC;                 // [4][NC] q0 partials
                if (t == 0) {
#pragma unroll
                    for (int b = 0; b < NB; ++b) scq[w * NC + b * 8 + g] = pq[b];
                }
                if (w >= 2) {
                    double2 *d2 = reinterpret_cast<double2 *>(scr) + (size_t)(w - 2) * TILES * 32 + lane;
#pragma unroll
                    for (int i = 0; i < TILES; ++i) d2[i * 32] = make_double2(acc[i][0], acc[i][1]);
                }
                group_sync(grp);
                if (w < 2) {
                    const double2 *d2 = reinterpret_cast<const double2 *>(scr) + (size_t)w * TILES * 32 + lane;
#pragma unroll
                    for (int i = 0; i < TILES; ++i) { const double2 v = d2[i * 32]; acc[i][0] += v.x; acc[i][1] += v.y; }
                }
                group_sync(grp);
                if (w == 1) {
                    double2 *d2 = reinterpret_cast<double2 *>(scr) + lane;
#pragma unroll
                    for (int i = 0; i < TILES; ++i) d2[i * 32] = make_double2(acc[i][0], acc[i][1]);
                }
                group_sync(grp);

                // ================= phase C: reduced solve (warp 0) =======================
                if (w == 0) {
                    const double2 *d2 = reinterpret_cast<const double2 *>(scr) + lane;
#pragma unroll
                    for (int i = 0; i < TILES; ++i) { const double2 v = d2[i * 32]; acc[i][0] += v.x; acc[i][1] += v.y; }
                    __syncwarp();
                    // spread the accumulator tiles into a dense (NC x (NC+1)) matrix; column n = br
                    double *mat = g_rows;   // over the reduction scratch (this warp has consumed it)
                    constexpr int LD = NC + 1;
#pragma unroll
                    for (int I = 0; I < NB; ++I) {
#pragma unroll
                        for (int J = LSPG ? I : 0; J < NB; ++J) {
                            const int ti = Tl::index(I, J);
                            const int r = I * 8 + g, c = J * 8 + 2 * t;
                            mat[r * LD + c] = acc[ti][0];
                            mat[r * LD + c + 1] = acc[ti][1];
                            if (LSPG && J > I) {
                                mat[c * LD + r] = acc[ti][0];
                                mat[(c + 1) * LD + r] = acc[ti][1];
                            }
                        }
                    }
                    __syncwarp();
                    double a[NC];
#pragma unroll
                    for (int j = 0; j < NC; ++j) a[j] = (lane < n && j < n) ? mat[lane * LD + j] : 0.0;
                    const double rhs = (lane < n) ? -mat[lane * LD + n] : 0.0;
                    const double dq = lu_solve_rows<NC>(a, rhs, n, lane);
                    double q0 = 0.0;
                    if (lane < n) q0 = (scq[lane] + scq[2 * NC + lane]) + (scq[NC + lane] + scq[3 * NC + lane]);
                    const double q = (lane < n) ? q0 + dq : 0.0;
                    const double sd = warp_sum((lane < n) ? dq * dq : 0.0);
                    const double sq = warp_sum(q * q);
                    const double err = sqrt(sd) / sqrt(sq);
                    g_q[lane] = q;
                    if (lane == 0) g_misc[1] = (err > p.tol) ? 1 : 0;
                }
                group_sync(grp);

                // ================= phase D: lift U_{k+1} = Phi q (:773) ===================
                {
                    double qv[NB];
#pragma unroll
                    for (int b = 0; b < NB; ++b) qv[b] = (b * 8 + g < 32) ? g_q[(b * 8 + g) & 31] : 0.0;
                    double *up = g_u + u_idx(run0, lgS);
#pragma unroll 4
                    for (int s = 0; s < S; ++s) {
                        const double *r0 = phi_run + (size_t)s * row_stride;
                        double v = 0.0;
#pragma unroll
                        for (int b = 0; b < NB; ++b) v = fma((colok[b] && run0 + s < nreal) ? r0[b * 8 + g] : 0.0, qv[b], v);
                        v += shfl_xor(v, 4);
                        v += shfl_xor(v, 8);
                        v += shfl_xor(v, 16);
                        if (g == 0 && run0 + s < N) up[s] = v;
                    }
                }
                ++k;
                const bool more = g_misc[1] != 0 && k < p.kmax;
                group_sync(grp);
                if (!more) break;
            }
            // ---- snapshot step+1 (:783) and iteration count --------------------------------
            double *Un1 = Us + (long long)(step + 1) * N;
            for (int i = gt; i < N; i += PGT) Un1[i] = g_u[u_idx(i, lgS)];
            if (gt == 0 && p.iters) p.iters[(long long)smp * p.nsteps + step] = k;
        }
        group_sync(grp);   // g_misc[0] is rewritten by the next queue pull
    }
}

template <int NB, int G, bool LSPG>
static int launch_prom_t(PromParams p, int phi_smem, cudaStream_t st, bool *fits) {
    const PromGeom geo = PromGeom::make(p.N);
    p.phi_smem = phi_smem;
    const PromSmem lay = PromSmem::make(geo.NP, geo.S, p.n, NB * 8, PromTiles<NB, LSPG>::COUNT, phi_smem);
    const size_t smem = sizeof(double) * ((size_t)lay.per_cta + (size_t)G * lay.per_group);
    int dev = 0, sms = 0, maxsm = 0;
    BX_CUDA(cudaGetDevice(&dev));
    BX_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    BX_CUDA(cudaDeviceGetAttribute(&maxsm, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    *fits = smem <= (size_t)maxsm;
    if (!*fits) return BX_OK;
    auto kern = prom_kernel<NB, G, LSPG>;
    BX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long ctas = sms;
    const long long need = ((long long)p.B + G - 1) / G;
    if (ctas > need) ctas = need;
    BX_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st));
    kern<<<(unsigned)ctas, G * PGT, smem, st>>>(p);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

template <int NB, bool LSPG>
static int launch_prom_nb(const PromParams &p, cudaStream_t st) {
    bool fits = false;
    int rc;
    // Phi in shared memory with as many trajectory groups per CTA as fit; else Phi from global / L2
    for (int phi_smem = 1; phi_smem >= 0; --phi_smem) {
        rc = launch_prom_t<NB, 3, LSPG>(p, phi_smem, st, &fits);
        if (rc != BX_OK || fits) return rc;
        rc = launch_prom_t<NB, 2, LSPG>(p, phi_smem, st, &fits);
        if (rc != BX_OK || fits) return rc;
        rc = launch_prom_t<NB, 1, LSPG>(p, phi_smem, st, &fits);
        if (rc != BX_OK || fits) return rc;
    }
    set_error("bx_prom_trajectories: N = %d, n = %d does not fit in shared memory", p.N, p.n);
    return BX_ERR_UNSUPPORTED;
}

template <bool LSPG>
static int launch_prom(const PromParams &p, cudaStream_t st) {
    if (p.n <= 15) return launch_prom_nb<2, LSPG>(p, st);
    if (p.n <= 23) return launch_prom_nb<3, LSPG>(p, st);
    if (p.n <= 31) return launch_prom_nb<4, LSPG>(p, st);
    set_error("bx_prom_trajectories: n = %d modes > 31 is not supported by this kernel", p.n);
    return BX_ERR_UNSUPPORTED;
}

}  // namespace bx

extern "C" int bx_prom_trajectories(int B, int N, int n, int nsteps, const double *X, const double *u0,
                                    int u0_batched, const double *mu1, const double *mu2, double At,
                                    double E, const double *Phi, int projection, double tol, int kmax,
                                    double *U, int *iters, void *workspace, void *stream) {
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
    PromParams p;
    p.B = B; p.N = N; p.n = n; p.nsteps = nsteps; p.kmax = kmax;
    p.X = X; p.u0 = u0; p.u0_stride = u0_batched ? N : 0;
    p.mu1 = mu1; p.mu2 = mu2; p.At = At; p.E = E; p.tol = tol;
    p.Phi = Phi; p.U = U; p.iters = iters; p.counter = static_cast<unsigned int *>(workspace);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return projection == BX_LSPG ? launch_prom<true>(p, st) : launch_prom<false>(p, st);
}
