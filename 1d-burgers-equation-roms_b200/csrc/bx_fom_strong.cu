// bx_fom_strong.cu — latency instance of the fused FOM trajectory kernel (FEM/fem_burgers.py:646-707):
// the one a sweep uses when it holds fewer trajectories than the GPU has warp slots for the throughput
// kernel (bx_fom.cu) — e.g. 4096 samples sharded over 8 GPUs = 512 per GPU = 3.5 per SM.  There the sweep
// time is the serial chain of one trajectory (~6000 Picard iterations), not FP64 issue, so this kernel
// shortens the chain of ONE iteration:
//   * one CTA = one trajectory = W warps (1, 2 or 4), lane l of the 32 W lanes owns M consecutive nodes;
//     CTAs are spread over all SMs by the hardware scheduler (3 or 4 trajectories per SM at B = 512);
//   * the arithmetic is the same as in bx_fom.cu (closed-form element integrals, continuant Thomas sweep per
//     lane, condensation to a 32 W-row system in the lanes' last unknowns);
//   * SPECULATIVE FRONT HALF: the convergence test of iteration k (two norms, a 7-shuffle reduction and,
//     for W > 1, an exchange through shared memory) is taken off the chain — right after the update the
//     lanes start assembling and sweeping iteration k+1, the reduction's shuffles are interleaved with that
//     work, the partial norms travel with the condensed rows of iteration k+1 behind the SAME barrier, and
//     the loop is left before the back-substitution of k+1 when k had converged (or reached kmax).  What is
//     discarded is one front half per time step (~4 % more instructions); the states and counts are those of
//     the reference loop (:663-698) exactly, because u is only ever updated by confirmed iterations;
//   * W = 2: the 64-row condensed system loses one level by a division-free cyclic-reduction step on the two
//     rows of a lane, then the 32-row shuffle PCR, both warps redundantly (no second exchange); the PCR is
//     renormalised by an exact power of two instead of a reciprocal;
//   * two named barriers per iteration for W > 1 (condensed rows + norms; halo), none for W = 1.
#include "bx_fom.cuh"
#include <cmath>

namespace bx {

// 2^-exponent(x): scaling by it is exact (x normal; the condensed pivots are O(1))
__device__ __forceinline__ double pow2_inverse_scale(double x) {
    const int e = (__double2hiint(x) >> 20) & 0x7ff;
    return __hiloint2double((2046 - e) << 20, 0);
}

// 32-row PCR as warp_pcr (bx_common.cuh), renormalised half-way by an exact power of two
__device__ __forceinline__ double warp_pcr_pow2(double al, double be, double ga, double rho) {
    pcr_step(al, be, ga, rho, 1);
    pcr_step(al, be, ga, rho, 2);
    {
        const double sc = pow2_inverse_scale(be);
        al *= sc; ga *= sc; rho *= sc; be *= sc;
    }
    pcr_step(al, be, ga, rho, 4);
    pcr_step(al, be, ga, rho, 8);
    pcr_step(al, be, ga, rho, 16);
    return rho * rcp(be);
}

// condensed row of group lane q as the lanes publish it: F, G, R of the FIRST row of lane q+1 and the
// couplings / right-hand side of the LAST row of lane q  (48 bytes, 16-byte aligned halves)
struct __align__(16) CondRow { double F, G, R, fl, cl, rl; };

template <int M, int W> struct StrongSmem {
    static constexpr int LN = 32 * W;
    static constexpr int EL = (M + 1) * LN;
    static constexpr int NL = M * LN;
    static constexpr int ROWS = W > 1 ? 6 * LN : 0;
    static constexpr int SOL = W > 2 ? W * LN : 0;
    static constexpr int MISC = W > 1 ? 4 * W + 2 : 0;   // halo (2 W) | norms (2 W) | sample
    static constexpr int TOTAL = 3 * EL + 3 * NL + ROWS + SOL + MISC;
    static constexpr size_t bytes() { return sizeof(double) * TOTAL; }
};

// resident CTAs per SM the register allocation is held to: 128 registers per thread up to 8 rows per lane,
// 255 above (the sweeps this kernel is for hold at most a handful of trajectories per SM)
template <int M, int W> struct StrongShape { static constexpr int MINB = (M > 8 ? 8 : 16) / W; };

template <int M, int W, bool EXACT>
__global__ void __launch_bounds__(32 * W, StrongShape<M, W>::MINB) fom_strong_kernel(const FomParams p) {
    using L = StrongSmem<M, W>;
    constexpr int LN = L::LN;
    extern __shared__ __align__(16) double smem[];
    double *s_off = smem;             // h/6 - At E / h   (off-diagonal constant of A; 0 where no element)
    double *s_h6 = s_off + L::EL;     // h/6
    double *s_dg = s_h6 + L::EL;      // diagonal constant of A (1 for padded rows)
    double *w_hfs = s_dg + L::NL;     // (h/6) (f_gp1 + f_gp2) per element
    double *w_atf = w_hfs + L::EL;    // At * F per node
    double *w_bn = w_atf + L::NL;     // M U^n + At F per node (per time step)
    CondRow *x_row = reinterpret_cast<CondRow *>(w_bn + L::NL);
    double *x_sol = w_bn + L::NL + L::ROWS;
    double *x_halo = x_sol + L::SOL;          // [2 w] = first node of warp w, [2 w + 1] = its last node
    double *x_nrm = x_halo + 2 * W;           // [2 w] = |delta|^2 of warp w, [2 w + 1] = |U|^2
    int *x_smp = reinterpret_cast<int *>(x_nrm + 2 * W);
    const int lane = threadIdx.x & 31;
    const int wq = threadIdx.x >> 5;
    const int gl = threadIdx.x;       // lane in the group
    const int N = p.N;
    const double At = p.At;

    // ---- mesh constants, once per CTA -------------------------------------------------------------
    for (int idx = threadIdx.x; idx < L::EL; idx += LN) {
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
    for (int idx = threadIdx.x; idx < L::NL; idx += LN) {
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
    if (W > 1 && gl == LN - 1) { x_row[gl].F = 0.0; x_row[gl].G = 0.0; x_row[gl].R = 0.0; }   // nothing right of the last lane
    __syncthreads();

    const int node0 = gl * M;
    const bool vec_ok = (node0 + M <= N) && ((N & 1) == 0) && ((M & 1) == 0);
    const int jl = N - 1 - node0;   // local index of the last real row (see bx_fom.cu)
    const double a6 = At / 6.0;
    const double c1 = 1.5 * At;   // (folding c1 into the stored forcing constant, as bx_fom.cu does, is slower here: 10.4 against 9.9 ms at 512 trajectories)
    const bool leftv = gl > 0 && jl >= 0;
    const double a6L = leftv ? a6 : 0.0;

    for (;;) {
        int s = 0;
        if (W == 1) {
            if (lane == 0) s = (int)atomicAdd(p.counter, 1u);
            s = __shfl_sync(FULL, s, 0);
        } else {
            if (gl == 0) *x_smp = (int)atomicAdd(p.counter, 1u);
            __syncthreads();
            s = *x_smp;
        }
        if (s >= p.B) break;
        const double mu1 = p.mu1[s];
        const double mu2 = p.mu2[s];

        // ---- per-sample forcing constants (FEM/fem_burgers.py:449-455, 565) -------------------------
        {
            double left = 0.0;
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
                    c0 = (f1 * BX_GA + f2 * BX_GB) * dV;
                    cr = (f1 * BX_GB + f2 * BX_GA) * dV;
                }
                w_hfs[(jj + 1) * LN + gl] = hfs;
                if (jj >= 0) w_atf[jj * LN + gl] = At * (left + c0);
                left = cr;
            }
        }

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
        double uL, uR;   // the neighbouring lanes' boundary nodes of the current iterate
        if (W > 1) {
            if (lane == 0) x_halo[2 * wq] = u[0];
            if (lane == 31) x_halo[2 * wq + 1] = u[M - 1];
            __syncthreads();
        }
        auto halo = [&]() {
            uL = shfl_up(u[M - 1], 1); uR = shfl_down(u[0], 1);
            if (W > 1) {
                if (lane == 0 && wq > 0) uL = x_halo[2 * (wq - 1) + 1];
                if (lane == 31 && wq < W - 1) uR = x_halo[2 * (wq + 1)];
            }
        };
        halo();

        double cs[M], rs[M], fs[M];
        // front half of a Picard iteration: assembly of A, b, r = b - A U_k from the current iterate and the
        // lane-local continuant Thomas sweep (:667-692)
        auto front = [&]() {
            Continuant<M, true> ct;
            double offL, qL, GL;
            {
                const double u1 = uL, u2 = u[0];
                const double sm = u1 + u2;
                qL = fma(a6L, u2, a6L * sm);
                const double z = fma(sm * (u2 - u1), 1.0 / 6.0, -w_hfs[gl]);
                GL = ((leftv ? c1 : 0.0) * rcp1(fmax(fabs(sm), 2.0e-10))) * z;
                offL = s_off[gl];
            }
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const int ei = (j + 1) * LN + gl;
                const double um = (j == 0) ? uL : u[j > 0 ? j - 1 : 0];
                const double up = (j == M - 1) ? uR : u[j < M - 1 ? j + 1 : j];
                const double u1 = u[j], u2 = up;
                const double sm = u1 + u2;
                const double t = a6 * sm;
                double pe = fma(a6, u1, t);
                double qe = fma(a6, u2, t);
                const double z = fma(sm * (u2 - u1), 1.0 / 6.0, -w_hfs[ei]);
                double Ge = (c1 * rcp1(fmax(fabs(sm), 2.0e-10))) * z;
                if ((!EXACT || j == M - 1) && j == jl) { pe = 0.0; qe = 0.0; Ge = 0.0; }
                const double off = s_off[ei];
                double a = offL - qL;
                double c = off + pe;
                double b = s_dg[j * LN + gl] + (qL - pe);
                const double rhs = w_bn[j * LN + gl] - GL + Ge;
                double r = rhs - fma(a, um, fma(b, u1, c * u2));   // (A U_k off the chain of the SUPG reciprocals)
                if (j == 0 && gl == 0) { a = 0.0; b = 1.0; c = 0.0; r = mu1 - u1; }   // Dirichlet row (:679-680, :686)
                ct.row(j, a, b, c, r, cs[j], rs[j], fs[j]);
                offL = off; qL = qe; GL = Ge;
            }
        };

        int n = 0;
        bool finite = true;
        for (; finite && n < p.nsteps; ++n) {
            // ---- bn = M U^n + At F (constant during the Picard loop, :683) ---------------------------
            {
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
            front();
            int k = 0;
            double sd = 0.0, su = 0.0;   // warp-level sums of the last confirmed iteration
            for (;;) {
                double xl, P;
                if constexpr (W == 1) {
                    {
                        double F, G, R;
                        partition_upward<M>(cs, rs, fs, F, G, R);
                        const double Fn = shfl_down(F, 1), Gn = shfl_down(G, 1), Rn = shfl_down(R, 1);
                        const double cl = (lane == 31) ? 0.0 : cs[M - 1];
                        const double al = (lane == 0) ? 0.0 : fs[M - 1];
                        xl = warp_pcr_pow2(al, fma(-cl, Fn, 1.0), -cl * Gn, fma(-cl, Rn, rs[M - 1]));
                        P = shfl_up(xl, 1);
                        if (lane == 0) P = 0.0;
                    }
                } else {
                    double F, G, R;
                    partition_upward<M>(cs, rs, fs, F, G, R);
                    if (gl > 0) {
                        *reinterpret_cast<double2 *>(&x_row[gl - 1].F) = make_double2(F, G);
                        x_row[gl - 1].R = R;
                    }
                    x_row[gl].fl = (gl == 0) ? 0.0 : fs[M - 1];
                    *reinterpret_cast<double2 *>(&x_row[gl].cl) = make_double2((gl == LN - 1) ? 0.0 : cs[M - 1], rs[M - 1]);
                    if (k > 0 && lane == 0) *reinterpret_cast<double2 *>(&x_nrm[2 * wq]) = make_double2(sd, su);
                    __syncthreads();
                    if (k > 0) {
                        double td = x_nrm[0], tu = x_nrm[1];
#pragma unroll
                        for (int w = 1; w < W; ++w) { td += x_nrm[2 * w]; tu += x_nrm[2 * w + 1]; }
                        sd = td; su = tu;   // every warp: the same sums in the same order
                        if (!(td > p.tol2 * tu) || k >= p.kmax) break;
                    }
                    if constexpr (W == 2) {
                        // rows 2 lane (e) and 2 lane + 1 (o) of the 64-row system; the even row of the next lane (n)
                        const CondRow re = x_row[2 * lane], ro = x_row[2 * lane + 1];
                        const double al_e = re.fl, be_e = fma(-re.cl, re.F, 1.0), ga_e = -re.cl * re.G, rho_e = fma(-re.cl, re.R, re.rl);
                        const double al_o = ro.fl, be_o = fma(-ro.cl, ro.F, 1.0), ga_o = -ro.cl * ro.G, rho_o = fma(-ro.cl, ro.R, ro.rl);
                        const double al_n = shfl_down(al_e, 1), be_n = shfl_down(be_e, 1), ga_n = shfl_down(ga_e, 1),
                                     rho_n = shfl_down(rho_e, 1);
                        const double inv_e = rcp(be_e);      // off the chain
                        // division-free cyclic-reduction step: odd rows in the odd unknowns only
                        const double bb = be_e * be_n, ab = al_o * be_n, gb = ga_o * be_e;
                        const double be2 = fma(be_o, bb, -fma(ab, ga_e, gb * al_n));
                        const double rho2 = fma(rho_o, bb, -fma(ab, rho_e, gb * rho_n));
                        const double x_o = warp_pcr_pow2(-ab * al_e, be2, -gb * ga_n, rho2);
                        const double x_p = shfl_up(x_o, 1);   // lane 0: al_e = 0
                        const double x_e = (rho_e - fma(al_e, x_p, ga_e * x_o)) * inv_e;
                        // unknown q lives in lane q / 2: even -> x_e, odd -> x_o
                        const int src = wq * 16 + (lane >> 1);
                        const double se = shfl_idx(x_e, src), so = shfl_idx(x_o, src);
                        const double sp = shfl_idx(x_o, src > 0 ? src - 1 : 0);
                        xl = (lane & 1) ? so : se;
                        P = (lane & 1) ? se : sp;
                        if (gl == 0) P = 0.0;
                    } else {
                        double cs2[W], rs2[W], fs2[W];
                        Continuant<W> c2;
#pragma unroll
                        for (int j = 0; j < W; ++j) {
                            const CondRow rq = x_row[W * lane + j];
                            c2.row(j, rq.fl, fma(-rq.cl, rq.F, 1.0), -rq.cl * rq.G, fma(-rq.cl, rq.R, rq.rl),
                                   cs2[j], rs2[j], fs2[j]);
                        }
                        double F2, G2, R2;
                        partition_upward<W>(cs2, rs2, fs2, F2, G2, R2);
                        const double Fn = shfl_down(F2, 1), Gn = shfl_down(G2, 1), Rn = shfl_down(R2, 1);
                        const double cl = (lane == 31) ? 0.0 : cs2[W - 1];
                        const double al = (lane == 0) ? 0.0 : fs2[W - 1];
                        double x2 = warp_pcr_pow2(al, fma(-cl, Fn, 1.0), -cl * Gn, fma(-cl, Rn, rs2[W - 1]));
                        double P2 = shfl_up(x2, 1);
                        if (lane == 0) P2 = 0.0;
                        double *sol = x_sol + wq * LN;
                        sol[W * lane + W - 1] = x2;
#pragma unroll
                        for (int j = W - 2; j >= 0; --j) {
                            x2 = fma(-cs2[j], x2, fma(-fs2[j], P2, rs2[j]));
                            sol[W * lane + j] = x2;
                        }
                        __syncwarp();
                        xl = sol[gl];
                        P = (gl > 0) ? sol[gl - 1] : 0.0;
                    }
                }
                // back-substitution fused with the update and the two norms (:695-698)
                double x = xl;
                double d0 = 0.0, d1 = 0.0, n0 = 0.0, n1 = 0.0;
#pragma unroll
                for (int j = M - 1; j >= 0; --j) {
                    if (j < M - 1) x = fma(-cs[j], x, fma(-fs[j], P, rs[j]));
                    u[j] += x;
                    if (j & 1) { d1 = fma(x, x, d1); n1 = fma(u[j], u[j], n1); }
                    else       { d0 = fma(x, x, d0); n0 = fma(u[j], u[j], n0); }
                }
                ++k;
                if (W > 1) {
                    if (lane == 0) x_halo[2 * wq] = u[0];
                    if (lane == 31) x_halo[2 * wq + 1] = u[M - 1];
                    __syncthreads();
                }
                halo();
                // the norms of iteration k ride along with the front half of iteration k+1
                sd = d0 + d1; su = n0 + n1;
                warp_sum2(sd, su, lane);
                front();
                if constexpr (W == 1) {   // (W > 1 decides behind the barrier that carries the next condensed rows)
                    if (!(sd > p.tol2 * su) || k >= p.kmax) break;
                }
            }

            // ---- snapshot n+1 (:705) and iteration count -------------------------------------------
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
        }
        if (!finite) {   // as bx_fom.cu: flag -1 and NaN snapshots from the first non-finite step on
            const double qnan = __longlong_as_double(0x7ff8000000000000ll);
            for (; n < p.nsteps; ++n) {
                double *Un1 = Us + (long long)(n + 1) * N;
#pragma unroll
                for (int j = 0; j < M; ++j) if (node0 + j < N) Un1[j] = qnan;
                if (gl == 0 && p.iters) p.iters[(long long)s * p.nsteps + n] = -1;
            }
        }
        if (W > 1) __syncthreads();   // the sample slot and the exchange area are reused by the next trajectory
    }
}

template <int M, int W, bool EXACT>
static int launch_strong_t(const FomParams &p, cudaStream_t st) {
    const size_t smem = StrongSmem<M, W>::bytes();
    auto kern = fom_strong_kernel<M, W, EXACT>;
    DeviceInfo di;
    int rc = device_info(di);
    if (rc != BX_OK) return rc;
    int occ = 0;
    rc = kernel_occupancy(reinterpret_cast<const void *>(kern), 32 * W, smem, &occ);
    if (rc != BX_OK) return rc;
    BX_REQUIRE(occ >= 1, "fom latency kernel does not fit on an SM (smem %zu B)", smem);
    long long ctas = (long long)di.sms * occ;
    if (ctas > p.B) ctas = p.B;
    BX_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st));
    kern<<<(unsigned)ctas, 32 * W, smem, st>>>(p);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

template <int M, int W>
static int launch_strong(const FomParams &p, cudaStream_t st) {
    return (p.N == 32 * W * M) ? launch_strong_t<M, W, true>(p, st) : launch_strong_t<M, W, false>(p, st);
}

int launch_fom_strong(const FomParams &p, int W, cudaStream_t st) {
    const int N = p.N;
    if (N <= 256) {
        if (W == 4) return launch_strong<2, 4>(p, st);
        if (W == 2) return launch_strong<4, 2>(p, st);
        return launch_strong<8, 1>(p, st);
    }
    if (N <= 512) {
        if (W == 4) return launch_strong<4, 4>(p, st);
        if (W == 2) return launch_strong<8, 2>(p, st);
        return launch_strong<16, 1>(p, st);
    }
    if (W == 4) return launch_strong<8, 4>(p, st);
    if (W == 2) return launch_strong<16, 2>(p, st);
    return launch_strong<32, 1>(p, st);
}

}  // namespace bx
