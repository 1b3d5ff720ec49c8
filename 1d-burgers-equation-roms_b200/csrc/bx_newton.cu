// bx_newton.cu — fused trajectory kernel for FEMBurgers.fom_burgers_newton AS CODED
// (FEM/fem_burgers.py:584-643), for B (mu1, mu2) samples at once.
//
// What the reference computes: at :616 the term `At * dC_dU @ U0` is a matrix-VECTOR
// product (operator precedence), so the "Jacobian" is J = T + 1 w^T with the tridiagonal
// T = M + At E K + At C(U_k) and the row vector w = At (dC_dU U_k) broadcast onto every row; row 0
// is then replaced by e0^T (:619-620).  Hence J = T' + e~ w^T, T' = T with the Dirichlet row,
// e~ = (0, 1, ..., 1), and J delta = -R is solved exactly by Sherman-Morrison with ONE tridiagonal
// factorisation and two right-hand sides:
//     T' y = -R,  T' z = e~,  delta = y - z (w.y) / (1 + w.z).
// R = T U_k - (M U^n + At F), R[0] = U_k[0] - mu1 (no SUPG term, :623-626);
// U_{k+1} = U_k + lambda delta, err = |U_{k+1} - U_k| / |U_{k+1}| (:632-635).
//
// Same thread mapping as bx_fom.cu: one warp per trajectory, lane l owns nodes l*M .. l*M+M-1 in
// registers, partition solve (continuant Thomas sweep per lane, 32-row PCR across lanes), here with
// two right-hand sides sharing the factorisation.  HBM traffic: the 8N-byte snapshot per step.
#include "bx_common.cuh"

namespace bx {

struct NewtonParams {
    int B, N, nsteps, kmax;
    const double *X;
    const double *u0;
    long long u0_stride;
    const double *mu1;
    const double *mu2;
    double At, E, lambda, tol2;
    double *U;
    int *iters;
    unsigned int *counter;
};

// division-free PCR step on rows with two right-hand sides (see pcr_step in bx_common.cuh)
__device__ __forceinline__ void pcr_step2(double &al, double &be, double &ga, double &r1, double &r2, int s) {
    const double al_m = shfl_up(al, s), be_m = shfl_up(be, s), ga_m = shfl_up(ga, s);
    const double r1_m = shfl_up(r1, s), r2_m = shfl_up(r2, s);
    const double al_p = shfl_down(al, s), be_p = shfl_down(be, s), ga_p = shfl_down(ga, s);
    const double r1_p = shfl_down(r1, s), r2_p = shfl_down(r2, s);
    const double bb = be_m * be_p, ab = al * be_p, gb = ga * be_m;
    be = fma(be, bb, -fma(ab, ga_m, gb * al_p));
    r1 = fma(r1, bb, -fma(ab, r1_m, gb * r1_p));
    r2 = fma(r2, bb, -fma(ab, r2_m, gb * r2_p));
    al = -ab * al_m;
    ga = -gb * ga_p;
}

__device__ __forceinline__ void warp_pcr2(double al, double be, double ga, double &r1, double &r2) {
    pcr_step2(al, be, ga, r1, r2, 1);
    pcr_step2(al, be, ga, r1, r2, 2);
    {
        const double inv = rcp(be);
        al *= inv; ga *= inv; r1 *= inv; r2 *= inv; be = 1.0;
    }
    pcr_step2(al, be, ga, r1, r2, 4);
    pcr_step2(al, be, ga, r1, r2, 8);
    pcr_step2(al, be, ga, r1, r2, 16);
    const double inv = rcp(be);
    r1 *= inv; r2 *= inv;
}

// the four swept-row arrays (cs, ys, zs, fs) of a lane: registers up to 16 rows per lane; at 32 rows per lane (N > 512)
// they would need 256 registers beside u[32] (round 1: 1640 bytes of stack, 3.5 KB of spill traffic per iteration), so
// they live in the warp's shared memory instead ([array][row][lane]: conflict-free)
template <int M> struct NewtonShape { static constexpr bool ROWS_IN_SMEM = M > 16; };

template <int M> struct NewtonSmem {
    static constexpr int EL = (M + 1) * 32;
    static constexpr int NL = M * 32;
    static constexpr int MESH = 2 * EL + NL;   // off, h6 | dg
    static constexpr int PER_WARP = 2 * NL + (NewtonShape<M>::ROWS_IN_SMEM ? 4 * NL : 0);   // AtF, bn | cs, ys, zs, fs
    static constexpr size_t bytes(int warps) { return sizeof(double) * (MESH + (size_t)warps * PER_WARP); }
};

template <int M, bool SMEM> struct RowStore {
    double r[SMEM ? 1 : M];
    double *s;   // SMEM: this lane's column of the warp's [row][lane] array
    __device__ __forceinline__ void set(int j, double v) { if (SMEM) s[j * 32] = v; else r[SMEM ? 0 : j] = v; }
    // volatile: keeps the 4 x 32 loads of an unrolled sweep where they are used instead of hoisted into registers
    __device__ __forceinline__ double get(int j) const { return SMEM ? *reinterpret_cast<volatile double *>(s + j * 32) : r[SMEM ? 0 : j]; }
};

template <int M, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) fom_newton_kernel(const NewtonParams p) {
    using L = NewtonSmem<M>;
    extern __shared__ __align__(16) double smem[];
    double *s_off = smem;
    double *s_h6 = s_off + L::EL;
    double *s_dg = s_h6 + L::EL;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    double *w_atf = s_dg + L::NL + warp * L::PER_WARP;
    double *w_bn = w_atf + L::NL;

    const int N = p.N;
    const double At = p.At;

    for (int idx = threadIdx.x; idx < L::EL; idx += WARPS * 32) {
        const int l = idx & 31, jj = (idx >> 5) - 1;
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
        const int l = idx & 31, j = idx >> 5;
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
    const int jl = N - 1 - node0;   // local index of the last real row
    const double a6 = At / 6.0;
    const double hAt = 0.5 * At;
    const bool leftv = lane > 0 && jl >= 0;
    const double a6L = leftv ? a6 : 0.0;

    for (;;) {
        int s = 0;
        if (lane == 0) s = (int)atomicAdd(p.counter, 1u);
        s = __shfl_sync(FULL, s, 0);
        if (s >= p.B) break;
        const double mu1 = p.mu1[s];
        const double mu2 = p.mu2[s];

        // At * F(mu2) per node (:427-461)
        {
            double left = 0.0;
#pragma unroll
            for (int jj = -1; jj < M; ++jj) {
                const int e = node0 + jj;
                double c0 = 0.0, cr = 0.0;
                if (e >= 0 && e <= N - 2) {
                    const double x1 = p.X[e], x2 = p.X[e + 1];
                    const double f1 = 0.02 * exp(mu2 * (BX_GA * x1 + BX_GB * x2));
                    const double f2 = 0.02 * exp(mu2 * (BX_GB * x1 + BX_GA * x2));
                    const double dV = 0.5 * (x2 - x1);
                    c0 = (f1 * BX_GA + f2 * BX_GB) * dV;
                    cr = (f1 * BX_GB + f2 * BX_GA) * dV;
                }
                if (jj >= 0) w_atf[jj * 32 + lane] = At * (left + c0);
                left = cr;
            }
        }

        double u[M];
        const double *u0p = p.u0 + (long long)s * p.u0_stride + node0;
        double *Us = p.U + (long long)s * (long long)(p.nsteps + 1) * N + node0;
#pragma unroll
        for (int j = 0; j < M; ++j) {
            u[j] = (node0 + j < N) ? u0p[j] : 0.0;
            if (node0 + j < N) Us[j] = u[j];
        }
        __syncwarp();

        for (int n = 0; n < p.nsteps; ++n) {
            {   // bn = M U^n + At F (:623)
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
                const double uL = shfl_up(u[M - 1], 1), uR = shfl_down(u[0], 1);
                constexpr bool RS = NewtonShape<M>::ROWS_IN_SMEM;
                RowStore<M, RS> cs, ys, zs, fs;
                cs.s = w_bn + L::NL + lane; ys.s = cs.s + L::NL; zs.s = ys.s + L::NL; fs.s = zs.s + L::NL;
                // continuant recurrences (see Continuant in bx_common.cuh), two right-hand sides
                double thm1 = 1.0, thm2 = 1.0, r1m = 0.0, r2m = 0.0, phm = 0.0, cm1 = 0.0;
                double offL = s_off[lane];
                double qL = fma(a6L, u[0], a6L * (uL + u[0]));
#pragma unroll
                for (int j = 0; j < M; ++j) {
                    const int ei = (j + 1) * 32 + lane;
                    const double um = (j == 0) ? uL : u[j > 0 ? j - 1 : 0];
                    const double up = (j == M - 1) ? uR : u[j < M - 1 ? j + 1 : j];
                    const double u1 = u[j], u2 = up;
                    const double t = a6 * (u1 + u2);
                    double pe = fma(a6, u1, t);
                    double qe = fma(a6, u2, t);
                    if (j == jl) { pe = 0.0; qe = 0.0; }
                    const double off = s_off[ei];
                    double a = offL - qL;
                    double c = off + pe;
                    double b = s_dg[j * 32 + lane] + (qL - pe);
                    double r = w_bn[j * 32 + lane] - fma(a, um, fma(b, u1, c * u2));   // -R
                    double e = (node0 + j < N) ? 1.0 : 0.0;                              // e~
                    if (j == 0 && lane == 0) { a = 0.0; b = 1.0; c = 0.0; r = mu1 - u1; e = 0.0; }
                    double th, r1, r2, ph, ct;
                    if (j == 0) {
                        th = b; r1 = r; r2 = e; ph = a; ct = c;
                    } else {
                        th = fma(b, thm1, -(a * cm1) * thm2);
                        r1 = fma(-a, r1m, r * thm1);
                        r2 = fma(-a, r2m, e * thm1);
                        ph = -a * phm;
                        ct = c * thm1;
                    }
                    const double inv = rcp(th);
                    cs.set(j, ct * inv); ys.set(j, r1 * inv); zs.set(j, r2 * inv); fs.set(j, ph * inv);
                    thm2 = thm1; thm1 = th; r1m = r1; r2m = r2; phm = ph; cm1 = c;
                    offL = off; qL = qe;
                }
                // condensation to one row per lane (rows: fs P + x_j + cs x_{j+1} = ys | zs)
                double F = 0.0, G = 0.0, Y = 0.0, Z = 0.0;
                if (M >= 2) {
                    F = fs.get(M - 2); G = cs.get(M - 2); Y = ys.get(M - 2); Z = zs.get(M - 2);
#pragma unroll
                    for (int j = M - 3; j >= 0; --j) {
                        const double cj = cs.get(j);
                        F = fma(-cj, F, fs.get(j));
                        G = -cj * G;
                        Y = fma(-cj, Y, ys.get(j));
                        Z = fma(-cj, Z, zs.get(j));
                    }
                }
                const double Fn = shfl_down(F, 1), Gn = shfl_down(G, 1), Yn = shfl_down(Y, 1), Zn = shfl_down(Z, 1);
                const double cl = (lane == 31) ? 0.0 : cs.get(M - 1);
                const double al = (lane == 0) ? 0.0 : fs.get(M - 1);
                double yl = fma(-cl, Yn, ys.get(M - 1)), zl = fma(-cl, Zn, zs.get(M - 1));
                warp_pcr2(al, fma(-cl, Fn, 1.0), -cl * Gn, yl, zl);
                double Py = shfl_up(yl, 1), Pz = shfl_up(zl, 1);
                if (lane == 0) { Py = 0.0; Pz = 0.0; }
                // back-substitution of both solutions, w.y and w.z on the fly (:616: w = At dC_dU U_k)
                double wy = 0.0, wz = 0.0;
                double y = yl, z = zl;
#pragma unroll
                for (int j = M - 1; j >= 0; --j) {
                    if (j < M - 1) {
                        const double cj = cs.get(j), fj = fs.get(j);
                        y = fma(-cj, y, fma(-fj, Py, ys.get(j)));
                        z = fma(-cj, z, fma(-fj, Pz, zs.get(j)));
                    }
                    ys.set(j, y); zs.set(j, z);
                    const int i = node0 + j;
                    const double um = (j == 0) ? uL : u[j > 0 ? j - 1 : 0];
                    const double up = (j == M - 1) ? uR : u[j < M - 1 ? j + 1 : j];
                    double w = 0.0;
                    if (i < N) w = hAt * ((i == N - 1 ? u[j] : up) - (i == 0 ? u[j] : um));
                    wy = fma(w, y, wy);
                    wz = fma(w, z, wz);
                }
                wy = warp_sum(wy);
                wz = warp_sum(wz);
                const double beta = wy / (1.0 + wz);
                sd = 0.0; su = 0.0;
#pragma unroll
                for (int j = 0; j < M; ++j) {
                    const double delta = fma(-zs.get(j), beta, ys.get(j));
                    const double un = fma(p.lambda, delta, u[j]);
                    const double d = un - u[j];
                    sd = fma(d, d, sd);
                    su = fma(un, un, su);
                    u[j] = un;
                }
                sd = warp_sum(sd);
                su = warp_sum(su);
                ++k;
            } while (sd > p.tol2 * su && k < p.kmax);

            double *Un1 = Us + (long long)(n + 1) * N;
#pragma unroll
            for (int j = 0; j < M; ++j) if (node0 + j < N) Un1[j] = u[j];
            const bool finite = su <= 1.7976931348623157e308;   // false for inf and NaN
            if (lane == 0 && p.iters) p.iters[(long long)s * p.nsteps + n] = finite ? k : -1;
            if (!finite) {   // flag -1 and NaN snapshots from here on; the batch goes on (include/burgers_b200.h)
                const double qnan = __longlong_as_double(0x7ff8000000000000ll);
                for (++n; n < p.nsteps; ++n) {
                    double *Un2 = Us + (long long)(n + 1) * N;
#pragma unroll
                    for (int j = 0; j < M; ++j) if (node0 + j < N) Un2[j] = qnan;
                    if (lane == 0 && p.iters) p.iters[(long long)s * p.nsteps + n] = -1;
                }
            }
        }
    }
}

template <int M, int WARPS>
static int launch_newton(const NewtonParams &p, cudaStream_t st) {
    using L = NewtonSmem<M>;
    const size_t smem = L::bytes(WARPS);
    auto kern = fom_newton_kernel<M, WARPS>;
    DeviceInfo di;
    int rc = device_info(di);
    if (rc != BX_OK) return rc;
    int occ = 0;
    rc = kernel_occupancy(reinterpret_cast<const void *>(kern), WARPS * 32, smem, &occ);
    if (rc != BX_OK) return rc;
    BX_REQUIRE(occ >= 1, "newton kernel does not fit on an SM (smem %zu B)", smem);
    long long ctas = (long long)di.sms * occ;
    const long long need = ((long long)p.B + WARPS - 1) / WARPS;
    if (ctas > need) ctas = need;
    BX_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st));
    kern<<<(unsigned)ctas, WARPS * 32, smem, st>>>(p);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

}  // namespace bx

extern "C" int bx_fom_newton_trajectories(int B, int N, int nsteps, const double *X, const double *u0,
                                          int u0_batched, const double *mu1, const double *mu2, double At,
                                          double E, double lambda, double tol, int kmax, double *U,
                                          int *iters, void *workspace, void *stream) {
    using namespace bx;
    BX_REQUIRE(B >= 0 && nsteps >= 0, "B and nsteps must be non-negative");
    BX_REQUIRE(N >= 2, "need at least 2 nodes (got N = %d)", N);
    BX_REQUIRE(kmax >= 1, "kmax must be >= 1 (the reference loop body defines U1)");
    if (B == 0) return BX_OK;
    BX_REQUIRE(X && u0 && mu1 && mu2 && U && workspace, "null device pointer");
    NewtonParams p;
    p.B = B; p.N = N; p.nsteps = nsteps; p.kmax = kmax;
    p.X = X; p.u0 = u0; p.u0_stride = u0_batched ? N : 0;
    p.mu1 = mu1; p.mu2 = mu2; p.At = At; p.E = E; p.lambda = lambda; p.tol2 = tol * tol;
    p.U = U; p.iters = iters; p.counter = queue_slot(workspace);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (N <= 64) return launch_newton<2, 4>(p, st);
    if (N <= 128) return launch_newton<4, 4>(p, st);
    if (N <= 256) return launch_newton<8, 4>(p, st);
    if (N <= 512) return launch_newton<16, 4>(p, st);
    if (N <= 1024) return launch_newton<32, 4>(p, st);
    set_error("bx_fom_newton_trajectories: N = %d > 1024 is not supported", N);
    return BX_ERR_UNSUPPORTED;
}
