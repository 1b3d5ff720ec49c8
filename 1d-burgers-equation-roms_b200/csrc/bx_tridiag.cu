// bx_tridiag.cu — K2: batched tridiagonal solve + update + convergence norms with the bands in
// HBM (FEM/fem_burgers.py:692-698: delta = spsolve(A, -R); U1 = U0 + delta;
// err = |delta| / |U1|).  One warp per system, N <= 1024.
//
// The warp streams its four input rows (dl, d, du, R) from HBM with unit-stride 16-byte loads
// into a padded shared-memory tile (row stride M+1 doubles, so that lane l reading its own M
// consecutive entries is bank-conflict free), then runs the same partition solve as the fused
// trajectory kernel (bx_fom.cu): normalised Thomas per lane, upward condensation, 32-row PCR
// across the lanes by shuffles, local back-substitution.  Algorithmic HBM traffic: 40 N bytes
// read + 8 N written per system.
#include "bx_common.cuh"

namespace bx {

template <int M> __device__ __forceinline__ int pad_pos(int i) { return (i / M) * (M + 1) + (i % M); }

// coalesced global -> padded smem tile; entries i >= N get `fill`
template <int M>
__device__ __forceinline__ void stage_in(double *tile, const double *__restrict__ g, int N, int lane,
                                         double fill, bool vec) {
    if (vec) {
#pragma unroll
        for (int k = 0; k < M / 2; ++k) {
            const int i = 2 * lane + 64 * k;
            double2 v = make_double2(fill, fill);
            if (i + 1 < N) v = *reinterpret_cast<const double2 *>(g + i);
            else if (i < N) v.x = g[i];
            const int pp = pad_pos<M>(i);
            tile[pp] = v.x; tile[pp + 1] = v.y;
        }
    } else {
#pragma unroll
        for (int k = 0; k < M; ++k) {
            const int i = lane + 32 * k;
            tile[pad_pos<M>(i)] = (i < N) ? g[i] : fill;
        }
    }
}

template <int M, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
tridiag_solve_update_kernel(int B, int N, const double *__restrict__ dl, const double *__restrict__ d,
                            const double *__restrict__ du, const double *__restrict__ R,
                            const double *Uk, double *Uk1, double *__restrict__ err) {
    constexpr int TILE = 32 * (M + 1);
    extern __shared__ __align__(16) double smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *t_a = smem + (size_t)warp * 4 * TILE;
    double *t_b = t_a + TILE, *t_c = t_b + TILE, *t_r = t_c + TILE;
    const int b = blockIdx.x * WARPS + warp;
    if (b >= B) return;
    const long long o = (long long)b * N;
    const bool vec = ((N & 1) == 0);  // rows start 16-byte aligned when N is even
    stage_in<M>(t_a, dl + o, N, lane, 0.0, vec);
    stage_in<M>(t_b, d + o, N, lane, 1.0, vec);
    stage_in<M>(t_c, du + o, N, lane, 0.0, vec);
    stage_in<M>(t_r, R + o, N, lane, 0.0, vec);
    __syncwarp();

    const int base = lane * (M + 1);
    double cs[M], rs[M], fs[M];
    Continuant<M> ct;
#pragma unroll
    for (int j = 0; j < M; ++j) {
        double a = t_a[base + j];
        const double bb = t_b[base + j];
        double c = t_c[base + j];
        const double r = -t_r[base + j];
        if (lane == 0 && j == 0) a = 0.0;            // dl[0] is unused by contract
        if (lane * M + j >= N - 1) c = 0.0;          // du[N-1] is unused by contract
        ct.row(j, a, bb, c, r, cs[j], rs[j], fs[j]);
    }
    __syncwarp();
    stage_in<M>(t_a, Uk + o, N, lane, 0.0, vec);  // reuse the dl tile for U_k

    double xl, P;
    partition_condense<M>(cs, rs, fs, lane, xl, P);
    __syncwarp();
    double x = xl, sd = 0.0, su = 0.0;
#pragma unroll
    for (int j = M - 1; j >= 0; --j) {
        if (j < M - 1) x = fma(-cs[j], x, fma(-fs[j], P, rs[j]));
        const double un = t_a[base + j] + x;
        sd = fma(x, x, sd);
        su = fma(un, un, su);
        t_b[base + j] = un;
    }
    sd = warp_sum(sd);
    su = warp_sum(su);
    __syncwarp();
    // padded tile -> HBM, unit stride
    if (vec) {
#pragma unroll
        for (int k = 0; k < M / 2; ++k) {
            const int i = 2 * lane + 64 * k;
            const int pp = pad_pos<M>(i);
            if (i + 1 < N) *reinterpret_cast<double2 *>(Uk1 + o + i) = make_double2(t_b[pp], t_b[pp + 1]);
            else if (i < N) Uk1[o + i] = t_b[pp];
        }
    } else {
#pragma unroll
        for (int k = 0; k < M; ++k) {
            const int i = lane + 32 * k;
            if (i < N) Uk1[o + i] = t_b[pad_pos<M>(i)];
        }
    }
    if (lane == 0 && err) err[b] = sqrt(sd) / sqrt(su);
}

template <int M>
static int launch_solve(int B, int N, const double *dl, const double *d, const double *du, const double *R,
                        const double *Uk, double *Uk1, double *err, cudaStream_t st) {
    constexpr int WARPS = 4;
    const size_t smem = sizeof(double) * WARPS * 4 * 32 * (M + 1);
    auto kern = tridiag_solve_update_kernel<M, WARPS>;
    {
        const int rco = kernel_occupancy(reinterpret_cast<const void *>(kern), WARPS * 32, smem, nullptr);
        if (rco != BX_OK) return rco;
    }
    kern<<<(B + WARPS - 1) / WARPS, WARPS * 32, smem, st>>>(B, N, dl, d, du, R, Uk, Uk1, err);
    BX_LAUNCH_CHECK();
    return BX_OK;
}

}  // namespace bx

extern "C" int bx_tridiag_solve_update(int B, int N, const double *dl, const double *d, const double *du,
                                       const double *R, const double *Uk, double *Uk1, double *err,
                                       void *stream) {
    using namespace bx;
    BX_REQUIRE(B >= 0, "B must be non-negative");
    BX_REQUIRE(N >= 2, "need at least 2 nodes (got %d)", N);
    if (B == 0) return BX_OK;
    BX_REQUIRE(dl && d && du && R && Uk && Uk1, "null device pointer");
    cudaStream_t st = (cudaStream_t)stream;
    if (N <= 64) return launch_solve<2>(B, N, dl, d, du, R, Uk, Uk1, err, st);
    if (N <= 128) return launch_solve<4>(B, N, dl, d, du, R, Uk, Uk1, err, st);
    if (N <= 256) return launch_solve<8>(B, N, dl, d, du, R, Uk, Uk1, err, st);
    if (N <= 512) return launch_solve<16>(B, N, dl, d, du, R, Uk, Uk1, err, st);
    if (N <= 1024) return launch_solve<32>(B, N, dl, d, du, R, Uk, Uk1, err, st);
    set_error("bx_tridiag_solve_update: N = %d > 1024 needs the multi-CTA cyclic-reduction path", N);
    return BX_ERR_UNSUPPORTED;
}
