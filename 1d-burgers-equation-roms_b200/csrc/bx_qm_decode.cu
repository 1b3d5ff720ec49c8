// bx_qm_decode.cu — decoder and tangent of the quadratic manifold for B reduced states at once:
//     u = Phi q + H sym(q x q)            (get_single_Q, FEM/fem_burgers.py:275-289, decoder at :1116-1118)
//     T = Phi + H dQ/dq                   (get_dQ_dq :292-312, tangent at :1148)
// With Hs_i the symmetric n x n matrix of row i of H (Hs_i[j][m] = H[i][idx(min, max)] (1 + delta_jm), monomial order
// i outer, j >= i inner) both are one symmetric matrix-vector product per node:  T_i = Phi_i + Hs_i q,
// u_i = (Phi_i + T_i) . q / 2.  This is the stand-alone form of the tangent pass that bx_qm.cu fuses into its
// trajectory kernel for n <= 31; it serves the stepwise quadratic-manifold path for larger bases (n <= 128), whose
// reduced system comes from bx_reduced_system with T as a per-sample tangent.
// One CTA per node: the packed row of H (n(n+1)/2 doubles) is staged in shared memory once and reused by all B
// samples; a warp takes one sample at a time, lane j (and j + 32, ...) owns mode j.
#include "bx_common.cuh"

namespace bx {

constexpr int QD_WARPS = 8;

__global__ void __launch_bounds__(QD_WARPS * 32) qm_decode_kernel(int B, int N, int n, const double *Phi, const double *H,
                                                                  const double *q, double *u, double *T) {
    extern __shared__ __align__(16) double smem[];
    const int k = n * (n + 1) / 2;
    double *s_h = smem;                        // packed row of H
    double *s_phi = s_h + k;                   // row of Phi
    double *s_q = s_phi + n + (threadIdx.x >> 5) * n;   // this warp's q, after the CTA-wide arrays
    const int i = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int c = threadIdx.x; c < k; c += blockDim.x) s_h[c] = H[(size_t)i * k + c];
    for (int c = threadIdx.x; c < n; c += blockDim.x) s_phi[c] = Phi[(size_t)i * n + c];
    __syncthreads();
    for (int s = warp; s < B; s += QD_WARPS) {
        __syncwarp();
        for (int c = lane; c < n; c += 32) s_q[c] = q[(size_t)s * n + c];
        __syncwarp();
        double usum = 0.0;
        for (int j = lane; j < n; j += 32) {
            // row j of Hs_i: entries m < j sit at idx(m, j) = m n - m (m - 1) / 2 + (j - m), entries m >= j are the
            // contiguous run starting at idx(j, j)
            double acc = 0.0;
            int idx = j;                           // idx(0, j)
            for (int m = 0; m < j; ++m) {
                acc = fma(s_h[idx], s_q[m], acc);
                idx += n - m - 1;                  // idx(m + 1, j) - idx(m, j)
            }
            // idx == idx(j, j) now
            acc = fma(2.0 * s_h[idx], s_q[j], acc);
            for (int m = j + 1; m < n; ++m) acc = fma(s_h[idx + (m - j)], s_q[m], acc);
            const double ph = s_phi[j];
            const double t = ph + acc;
            if (T) T[((size_t)s * N + i) * n + j] = t;
            usum = fma(ph + t, s_q[j], usum);
        }
        usum = warp_sum(usum);
        if (lane == 0) u[(size_t)s * N + i] = 0.5 * usum;
    }
}

}  // namespace bx

extern "C" int bx_qm_decode(int B, int N, int n, const double *Phi, const double *H, const double *q, double *u,
                            double *T, void *stream) {
    using namespace bx;
    BX_REQUIRE(B >= 0 && N >= 1, "B >= 0 and N >= 1");
    BX_REQUIRE(n >= 1 && n <= 128, "bx_qm_decode: 1 <= n <= 128 (got %d)", n);
    if (B == 0) return BX_OK;
    BX_REQUIRE(Phi && H && q && u, "null device pointer");
    const size_t smem = sizeof(double) * ((size_t)n * (n + 1) / 2 + n + (size_t)QD_WARPS * n);
    const int rc = kernel_occupancy(reinterpret_cast<const void *>(qm_decode_kernel), QD_WARPS * 32, smem, nullptr);
    if (rc != BX_OK) return rc;
    qm_decode_kernel<<<(unsigned)N, QD_WARPS * 32, smem, static_cast<cudaStream_t>(stream)>>>(B, N, n, Phi, H, q, u, T);
    BX_LAUNCH_CHECK();
    return BX_OK;
}
