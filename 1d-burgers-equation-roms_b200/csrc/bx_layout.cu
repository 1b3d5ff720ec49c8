// bx_layout.cu — trajectory buffer (B, rows, N) -> the reference's per-sample layout (B, N, rows).
// FEMBurgers.fom_burgers returns U of shape (N, nTimeSteps+1), C order (FEM/fem_burgers.py:650), and the
// sweep drivers np.save exactly that array (FEM/paper_training_stage.py:52-53); POD/pod.py:80-82 hstacks
// the files.  The trajectory kernels write one snapshot per row (coalesced stores of the hot loop), so the
// files are produced by one batched tiled transpose on the device followed by a straight D2H copy.
// HBM bound: 8 rows N bytes read + the same written per sample; 32 x 32 tiles through padded shared memory,
// both sides unit-stride across the warp.
#include "bx_common.cuh"

namespace bx {

__global__ void __launch_bounds__(256) transpose_snapshots_kernel(int rows, int N, const double *__restrict__ U,
                                                                  double *__restrict__ Ur) {
    __shared__ double tile[32][33];
    const long long b = blockIdx.z;
    const double *src = U + b * (long long)rows * N;
    double *dst = Ur + b * (long long)rows * N;
    const int i0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int r = r0 + ty + 8 * k, i = i0 + tx;
        if (r < rows && i < N) tile[ty + 8 * k][tx] = src[(long long)r * N + i];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int i = i0 + ty + 8 * k, r = r0 + tx;
        if (i < N && r < rows) dst[(long long)i * rows + r] = tile[tx][ty + 8 * k];
    }
}

}  // namespace bx

extern "C" int bx_snapshots_to_reference_layout(int B, int rows, int N, const double *U, double *Uref,
                                                void *stream) {
    using namespace bx;
    BX_REQUIRE(B >= 0 && rows >= 0 && N >= 0, "B, rows and N must be non-negative");
    if (B == 0 || rows == 0 || N == 0) return BX_OK;
    BX_REQUIRE(U && Uref && U != Uref, "null or aliased device pointer");
    BX_REQUIRE(B <= 65535, "at most 65535 samples per call (got %d)", B);
    dim3 grid((unsigned)((N + 31) / 32), (unsigned)((rows + 31) / 32), (unsigned)B);
    transpose_snapshots_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(rows, N, U, Uref);
    BX_LAUNCH_CHECK();
    return BX_OK;
}
