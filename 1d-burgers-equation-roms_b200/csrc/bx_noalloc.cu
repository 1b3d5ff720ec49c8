// bx_noalloc.cu — bx_fom_picard_trajectories_noalloc: the batched FOM sweep for callers that cannot allocate device
// memory (code under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The sweep kernels need scratch: a 256-byte head of work-queue counters and one progress word per trajectory
// (bx_fom_workspace_bytes).  Here that scratch is a file-scope device array of fixed size, which every GPU that loads
// the library holds until the process exits: 256 + 4 * BX_NOALLOC_MAX_TRAJECTORIES bytes = 262400 bytes.
//   * In HOST code the name g_fom_scratch is the host-side stand-in of the symbol, not a device address: the address
//     the kernels get comes from cudaGetSymbolAddress, asked once per device with that device current.
//   * Every call on a device shares the array.  A call clears what has to start at zero (the head and the progress words
//     of its B trajectories) on the caller's stream before the sweep; calls that overlap on the GPU (two streams, or
//     two host threads on one device) race on the counters and must not be issued — use the workspace form there.
//   * Nothing here, and nothing in bx_fom_picard_trajectories below it, calls cudaMalloc / cudaFree or an allocating
//     library routine: the set-up calls are cudaGetDevice, cudaGetSymbolAddress, cudaFuncSetAttribute and the
//     occupancy query; the clears are a kernel launch and cudaMemsetAsync.
#include "bx_common.cuh"
#include <mutex>

namespace bx {

constexpr size_t NOALLOC_HEAD = 256;
constexpr size_t NOALLOC_BYTES = NOALLOC_HEAD + 4u * (size_t)BX_NOALLOC_MAX_TRAJECTORIES;

static __device__ __align__(256) unsigned char g_fom_scratch[NOALLOC_BYTES];

// device code names the array directly: zero the head and the first `words` progress words
__global__ void noalloc_clear_kernel(int words) {
    unsigned int *w = reinterpret_cast<unsigned int *>(g_fom_scratch);
    const int total = (int)(NOALLOC_HEAD / 4) + words;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) w[i] = 0u;
}

constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mutex;
static void *g_noalloc_ptr[NOALLOC_MAX_DEVICES];   // device address of g_fom_scratch on device d, once asked

// device address of the scratch array on the CURRENT device
static int noalloc_scratch(void **out) {
    int dev = 0;
    BX_CUDA(cudaGetDevice(&dev));
    BX_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "bx_fom_picard_trajectories_noalloc: device ordinal %d >= %d", dev,
               NOALLOC_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[dev]) {
        void *ptr = nullptr;
        BX_CUDA(cudaGetSymbolAddress(&ptr, g_fom_scratch));
        BX_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned a null address for the scratch array");
        g_noalloc_ptr[dev] = ptr;
    }
    *out = g_noalloc_ptr[dev];
    return BX_OK;
}

}  // namespace bx

extern "C" int bx_fom_picard_trajectories_noalloc(int B, int N, int nsteps, const double *X, const double *u0,
                                                  int u0_batched, const double *mu1, const double *mu2, double At,
                                                  double E, double tol, int kmax, double *U, int *iters,
                                                  double *err_hist, void *stream) {
    using namespace bx;
    BX_REQUIRE(B >= 0, "B must be non-negative");
    if (B > BX_NOALLOC_MAX_TRAJECTORIES) {
        set_error("bx_fom_picard_trajectories_noalloc: B = %d exceeds the static scratch (BX_NOALLOC_MAX_TRAJECTORIES = %d); "
                  "use bx_fom_picard_trajectories with a caller workspace",
                  B, BX_NOALLOC_MAX_TRAJECTORIES);
        return BX_ERR_UNSUPPORTED;
    }
    void *scratch = nullptr;
    const int rc = noalloc_scratch(&scratch);
    if (rc != BX_OK) return rc;
    if (B > 0) {
        // the array keeps the last call's contents: counters and progress words start at zero
        const int words = B;
        const int threads = 256;
        const int blocks = (int)((NOALLOC_HEAD / 4 + (size_t)words + threads - 1) / threads);
        noalloc_clear_kernel<<<blocks < 148 ? blocks : 148, threads, 0, static_cast<cudaStream_t>(stream)>>>(words);
        BX_LAUNCH_CHECK();
    }
    return bx_fom_picard_trajectories(B, N, nsteps, X, u0, u0_batched, mu1, mu2, At, E, tol, kmax, U, iters, err_hist,
                                      scratch, stream);
}
