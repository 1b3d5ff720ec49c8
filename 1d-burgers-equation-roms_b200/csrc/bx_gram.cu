// bx_gram.cu — snapshot Gram matrix G (+)= S S^T for the POD basis (replaces the SVD of the
// 512 x (n_traj * 501) snapshot matrix at POD/pod.py:80-84: the left singular vectors of S are the
// eigenvectors of S S^T, and S S^T is the only part that touches the snapshots).
//
// S is taken as the trajectory buffer itself: `cols` snapshots of N nodes, one snapshot per row
// ((B, nsteps+1, N) flattened), so G[i][j] = sum_r S[r][i] S[r][j].
//
//   * FP64 tensor pipe: mma.sync.m8n8k4.f64 (DMMA); tcgen05 has no FP64 kind.  The k index of the
//     MMA is the snapshot index, so the A fragment (row-major, A[i][k] = S[k][i]) and the B
//     fragment (col-major, B[k][j] = S[k][j]) are both "4 snapshots x 8 consecutive nodes" reads of
//     the same shared-memory tile: no transpose anywhere.
//   * only the upper block triangle of 128 x 128 tiles is computed (10 of 16 for N = 512).
//   * the work list (tile, 16-snapshot chunk) is split evenly over one persistent CTA per SM;
//     a CTA keeps its tile accumulators in registers across its whole range and spills one
//     partial tile per tile it touched (at most ceil(T/P)+1).  A second kernel adds the partials
//     in CTA order (fixed order => bit-reproducible G, so every rank derives the same basis after
//     the all-reduce) and mirrors the result into the lower triangle.
//   * a producer warp streams the snapshot rows with 1-D bulk copies (cp.async.bulk, TMA engine)
//     into a 4-stage ring of padded shared-memory tiles; full/empty mbarriers, no __syncthreads in
//     the main loop.  Row stride 132 doubles (= 4 mod 16) makes every fragment load conflict free.
// HBM traffic: every snapshot row is read once per tile column block it feeds (L2 absorbs the
// re-reads: concurrently running CTAs walk the same snapshot range); bound: FP64 tensor pipe.
#include "bx_common.cuh"
#include "bx_async.cuh"

namespace bx {

constexpr int GT = 128;          // tile edge (nodes)
constexpr int GKC = 16;          // snapshots per pipeline stage
constexpr int GLD = GT + 4;      // padded row stride in doubles
constexpr int GSTAGES = 4;
constexpr int GWARPS = 8;        // consumer warps, 4 (rows) x 2 (cols); warp GWARPS is the producer
constexpr int GTHREADS = (GWARPS + 1) * 32;
constexpr size_t G_TILE_DOUBLES = (size_t)GSTAGES * 2 * GKC * GLD;
constexpr size_t G_SMEM = G_TILE_DOUBLES * sizeof(double) + 2 * GSTAGES * sizeof(uint64_t);

struct GramParams {
    int N, nb, T, segmax, accumulate, bulk_ok;
    long long cols, C, items;   // C = chunks per tile, items = T * C
    const double *S;
    double *part;               // [P][segmax][GT][GT]
    double *G;                  // (N, N) row-major
};

__device__ __forceinline__ void dmma884g(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// upper-triangle tile index -> block row I <= block column J
__device__ __forceinline__ void tile_ij(int t, int nb, int &I, int &J) {
    I = 0;
    while (t >= nb - I) { t -= nb - I; ++I; }
    J = I + t;
}

__global__ void __launch_bounds__(GTHREADS, 1) gram_kernel(const GramParams p) {
    extern __shared__ __align__(128) unsigned char g_raw[];
    double *tiles = reinterpret_cast<double *>(g_raw);                       // [stage][blk][row][GLD]
    uint64_t *full = reinterpret_cast<uint64_t *>(g_raw + G_TILE_DOUBLES * sizeof(double));
    uint64_t *empty = full + GSTAGES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // pads and columns beyond N are never written by the copies: they must read as zero
    for (int i = threadIdx.x; i < (int)G_TILE_DOUBLES; i += GTHREADS) tiles[i] = 0.0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < GSTAGES; ++s) { mbar_init(&full[s], 32); mbar_init(&empty[s], GWARPS); }
        fence_mbar_init();
    }
    fence_proxy_async();
    __syncthreads();

    const long long P = gridDim.x;
    const long long i0 = p.items * blockIdx.x / P, i1 = p.items * (blockIdx.x + 1) / P;

    if (warp == GWARPS) {
        // ---------------- producer: lane -> (snapshot row of the chunk, column block) ----------
        const int row = lane & 15, blk = lane >> 4;
        for (long long it = i0; it < i1; ++it) {
            const long long nth = it - i0;
            const int s = (int)(nth % GSTAGES);
            const uint32_t ph = (uint32_t)((nth / GSTAGES) & 1);
            mbar_wait(&empty[s], ph ^ 1u);
            const int tile = (int)(it / p.C);
            const long long chunk = it - (long long)tile * p.C;
            int I, J;
            tile_ij(tile, p.nb, I, J);
            const long long r = chunk * GKC + row;
            const int c0 = (blk ? J : I) * GT;
            const int valid = min(GT, p.N - c0);
            double *dst = tiles + ((size_t)(s * 2 + blk) * GKC + row) * GLD;
            if (blk == 1 && I == J) {
                mbar_arrive(&full[s]);   // diagonal tile: both operands come from block 0
            } else if (r < p.cols) {
                const double *src = p.S + r * (long long)p.N + c0;
                if (p.bulk_ok) {
                    mbar_arrive_expect_tx(&full[s], (uint32_t)valid * 8u);
                    bulk_g2s(dst, src, (uint32_t)valid * 8u, &full[s]);
                } else {   // odd N / misaligned S: plain loads (slow path, same arithmetic)
                    for (int c = 0; c < valid; ++c) dst[c] = src[c];
                    fence_proxy_async();
                    mbar_arrive(&full[s]);
                }
            } else {       // past the last snapshot: the row contributes nothing
                for (int c = 0; c < GT; ++c) dst[c] = 0.0;
                fence_proxy_async();
                mbar_arrive(&full[s]);
            }
        }
        return;
    }

    // ---------------- consumers -----------------------------------------------------------------
    const int wi = warp >> 1, wj = warp & 1;
    const int g = lane >> 2, t = lane & 3;
    double acc[4][8][2];
    int cur = -1, seg = 0, I = 0, J = 0;

    auto flush = [&]() {
        double *dst = p.part + ((size_t)blockIdx.x * p.segmax + seg) * (size_t)(GT * GT);
#pragma unroll
        for (int m = 0; m < 4; ++m)
#pragma unroll
            for (int nn = 0; nn < 8; ++nn) {
                const int r = wi * 32 + m * 8 + g, c = wj * 64 + nn * 8 + 2 * t;
                *reinterpret_cast<double2 *>(dst + r * GT + c) = make_double2(acc[m][nn][0], acc[m][nn][1]);
            }
        ++seg;
    };

    for (long long it = i0; it < i1; ++it) {
        const long long nth = it - i0;
        const int s = (int)(nth % GSTAGES);
        const uint32_t ph = (uint32_t)((nth / GSTAGES) & 1);
        const int tile = (int)(it / p.C);
        if (tile != cur) {
            if (cur >= 0) flush();
#pragma unroll
            for (int m = 0; m < 4; ++m)
#pragma unroll
                for (int nn = 0; nn < 8; ++nn) { acc[m][nn][0] = 0.0; acc[m][nn][1] = 0.0; }
            cur = tile;
            tile_ij(tile, p.nb, I, J);
        }
        mbar_wait(&full[s], ph);
        const double *sA = tiles + (size_t)(s * 2) * GKC * GLD + wi * 32 + g;
        const double *sB = tiles + (size_t)(s * 2 + (I == J ? 0 : 1)) * GKC * GLD + wj * 64 + g;
#pragma unroll
        for (int k0 = 0; k0 < GKC; k0 += 4) {
            double a[4], b[8];
#pragma unroll
            for (int m = 0; m < 4; ++m) a[m] = sA[(k0 + t) * GLD + m * 8];
#pragma unroll
            for (int nn = 0; nn < 8; ++nn) b[nn] = sB[(k0 + t) * GLD + nn * 8];
#pragma unroll
            for (int m = 0; m < 4; ++m)
#pragma unroll
                for (int nn = 0; nn < 8; ++nn) dmma884g(acc[m][nn][0], acc[m][nn][1], a[m], b[nn]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
    if (cur >= 0) flush();
}

// G (+)= sum of the partial tiles, in CTA order; mirrors into the lower triangle.
__global__ void __launch_bounds__(256) gram_reduce_kernel(const GramParams p, int P) {
    __shared__ int s_lo, s_hi;
    const int tile = blockIdx.y;
    const long long t0 = (long long)tile * p.C, t1 = t0 + p.C;
    if (threadIdx.x == 0) {
        int lo = P, hi = -1;
        for (int c = 0; c < P; ++c) {
            const long long a = p.items * c / P, b = p.items * (c + 1) / P;
            if (a < b && a < t1 && b > t0) { if (c < lo) lo = c; hi = c; }
        }
        s_lo = lo; s_hi = hi;
    }
    __syncthreads();
    int I, J;
    tile_ij(tile, p.nb, I, J);
    const int e = blockIdx.x * 256 + threadIdx.x;   // element of the tile
    const int r = e / GT, c = e % GT;
    const int row = I * GT + r, col = J * GT + c;
    if (row >= p.N || col >= p.N) return;
    if (I == J && r > c) return;                     // the (c, r) thread writes both
    double sum = 0.0;
    for (int cta = s_lo; cta <= s_hi; ++cta) {
        const long long a = p.items * cta / P;
        const int sgm = tile - (int)(a / p.C);
        sum += p.part[((size_t)cta * p.segmax + sgm) * (size_t)(GT * GT) + e];
    }
    double *g1 = p.G + (size_t)row * p.N + col;
    double *g2 = p.G + (size_t)col * p.N + row;
    if (p.accumulate) sum += *g1;
    *g1 = sum;
    if (row != col) *g2 = sum;
}

static int gram_plan(int N, long long cols, GramParams &p, int &P) {
    DeviceInfo di;
    {
        const int rcd = device_info(di);
        if (rcd != BX_OK) return rcd;
    }
    const int sms = di.sms;
    p.N = N; p.cols = cols;
    p.nb = (N + GT - 1) / GT;
    p.T = p.nb * (p.nb + 1) / 2;
    p.C = (cols + GKC - 1) / GKC;
    p.items = (long long)p.T * p.C;
    P = (int)((p.items < sms) ? p.items : sms);
    if (P < 1) P = 1;
    const long long per = (p.items + P - 1) / P;
    long long sm = (per + p.C - 1) / p.C + 1;
    if (sm > p.T) sm = p.T;
    p.segmax = (int)sm;
    return BX_OK;
}

}  // namespace bx

extern "C" long long bx_gram_workspace_bytes(int N, long long cols) {
    using namespace bx;
    if (N < 1 || cols < 1) return 0;
    GramParams p; int P = 1;
    if (gram_plan(N, cols, p, P) != BX_OK) return -1;
    return (long long)P * p.segmax * GT * GT * (long long)sizeof(double);
}

extern "C" int bx_gram_accumulate(int N, long long cols, const double *S, double *G, int accumulate,
                                  void *workspace, void *stream) {
    using namespace bx;
    BX_REQUIRE(N >= 1 && cols >= 0, "need N >= 1 nodes and cols >= 0 snapshots");
    BX_REQUIRE(G, "null device pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (cols == 0) {
        if (!accumulate) BX_CUDA(cudaMemsetAsync(G, 0, sizeof(double) * (size_t)N * N, st));
        return BX_OK;
    }
    BX_REQUIRE(S && workspace, "null device pointer");
    GramParams p; int P = 1;
    int rc = gram_plan(N, cols, p, P);
    if (rc != BX_OK) return rc;
    p.S = S; p.G = G; p.part = static_cast<double *>(workspace); p.accumulate = accumulate ? 1 : 0;
    p.bulk_ok = ((N & 1) == 0) && ((reinterpret_cast<uintptr_t>(S) & 15u) == 0);
    {
        const int rco = kernel_occupancy(reinterpret_cast<const void *>(gram_kernel), GTHREADS, G_SMEM, nullptr);
        if (rco != BX_OK) return rco;
    }
    gram_kernel<<<P, GTHREADS, G_SMEM, st>>>(p);
    BX_LAUNCH_CHECK();
    gram_reduce_kernel<<<dim3(GT * GT / 256, p.T), 256, 0, st>>>(p, P);
    BX_LAUNCH_CHECK();
    return BX_OK;
}
