// bx_fine.cu — fine-mesh FOM trajectory: FEMBurgers.fom_burgers (FEM/fem_burgers.py:646-707) for a
// mesh far larger than one warp can hold (BASELINE config 5: 2^20 nodes, one trajectory).
//
// ONE persistent cooperative kernel integrates the whole trajectory.  The mesh is cut into
// contiguous partitions three levels deep — FM nodes per thread (16, 8 or 4: the fewest whose CTAs are all
// co-resident), 32 threads per warp, 16 warps per CTA, G = ceil(N / (512 FM)) CTAs — and each Picard iteration
// solves the N x N tridiagonal system A delta = -R by a partition (SPIKE / cyclic-reduction type) method on that
// hierarchy:
//   thread  assembles its FM rows from the closed forms of the 2-point-Gauss element integrals
//           (as bx_fom.cu), sweeps them in continuant form carrying the coupling to the previous
//           partition's last unknown P, and condenses to
//              row:  al P + be x_last + ga E = rho      (E = first unknown of the next partition)
//              fgr:  x_first = R - F P - G x_last
//   warp / CTA   "group_up": neighbours' fgr close the interior couplings (E -> x_last of the next
//           partition), a parallel cyclic reduction with three right-hand sides solves the
//           group's rows for  x_last(j) = y_j - P z1_j + X z2_j  (P, X: the group's own boundary
//           unknowns), and the group's row / fgr follow by substitution;
//   grid    every CTA publishes its row + fgr, waits for those of all CTAs, solves the G-row closed system
//           redundantly (5 rows per lane + warp PCR) and walks back down.
// Then local back-substitution, update, |delta|^2 and |U|^2.
//
// ONE synchronisation of the grid per Picard iteration, and no grid barrier:
//   * the norm partials and the partition-edge states of update k travel WITH the rows of iteration k + 1: the
//     convergence test runs one synchronisation late, so the front half of k + 1 (assembly, sweeps, condensation) is
//     speculative and is discarded when update k had converged or reached kmax (one front half per time step;
//     u is only changed by confirmed iterations: states and counts are those of the reference loop);
//   * the states just outside a CTA need no exchange of their own: they are the neighbours' published edge states plus
//     the interface solution every CTA has anyway (x_last of the CTA before, x_first = R - F P - G x_last of the one after);
//   * what a CTA publishes (11 doubles) travels as 22 epoch-tagged 8-byte words; every thread polls its share of the
//     G x 22 words into shared memory — no fence, no atomic counter, no second read of L2 (see "synchronisation" below).
// Measured on B200, 2^20 nodes: 19.3 us per Picard iteration with two grid barriers -> 13.8 us (profiles/r4*).
//
// U_k lives in registers and the swept rows (cs, rs, fs) in shared memory (3 x 8 x 512 FM bytes) for the whole
// trajectory; mesh / forcing constants and M U^n + At F stream from an L2-resident scratch in a [row][thread]
// interleaved layout (coalesced) — at FM = 16 through a two-stage cp.async ring.  HBM traffic: the 8N-byte snapshot
// per time step.
//
// MULTI (bx_fom_picard_finemesh_ranks): the trajectory is cut once more, into R <= 8 contiguous node blocks, one per
// GPU (one process per GPU), and the solve gets a FOURTH level: every CTA of a rank folds the rank's CTA rows into
// ONE row + fgr (two group_up stages: 32 CTAs per warp, then the <= 5 groups), CTA 0 stores it, with the rank's norm
// partials and block-edge states (11 doubles as 22 tagged words), straight into every rank's exchange buffer through
// peer memory (NVLink; cudaIpc-mapped), every CTA polls the R slots in its OWN buffer, solves the R-row closed system
// redundantly and walks back down.  One NVLink crossing per Picard iteration, no collective library on the data path,
// no host round trip; the state of a rank never leaves its GPU.  Meshes up to R x 148 x 8192 nodes.
#include "bx_common.cuh"
#include <cstdint>
#include <cstring>
#include <type_traits>

namespace bx {

constexpr int FM_MAX = 16;        // nodes per thread: 16 (8192 nodes per CTA), 8 or 4 — the fewest that keep the CTAs co-resident
constexpr int FT = 512;           // threads per CTA
constexpr int FW = FT / 32;       // warps per CTA
constexpr int FBLOCK = FM_MAX * FT;   // granularity of the node blocks of a partitioned trajectory
constexpr int FGMAX = 160;        // CTAs the top-level solve covers (5 rows per lane)
constexpr int FRMAX = 8;          // ranks the rank-level solve covers

// exchange buffer of a rank (peer-mapped by every other rank), byte offsets
constexpr size_t PX_EPOCH = 0;        // u64: exchanges done on these buffers so far (kept by the kernels, the same on every rank)
constexpr size_t PX_FAIL = 128;       // u64: non-zero once any rank gave up waiting
constexpr size_t PX_PAYLOAD = 256;    // u64 [2 parities][FRMAX sources][32 words]
constexpr int PX_VALUES = 11;         // a rank's row + fgr (7), norm partials (2), edge states (2)
constexpr int PX_WORDS = 2 * PX_VALUES;   // each double as 2 words of (32-bit epoch tag << 32 | 32 data bits)
constexpr size_t PX_BYTES = 256 + 2 * FRMAX * 32 * 8;
constexpr unsigned long long PX_TIMEOUT_NS = 10ull * 1000ull * 1000ull * 1000ull;

struct FineParams {
    int B, N, nsteps, kmax, G;
    const double *X;
    const double *u0;
    long long u0_stride;
    const double *mu1;
    const double *mu2;
    double At, E, tol2;
    double *U;
    int *iters;
    double *g_thr;            // per CTA: h6 | hfs ((FM+1) x FT each, element jj at [(jj+1) * FT + tid]) | atf | bn (FM x FT each, node j at [j * FT + tid])
    unsigned long long *g_ll; // [2 parities][G][32]: a CTA's row + fgr (7), norm partials (2), edge states (2) as 22 tagged words
    unsigned long long *bar;  // [16] = abort word of the rank (MULTI)
    // MULTI only: this rank's node block [off, off + Nloc) of the N-node mesh; u0 / U / iters are the block's
    int R, rank, off, Nloc;
    char *peer[FRMAX];        // exchange buffer of every rank as mapped here ([rank] = own)
};

struct Lvl { double y, z1, z2; };

__device__ __forceinline__ void cp_async8(unsigned int smem_addr, const double *gptr) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_addr), "l"(gptr) : "memory");
}

// BX_FINE_TRACE (experiments, tools/fine_trace.py): cycles of thread 0 of CTA 0 between marks, summed per mark
#ifdef BX_FINE_TRACE
#ifndef FINE_TRACE_RANK
#define FINE_TRACE_RANK 0   // which rank of a partitioned trajectory is traced (clock64 is per SM: one CTA only)
#endif
__device__ unsigned long long g_fine_trace[32];
__device__ unsigned long long g_fine_tlast;
#define FINE_MARK(i)                                                      \
    do {                                                                  \
        if (threadIdx.x == 0 && blockIdx.x == 0 && p.rank == FINE_TRACE_RANK) { \
            const unsigned long long t_ = clock64();                      \
            g_fine_trace[i] += t_ - g_fine_tlast;                         \
            g_fine_trace[16 + (i)] += 1ull;                               \
            g_fine_tlast = t_;                                            \
        }                                                                 \
    } while (0)
#else
#define FINE_MARK(i) do { } while (0)
#endif

__device__ __forceinline__ void warp_sum_pair(double &a, double &b) {
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) { a += shfl_xor(a, s); b += shfl_xor(b, s); }
}

// one division-free PCR step on a row with three right-hand sides (see pcr_step in bx_common.cuh)
__device__ __forceinline__ void pcr_step3(double &al, double &be, double &ga, double &r0, double &r1, double &r2, int s) {
    const double al_m = shfl_up(al, s), be_m = shfl_up(be, s), ga_m = shfl_up(ga, s);
    const double al_p = shfl_down(al, s), be_p = shfl_down(be, s), ga_p = shfl_down(ga, s);
    const double r0m = shfl_up(r0, s), r1m = shfl_up(r1, s), r2m = shfl_up(r2, s);
    const double r0p = shfl_down(r0, s), r1p = shfl_down(r1, s), r2p = shfl_down(r2, s);
    const double bb = be_m * be_p, ab = al * be_p, gb = ga * be_m;
    be = fma(be, bb, -fma(ab, ga_m, gb * al_p));
    r0 = fma(r0, bb, -fma(ab, r0m, gb * r0p));
    r1 = fma(r1, bb, -fma(ab, r1m, gb * r1p));
    r2 = fma(r2, bb, -fma(ab, r2m, gb * r2p));
    al = -ab * al_m;
    ga = -gb * ga_p;
}

// Group m (<= 32) consecutive partitions, one per lane (lanes >= m must hold identity rows:
// al = ga = 0, be = 1, rho = 0, F = G = R = 0).  In: the lane's row (al, be, gaE, rho) and fgr.
// Out: lv (x_last(lane) = y - P z1 + X z2) and, valid on EVERY lane, the group's row and fgr.
// STEPS: PCR steps, 2^STEPS >= m (after s steps a row only couples to rows 2^s away: none left in the group)
template <int STEPS = 5>
__device__ __forceinline__ Lvl group_up(int m, int lane, double &al, double &be, double &ga, double &rho,
                                        double &F, double &G, double &R) {
    // close the interior couplings with the next partition's fgr
    const double Fn = shfl_down(F, 1), Gn = shfl_down(G, 1), Rn = shfl_down(R, 1);
    const bool last = lane == m - 1;
    const double gaE = ga;
    double a_ = al, b_ = be, g_ = 0.0, r0 = rho, r1 = 0.0, r2 = 0.0;
    if (lane < m - 1) {
        b_ = fma(-gaE, Fn, be);
        g_ = -gaE * Gn;
        r0 = fma(-gaE, Rn, rho);
    }
    if (lane == 0) { r1 = a_; a_ = 0.0; }                    // coupling to the group's P moves to the right
    if (last) { a_ = 0.0; b_ = 1.0; g_ = 0.0; r0 = 0.0; r1 = 0.0; r2 = 1.0; }   // x_last(m-1) = X
    if (lane >= m) { a_ = 0.0; b_ = 1.0; g_ = 0.0; r0 = 0.0; r1 = 0.0; r2 = 0.0; }
    pcr_step3(a_, b_, g_, r0, r1, r2, 1);
    pcr_step3(a_, b_, g_, r0, r1, r2, 2);
    {
        const double inv = rcp(b_);
        a_ *= inv; g_ *= inv; r0 *= inv; r1 *= inv; r2 *= inv; b_ = 1.0;
    }
    if (STEPS > 2) pcr_step3(a_, b_, g_, r0, r1, r2, 4);
    if (STEPS > 3) pcr_step3(a_, b_, g_, r0, r1, r2, 8);
    if (STEPS > 4) pcr_step3(a_, b_, g_, r0, r1, r2, 16);
    const double inv = rcp(b_);
    Lvl lv;
    lv.y = r0 * inv; lv.z1 = r1 * inv; lv.z2 = r2 * inv;
    // the group's row: the true last row with x_last(m-2) substituted (lane -1 stands for P itself)
    double ym = shfl_up(lv.y, 1), z1m = shfl_up(lv.z1, 1), z2m = shfl_up(lv.z2, 1);
    if (m == 1) { ym = 0.0; z1m = -1.0; z2m = 0.0; }
    const double gal = -al * z1m, gbe = fma(al, z2m, be), grho = fma(-al, ym, rho);
    // the group's fgr from partition 0
    const double gF = fma(-G, lv.z1, F), gG = G * lv.z2, gR = fma(-G, lv.y, R);
    al = shfl_idx(gal, m - 1); be = shfl_idx(gbe, m - 1); ga = shfl_idx(gaE, m - 1); rho = shfl_idx(grho, m - 1);
    F = shfl_idx(gF, 0); G = shfl_idx(gG, 0); R = shfl_idx(gR, 0);
    return lv;
}

// ---- synchronisation of the CTAs of the (co-resident, cooperative) grid without a barrier --------------------------
// What a CTA publishes per Picard iteration — its row + fgr, the norm partials and edge states of its last update:
// LL_VALUES doubles — travels as 2 x LL_VALUES 8-byte words of (32-bit epoch tag << 32 | 32 data bits).  An aligned
// 8-byte store is seen whole, so a word whose tag equals the epoch IS the data: no fence, no counter, no second read;
// every thread of every CTA polls its share of the G x 22 words (all loads of a round in flight together) and drops
// the data halves into shared memory.  Slots alternate with the parity of the epoch: a CTA can publish epoch e + 2 only
// after it has seen epoch e + 1 of every CTA, which they publish after all their threads have read epoch e.
constexpr int LL_VALUES = 11;
constexpr int LL_WORDS = 2 * LL_VALUES;
constexpr int LL_STRIDE = LL_WORDS;   // words per CTA slot: packed, word 22 c + w of a synchronisation sits at [22 c + w]
__device__ __forceinline__ void st_gpu_u64(unsigned long long *a, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(a), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_gpu_u64(const unsigned long long *a) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(a) : "memory");
    return v;
}

// ---- peer-memory exchange between the ranks of a partitioned trajectory (MULTI) -------------------------
// Each double travels as two 8-byte words that carry 32 data bits and the 32-bit epoch of the exchange: an aligned
// 8-byte store arrives whole, so a word whose tag equals the expected epoch IS the data — no fence, no separate flag,
// one NVLink crossing per exchange (the scheme NCCL calls LL).  Slots alternate with the parity of the epoch; a slot is
// rewritten two exchanges of its kind later, which a rank can only reach after every rank has published the exchange in
// between, i.e. after every CTA of every rank has read this one (a grid barrier lies between reading and publishing).
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void st_sys_u64(unsigned long long *a, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(a), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_sys_u64(const unsigned long long *a) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(a) : "memory");
    return v;
}
__device__ __forceinline__ size_t px_slot(unsigned int ep, int src) {
    return PX_PAYLOAD + 8 * 32 * ((size_t)(ep & 1u) * FRMAX + src);
}

// ONE warp of the rank (all 32 lanes, v identical on every lane): word w of this rank's 7 values goes to every rank's
// buffer, one store per (destination, word) spread over the lanes
__device__ __forceinline__ void peer_publish(const FineParams &p, unsigned int ep, int lane, const double (&v)[PX_VALUES]) {
    const int total = p.R * PX_WORDS;
    for (int idx = lane; idx < total; idx += 32) {
        const int dst = idx / PX_WORDS, w = idx - dst * PX_WORDS;
        double d = v[0];
#pragma unroll
        for (int i = 1; i < PX_VALUES; ++i) d = (w >> 1) == i ? v[i] : d;
        const unsigned int half = (w & 1) ? (unsigned int)__double2hiint(d) : (unsigned int)__double2loint(d);
        unsigned long long *slot = reinterpret_cast<unsigned long long *>(p.peer[dst] + px_slot(ep, p.rank));
        st_sys_u64(slot + w, ((unsigned long long)ep << 32) | half);
    }
}

// one warp of every CTA: lane s < R waits until the 14 words of rank s in the OWN buffer carry this epoch and unpacks its
// 7 values (v[7] = 0; lanes >= R: zeros).  A wait that exceeds PX_TIMEOUT_NS, or a failure word set by anybody, raises the
// failure word of every rank and this rank's abort word; the values are then meaningless and the next grid barrier
// ends the kernel.
__device__ __forceinline__ void peer_gather(const FineParams &p, unsigned int ep, int lane, double (&v)[PX_VALUES]) {
#pragma unroll
    for (int i = 0; i < PX_VALUES; ++i) v[i] = 0.0;
    if (lane < p.R) {
        char *own = p.peer[p.rank];
        const unsigned long long *slot = reinterpret_cast<const unsigned long long *>(own + px_slot(ep, lane));
        volatile unsigned long long *failw = reinterpret_cast<volatile unsigned long long *>(own + PX_FAIL);
        volatile unsigned long long *abortw = reinterpret_cast<volatile unsigned long long *>(p.bar + 16);
        unsigned long long t0 = 0ull;
        bool ok = true;
        unsigned spins = 0;
        unsigned long long wd[PX_WORDS];
        for (;;) {
            bool all = true;
#pragma unroll
            for (int w = 0; w < PX_WORDS; ++w) wd[w] = ld_sys_u64(slot + w);
#pragma unroll
            for (int w = 0; w < PX_WORDS; ++w) all = all && (unsigned int)(wd[w] >> 32) == ep;
            if (all) break;
            if ((++spins & 63u) == 0u) {
                if (*failw != 0ull || *abortw != 0ull) { ok = false; break; }
                const unsigned long long t = global_timer_ns();
                if (t0 == 0ull) t0 = t;
                if (t - t0 > PX_TIMEOUT_NS) { ok = false; break; }
            }
        }
        if (!ok) {
            for (int s = 0; s < p.R; ++s) *reinterpret_cast<volatile unsigned long long *>(p.peer[s] + PX_FAIL) = 1ull;
            *abortw = 1ull;
            __threadfence_system();
        } else {
#pragma unroll
            for (int i = 0; i < PX_VALUES; ++i) v[i] = __hiloint2double((int)(unsigned int)wd[2 * i + 1], (int)(unsigned int)wd[2 * i]);
        }
    }
    __syncwarp();
}

template <int FM, bool MULTI> __global__ void __launch_bounds__(FT, 1) fine_fom_kernel(const FineParams p) {
    constexpr int FNODES = FM * FT;   // nodes per CTA
    extern __shared__ __align__(16) double smem[];
    double *s_cs = smem;                    // swept rows of this thread: [j * FT + tid]
    double *s_rs = s_cs + FNODES;
    double *s_fs = s_rs + FNODES;
    double *s_x = s_fs + FNODES;            // small exchange area (256 doubles)
    double *s_wrow = s_x;                   // [FW][8] warp rows + fgr
    double *s_top = s_x + 8 * FW;           // [FGMAX + 1]: x_last of CTA c at [c + 1], [0] = 0
    double *s_wx = s_top + FGMAX + 2;       // [FW + 1]: x_last of warp w at [w + 1], [0] = CTA's P
    double *s_red = s_wx + FW + 2;          // [FW][4] norm partials / warp edge states
    double *s_dec = s_red + 4 * FW;         // [10]: sd, su | states outside the CTA: corrected L, R, current L, R | rank: L, R, dR
    double *s_grp = s_dec + 10;             // MULTI: [5][8] rows + fgr of the groups of 32 CTAs
    double *s_gx = s_grp + 40;              // MULTI: [6]: x_last of group g at [g + 1], [0] = the rank's P
    int *s_abort = reinterpret_cast<int *>(s_gx + 8);
    double *s_meta = s_gx + 10;             // [4]: norm partials of the CTA's last update, its edge states (travel with the next rows)
    double *s_lc = s_meta + 4;              // [FW][3]: warp 0's CTA-level y, z1, z2 (kept out of the registers of 15 other warps)
    double *s_rows = s_x + 512;             // [G][LL_VALUES]: what every CTA published for this synchronisation
    unsigned int *s_rows32 = reinterpret_cast<unsigned int *>(s_rows);
    // the same bytes, between synchronisations: ring of two row stages [stage][h6R | hfs | bn][tid] that streams the
    // thread's per-row constants from the L2-resident scratch two rows ahead of the sweep (cp.async: no registers held)
    double *s_ring = s_rows + threadIdx.x;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cta = blockIdx.x, G = p.G, N = p.N;
    const int off = MULTI ? p.off : 0;       // first node of this rank's block
    const int Nl = MULTI ? p.Nloc : N;       // nodes of the block (row length of u0 / U)
    const int node0 = off + (cta * FT + tid) * FM;   // global node of the thread's first row
    const int jl = N - 1 - node0;           // local index of the last real row (may be outside [0, FM))
    const double At = p.At;
    const double a6 = At / 6.0, c1 = 1.5 * At;
    const bool has_e = p.E != 0.0;
    const double ate = At * p.E;
    unsigned int nsync = 0u;   // synchronisations so far = epoch of the last one
    // epoch of the exchanges between ranks: continues where the last launch on these buffers stopped
    unsigned int xe = 0u;
    if (MULTI) xe = (unsigned int)*reinterpret_cast<const volatile unsigned long long *>(p.peer[p.rank] + PX_EPOCH);

    // one pointer, constant offsets: the four per-thread scratch arrays of this CTA are contiguous
    double *t_h6 = p.g_thr + (size_t)cta * ((4 * FM + 2) * FT) + tid;
    double *t_hfs = t_h6 + (FM + 1) * FT;
    double *t_atf = t_hfs + (FM + 1) * FT;
    double *t_bn = t_atf + FNODES;
    double *t_cs = s_cs + tid, *t_rs = s_rs + tid, *t_fs = s_fs + tid;

    // ---- mesh constants of this thread's elements (jj = -1 .. FM-1), once ----------------------
    for (int jj = -1; jj < FM; ++jj) {
        const int e = node0 + jj;
        double h6 = 0.0;
        if (e >= 0 && e <= N - 2) h6 = (p.X[e + 1] - p.X[e]) / 6.0;
        t_h6[(jj + 1) * FT] = h6;
    }

    if (MULTI && tid == 0) *s_abort = 0;
    // (measured: the ring pays at 16 rows per thread, 15.4 -> 14.1 us per iteration; with 8 or 4 rows the compiler has the
    // registers to hoist the plain loads itself and the ring only adds instructions: 6.3 -> 6.8 us at 4 rows)
    constexpr bool RING = FM >= 16;
    const unsigned int ring_sa = (unsigned int)__cvta_generic_to_shared(s_ring);
    auto ring_issue = [&](int j) {   // row j -> stage j & 1
        const unsigned int dst = ring_sa + (unsigned int)((j & 1) * 3 * FT * 8);
        cp_async8(dst, t_h6 + (j + 1) * FT);
        cp_async8(dst + FT * 8, t_hfs + (j + 1) * FT);
        cp_async8(dst + 2 * FT * 8, t_bn + j * FT);
    };
    auto ring_start = [&]() {   // rows 0 and 1 of the next sweep
        if (!RING) return;
        asm volatile("cp.async.wait_all;" ::: "memory");
        ring_issue(0);
        asm volatile("cp.async.commit_group;" ::: "memory");
        ring_issue(1);
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    const bool fast_thread = !has_e && node0 > 0 && jl >= FM;
    const bool leftv = node0 > 0 && jl >= 0;
    const double a6L = leftv ? a6 : 0.0, c1L = leftv ? c1 : 0.0;

    for (int smp = 0; smp < p.B; ++smp) {
        const double mu1 = p.mu1[smp], mu2 = p.mu2[smp];
        // ---- per-sample forcing constants (FEM/fem_burgers.py:449-455, 565) ---------------------
        {
            double left = 0.0;
            for (int jj = -1; jj < FM; ++jj) {
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
                t_hfs[(jj + 1) * FT] = hfs;
                if (jj >= 0) t_atf[jj * FT] = At * (left + c0);
                left = cr;
            }
        }
        // ---- initial condition -> registers + snapshot 0 ---------------------------------------
        double u[FM];
        double *Us = p.U + (long long)smp * (long long)(p.nsteps + 1) * Nl;
        {
            const double *u0p = p.u0 + (long long)smp * p.u0_stride;
#pragma unroll
            for (int j = 0; j < FM; ++j) {
                u[j] = (node0 + j < N) ? u0p[node0 - off + j] : 0.0;
                if (node0 + j < N) Us[node0 - off + j] = u[j];
            }
        }
        double sd = 0.0, su = 0.0, uL = 0.0, uR = 0.0;

        // After an update, inside the CTA: the neighbours' edge states through shuffles / shared memory, the CTA's norm
        // partials and own edge states into warp 0 (they travel with the next rows); the states just outside the CTA are
        // s_dec[slotL], s_dec[slotR]
        auto cta_edges = [&](double psd, double psu, int slotL, int slotR) {
            warp_sum_pair(psd, psu);
            if (lane == 0) { s_red[4 * warp + 0] = psd; s_red[4 * warp + 1] = psu; s_red[4 * warp + 2] = u[0]; }
            if (lane == 31) s_red[4 * warp + 3] = u[FM - 1];
            __syncthreads();
            if (tid == 0) {
                double a = 0.0, b = 0.0;
#pragma unroll
                for (int w = 0; w < FW; ++w) { a += s_red[4 * w]; b += s_red[4 * w + 1]; }
                s_meta[0] = a; s_meta[1] = b; s_meta[2] = s_red[2]; s_meta[3] = s_red[4 * (FW - 1) + 3];
            }
            uL = shfl_up(u[FM - 1], 1);
            uR = shfl_down(u[0], 1);
            if (lane == 0) uL = warp > 0 ? s_red[4 * (warp - 1) + 3] : s_dec[slotL];
            if (lane == 31) uR = warp < FW - 1 ? s_red[4 * (warp + 1) + 2] : s_dec[slotR];
        };

        // ONE synchronisation of the whole mesh per Picard iteration.  In: the CTA's row + fgr (warp 0) and lc.  They are
        // published together with the norm partials and edge states of the PREVIOUS update (double-buffered by the parity
        // of the synchronisation: a CTA can run at most one synchronisation ahead of another).  Out: sd, su of the
        // previous update (summed in CTA order, then rank order: every CTA takes the same decision), x_last of every warp
        // of this CTA in s_wx, and the states just outside the CTA: s_dec[4], s_dec[5] as the neighbours hold them now,
        // s_dec[2], s_dec[3] with this solve's correction added (x_last of the CTA before, x_first of the CTA after —
        // what the neighbours will hold after applying this solve).
        auto sync_step = [&](double a2, double b2, double g2, double r2, double F2, double G2, double R2) -> bool {
            const unsigned int ep = ++nsync;   // tags start at 1: the buffer is zeroed at launch
            unsigned long long *slots = p.g_ll + (size_t)(ep & 1u) * ((size_t)G * LL_STRIDE);
            if (warp == 0 && lane < LL_WORDS) {
                const int vi = lane >> 1;
                double d = a2;
                d = vi == 1 ? b2 : d; d = vi == 2 ? g2 : d; d = vi == 3 ? r2 : d; d = vi == 4 ? F2 : d; d = vi == 5 ? G2 : d;
                d = vi == 6 ? R2 : d;
                if (vi >= 7) d = s_meta[vi - 7];
                const unsigned int half = (lane & 1) ? (unsigned int)__double2hiint(d) : (unsigned int)__double2loint(d);
                st_gpu_u64(slots + (size_t)cta * LL_STRIDE + lane, ((unsigned long long)ep << 32) | half);
            }
            FINE_MARK(4);
            {
                // word idx = 22 c + w of the G x 22 goes to thread idx mod FT; its data half lands at s_rows32[idx]
                constexpr int MAXW = (FGMAX * LL_WORDS + FT - 1) / FT;
                const int total = G * LL_WORDS;
                volatile unsigned long long *abortw = reinterpret_cast<volatile unsigned long long *>(p.bar + 16);
                unsigned long long wd[MAXW];
                unsigned spins = 0;
                for (;;) {
                    bool all = true;
#pragma unroll
                    for (int q = 0; q < MAXW; ++q) {
                        const int idx = tid + q * FT;
                        if (idx < total) wd[q] = ld_gpu_u64(slots + idx);
                    }
#pragma unroll
                    for (int q = 0; q < MAXW; ++q)
                        if (tid + q * FT < total) all = all && (unsigned int)(wd[q] >> 32) == ep;
                    if (all) break;
                    if (MULTI && (++spins & 255u) == 0u && *abortw != 0ull) { *s_abort = 1; break; }
                }
#pragma unroll
                for (int q = 0; q < MAXW; ++q)
                    if (tid + q * FT < total) s_rows32[tid + q * FT] = (unsigned int)wd[q];
            }
            FINE_MARK(5);
            __syncthreads();
            FINE_MARK(6);
            if (MULTI && *s_abort != 0) return false;
            if (MULTI) {
                // ---- rank level: fold the G rows of this rank, exchange with the other ranks, unfold --------------
                const int ng = (G + 31) >> 5;   // groups of 32 CTAs, one per warp
                Lvl l1; l1.y = 0.0; l1.z1 = 0.0; l1.z2 = 0.0;
                if (warp < ng) {
                    const int c = 32 * warp + lane;
                    double a3 = 0.0, b3 = 1.0, g3 = 0.0, r3 = 0.0, F3 = 0.0, G3 = 0.0, R3 = 0.0;
                    if (c < G) {
                        const double *d = s_rows + LL_VALUES * c;
                        a3 = d[0]; b3 = d[1]; g3 = d[2]; r3 = d[3]; F3 = d[4]; G3 = d[5]; R3 = d[6];
                    }
                    const int m = G - 32 * warp < 32 ? G - 32 * warp : 32;
                    l1 = group_up(m, lane, a3, b3, g3, r3, F3, G3, R3);
                    if (lane == 0) {
                        double *d = s_grp + 8 * warp;
                        d[0] = a3; d[1] = b3; d[2] = g3; d[3] = r3; d[4] = F3; d[5] = G3; d[6] = R3;
                    }
                } else if (warp == ng) {
                    // beside the fold, an idle warp: the rank's norm partials, fixed order (lane l takes CTAs 5 l .. 5 l + 4,
                    // then a butterfly)
                    double ra = 0.0, rb = 0.0;
#pragma unroll
                    for (int j = 0; j < 5; ++j) {
                        const int c = 5 * lane + j;
                        if (c < G) { ra += s_rows[LL_VALUES * c + 7]; rb += s_rows[LL_VALUES * c + 8]; }
                    }
                    warp_sum_pair(ra, rb);
                    if (lane == 0) { s_dec[6] = ra; s_dec[7] = rb; }
                }
                __syncthreads();
                if (warp == 0) {
                    double a4 = 0.0, b4 = 1.0, g4 = 0.0, r4 = 0.0, F4 = 0.0, G4 = 0.0, R4 = 0.0;
                    if (lane < ng) {
                        const double *d = s_grp + 8 * lane;
                        a4 = d[0]; b4 = d[1]; g4 = d[2]; r4 = d[3]; F4 = d[4]; G4 = d[5]; R4 = d[6];
                    }
                    const Lvl l2 = group_up<3>(ng, lane, a4, b4, g4, r4, F4, G4, R4);   // -> the rank's row + fgr
                    const double ra = s_dec[6], rb = s_dec[7];   // (s_dec[6 .. 8] are rewritten below, by this warp)
                    ++xe;
                    double v[PX_VALUES];
                    if (cta == 0) {
                        v[0] = a4; v[1] = b4; v[2] = g4; v[3] = r4; v[4] = F4; v[5] = G4; v[6] = R4;
                        v[7] = ra; v[8] = rb; v[9] = s_rows[9]; v[10] = s_rows[LL_VALUES * (G - 1) + 10];
                        peer_publish(p, xe, lane, v);
                    }
                    peer_gather(p, xe, lane, v);
                    double a = 0.0, b = 0.0;
                    for (int r = 0; r < p.R; ++r) { a += shfl_idx(v[7], r); b += shfl_idx(v[8], r); }
                    // closed R-row system: one more group_up (lanes >= R hold identity rows), then the group's row
                    // with nothing left (P = 0) or right (ga = 0) of the mesh gives x_last of the last rank
                    double a5 = 0.0, b5 = 1.0, g5 = 0.0, r5 = 0.0, F5 = 0.0, G5 = 0.0, R5 = 0.0;
                    if (lane < p.R) { a5 = v[0]; b5 = v[1]; g5 = v[2]; r5 = v[3]; F5 = v[4]; G5 = v[5]; R5 = v[6]; }
                    const Lvl l3 = group_up<3>(p.R, lane, a5, b5, g5, r5, F5, G5, R5);
                    static_assert(FRMAX <= 8 && FGMAX <= 8 * 32, "rank-level groups: three PCR steps");
                    const double Xtop = r5 * rcp(b5);
                    const double xr = fma(Xtop, l3.z2, l3.y);          // x_last of rank `lane`
                    double Pr = shfl_idx(xr, p.rank > 0 ? p.rank - 1 : 0);
                    if (p.rank == 0) Pr = 0.0;
                    const double Xr = shfl_idx(xr, p.rank);
                    // states just outside the rank's block, and the correction of the one on the right: x_first of the next rank
                    const int rn = p.rank < p.R - 1 ? p.rank + 1 : p.rank;
                    double tLr = shfl_idx(v[10], p.rank > 0 ? p.rank - 1 : 0), tRr = shfl_idx(v[9], rn);
                    double dRr = fma(-shfl_idx(v[5], rn), shfl_idx(xr, rn), fma(-shfl_idx(v[4], rn), Xr, shfl_idx(v[6], rn)));
                    if (p.rank == 0) tLr = 0.0;
                    if (p.rank == p.R - 1) { tRr = 0.0; dRr = 0.0; }
                    if (lane < ng) s_gx[lane + 1] = fma(Xr, l2.z2, fma(-Pr, l2.z1, l2.y));
                    if (lane == 0) {
                        s_gx[0] = Pr;
                        s_dec[0] = a; s_dec[1] = b; s_dec[6] = tLr; s_dec[7] = tRr; s_dec[8] = dRr;
                    }
                }
                __syncthreads();
                if (warp < ng) {
                    const double Pg = s_gx[warp], Xg = s_gx[warp + 1];
                    if (32 * warp + lane < G) s_top[32 * warp + lane + 1] = fma(Xg, l1.z2, fma(-Pg, l1.z1, l1.y));
                    if (warp == 0 && lane == 0) s_top[0] = Pg;
                }
                __syncthreads();
            }
            if (!MULTI && warp == 0) {   // ---- grid level: closed G-row system, solved by every CTA
                constexpr int TM = FGMAX / 32;   // 5 rows per lane
                double cs5[TM], rs5[TM], fs5[TM];
                Continuant<TM> ct;
#pragma unroll
                for (int j = 0; j < TM; ++j) {
                    const int c = lane * TM + j;
                    double a = 0.0, b = 1.0, cc = 0.0, r = 0.0;
                    if (c < G) {
                        const double *d = s_rows + LL_VALUES * c;
                        a = d[0]; b = d[1]; r = d[3];
                        if (c < G - 1) {   // E of CTA c = x_first of CTA c+1
                            const double gE = d[2];
                            const double Fn = d[LL_VALUES + 4], Gn = d[LL_VALUES + 5], Rn = d[LL_VALUES + 6];
                            b = fma(-gE, Fn, b); cc = -gE * Gn; r = fma(-gE, Rn, r);
                        }
                        if (c == 0) a = 0.0;
                    }
                    ct.row(j, a, b, cc, r, cs5[j], rs5[j], fs5[j]);
                }
                double xl, P;
                partition_condense<TM>(cs5, rs5, fs5, lane, xl, P);
                double x = xl;
#pragma unroll
                for (int j = TM - 1; j >= 0; --j) {
                    if (j < TM - 1) x = fma(-cs5[j], x, fma(-fs5[j], P, rs5[j]));
                    s_top[lane * TM + j + 1] = x;
                }
                if (lane == 0) s_top[0] = 0.0;
                __syncwarp();
                FINE_MARK(8);
                if (lane == 0) { s_dec[6] = 0.0; s_dec[7] = 0.0; s_dec[8] = 0.0; }
                __syncwarp();
            }
            if (!MULTI && warp == 1) {
                // beside it, warp 1: norms of the previous update, fixed order: lane l takes CTAs 5 l .. 5 l + 4, then a butterfly
                double a = 0.0, b = 0.0;
#pragma unroll
                for (int j = 0; j < 5; ++j) {
                    const int c = 5 * lane + j;
                    if (c < G) { a += s_rows[LL_VALUES * c + 7]; b += s_rows[LL_VALUES * c + 8]; }
                }
                warp_sum_pair(a, b);
                if (lane == 0) { s_dec[0] = a; s_dec[1] = b; }
            }
            if (warp == 0) {
                // CTA level, downwards: x_last of every warp; the states just outside the CTA
                const double Pc = s_top[cta], Xc = s_top[cta + 1];
                if (lane < FW) s_wx[lane + 1] = fma(Xc, s_lc[3 * lane + 2], fma(-Pc, s_lc[3 * lane + 1], s_lc[3 * lane]));
                if (lane == 0) {
                    s_wx[0] = Pc;
                    const double tL = cta > 0 ? s_rows[LL_VALUES * (cta - 1) + 10] : s_dec[6];
                    double tR = s_dec[7], dR = s_dec[8];
                    if (cta < G - 1) {   // x_first of the next CTA = R - F x_last(this CTA) - G x_last(next CTA)
                        const double *d = s_rows + LL_VALUES * (cta + 1);
                        tR = d[9];
                        dR = fma(-d[5], s_top[cta + 2], fma(-d[4], Xc, d[6]));
                    }
                    s_dec[2] = tL + Pc; s_dec[3] = tR + dR; s_dec[4] = tL; s_dec[5] = tR;
                }
            }
            FINE_MARK(9);
            __syncthreads();
            FINE_MARK(10);
            sd = s_dec[0]; su = s_dec[1];
            return true;
        };
        // the states just outside the CTA as the neighbours hold them (sample start; after a discarded front half)
        auto adopt_edges = [&]() {
            if (tid == 0) uL = s_dec[4];
            if (tid == FT - 1) uR = s_dec[5];
        };

        // sample start: one synchronisation with identity rows brings the neighbours' edge states
        cta_edges(0.0, 0.0, 4, 5);
        {
            if (tid < FW) { s_lc[3 * tid] = 0.0; s_lc[3 * tid + 1] = 0.0; s_lc[3 * tid + 2] = 0.0; }
            if (!sync_step(0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 0.0)) return;
            adopt_edges();
        }

        for (int n = 0; n < p.nsteps; ++n) {
            // ---- bn = M U^n + At F (constant during the Picard loop, :683) ---------------------
            {
                double h6L = t_h6[0];
#pragma unroll
                for (int j = 0; j < FM; ++j) {
                    const double h6R = t_h6[(j + 1) * FT];
                    const double um = (j == 0) ? uL : u[j > 0 ? j - 1 : 0];
                    const double up = (j == FM - 1) ? uR : u[j < FM - 1 ? j + 1 : j];
                    t_bn[j * FT] = h6L * um + (2.0 * (h6L + h6R)) * u[j] + h6R * up + t_atf[j * FT];
                    h6L = h6R;
                }
            }
            // Picard loop (:663-698) with the convergence test one synchronisation late: the norms of update k travel with the
            // rows of iteration k + 1, whose front half (assembly, sweeps, condensation) is therefore speculative — it is
            // discarded when update k turns out to have converged (or reached kmax).  u is only ever changed by confirmed
            // iterations, so states and counts are those of the reference loop.
            int k = 0;
            if (RING) __threadfence_block();   // this thread's bn stores before its asynchronous copies of them
            ring_start();
            FINE_MARK(13);
            for (;;) {
                // ================= thread level: assemble + sweep (:667-692) ======================
                double al, be, ga, rho, F, Gc, R;
                // FAST: an interior thread (no Dirichlet row, no last element, no padded row) with E = 0 — all but a handful of
                // threads of the grid: the same arithmetic without the 12 selects per row and the diffusion terms (which are
                // exact zeros there: identical results)
                auto sweep_rows = [&](auto fast_c) {
                    constexpr bool FAST = decltype(fast_c)::value;
                    Continuant<FM, true> ct;   // 1-ulp reciprocals (3 DFMA), as in the FOM sweep kernels
                    double h6L = t_h6[0];
                    double qL, GL;
                    {
                        const double u1 = uL, u2 = u[0];
                        const double sm = u1 + u2;
                        qL = fma(a6L, u2, a6L * sm);
                        const double z = fma(sm * (u2 - u1), 1.0 / 6.0, -t_hfs[0]);
                        GL = (c1L * rcp1(fmax(fabs(sm), 2.0e-10))) * z;
                    }
                    double csj = 0.0, rsj = 0.0, fsj = 0.0;
#pragma unroll
                    for (int j = 0; j < FM; ++j) {
                        double h6R, hfsj, bnj;
                        if (RING) {
                            asm volatile("cp.async.wait_group 1;" ::: "memory");   // row j has landed (row j + 1 may be in flight)
                            h6R = s_ring[(j & 1) * 3 * FT]; hfsj = s_ring[(j & 1) * 3 * FT + FT]; bnj = s_ring[(j & 1) * 3 * FT + 2 * FT];
                            if (j + 2 < FM) ring_issue(j + 2);
                            asm volatile("cp.async.commit_group;" ::: "memory");
                        } else {
                            h6R = t_h6[(j + 1) * FT]; hfsj = t_hfs[(j + 1) * FT]; bnj = t_bn[j * FT];
                        }
                        const double um = (j == 0) ? uL : u[j > 0 ? j - 1 : 0];
                        const double up = (j == FM - 1) ? uR : u[j < FM - 1 ? j + 1 : j];
                        const double u1 = u[j], u2 = up;
                        const double sm = u1 + u2;
                        const double tt = a6 * sm;
                        double pe = fma(a6, u1, tt);
                        double qe = fma(a6, u2, tt);
                        const double z = fma(sm * (u2 - u1), 1.0 / 6.0, -hfsj);
                        double Ge = (c1 * rcp1(fmax(fabs(sm), 2.0e-10))) * z;
                        double a, b, c;
                        if (FAST) {
                            a = h6L - qL;
                            c = h6R + pe;
                            b = 2.0 * (h6L + h6R) + (qL - pe);
                        } else {
                            if (j >= jl) { pe = 0.0; qe = 0.0; Ge = 0.0; }   // no element to the right of node N-1
                            double keL = 0.0, keR = 0.0;
                            if (has_e) {
                                keL = h6L > 0.0 ? ate / (6.0 * h6L) : 0.0;
                                keR = h6R > 0.0 ? ate / (6.0 * h6R) : 0.0;
                            }
                            a = (h6L - keL) - qL;
                            c = (h6R - keR) + pe;
                            b = (2.0 * (h6L + h6R) + (keL + keR)) + (qL - pe);
                        }
                        double r = (bnj - GL + Ge) - fma(a, um, fma(b, u1, c * u2));
                        if (!FAST) {
                            if (j > jl) { a = 0.0; b = 1.0; c = 0.0; r = 0.0; }              // padded row
                            if (j == 0 && node0 == 0) { a = 0.0; b = 1.0; c = 0.0; r = mu1 - u1; }   // Dirichlet row (:679-686)
                        }
                        ct.row(j, a, b, c, r, csj, rsj, fsj);
                        t_cs[j * FT] = csj; t_rs[j * FT] = rsj; t_fs[j * FT] = fsj;
                        h6L = h6R; qL = qe; GL = Ge;
                    }
                    // row of the last unknown, and x_first = R - F P - G x_last by an upward sweep
                    al = fsj; be = 1.0; ga = csj; rho = rsj;
                    F = t_fs[(FM - 2) * FT]; Gc = t_cs[(FM - 2) * FT]; R = t_rs[(FM - 2) * FT];
#pragma unroll
                    for (int j = FM - 3; j >= 0; --j) {
                        const double cj = t_cs[j * FT];
                        F = fma(-cj, F, t_fs[j * FT]);
                        Gc = -cj * Gc;
                        R = fma(-cj, R, t_rs[j * FT]);
                    }
                };
                // (measured: 8.7 -> 8.3 us per iteration at 8 rows per thread, 6.5 -> 6.3 at 4; at 16 rows the second unrolled
                // copy of the sweep makes the kernel slower, 13.4 -> 14.7 us, so that instance keeps the general form only)
                if (FM <= 8 && fast_thread) sweep_rows(std::true_type{}); else sweep_rows(std::false_type{});
                // ================= warp level ====================================================
                FINE_MARK(0);
                const Lvl lw = group_up(32, lane, al, be, ga, rho, F, Gc, R);
                if (lane == 0) {
                    double *d = s_wrow + 8 * warp;
                    d[0] = al; d[1] = be; d[2] = ga; d[3] = rho; d[4] = F; d[5] = Gc; d[6] = R;
                }
                FINE_MARK(1);
                __syncthreads();
                FINE_MARK(2);
                // ================= CTA level (warp 0) ==============================================
                double a2 = 0.0, b2 = 1.0, g2 = 0.0, r2 = 0.0, F2 = 0.0, G2 = 0.0, R2 = 0.0;
                if (warp == 0) {
                    if (lane < FW) {
                        const double *d = s_wrow + 8 * lane;
                        a2 = d[0]; b2 = d[1]; g2 = d[2]; r2 = d[3]; F2 = d[4]; G2 = d[5]; R2 = d[6];
                    }
                    const Lvl lc = group_up<4>(FW, lane, a2, b2, g2, r2, F2, G2, R2);
                    if (lane < FW) { s_lc[3 * lane] = lc.y; s_lc[3 * lane + 1] = lc.z1; s_lc[3 * lane + 2] = lc.z2; }
                    static_assert(FW <= 16, "CTA-level group: four PCR steps");
                }
                FINE_MARK(3);
                if (!sync_step(a2, b2, g2, r2, F2, G2, R2)) return;
                if (k > 0 && !(sd > p.tol2 * su && k < p.kmax)) { adopt_edges(); break; }
                ring_start();   // for the next sweep, while this iteration is finished (bn does not change inside a step)
                // ================= warp level, downwards; back-substitution, update, norms ========
                double psd = 0.0, psu = 0.0;
                {
                    const double Pw = s_wx[warp], Xw = s_wx[warp + 1];
                    const double xl = fma(Xw, lw.z2, fma(-Pw, lw.z1, lw.y));
                    double P = shfl_up(xl, 1);
                    if (lane == 0) P = Pw;
                    double x = xl;
#pragma unroll
                    for (int j = FM - 1; j >= 0; --j) {
                        if (j < FM - 1) x = fma(-t_cs[j * FT], x, fma(-t_fs[j * FT], P, t_rs[j * FT]));
                        psd = fma(x, x, psd);
                        u[j] += x;
                        psu = fma(u[j], u[j], psu);
                    }
                }
                FINE_MARK(11);
                cta_edges(psd, psu, 2, 3);
                FINE_MARK(12);
                ++k;
            }

            // ---- snapshot n+1 (:705) and iteration count -------------------------------------------
            double *Un1 = Us + (long long)(n + 1) * Nl + (node0 - off);
            if (node0 + FM <= N && ((Nl & 1) == 0)) {
#pragma unroll
                for (int j = 0; j < FM; j += 2) *reinterpret_cast<double2 *>(Un1 + j) = make_double2(u[j], u[j + 1]);
            } else {
#pragma unroll
                for (int j = 0; j < FM; ++j) if (node0 + j < N) Un1[j] = u[j];
            }
            // non-finite state (config 5b: At = 0.05 on this mesh overflows): flag -1; the following steps stay
            // non-finite and carry the flag as well (include/burgers_b200.h)
            if (cta == 0 && tid == 0 && p.iters)
                p.iters[(long long)smp * p.nsteps + n] = (su <= 1.7976931348623157e308) ? k : -1;
        }
    }
    if (MULTI && cta == 0 && tid == 0)   // the next launch on these buffers continues the epochs
        *reinterpret_cast<volatile unsigned long long *>(p.peer[p.rank] + PX_EPOCH) = xe;
}

// swept rows (3 x nodes) | exchange area (512) | the rows of all CTAs, aliased with the two-stage constant ring (2 x 3 x FT)
constexpr size_t fine_smem(int fm) { return sizeof(double) * (3 * (size_t)fm * FT + 512 + 2 * 3 * (size_t)FT); }
static_assert(2 * 3 * FT >= FGMAX * LL_VALUES, "the ring covers the row area");

struct FinePlan {
    int G;
    size_t off_thr, off_ll, ll_bytes, total;
};

static int fine_plan(int N, int fm, FinePlan &pl) {
    const int nodes = fm * FT;
    pl.G = (N + nodes - 1) / nodes;
    size_t o = 256;
    auto take = [&](size_t doubles) { size_t at = o; o += sizeof(double) * doubles; o = (o + 255) & ~(size_t)255; return at; };
    pl.off_thr = take((size_t)pl.G * (4 * fm + 2) * FT);
    pl.off_ll = take(2 * (size_t)pl.G * LL_STRIDE);   // double-buffered by the parity of the synchronisation
    pl.ll_bytes = sizeof(unsigned long long) * 2 * (size_t)pl.G * LL_STRIDE;
    pl.total = o;
    return BX_OK;
}

}  // namespace bx

extern "C" long long bx_fine_workspace_bytes(int N) {
    using namespace bx;
    if (N < 2) return -1;
    size_t most = 0;
    for (int fm = 4; fm <= FM_MAX; fm *= 2) {
        FinePlan pl;
        fine_plan(N, fm, pl);
        most = pl.total > most ? pl.total : most;
    }
    return (long long)most;
}

namespace bx {

// one instance: MULTI = false integrates the whole mesh on this device, MULTI = true this rank's node block.
// probe: return 1 instead of an error when the CTAs of this instance cannot all be resident
template <int FM, bool MULTI>
static int fine_launch_rows(FineParams &p, void *workspace, cudaStream_t st, int limit, bool probe) {
    FinePlan pl;
    fine_plan(MULTI ? p.Nloc : p.N, FM, pl);
    int occ = 0;
    {
        const int rco = kernel_occupancy(reinterpret_cast<const void *>(fine_fom_kernel<FM, MULTI>), FT, fine_smem(FM), &occ);
        if (rco != BX_OK) return rco;
    }
    if (occ < 1 || pl.G > limit || pl.G > FGMAX) {
        if (probe) return 1;
        set_error("bx_fom_picard_finemesh: %d nodes need %d co-resident CTAs of %d nodes; this device runs %d",
                  MULTI ? p.Nloc : p.N, pl.G, FM * FT, occ >= 1 ? (limit < FGMAX ? limit : FGMAX) : 0);
        return BX_ERR_UNSUPPORTED;
    }
    char *ws = static_cast<char *>(workspace);
    p.G = pl.G;
    p.bar = reinterpret_cast<unsigned long long *>(ws);
    p.g_thr = reinterpret_cast<double *>(ws + pl.off_thr);
    p.g_ll = reinterpret_cast<unsigned long long *>(ws + pl.off_ll);
    BX_CUDA(cudaMemsetAsync(p.bar, 0, 256, st));
    BX_CUDA(cudaMemsetAsync(p.g_ll, 0, pl.ll_bytes, st));
    void *args[] = {&p};
    BX_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void *>(fine_fom_kernel<FM, MULTI>), dim3(pl.G), dim3(FT), args,
                                        fine_smem(FM), st));
    return BX_OK;
}

// rows per thread: the fewest (4, 8, 16) whose CTAs are all co-resident, one per SM — a Picard iteration is a chain of
// latencies (two grid barriers, three group solves) plus the thread-level sweep, which shrinks with the rows per thread;
// BX_TUNE_FINE_ROWS forces a value
template <bool MULTI>
static int fine_launch(FineParams &p, void *workspace, cudaStream_t st) {
    DeviceInfo di;
    {
        const int rcd = device_info(di);
        if (rcd != BX_OK) return rcd;
    }
    if (!di.cooperative) {
        set_error("bx_fom_picard_finemesh: the device does not support cooperative launches");
        return BX_ERR_UNSUPPORTED;
    }
    const int forced = tuning(BX_TUNE_FINE_ROWS);
    if (forced == 4) return fine_launch_rows<4, MULTI>(p, workspace, st, di.sms, false);
    if (forced == 8) return fine_launch_rows<8, MULTI>(p, workspace, st, di.sms, false);
    if (forced == 16) return fine_launch_rows<16, MULTI>(p, workspace, st, di.sms, false);
    int rc = fine_launch_rows<4, MULTI>(p, workspace, st, di.sms, true);
    if (rc == 1) rc = fine_launch_rows<8, MULTI>(p, workspace, st, di.sms, true);
    if (rc == 1) rc = fine_launch_rows<16, MULTI>(p, workspace, st, di.sms, false);
    return rc;
}

}  // namespace bx

extern "C" int bx_fom_picard_finemesh(int B, int N, int nsteps, const double *X, const double *u0, int u0_batched,
                                      const double *mu1, const double *mu2, double At, double E, double tol,
                                      int kmax, double *U, int *iters, void *workspace, void *stream) {
    using namespace bx;
    BX_REQUIRE(B >= 0 && nsteps >= 0, "B and nsteps must be non-negative");
    BX_REQUIRE(N >= 2, "need at least 2 nodes (got N = %d)", N);
    BX_REQUIRE(kmax >= 1, "kmax must be >= 1 (the reference loop body defines U1)");
    if (B == 0) return BX_OK;
    BX_REQUIRE(X && u0 && mu1 && mu2 && U && workspace, "null device pointer");
    FineParams p = {};
    p.B = B; p.N = N; p.nsteps = nsteps; p.kmax = kmax;
    p.X = X; p.u0 = u0; p.u0_stride = u0_batched ? N : 0;
    p.mu1 = mu1; p.mu2 = mu2; p.At = At; p.E = E; p.tol2 = tol * tol;
    p.U = U; p.iters = iters;
    p.R = 1; p.rank = 0; p.off = 0; p.Nloc = N;
    return fine_launch<false>(p, workspace, static_cast<cudaStream_t>(stream));
}

// ---- the trajectory partitioned over R ranks (one process per GPU) ------------------------------------------
extern "C" long long bx_fine_peer_bytes(void) { return (long long)bx::PX_BYTES; }

// set-up calls (they allocate / map; not for the time-stepping loop): the exchange buffer has to come from cudaMalloc
// itself so that its IPC handle maps the buffer and nothing else
extern "C" int bx_peer_alloc(void **buf, void *ipc_handle_64_bytes) {
    using namespace bx;
    BX_REQUIRE(buf && ipc_handle_64_bytes, "null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    void *ptr = nullptr;
    BX_CUDA(cudaMalloc(&ptr, PX_BYTES));
    BX_CUDA(cudaMemset(ptr, 0, PX_BYTES));
    cudaIpcMemHandle_t h;
    BX_CUDA(cudaIpcGetMemHandle(&h, ptr));
    memcpy(ipc_handle_64_bytes, &h, sizeof h);
    *buf = ptr;
    return BX_OK;
}
extern "C" int bx_peer_open(const void *ipc_handle_64_bytes, void **buf) {
    using namespace bx;
    BX_REQUIRE(buf && ipc_handle_64_bytes, "null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle_64_bytes, sizeof h);
    BX_CUDA(cudaIpcOpenMemHandle(buf, h, cudaIpcMemLazyEnablePeerAccess));
    return BX_OK;
}
extern "C" int bx_peer_close(void *buf) {
    using namespace bx;
    BX_CUDA(cudaIpcCloseMemHandle(buf));
    return BX_OK;
}
extern "C" int bx_peer_free(void *buf) {
    using namespace bx;
    BX_CUDA(cudaFree(buf));
    return BX_OK;
}

extern "C" int bx_fom_picard_finemesh_ranks(int B, int N, int node_offset, int N_local, int nsteps, const double *X,
                                            const double *u0_local, int u0_batched, const double *mu1, const double *mu2,
                                            double At, double E, double tol, int kmax, double *U_local, int *iters,
                                            void *workspace, int R, int rank, void *const *peer_bufs, void *stream) {
    using namespace bx;
    BX_REQUIRE(B >= 0 && nsteps >= 0, "B and nsteps must be non-negative");
    BX_REQUIRE(N >= 2, "need at least 2 nodes (got N = %d)", N);
    BX_REQUIRE(kmax >= 1, "kmax must be >= 1 (the reference loop body defines U1)");
    BX_REQUIRE(R >= 1 && R <= FRMAX && rank >= 0 && rank < R, "1 <= R <= %d ranks, 0 <= rank < R (got R = %d, rank = %d)", FRMAX, R, rank);
    BX_REQUIRE(N_local >= 1 && node_offset >= 0 && (long long)node_offset + N_local <= N, "node block [%d, %d + %d) outside the mesh", node_offset, node_offset, N_local);
    BX_REQUIRE(node_offset % FBLOCK == 0, "node_offset must be a multiple of %d", FBLOCK);
    BX_REQUIRE(node_offset + N_local == N || N_local % FBLOCK == 0, "every block but the last holds a multiple of %d nodes", FBLOCK);
    BX_REQUIRE((rank == 0) == (node_offset == 0) && (rank == R - 1) == (node_offset + N_local == N), "blocks must be in rank order and cover the mesh");
    if (B == 0) return BX_OK;
    BX_REQUIRE(X && u0_local && mu1 && mu2 && U_local && workspace && peer_bufs, "null pointer");
    FineParams p = {};
    p.B = B; p.N = N; p.nsteps = nsteps; p.kmax = kmax;
    p.X = X; p.u0 = u0_local; p.u0_stride = u0_batched ? N_local : 0;
    p.mu1 = mu1; p.mu2 = mu2; p.At = At; p.E = E; p.tol2 = tol * tol;
    p.U = U_local; p.iters = iters;
    p.R = R; p.rank = rank; p.off = node_offset; p.Nloc = N_local;
    for (int r = 0; r < R; ++r) {
        BX_REQUIRE(peer_bufs[r] != nullptr, "peer_bufs[%d] is null", r);
        p.peer[r] = static_cast<char *>(peer_bufs[r]);
    }
    return fine_launch<true>(p, workspace, static_cast<cudaStream_t>(stream));
}

// 1 when a rank gave up waiting for another one during a launch on this exchange buffer (the outputs of that launch
// are then invalid on every rank), 0 otherwise; synchronises the stream
extern "C" int bx_fine_peer_failed(const void *own_buf, void *stream) {
    using namespace bx;
    BX_REQUIRE(own_buf, "null pointer");
    unsigned long long w = 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    BX_CUDA(cudaMemcpyAsync(&w, static_cast<const char *>(own_buf) + PX_FAIL, sizeof w, cudaMemcpyDeviceToHost, st));
    BX_CUDA(cudaStreamSynchronize(st));
    return w != 0ull ? 1 : 0;
}

#ifdef BX_FINE_TRACE
extern "C" int bx_fine_trace_read(unsigned long long *out32, int reset) {
    using namespace bx;
    BX_CUDA(cudaDeviceSynchronize());
    BX_CUDA(cudaMemcpyFromSymbol(out32, g_fine_trace, sizeof(unsigned long long) * 32));
    if (reset) {
        unsigned long long z[32] = {};
        BX_CUDA(cudaMemcpyToSymbol(g_fine_trace, z, sizeof z));
    }
    return BX_OK;
}
#endif
