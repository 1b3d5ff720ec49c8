// bx_fom.cuh — parameter block, shared-memory carve-up and small helpers shared by the two FOM trajectory
// kernels: the throughput kernel (bx_fom.cu) and the latency kernel (bx_fom_strong.cu).
#pragma once
#include "bx_common.cuh"

namespace bx {

struct FomParams {
    int B, N, nsteps, kmax;
    const double *X;
    const double *u0;
    long long u0_stride;  // 0: shared initial condition, N: per-sample
    const double *mu1;
    const double *mu2;
    double At, E, tol2;
    double a6, c1, c16;   // At / 6 (convection), 1.5 At and At / 4 (SUPG): kernel parameters so that they can stay in the constant bank
    double *U;
    int *iters;
    double *err_hist;     // optional (B, nsteps, kmax): |delta| / |U_{k+1}| after every iteration (:698)
    unsigned int *counter;
    int *progress;        // (B,) segments completed per trajectory (-1: not finite any more); used when nseg > 1
    int nseg, seg;        // throughput kernel: the nsteps time steps are cut into nseg segments of seg steps
};

// shared-memory carve-up of the throughput kernel (doubles): element arrays have (M+1)*32 entries, indexed
// (j+1)*32 + l for the element between local nodes j and j+1 of lane l (j = -1 is the element shared with
// the previous lane); node arrays have M*32 entries indexed j*32 + l.  Both are bank-conflict free for
// 64-bit accesses.
template <int M> struct FomSmem {
    static constexpr int EL = (M + 1) * 32;
    static constexpr int NL = M * 32;
    static constexpr int MESH = 2 * EL + NL;                 // off, h6 | dg
    static constexpr int PER_WARP = EL + 2 * NL + 2;         // hfs | AtF, bn | mu1
    static constexpr size_t bytes(int warps) { return sizeof(double) * (MESH + (size_t)warps * PER_WARP); }
};

// Sum two values over the warp with 7 instead of 10 64-bit shuffles: the two half-warps first
// trade the value they do not keep, each half reduces one quantity, lanes 0 / 16 broadcast.
__device__ __forceinline__ void warp_sum2(double &a, double &b, int lane) {
    const bool lo = lane < 16;
    double mine = lo ? a : b;
    const double give = lo ? b : a;
    mine += shfl_xor(give, 16);
#pragma unroll
    for (int s = 8; s > 0; s >>= 1) mine += shfl_xor(mine, s);
    a = shfl_idx(mine, 0);
    b = shfl_idx(mine, 16);
}

// latency kernel (bx_fom_strong.cu): W = 1, 2 or 4 warps per trajectory, speculative next-iteration front half
int launch_fom_strong(const FomParams &p, int W, cudaStream_t st);

}  // namespace bx
