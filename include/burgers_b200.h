/*
 * burgers_b200.h — C ABI of libburgers_b200.so (sm_100a).
 *
 * Drop-in boundary for ONE hot path of SADPR/1D-Burgers-Equation-ROMs: the implicit
 * linear-FEM time stepping of the 1D inviscid Burgers FOM over many (mu1, mu2) samples and
 * the POD Galerkin/LSPG (and quadratic-manifold) PROM step that reuses the same operators.
 * Each entry point cites the reference interface (file:line under the reference root) it
 * replaces.  INTEGRATION.md shows the ctypes binding a maintainer of the reference adds.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; no torch / C++ types.
 *   - every pointer is a DEVICE pointer (FP64 unless stated) owned by the caller; nothing is
 *     allocated, freed or synchronised inside; work is enqueued on `stream`
 *     (a cudaStream_t passed as void*, NULL = legacy default stream).
 *   - return value: BX_OK (0) or a negative BX_ERR_* code; bx_last_error() gives the text.
 *   - "bands" of a tridiagonal N x N matrix are three length-N arrays dl (sub-diagonal,
 *     dl[0] unused = 0), d (diagonal), du (super-diagonal, du[N-1] unused = 0).  Batched
 *     arrays are sample-major: element (b, i) at [b*N + i].
 *   - the mesh is the reference's structured 2-node mesh T[e] = (e+1, e+2)
 *     (FEM/fom.py:16); X (N nodes, strictly increasing, possibly non-uniform) is the only
 *     mesh input.
 *   - trajectories are written as U[b][step][node], (B, nsteps+1, N); the reference's
 *     per-sample layout (N, nsteps+1) (FEM/fem_burgers.py:650) is its transpose.
 *   - iteration counts: iters[b][step], int32, the number of nonlinear iterations the
 *     reference's while-loop would have executed for that step.
 *   - non-finite states never abort a batch: from the first step of sample b whose state is not
 *     finite (overflowing Picard iteration, singular reduced system — where the reference prints nan
 *     or raises numpy.linalg.LinAlgError, FEM/fem_burgers.py:767), iters[b][step] = -1 and the
 *     snapshots are NaN; the other samples are unaffected.
 *   - a workspace must not be shared by launches that may run concurrently (one per stream).
 */
#ifndef BURGERS_B200_H
#define BURGERS_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define BX_OK                0
#define BX_ERR_INVALID      -1   /* bad argument (size, null pointer, unsupported N)     */
#define BX_ERR_CUDA         -2   /* a CUDA runtime call / launch failed                  */
#define BX_ERR_UNSUPPORTED  -3   /* configuration outside what the kernels cover         */
#define BX_ERR_PROJECTION   -4   /* projection is not BX_GALERKIN / BX_LSPG
                                    (reference: ValueError, FEM/fem_burgers.py:764,1158)  */

#define BX_GALERKIN 0
#define BX_LSPG     1

int         bx_version(void);
const char *bx_last_error(void);
/* number of SMs / device name of the current device (diagnostics for bench.py) */
int         bx_device_sm_count(void);

/* ---- explicit tuning knobs -------------------------------------------------------------------
 * Launch-shape choices are automatic; these knobs force one for A/B measurements and for tests that
 * must reach every instance.  Process-wide, value < 0 (the default) = automatic.  They replace the
 * environment switches of round 1: nothing on the launch path reads the environment.               */
#define BX_TUNE_FOM_WARPS_PER_TRAJECTORY 0  /* 1, 2 or 4 warps integrate one FOM trajectory (N <= 512)   */
#define BX_TUNE_PROM_CTA_WARPS           1  /* 8 or 16 warps per trajectory, POD PROM with n >= 32        */
#define BX_TUNE_PROM_CTA_CHUNK           2  /* 8, 12, 16, 24 or 32 nodes per staged Phi chunk, n >= 32   */
#define BX_TUNE_QM_SCHEDULE              3  /* tangent rows of the quadratic-manifold PROM:
                                               0 global scratch, 1 shared-memory tile, 2 chunked global   */
#define BX_TUNE_PROM_WARPS_PER_TRAJECTORY 4 /* POD PROM with 32 <= n <= 48: 1 = one warp per trajectory (default), 8 = the CTA kernel */
#define BX_TUNE_FOM_KERNEL               5  /* 0 throughput kernel, 1 latency kernel (speculative front half) */
#define BX_TUNE_FOM_SEGMENT_STEPS        6  /* time steps per round-robin work unit of the FOM / POD-PROM / quadratic-manifold sweeps (0: whole trajectories) */
#define BX_TUNE_QM_GROUP                 7  /* 1 .. 8 trajectories per lock-step group of the quadratic-manifold PROM */
#define BX_TUNE_FINE_ROWS                8  /* 4, 8 or 16 mesh rows per thread of the fine-mesh kernels (2048 / 4096 / 8192 nodes per CTA) */
#define BX_TUNE_COUNT                    9
int bx_set_tuning(int knob, int value);
int bx_get_tuning(int knob);

/* ---- FP64 peak probe (roofline denominators; MEASURED_PEAKS.json has no FP64 entry) --------------
 * Launches a saturating stream of independent DFMA (kind BX_PROBE_DFMA) or mma.sync.m8n8k4.f64
 * (BX_PROBE_DMMA) instructions on `stream` (16 warps per SM, `iters` rounds) and returns the number of
 * floating-point operations the launch executes (negative BX_ERR_* on error).  The caller times it
 * with events on the same stream: peak = flops / seconds.  sink: one device double (never written). */
#define BX_PROBE_DFMA 0
#define BX_PROBE_DMMA 1
long long bx_probe_fp64_peak(int kind, int iters, double *sink, void *stream);

/* ---- L1 operators -------------------------------------------------------------------- */
/* FEMBurgers.compute_mass_matrix, FEM/fem_burgers.py:324-353 */
int bx_mass_bands(int N, const double *X, double *dl, double *d, double *du, void *stream);
/* FEMBurgers.compute_diffusion_matrix, FEM/fem_burgers.py:355-387 */
int bx_diffusion_bands(int N, const double *X, double *dl, double *d, double *du, void *stream);
/* FEMBurgers.compute_convection_matrix(U_n), FEM/fem_burgers.py:389-425 ; U (B, N) */
int bx_convection_bands(int B, int N, const double *X, const double *U,
                        double *dl, double *d, double *du, void *stream);
/* FEMBurgers.compute_convection_matrix_derivative(U_n), FEM/fem_burgers.py:463-498 */
int bx_convection_derivative_bands(int N, const double *X, double *dl, double *d, double *du,
                                   void *stream);
/* FEMBurgers.compute_forcing_vector(mu2), FEM/fem_burgers.py:427-461 ; mu2 (B,), F (B, N) */
int bx_forcing_vector(int B, int N, const double *X, const double *mu2, double *F, void *stream);
/* FEMBurgers.compute_supg_term(U_n, mu2), FEM/fem_burgers.py:500-581 ; S (B, N) */
int bx_supg_term(int B, int N, const double *X, const double *U, const double *mu2, double *S,
                 void *stream);

/* ---- one Picard linearisation, bands materialised in HBM ----------------------------- */
/* FEM/fem_burgers.py:667-689 (= :730-752): A = M + At C(Uk) + At E K, row 0 <- e0^T;
 * b = M Un + At F(mu2) - At S(Uk, mu2) (S dropped when supg == 0, as :1141), b[0] = mu1;
 * R = A Uk - b.  Writes dl, d, du, R (and b when b != NULL), all (B, N).                */
int bx_assemble_picard(int B, int N, const double *X, const double *Uk, const double *Un,
                       const double *mu1, const double *mu2, double At, double E, int supg,
                       double *dl, double *d, double *du, double *b, double *R, void *stream);
/* FEM/fem_burgers.py:692-698: delta = A^-1 (-R); Uk1 = Uk + delta;
 * err[b] = |delta|_2 / |Uk1|_2.  One warp per system; N <= 1024.  Uk1 may alias Uk.      */
int bx_tridiag_solve_update(int B, int N, const double *dl, const double *d, const double *du,
                            const double *R, const double *Uk, double *Uk1, double *err,
                            void *stream);

/* ---- whole trajectories (the sweep of FEM/paper_training_stage.py:28-53) ------------- */
/* bytes of zero-initialisable scratch the trajectory kernels need (work-queue counters). */
long long bx_workspace_bytes(void);

/* FEMBurgers.fom_burgers, FEM/fem_burgers.py:646-707, for B samples at once.
 * u0: (N,) shared by all samples when u0_batched == 0, else (B, N).
 * tol, kmax: the literals 1e-6 / 20 of :663.  U: (B, nsteps+1, N).  iters: (B, nsteps) or NULL.
 * err_hist: NULL, or (B, nsteps, kmax): entry [b][step][k] receives |delta_U| / |U1| after iteration k
 * (the value the reference prints as "Iteration: k+1, Error: ..." at :664); entries of iterations that
 * did not run are left untouched.
 * N <= 1024.  Sweeps that fill the device run one warp per trajectory in round-robin segments of
 * BX_TUNE_FOM_SEGMENT_STEPS time steps (csrc/bx_fom.cu); sweeps with at most 8 trajectories per SM run the
 * latency kernel (csrc/bx_fom_strong.cu) with one, two or four warps per trajectory
 * (BX_TUNE_FOM_KERNEL / BX_TUNE_FOM_WARPS_PER_TRAJECTORY force the choice).
 * workspace: bx_fom_workspace_bytes(B) bytes (contents ignored, overwritten).             */
long long bx_fom_workspace_bytes(int B);
/* the launch a sweep of B samples on N nodes gets on the current device: 16 * kernel + warps per trajectory
 * (kernel 0 = throughput kernel, 1 = latency kernel); negative = error code                            */
int bx_fom_plan(int B, int N);
int bx_fom_picard_trajectories(int B, int N, int nsteps, const double *X, const double *u0,
                               int u0_batched, const double *mu1, const double *mu2,
                               double At, double E, double tol, int kmax,
                               double *U, int *iters, double *err_hist, void *workspace, void *stream);

/* No counterpart in the reference.  bx_fom_picard_trajectories for
 * callers that cannot allocate device memory (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc):
 * same arguments, outputs and error behaviour, without the workspace — the scratch is a static device array inside
 * the library (256 + 4 * BX_NOALLOC_MAX_TRAJECTORIES bytes, held by every GPU that loads the library until exit).
 * B > BX_NOALLOC_MAX_TRAJECTORIES is rejected with BX_ERR_UNSUPPORTED.  The call allocates nothing (no cudaMalloc /
 * cudaFree, no allocating library routine).  Every call on a device shares the array: calls that overlap on the GPU
 * (two streams, two host threads) race — issue them one after the other on one stream, or use the workspace form. */
#define BX_NOALLOC_MAX_TRAJECTORIES 65536
int bx_fom_picard_trajectories_noalloc(int B, int N, int nsteps, const double *X, const double *u0,
                                       int u0_batched, const double *mu1, const double *mu2,
                                       double At, double E, double tol, int kmax,
                                       double *U, int *iters, double *err_hist, void *stream);

/* FEMBurgers.fom_burgers, FEM/fem_burgers.py:646-707, on a mesh too large for one warp (N > 1024;
 * BASELINE config 5: 2^20 nodes): one persistent cooperative kernel per call, the N x N tridiagonal solve
 * of every Picard iteration done by a three-level partition / cyclic-reduction method across threads,
 * warps and CTAs (one synchronisation of the grid per iteration, by epoch-tagged words instead of a barrier), 4, 8
 * or 16 mesh rows per thread by mesh size (BX_TUNE_FINE_ROWS), state on chip for the whole trajectory.  The B samples
 * are integrated one after the other.  Same arguments and outputs as bx_fom_picard_trajectories.
 * N <= 8192 * min(SM count, 160).  workspace: bx_fine_workspace_bytes(N) bytes, contents ignored.       */
long long bx_fine_workspace_bytes(int N);
int bx_fom_picard_finemesh(int B, int N, int nsteps, const double *X, const double *u0, int u0_batched,
                           const double *mu1, const double *mu2, double At, double E, double tol, int kmax,
                           double *U, int *iters, void *workspace, void *stream);

/* The same trajectory (FEM/fem_burgers.py:646-707) with the mesh partitioned over R <= 8 GPUs, one process per GPU
 * (meshes up to R x 8192 x SM count nodes).  Rank r integrates the node block
 * [node_offset, node_offset + N_local): node_offset is a multiple of 8192, every block but the last holds a multiple
 * of 8192 nodes, blocks are in rank order and cover the mesh.  Every Picard iteration solves the N x N system by the
 * partition method of bx_fom_picard_finemesh with a fourth level: each rank condenses its block to one interface
 * row, stores it into every rank's exchange buffer through peer memory (NVLink) and solves the R-row closed system
 * redundantly; a second exchange carries the norm partials and the block-edge states.  No host round trip and no
 * collective library inside the launch: all ranks must launch (same B, N, nsteps, tol, kmax, the same sequence of
 * launches on a set of exchange buffers), the kernels wait for each other on the device.
 * X: the WHOLE mesh (N).  u0_local: (N_local,) or (B, N_local).  U_local: (B, nsteps+1, N_local).  iters: (B, nsteps) or
 * NULL, identical on every rank.  workspace: bx_fine_workspace_bytes(N_local) bytes.
 * peer_bufs: host array of R device pointers, [r] = rank r's exchange buffer as mapped in this process ([rank] = own).
 * The exchange buffers carry an epoch counter from launch to launch (2^32 exchanges = 2^31 Picard iterations per
 * set of buffers).  A rank that waits more than 10 s for another one gives up: every rank's launch then ends early
 * with invalid outputs, bx_fine_peer_failed returns 1 (nothing hangs) and the buffers have to be re-created.        */
long long bx_fine_peer_bytes(void);
/* set-up (these four allocate / map device memory; call them outside the time-stepping loop): an exchange buffer of
 * bx_fine_peer_bytes() zeroed bytes from cudaMalloc + its 64-byte cudaIpcMemHandle_t; map / unmap another rank's.  */
int bx_peer_alloc(void **buf, void *ipc_handle_64_bytes);
int bx_peer_open(const void *ipc_handle_64_bytes, void **buf);
int bx_peer_close(void *buf);
int bx_peer_free(void *buf);
int bx_fom_picard_finemesh_ranks(int B, int N, int node_offset, int N_local, int nsteps, const double *X,
                                 const double *u0_local, int u0_batched, const double *mu1, const double *mu2,
                                 double At, double E, double tol, int kmax, double *U_local, int *iters,
                                 void *workspace, int R, int rank, void *const *peer_bufs, void *stream);
/* 1 when a launch on this exchange buffer gave up waiting for another rank, else 0; synchronises the stream */
int bx_fine_peer_failed(const void *own_buf, void *stream);

/* FEMBurgers.pod_prom_burgers, FEM/fem_burgers.py:709-785 (north_star alias pod_prom_newton), for B
 * samples at once.  Phi: (N, n) row-major (C order) reduced basis shared by all samples.
 * projection: BX_GALERKIN (Ar = Phi^T A Phi, br = Phi^T R, :756-757) or BX_LSPG (J = A Phi,
 * Ar = J^T J, br = J^T R, :760-762); anything else returns BX_ERR_PROJECTION (ValueError, :764).
 * The reduced solve is LU with partial pivoting (np.linalg.solve -> LAPACK gesv, :767).
 * tol, kmax: the literals 1e-6 / 20 of :725.  err_hist: NULL or (B, nsteps, kmax), |dq| / |q| after every
 * iteration (the value printed at :727).  1 <= n <= 240 (the reference's stored truncations have
 * 9 ... 227 modes, POD/modes/): n <= 31 runs one warp per trajectory with Phi in shared memory, n >= 32
 * one CTA per trajectory with Phi streamed from L2 and a blocked LU of the reduced system.
 * workspace: bx_prom_workspace_bytes(B, N, n) bytes (work queue, per-sample forcing data and, for
 * large n, one reduced system per resident CTA; contents ignored, overwritten).                  */
long long bx_prom_workspace_bytes(int B, int N, int n);
int bx_prom_trajectories(int B, int N, int n, int nsteps, const double *X, const double *u0,
                         int u0_batched, const double *mu1, const double *mu2, double At, double E,
                         const double *Phi, int projection, double tol, int kmax,
                         double *U, int *iters, double *err_hist, void *workspace, void *stream);

/* Decoder and tangent of the quadratic manifold for B reduced states (FEM/fem_burgers.py:263-312, 1116-1118, 1148):
 * u = Phi q + H sym(q x q) -> u (B, N);  T = Phi + H dQ/dq -> T (B, N, n) (NULL: decoder only).  q: (B, n).
 * n <= 128.  The stepwise quadratic-manifold path for n > 31 (bx_qm_prom_trajectories covers n <= 31) is built
 * from this, bx_reduced_system (per-sample tangent, supg = 0) and a batched dense solve.                   */
int bx_qm_decode(int B, int N, int n, const double *Phi, const double *H, const double *q, double *u, double *T,
                 void *stream);

/* The reduced system of ONE nonlinear iteration for a caller-supplied tangent: the block every
 * PROM variant of the reference repeats with its own decoder tangent in the place of Phi — pod_prom_burgers
 * FEM/fem_burgers.py:730-762, local_prom_burgers :1022-1058, pod_ann_prom :1225-1262, pod_rbf_prom :1330-1375,
 * lie_prom, ae_prom.  Native precedent with the same in-place-output shape: compute_Ar_br(A, Phi, R, Ar, br,
 * echo_level), Stanford_2D/Burgers_2D/cpp_files/mkl_sparse_dense_operations.cpp:13-20.
 *   A = M + At C(Uk) + At E K (row 0 <- e0^T);  b = M Un + At F(mu2) - At S(Uk, mu2) [supg != 0], b[0] = mu1;
 *   R = A Uk - b;   Galerkin: Ar = T^T A T, br = T^T R;   LSPG: Ar = (A T)^T (A T), br = (A T)^T R.
 * Uk, Un: (B, N).  T: (N, n) row-major, shared by all samples (T_stride = 0) or one per sample, T_stride doubles
 * apart (>= N n).  Ar: (B, n, n) row-major, br: (B, n).  n <= 240, N <= 2048.  No workspace.               */
int bx_reduced_system(int B, int N, int n, const double *X, const double *Uk, const double *Un,
                      const double *mu1, const double *mu2, double At, double E, int supg,
                      const double *T, long long T_stride, int projection, double *Ar, double *br, void *stream);

/* FEMBurgers.pod_quadratic_manifold, FEM/fem_burgers.py:1081-1175, for B samples at once.
 * Decoder u = Phi q + H sym(q x q), tangent Phi + H dQ/dq (get_single_Q / get_dQ_dq, :275-312; monomial
 * order i outer, j >= i inner).  Phi: (N, n), H: (N, n(n+1)/2), both row-major, shared by all samples.
 * uxa: (B,) left Dirichlet value per sample (the reference's scalar `uxa`; the drivers pass mu1).
 * No SUPG term (:1141); q accumulates over the Gauss-Newton iterations and is re-projected from the full
 * state at every time step (:1129); stops when |dq| / max(1e-14, |q|) < tol (strict, :1165-1169) or after
 * itmax iterations; iters[b][step] = iterations executed.  projection: BX_GALERKIN / BX_LSPG, anything
 * else BX_ERR_PROJECTION (ValueError, :1158).  err_hist: NULL or (B, nsteps, itmax), the value printed at :1166.
 * 1 <= n <= 31 in this build.
 * workspace: bx_qm_workspace_bytes(B, N, n) bytes, 16-byte aligned, contents ignored.               */
long long bx_qm_workspace_bytes(int B, int N, int n);
int bx_qm_prom_trajectories(int B, int N, int n, int nsteps, const double *X, const double *u0,
                            int u0_batched, const double *uxa, const double *mu2, double At, double E,
                            const double *Phi, const double *H, int projection, double tol, int itmax,
                            double *U, int *iters, double *err_hist, void *workspace, void *stream);
/* 1 if tangent-row schedule `schedule` (BX_TUNE_QM_SCHEDULE) can run at (N, n) on the current device */
int bx_qm_schedule_available(int N, int n, int schedule);

/* FEMBurgers.fom_burgers_newton AS CODED, FEM/fem_burgers.py:584-643, for B samples at once.
 * At :616 `At * dC_dU @ U0` is a matrix-vector product whose result numpy broadcasts onto every row:
 * J = T + 1 w^T, T = M + At E K + At C(U_k), w = At dC_dU U_k, row 0 <- e0^T (:619-620).  Solved exactly
 * with one tridiagonal factorisation and two right-hand sides (Sherman-Morrison).  R has no SUPG term
 * (:623-626).  U_{k+1} = U_k + lambda delta; lambda, tol, kmax: the literals 0.1 / 1e-6 / 100 of
 * :597-598, :606.  N <= 1024.                                                                */
int bx_fom_newton_trajectories(int B, int N, int nsteps, const double *X, const double *u0,
                               int u0_batched, const double *mu1, const double *mu2,
                               double At, double E, double lambda, double tol, int kmax,
                               double *U, int *iters, void *workspace, void *stream);

/* ---- POD basis: snapshot Gram matrix (replaces np.linalg.svd of POD/pod.py:80-84) ------------ */
/* G (+)= S S^T with S given as `cols` snapshots of N nodes, one snapshot per row -- i.e. the trajectory
 * buffer U (B, nsteps+1, N) itself with cols = B*(nsteps+1).  G: (N, N) row-major, both triangles
 * written.  accumulate != 0 adds to the G already there (several trajectory buffers, one G per GPU);
 * the sum over GPUs is the caller's all-reduce (ncclAllReduce on N*N doubles).  FP64 tensor pipe (DMMA),
 * fixed summation order (bit-reproducible for a given N, cols and SM count).
 * workspace: bx_gram_workspace_bytes(N, cols) bytes, contents ignored.                                */
long long bx_gram_workspace_bytes(int N, long long cols);
int bx_gram_accumulate(int N, long long cols, const double *S, double *G, int accumulate,
                       void *workspace, void *stream);

/* ---- on-disk snapshot format of the sweep drivers (FEM/paper_training_stage.py:52-53) ------------------ */
/* Trajectory buffer U (B, rows, N) (one snapshot per row, as the trajectory entry points write it) ->
 * Uref (B, N, rows): per sample the array FEMBurgers.fom_burgers returns, (N, nTimeSteps+1) C order
 * (FEM/fem_burgers.py:650, 705), i.e. the payload of one `fem_simulation_mu1_*_mu2_*.npy` file that
 * POD/pod.py:80-82 loads and hstacks.  Batched tiled transpose; U and Uref must not alias; B <= 65535.   */
int bx_snapshots_to_reference_layout(int B, int rows, int N, const double *U, double *Uref, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* BURGERS_B200_H */
