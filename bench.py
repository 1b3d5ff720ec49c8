#!/usr/bin/env python3
"""bench.py — FOM time-steps/s on BASELINE.json config 2 (64 x 64 (mu1, mu2) grid = 4096 samples, 512 nodes,
dt = 0.05, 500 steps), STRONG scaling: the 4096 samples are sharded over the N GPUs (the loop being batched is
FEM/paper_training_stage.py:28-53; north_star: "samples are sharded across the 8 GPUs").

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
  python bench.py --impl reference [...]                         # reference algorithm on host cores

A "step" is one full sweep: 4096 samples x 500 time steps through the fused trajectory kernel (each rank its
shard).  One JSON line on stdout (rank 0).  Beside the headline the line carries
  * `weak`             the round-1 figure (every GPU integrates its own 4096-sample grid), N > 1 only;
  * `e2e`              the same strong sweep with HOST parameters in and HOST snapshots out (+ the raw pinned D2H floor);
  * `roofline`         bound = FP64 pipe: executed FP64 flop/s of the FOM kernel against the DFMA peak measured
                       in this run (bx_probe_fp64_peak); the 88*N byte model of SURVEY 8d as `model_frac`;
  * `extra_workloads`  configs 3, 4 and 5a under the same clock: snapshot Gram + all-reduce, POD-LSPG / Galerkin
                       PROM with 28 modes, quadratic-manifold builder + Galerkin PROM with 28 modes, fine mesh (one
                       2^20-node trajectory per GPU, and at N > 1 the same trajectory PARTITIONED over the N ranks);
  * `cpu_baseline`     the oracle port on one host core over a bounded sample (rank 0, N = 1).
See DESIGN.md "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "FOM time-steps/sec (512 nodes, batched mu)"
UNIT = "time-steps/s"
N_NODES, NSTEPS, AT, E = 512, 500, 0.05, 0.0
GRID = (64, 64)
N_MODES = 28


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--samples", type=int, default=GRID[0] * GRID[1], help="samples of the whole job (default 4096 = config 2)")
    ap.add_argument("--nsteps", type=int, default=NSTEPS, help="time steps per trajectory")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the config 3 / 4 / 5a workloads")
    ap.add_argument("--no-weak", action="store_true")
    ap.add_argument("--cpu-samples", type=int, default=16, help="trajectories in the bounded CPU sample")
    return ap.parse_args()


def config2_grid(samples):
    """The 64 x 64 grid, mu1-major (FEM/paper_training_stage.py:8-10); `samples` != 4096 keeps 64 mu2 columns."""
    n2 = GRID[1]
    n1 = (samples + n2 - 1) // n2
    mu1 = np.linspace(4.25, 5.5, n1)
    mu2 = np.linspace(0.015, 0.03, n2)
    return np.repeat(mu1, n2)[:samples].copy(), np.tile(mu2, n1)[:samples].copy()


def weak_grid_for_rank(samples, rank, world):
    """Round-1 weak-scaling workload: global grid (64*world) x 64, rank r takes the r-th block of `samples` points."""
    m1, m2 = config2_grid(samples * world)
    return m1[rank * samples:(rank + 1) * samples].copy(), m2[rank * samples:(rank + 1) * samples].copy()


def config_dict(samples, nsteps, world):
    return {"workload": f"configs[1]: batched FOM snapshot generation, {GRID[0]}x{GRID[1]} (mu1 in [4.25,5.5], "
                        f"mu2 in [0.015,0.03]) grid = {samples} samples in total, {N_NODES} nodes on [0,100], dt={AT}, "
                        f"{nsteps} steps, u0=1, E=0 (fom_burgers, Picard tol 1e-6 / cap 20)",
            "samples_total": samples, "samples_per_gpu": [int(b - a) for a, b in _shards(samples, world)],
            "nodes": N_NODES, "time_steps": nsteps, "dt": AT,
            "sharding": f"strong: the {samples} samples in contiguous blocks over {world} rank(s), no collective on the data path",
            "l2": "outputs (8.4 GB of snapshots per sweep over all ranks) exceed L2; inputs are 64 KB of parameters"}


def _shards(samples, world):
    out = []
    base, rem = divmod(samples, world)
    lo = 0
    for r in range(world):
        hi = lo + base + (1 if r < rem else 0)
        out.append((lo, hi))
        lo = hi
    return out


# ----------------------------------------------------------------------------------------------
# clocks sampling (nvidia-smi during the timed region)
# ----------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_capture(variant):
    """Static numbers of the committed `ncu --set full` capture of the FOM kernel instance `variant`
    (profiles/roofline_traffic.json): DRAM bytes per launch, FP64 warp instructions per Picard iteration, FP64-pipe
    utilisation — each tagged with the configuration and commit it was captured on."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("kernels", {}).get(variant)
        except Exception:
            return None
    return None


# ----------------------------------------------------------------------------------------------
# CPU legs (oracle port; the only places bench.py executes oracle/)
# ----------------------------------------------------------------------------------------------
def _cpu_worker(args):
    os.environ["OMP_NUM_THREADS"] = "1"
    from oracle import burgers_oracle as O
    mu1, mu2, nsteps = args
    X, _ = O.mesh(N_NODES)
    t0 = time.perf_counter()
    U, it = O.fom_burgers(X, AT, nsteps, np.ones(N_NODES), mu1, E, mu2)
    return time.perf_counter() - t0, int(it.sum())


def cpu_sample(n_traj, nsteps, procs):
    """n_traj trajectories spread evenly over the config-2 grid, on `procs` host processes."""
    import multiprocessing as mp
    m1, m2 = config2_grid(GRID[0] * GRID[1])
    sel = np.linspace(0, m1.size - 1, n_traj).astype(int)
    jobs = [(float(m1[i]), float(m2[i]), nsteps) for i in sel]
    t0 = time.perf_counter()
    if procs == 1:
        res = [_cpu_worker(j) for j in jobs]
    else:
        with mp.get_context("fork").Pool(procs) as pool:
            res = pool.map(_cpu_worker, jobs, chunksize=1)
    wall = time.perf_counter() - t0
    return wall, n_traj * nsteps / wall, sum(r[1] for r in res)


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's algorithm (oracle port, numpy + LAPACK gtsv, pinned to the unmodified
    reference by tests/test_oracle_golden.py) on all host cores; one "step" = a bounded sample of config 2 (one
    trajectory per core).  The scipy reference itself is ~640 x slower per core and does not travel."""
    if rank != 0:
        return
    procs = os.cpu_count() or 1
    n_traj = procs
    for _ in range(args.warmup):
        cpu_sample(min(n_traj, procs), 10, procs)
    walls, its = [], 0
    for _ in range(args.steps):
        w, _, it = cpu_sample(n_traj, args.nsteps, procs)
        walls.append(w); its += it
    wall = float(np.sum(walls))
    value = n_traj * args.nsteps * args.steps / wall
    sample = (f"{n_traj} trajectories per step spread evenly over the 64x64 grid, {args.nsteps} time steps each, "
              f"one process per core ({procs}); numpy restatement of fom_burgers pinned to the reference")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(args.samples, args.nsteps, world),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "picard_iterations": its}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# the CUDA arm
# ----------------------------------------------------------------------------------------------
class Bench:
    def __init__(self, args):
        import torch
        import torch.distributed as dist
        import burgers_b200
        self.torch, self.dist, self.bx = torch, dist, burgers_b200
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.fem = burgers_b200.FEMBurgers(np.linspace(0.0, 100.0, N_NODES))
        self.u0 = torch.ones(N_NODES, dtype=torch.float64, device=self.dev)
        self.launches = 0

    # -- helpers ---------------------------------------------------------------------------------
    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, x, op="max"):
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return float(t.item())

    def timed(self, fn, steps, warmup):
        """`warmup` untimed + `steps` timed calls of fn between barriers; CUDA events on the launching stream.
        Returns (total ms = max over ranks, per-call ms of this rank)."""
        torch = self.torch
        for _ in range(warmup):
            fn()
        self.barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        ev[0].record()
        for k in range(steps):
            fn()
            ev[k + 1].record()
        self.barrier()
        per = [ev[k].elapsed_time(ev[k + 1]) for k in range(steps)]
        return self.reduce(ev[0].elapsed_time(ev[steps])), per

    def fp64_peak(self, kind):
        """DFMA / DMMA peak of this GPU, measured now (TFLOP/s): bx_probe_fp64_peak timed with events."""
        torch = self.torch
        lib = self.bx._lib.load()
        sink = torch.zeros(8, dtype=torch.float64, device=self.dev)
        st = torch.cuda.current_stream().cuda_stream
        best = 0.0
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            flops = lib.bx_probe_fp64_peak(kind, 4000, sink.data_ptr(), st)
            e1.record()
            torch.cuda.synchronize()
            if flops < 0:
                raise RuntimeError("bx_probe_fp64_peak failed: " + lib.bx_last_error().decode())
            if rep:
                best = max(best, flops / (e0.elapsed_time(e1) * 1e-3) / 1e12)
        return best

    # -- headline: config 2, strong ---------------------------------------------------------------
    def run(self):
        torch, args = self.torch, self.args
        world, rank, dev, fem = self.world, self.rank, self.dev, self.fem
        nT = args.nsteps
        g1, g2 = config2_grid(args.samples)
        lo, hi = _shards(args.samples, world)[rank]
        B = hi - lo
        mu1 = torch.from_numpy(g1[lo:hi].copy()).to(dev); mu2 = torch.from_numpy(g2[lo:hi].copy()).to(dev)
        cap = args.samples if (world > 1 and not args.no_weak) else B     # the weak leg reuses the buffers
        Ubuf = torch.empty((cap, nT + 1, N_NODES), dtype=torch.float64, device=dev)
        itbuf = torch.empty((cap, nT), dtype=torch.int32, device=dev)
        U, iters = Ubuf[:B], itbuf[:B]

        def step():
            fem.fom_burgers_batch(AT, nT, self.u0, mu1, E, mu2, out=U, iters_out=iters)

        warm = max(args.warmup, 3)
        sampler = ClockSampler(self.local_rank)
        for _ in range(warm):
            step()
        self.barrier()
        if rank == 0:
            sampler.start()
        total_ms, per_launch_ms = self.timed(step, args.steps, 0)
        clocks = sampler.stop() if rank == 0 else None
        self.launches += args.steps
        total_iters = int(iters.sum().item())
        iters_all = self.reduce(total_iters, "sum")
        value = args.samples * nT * args.steps / (total_ms * 1e-3)
        lib = self.bx._lib.load()
        kern_used, W_used = self.fom_variant(B)

        # ---- weak leg (round-1 definition), N > 1 only --------------------------------------------
        weak = None
        if world > 1 and not args.no_weak:
            w1, w2 = weak_grid_for_rank(args.samples, rank, world)
            wm1 = torch.from_numpy(w1).to(dev); wm2 = torch.from_numpy(w2).to(dev)
            ksteps = min(args.steps, 5)
            wms, _ = self.timed(lambda: fem.fom_burgers_batch(AT, nT, self.u0, wm1, E, wm2, out=Ubuf, iters_out=itbuf),
                                ksteps, 1)
            self.launches += ksteps
            weak = {"value": world * args.samples * nT * ksteps / (wms * 1e-3), "unit": UNIT, "ms_per_step": wms / ksteps,
                    "samples_per_gpu": args.samples, "steps": ksteps,
                    "note": f"every rank integrates its own {args.samples}-sample block of a ({GRID[0]}*{world}) x {GRID[1]} grid"}

        # ---- end-to-end leg: host parameters in, host snapshots out, through the public sweep API ----
        e2e = None
        if not args.no_e2e:
            e2e = self.e2e_leg(g1[lo:hi], g2[lo:hi], B, nT)

        # ---- FP64 peaks of this GPU, measured in this run ---------------------------------------------
        dfma_peak = self.fp64_peak(self.bx._lib.BX_PROBE_DFMA)
        dmma_peak = self.fp64_peak(self.bx._lib.BX_PROBE_DMMA)

        # ---- configs 3, 4, 5a under the same clock ---------------------------------------------------
        extra = None
        if not args.no_extra:
            extra = self.extra_workloads(Ubuf, itbuf, mu1, mu2, B, nT, dmma_peak, dfma_peak)

        if rank != 0:
            if world > 1:
                self.dist.destroy_process_group()
            return

        # ---- roofline of the dominant (only) kernel of the headline --------------------------------------
        hbm_peak, hbm_src = measured_peak_hbm()
        k_ms = float(np.mean(per_launch_ms))
        variant = f"fom_{kern_used}_w{W_used}"
        cap_ = ncu_capture(variant) or {}
        instr = cap_.get("fp64_warp_instructions_per_picard_iteration")
        model_bytes = 88.0 * N_NODES * total_iters + 8.0 * N_NODES * B * nT     # SURVEY 8(d), this rank's launch
        model_gbs = model_bytes / (k_ms * 1e-3) / 1e9
        achieved = (instr * 64.0 * total_iters / (k_ms * 1e-3) / 1e12) if instr else None
        same_cfg = cap_.get("samples") == B and cap_.get("nsteps") == nT
        roofline = {
            "bound": "fp64", "unit": "TFLOP/s", "achieved": achieved, "peak": dfma_peak,
            "frac": (achieved / dfma_peak) if achieved else None,
            "peak_source": "DFMA issue peak of this GPU measured in this run (bx_probe_fp64_peak, 16 warps/SM of independent chains); "
                           "MEASURED_PEAKS.json has no FP64 entry",
            "achieved_source": (f"{instr} FP64 warp instructions per Picard iteration (ncu sm__inst_executed_pipe_fp64 of the committed "
                                f"capture {cap_.get('capture')}) x 64 flop x {total_iters} iterations of this launch / kernel time")
            if instr else "no committed ncu capture for this kernel instance",
            "traffic": cap_.get("dram_bytes_per_launch") if same_cfg else None,
            "traffic_source": (f"ncu --set full dram__bytes_read.sum + dram__bytes_write.sum, captured on {cap_.get('capture')}"
                               if same_cfg else "null: the committed capture is for another samples/nsteps"),
            "fp64_pipe_active_pct_ncu": cap_.get("fp64_pipe_active_pct") if same_cfg else None,
            "kernel": (f"bx::fom_picard_kernel (throughput instance, one warp per trajectory, round-robin segments)"
                       if kern_used == "throughput" else f"bx::fom_strong_kernel (latency instance, {W_used} warp(s) per trajectory)"),
            "kernel_ms": k_ms,
            "picard_iterations_per_launch": total_iters,
            "model_frac": model_gbs / hbm_peak, "model_gbs": model_gbs, "model_peak_gbs": hbm_peak, "model_peak_source": hbm_src,
            "model": "banded-dataflow equivalent of SURVEY 8d: 88*N B per Picard iteration + 8*N B per time step per sample; the "
                     "fused kernel keeps the bands in registers, so this is NOT a DRAM throughput (see traffic) and exceeds 1",
            "real_snapshot_bytes_per_launch": 8.0 * N_NODES * B * (nT + 1) + 4.0 * B * nT}

        cpu = None
        if not args.no_cpu_baseline and world == 1:
            wall, v, its = cpu_sample(args.cpu_samples, nT, 1)
            cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
                   "sample": f"{args.cpu_samples} of the 4096 config-2 trajectories (evenly spaced), {nT} steps each, "
                             f"{its} Picard iterations, {wall:.1f} s on one core (numpy + LAPACK gtsv oracle, pinned to the "
                             f"unmodified reference; the scipy reference itself runs ~640x slower per core, SURVEY 6)",
                   "host_cores_available": os.cpu_count()}

        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": warm, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": config_dict(args.samples, nT, world),
                "trajectories_per_s": args.samples * args.steps / (total_ms * 1e-3),
                "picard_iterations_per_s": iters_all * args.steps / (total_ms * 1e-3),
                "strong": {"value": value, "unit": UNIT, "ms_per_step": total_ms / args.steps, "samples_total": args.samples,
                           "kernel": kern_used, "warps_per_trajectory": W_used},
                "weak": weak, "roofline": roofline,
                "fp64_peaks_measured": {"dfma_tflops": dfma_peak, "dmma_tflops": dmma_peak},
                "extra_workloads": extra, "cpu_baseline": cpu, "e2e": e2e, "clocks": clocks,
                "gpu_launches": self.launches, "gpu": torch.cuda.get_device_name(dev)}
        print(json.dumps(line), flush=True)
        if world > 1:
            self.dist.destroy_process_group()

    def fom_variant(self, B):
        """(kernel, warps per trajectory) the library picks for B samples of 512 nodes: bx_fom_plan."""
        plan = self.bx._lib.load().bx_fom_plan(int(B), N_NODES)
        if plan < 0:
            raise RuntimeError("bx_fom_plan failed")
        return ("latency" if plan >= 16 else "throughput"), plan & 15

    # -- e2e ------------------------------------------------------------------------------------------
    def e2e_leg(self, m1, m2, B, nT):
        torch, args, fem = self.torch, self.args, self.fem
        from burgers_b200 import sweep as bsweep
        chunk = min(512, max(64, -(-B // 4)))     # D2H of a chunk hides the next chunk's kernel (tools/e2e_time.py)
        chunk = min(chunk, B)
        bufs = bsweep.HostSweepBuffers(fem, chunk, nT)
        h1 = torch.from_numpy(m1.copy()).pin_memory(); h2 = torch.from_numpy(m2.copy()).pin_memory()
        check = []

        def consumer(lo, hi, Uh, Ih):
            check.append(float(Uh[0, -1, -1]))

        for _ in range(2):
            bsweep.fom_sweep_to_host(fem, AT, nT, self.u0, h1, h2, E, chunk, consumer, bufs)
        self.barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            bsweep.fom_sweep_to_host(fem, AT, nT, self.u0, h1, h2, E, chunk, consumer, bufs)
        torch.cuda.synchronize()
        dt = self.reduce(time.perf_counter() - t0)
        nchunks = len(bsweep.chunk_schedule(B, chunk))
        self.launches += (2 + args.steps) * nchunks
        # raw floor of the same device->host traffic: every rank copies its snapshot bytes into the same pinned
        # buffers at the same time, no kernel (what the box's host path gives N concurrent ranks)
        self.barrier()
        t0 = time.perf_counter()
        reps = 2
        for _ in range(reps):
            left = B
            i = 0
            while left > 0:
                n = min(chunk, left)
                bufs.hU[i & 1][:n].copy_(bufs.dU[i & 1][:n], non_blocking=True)
                left -= n; i += 1
        torch.cuda.synchronize()
        floor_dt = self.reduce(time.perf_counter() - t0) / reps
        d2h = int(bufs.d2h_bytes)
        e2e = {"value": args.samples * nT * args.steps / dt, "unit": UNIT,
               "h2d_bytes_per_step": int(bufs.h2d_bytes), "d2h_bytes_per_step": d2h,
               "ms_per_step": 1e3 * dt / args.steps, "chunk_samples": chunk,
               "d2h_floor_ms": 1e3 * floor_dt, "d2h_floor_gbs_per_rank": d2h / floor_dt / 1e9,
               "d2h_floor_gbs_all_ranks": self.world * d2h / floor_dt / 1e9,
               "note": "pinned host params -> H2D -> fused kernel -> D2H of all snapshots + iteration counts, double-buffered "
                       "(burgers_b200.sweep.fom_sweep_to_host); bytes are per rank; the whole job's 4096 samples over all ranks; "
                       "d2h_floor_* = the same snapshot bytes copied by all ranks at once with no kernel (the host path's own limit)"}
        del bufs
        return e2e

    # -- configs 3, 4, 5a -------------------------------------------------------------------------------
    def extra_workloads(self, Ubuf, itbuf, mu1, mu2, B, nT, dmma_peak, dfma_peak):
        """Every other config of BASELINE.json under the same clock, sharded like the headline (strong).  Times are
        max over ranks; iteration totals are summed over ranks; TFLOP/s are ALGORITHMIC flops of the
        whole job / time / N GPUs, `frac` against this run's measured DMMA peak per GPU."""
        torch, fem, dev, world = self.torch, self.fem, self.dev, self.world
        from burgers_b200 import pod, quadratic
        out = {}
        U, iters = Ubuf[:B], itbuf[:B]
        N, n, k = N_NODES, N_MODES, N_MODES * (N_MODES + 1) // 2
        total = self.args.samples

        def job(name, fn, its_fn, flops_per_it, steps, warmup, peak, note):
            ms, _ = self.timed(fn, steps, warmup)
            its = self.reduce(its_fn(), "sum")
            ms1 = ms / steps
            tf = flops_per_it * its / (ms1 * 1e-3) / 1e12 / world
            out[name] = {"ms": ms1, "iterations": its, "iterations_per_s": its / (ms1 * 1e-3),
                         "time_steps_per_s": total * nT / (ms1 * 1e-3), "algorithmic_tflops_per_gpu": tf,
                         "frac_of_measured_peak": tf / peak if peak else None, "steps": steps, "note": note}

        # config 3 (a): snapshot Gram of this rank's FOM snapshots + the all-reduce + eigh -> Phi (28 modes)
        fem.fom_burgers_batch(AT, nT, self.u0, mu1, E, mu2, out=U, iters_out=iters)
        acc = pod.GramAccumulator(N, dev)

        def gram():
            acc.G.zero_(); acc.cols = 0
            acc.add(U)

        ms, _ = self.timed(gram, 3, 1)
        cols = U.shape[0] * U.shape[1]
        cols_all = self.reduce(cols, "sum")
        exe = N * (N + 128.0) * cols_all
        out["config3_gram"] = {"ms": ms / 3, "snapshots": cols_all,
                               "executed_tflops_per_gpu": exe / (ms / 3 * 1e-3) / 1e12 / world,
                               "frac_of_measured_peak": exe / (ms / 3 * 1e-3) / 1e12 / world / dmma_peak,
                               "algorithmic_flops": 2.0 * N * N * cols_all,
                               "note": "bx_gram_accumulate (DMMA SYRK, upper block triangle: N(N+128) cols executed flops)"}
        ms, _ = self.timed(lambda: pod.allreduce_gram(acc.G), 3, 1)
        out["config3_allreduce"] = {"ms": ms / 3, "bytes": N * N * 8, "note": "NCCL all-reduce(sum) of the Gram; no-op at 1 GPU"}
        gram(); pod.allreduce_gram(acc.G)
        pod.basis_from_gram(acc.G, n)      # untimed first call: cuSOLVER handle / workspace creation
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        Phi, s = pod.basis_from_gram(acc.G, n)
        e1.record(); torch.cuda.synchronize()
        out["config3_eigh"] = {"ms": e0.elapsed_time(e1), "note": "torch.linalg.eigh (cuSOLVER) on 512 x 512, every rank"}
        Phi = Phi.contiguous()

        # manifold for config 4 from the reference's 3 x 3 training grid (FEM/paper_training_stage.py:8-10), as
        # Quadratic_manifold/build_quadratic_manifold.py does (alpha = 1e-2), on this GPU
        from burgers_b200 import sweep as bsweep
        t1, t2 = bsweep.parameter_grid(3, 3)
        Ut, _ = fem.fom_burgers_batch(AT, nT, self.u0, t1, E, t2)
        e0.record()
        H = quadratic.build_manifold(Ut, Phi, alpha=1e-2)
        e1.record(); torch.cuda.synchronize()
        out["config4_build_H"] = {"ms": e0.elapsed_time(e1), "snapshots": Ut.shape[0] * Ut.shape[1], "monomials": k,
                                  "note": "burgers_b200.quadratic.build_manifold on the 3x3 training grid (cuSOLVER/cuBLAS library path)"}
        del Ut

        # config 3 (b): POD PROM sweeps with 28 modes
        prom_flops = 2.0 * N * n * n + 12.0 * N * n + (2.0 / 3.0) * n ** 3
        for proj in ("LSPG", "Galerkin"):
            job(f"config3_prom_{proj.lower()}28",
                lambda: fem.pod_prom_burgers_batch(AT, nT, self.u0, mu1, E, mu2, Phi, proj, out=U, iters_out=iters),
                lambda: int(iters.sum().item()), prom_flops, 3, 1, dmma_peak,
                f"pod_prom_burgers {proj}, 28 modes from this run's Gram basis, {total} samples x {nT} steps; flops/iteration = 2Nn^2+12Nn+2n^3/3")
            self.launches += 4 * 2
        # config 4: quadratic-manifold PROM, Galerkin, 28 modes
        qm_flops = prom_flops + 4.0 * N * k + 2.0 * N * (n + k)
        qsteps = min(nT, 20)
        fem.pod_quadratic_manifold_batch(AT, qsteps, self.u0, mu1, E, mu2, Phi, H, "Galerkin")   # warm-up pass
        job("config4_qm_galerkin28",
            lambda: fem.pod_quadratic_manifold_batch(AT, nT, self.u0, mu1, E, mu2, Phi, H, "Galerkin", out=U, iters_out=iters),
            lambda: int(iters.clamp(min=0).sum().item()), qm_flops, 1, 0, dmma_peak,
            f"pod_quadratic_manifold Galerkin, 28 modes / 406 monomials, {total} samples x {nT} steps (one timed pass after a "
            f"{qsteps}-step warm-up pass); flops/iteration = PROM + 4Nk + 2N(n+k)")
        out["config4_qm_galerkin28"]["flagged_nonfinite_steps"] = self.reduce(int((iters < 0).sum().item()), "sum")
        self.launches += 2 * 2

        # config 5a: one trajectory on 2^20 nodes at constant CFL; every rank integrates its own replica (max over ranks)
        NF = 1 << 20
        Xf = np.linspace(0.0, 100.0, NF)
        femf = self.bx.FEMBurgers(Xf)
        Atf = AT * 511.0 / (NF - 1)
        Uf = torch.empty((1, nT + 1, NF), dtype=torch.float64, device=dev)
        itf = torch.empty((1, nT), dtype=torch.int32, device=dev)
        u0f = torch.ones(NF, dtype=torch.float64, device=dev)
        ms, _ = self.timed(lambda: femf.fom_burgers_batch(Atf, nT, u0f, [4.75], E, [0.02], out=Uf, iters_out=itf), 3, 1)
        its = int(itf.sum().item())
        ms1 = ms / 3
        hbm_peak, _ = measured_peak_hbm()
        out["config5a_finemesh"] = {"ms": ms1, "iterations": its, "us_per_iteration": 1e3 * ms1 / max(its, 1),
                                    "time_steps_per_s": nT / (ms1 * 1e-3),
                                    "model_gbs": 88.0 * NF * its / (ms1 * 1e-3) / 1e9,
                                    "model_frac_of_hbm": 88.0 * NF * its / (ms1 * 1e-3) / 1e9 / hbm_peak,
                                    "note": "2^20 nodes, one trajectory per GPU, constant CFL At = 0.05*511/(N-1) (5a); 88*N B per iteration is the "
                                            "banded-dataflow model, the kernel keeps state on chip (latency chain, see DESIGN K-Fine)"}
        self.launches += 4
        del Uf, femf
        # the same trajectory PARTITIONED over the ranks: interface rows through peer memory
        if 1 < world <= 8:
            out["config5a_finemesh_ranks"] = self.finemesh_ranks(Xf, Atf, nT, its)
        out["peaks"] = {"dmma_tflops": dmma_peak, "dfma_tflops": dfma_peak}
        return out

    def finemesh_ranks(self, Xf, Atf, nT, its_single):
        """One 2^20-node trajectory over all ranks (bx_fom_picard_finemesh_ranks); a failure on any rank is reported by all."""
        torch, dev = self.torch, self.dev
        from burgers_b200.finemesh_dist import PeerExchange, fom_burgers_partitioned, node_partition
        ex = None
        try:
            ex = PeerExchange.create()
            off, cnt = node_partition(Xf.size, self.world)[self.rank]
            outb = torch.empty((1, nT + 1, cnt), dtype=torch.float64, device=dev)
            u0l = np.ones(cnt)
            best = None
            for rep in range(3):   # the first one is the warm-up
                _, it = fom_burgers_partitioned(Xf, Atf, nT, u0l, 4.75, E, 0.02, ex, out=outb)
                ms = self.reduce(ex.last_kernel_ms)
                if rep > 0:
                    best = ms if best is None else min(best, ms)
            self.launches += 3
            its = int(it.sum().item())
            return {"ms": best, "iterations": its, "us_per_iteration": 1e3 * best / max(its, 1), "ranks": self.world,
                    "nodes_per_rank": [c for _, c in node_partition(Xf.size, self.world)],
                    "iterations_equal_single_gpu": its == its_single, "finite": bool(torch.isfinite(outb).all().item()),
                    "note": "kernel time by CUDA events, max over ranks, best of 2 after a warm-up launch; every Picard iteration "
                            "exchanges one interface row per rank through NVLink peer memory (no collective call inside the launch)"}
        except Exception as e:   # noqa: BLE001 - the bench line must survive a box without IPC / peer access
            return {"error": f"{type(e).__name__}: {e}"[:300]}
        finally:
            if ex is not None:
                ex.close()


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    Bench(args).run()


if __name__ == "__main__":
    main()
