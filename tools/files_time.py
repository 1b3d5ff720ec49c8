"""Time the on-disk side of the sweep on cuda:0: python tools/files_time.py [B] [steps] [dir]
(1) the device transpose (B, steps+1, N) -> (B, N, steps+1) against the HBM roofline,
(2) sweep_to_npy end to end: FOM sweep + transpose + D2H + one .npy per sample."""
import os, sys, time, tempfile, shutil, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import burgers_b200
from burgers_b200 import sweep

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 500
root = sys.argv[3] if len(sys.argv) > 3 else tempfile.mkdtemp(prefix="bx_files_")
fem = burgers_b200.FEMBurgers(np.linspace(0, 100, 512))
g = int(round(B ** 0.5))
mu1, mu2 = sweep.parameter_grid(g, B // g)
U, _ = fem.fom_burgers_batch(0.05, steps, np.ones(512), mu1, 0.0, mu2)
Ur = torch.empty((U.shape[0], 512, steps + 1), dtype=torch.float64, device="cuda")
peak = None
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
for r in range(4):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); sweep.to_reference_layout(fem, U, out=Ur); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    gbs = 2 * U.numel() * 8 / ms / 1e6
    print(f"transpose B={U.shape[0]} x {steps + 1} x 512: {ms:.3f} ms, {gbs:.0f} GB/s (read + write)"
          + (f", {100 * gbs / peak:.0f} % of the measured HBM copy peak {peak:.0f} GB/s" if peak else ""))
assert torch.equal(Ur, U.permute(0, 2, 1))
for r in range(2):
    d = os.path.join(root, f"run{r}")
    t0 = time.perf_counter()
    paths, iters = sweep.sweep_to_npy(fem, d, 0.05, steps, np.ones(512), mu1, mu2, chunk=256, writers=8)
    dt = time.perf_counter() - t0
    nbytes = sum(os.path.getsize(p) for p in paths)
    print(f"sweep_to_npy B={len(paths)} steps={steps} -> {d}: {dt * 1e3:.0f} ms wall, {len(paths) / dt:.0f} files/s, "
          f"{nbytes / dt / 1e9:.2f} GB/s to the file system, {len(paths) * steps / dt:.3e} time-steps/s")
    A = np.load(paths[-1]); assert A.shape == (512, steps + 1)
    shutil.rmtree(d)
if len(sys.argv) <= 3:
    shutil.rmtree(root, ignore_errors=True)
