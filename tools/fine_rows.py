"""Single-GPU fine-mesh kernel, us per Picard iteration by mesh size and rows per thread (BX_TUNE_FINE_ROWS):
python tools/fine_rows.py [steps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import burgers_b200

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 100
sms = burgers_b200._lib.load().bx_device_sm_count()
print(f"rows per thread x mesh size: us per Picard iteration ({steps} steps, constant CFL, best of 3), {sms} SMs")
for N in [2049, 8192, 32768, 131072, 262144, 524288, 148 * 2048, 148 * 4096, 1 << 20]:
    X = np.linspace(0, 100, N)
    At = 0.05 * 511.0 / (N - 1)
    fem = burgers_b200.FEMBurgers(X)
    U = torch.empty((1, steps + 1, N), dtype=torch.float64, device="cuda")
    it = torch.empty((1, steps), dtype=torch.int32, device="cuda")
    u0 = torch.ones(N, dtype=torch.float64, device="cuda")
    mu1 = torch.tensor([4.75], dtype=torch.float64, device="cuda"); mu2 = torch.tensor([0.02], dtype=torch.float64, device="cuda")
    line = f"N={N:8d}:"
    ref = None
    for rows in (16, 8, 4, -1):
        burgers_b200.set_tuning("fine_rows", rows)
        if rows > 0 and -(-N // (rows * 512)) > min(sms, 160):
            line += f"  rows {rows:2d}: does not fit"
            continue
        best = 1e30
        for r in range(3):
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            fem.fom_burgers_batch(At, steps, u0, mu1, 0.0, mu2, out=U, iters_out=it)
            e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        tot = int(it.sum())
        last = U[0, -1].clone()
        same = "" if ref is None else (" (bitwise = rows 16)" if torch.equal(last, ref[0]) and tot == ref[1] else f" (max rel diff {float(((last - ref[0]).abs() / ref[0].abs()).max()):.1e}, its {tot} vs {ref[1]})")
        if ref is None:
            ref = (last, tot)
        line += f"  rows {'auto' if rows < 0 else format(rows, '2d')}: {best / tot * 1e3:6.2f} us{same}"
    print(line, flush=True)
burgers_b200.set_tuning("fine_rows", -1)
