"""Critical-path breakdown of the fine-mesh kernel (thread 0 of CTA 0, cycles between marks); needs a library built with
tools/build_variant.sh trace bx_fine.cu -DBX_FINE_TRACE:  BX_VARIANT_LIB=tools/bin/lib_trace.so python tools/fine_trace.py [N] [steps] [rows]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import burgers_b200
burgers_b200._lib.LIB_PATH = os.path.abspath(os.environ["BX_VARIANT_LIB"])
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 100
rows = int(sys.argv[3]) if len(sys.argv) > 3 else -1
NAMES = {0: "thread-level assembly + sweep + upward", 1: "warp group_up", 2: "__syncthreads (warp rows)", 3: "CTA group_up (warp 0)",
         4: "publish (22 tagged words)", 5: "poll + unpack to shared memory", 6: "__syncthreads (rows of all CTAs)",
         8: "top-level solve (5-row sweep, PCR)", 9: "norm sum, s_wx, edge states", 10: "__syncthreads (solution)",
         11: "warp down + back-substitution + update", 12: "cta_edges (norm partials, edges)", 13: "step boundary: snapshot, bn"}
lib = burgers_b200._lib.load()
raw = ctypes.CDLL(burgers_b200._lib.LIB_PATH)
X = np.linspace(0, 100, N); At = 0.05 * 511.0 / (N - 1)
fem = burgers_b200.FEMBurgers(X)
burgers_b200.set_tuning("fine_rows", rows)
U = torch.empty((1, steps + 1, N), dtype=torch.float64, device="cuda"); it = torch.empty((1, steps), dtype=torch.int32, device="cuda")
u0 = torch.ones(N, dtype=torch.float64, device="cuda")
mu1 = torch.tensor([4.75], dtype=torch.float64, device="cuda"); mu2 = torch.tensor([0.02], dtype=torch.float64, device="cuda")
buf = (ctypes.c_ulonglong * 32)()
fem.fom_burgers_batch(At, steps, u0, mu1, 0.0, mu2, out=U, iters_out=it)
raw.bx_fine_trace_read(buf, 1)
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record(); fem.fom_burgers_batch(At, steps, u0, mu1, 0.0, mu2, out=U, iters_out=it); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1); tot = int(it.sum())
raw.bx_fine_trace_read(buf, 1)
cyc = sum(buf[i] for i in range(16))
print(f"N={N} rows={rows} steps={steps}: {ms:.2f} ms, {tot} Picard iterations, {ms / tot * 1e3:.2f} us/iteration; traced cycles {cyc} = {cyc / (ms * 1e-3) / 1e9:.2f} GHz-equivalent")
for i in range(14):
    if buf[16 + i]:
        print(f"  [{i:2d}] {NAMES[i]:52s} {100.0 * buf[i] / cyc:5.1f} %  {buf[i] / buf[16 + i]:8.0f} cycles x {buf[16 + i]:6d}  = {buf[i] / cyc * ms / tot * 1e3:6.2f} us per iteration")
