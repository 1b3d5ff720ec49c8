"""Time the snapshot Gram kernel on a config-2-sized buffer: python tools/gram_time.py [B] [steps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import burgers_b200
from burgers_b200 import pod

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 500
N = 512
U = torch.rand((B, steps + 1, N), dtype=torch.float64, device="cuda") + 1.0
acc = pod.GramAccumulator(N)
cols = B * (steps + 1)
for r in range(4):
    acc.G.zero_()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); acc.add(U); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"gram N={N} cols={cols}: {ms:.2f} ms, {2.0 * N * N * cols / ms / 1e9:.2f} TFLOP/s (full 2 N^2 cols), "
          f"{(N * (N + 128)) * cols / ms / 1e9:.2f} TFLOP/s executed (10 of 16 tiles), read {8.0 * N * cols / ms / 1e6:.0f} GB/s algorithmic")
# cuBLAS reference (library dgemm/syrk on the same data): the FP64 peak this box sustains
S = U.reshape(-1, N)
for r in range(3):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); G2 = S.T @ S; e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"cuBLAS dgemm S^T S: {ms:.2f} ms, {2.0 * N * N * cols / ms / 1e9:.2f} TFLOP/s")
print("max rel diff vs cuBLAS:", float((acc.G - G2).abs().max() / G2.abs().max()))
