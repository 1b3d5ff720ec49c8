"""R rank kernels of the partitioned fine-mesh trajectory side by side on ONE device (as tests/test_zy_finemesh_ranks.py):
python tools/finemesh_ranks_local.py [N] [R] [steps] [rows]   -> ms per launch set, us per Picard iteration"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import burgers_b200
if os.environ.get("BX_VARIANT_LIB"):
    burgers_b200._lib.LIB_PATH = os.path.abspath(os.environ["BX_VARIANT_LIB"])
from burgers_b200 import _lib
from burgers_b200.finemesh_dist import PeerExchange, node_partition
N = int(sys.argv[1]) if len(sys.argv) > 1 else 524288
R = int(sys.argv[2]) if len(sys.argv) > 2 else 2
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 100
rows = int(sys.argv[4]) if len(sys.argv) > 4 else -1
burgers_b200.set_tuning("fine_rows", rows)
lib = _lib.load()
dev = torch.device("cuda:0")
X = torch.linspace(0.0, 100.0, N, dtype=torch.float64, device=dev)
At = 0.05 * 511.0 / (N - 1)
parts = node_partition(N, R)
ex = PeerExchange.local(R)
m1 = torch.tensor([4.75], dtype=torch.float64, device=dev); m2 = torch.tensor([0.02], dtype=torch.float64, device=dev)
streams = [torch.cuda.Stream(dev) for _ in range(R)]
keep = []
for off, cnt in parts:
    keep.append((torch.ones(cnt, dtype=torch.float64, device=dev), torch.empty((1, steps + 1, cnt), dtype=torch.float64, device=dev),
                 torch.empty((1, steps), dtype=torch.int32, device=dev),
                 torch.empty(int(lib.bx_fine_workspace_bytes(cnt)), dtype=torch.uint8, device=dev)))
for rep in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for r, (off, cnt) in enumerate(parts):
        u0, U, it, ws = keep[r]
        _lib.check(lib.bx_fom_picard_finemesh_ranks(1, N, off, cnt, steps, X.data_ptr(), u0.data_ptr(), 0, m1.data_ptr(), m2.data_ptr(),
                                                    At, 0.0, 1e-6, 20, U.data_ptr(), it.data_ptr(), ws.data_ptr(), R, r,
                                                    ex[r].pointer_array(), streams[r].cuda_stream), "ranks")
    torch.cuda.synchronize()
    ms = 1e3 * (time.perf_counter() - t0)
    tot = int(keep[0][2].sum())
    print(f"N={N} R={R} rows={rows} CTAs/rank={[-(-c // (512 * (rows if rows > 0 else 4))) for _, c in parts]}: {ms:.2f} ms, {tot} iterations, "
          f"{1e3 * ms / tot:.2f} us per iteration, failed={any(e.failed() for e in ex)}", flush=True)
