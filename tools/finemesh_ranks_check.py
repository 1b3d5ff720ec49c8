"""torchrun --nproc-per-node R tools/finemesh_ranks_check.py <N_parity> <N_timed> <steps>
One process per GPU: (1) parity of the partitioned fine-mesh trajectory (bx_fom_picard_finemesh_ranks, peer-memory
exchange) against the oracle and the single-GPU kernel at N_parity nodes x 12 steps; (2) kernel time (CUDA events, max
over ranks) at N_timed nodes x steps, beside the single-GPU kernel on rank 0 when the mesh fits one GPU."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import burgers_b200  # noqa: E402
from burgers_b200.finemesh_dist import PeerExchange, fom_burgers_partitioned, node_partition  # noqa: E402


def main():
    Np, Nt, steps = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    rank, R = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    gl = dist.new_group(backend="gloo")
    ex = PeerExchange.create(gl)
    dev = torch.device("cuda", torch.cuda.current_device())

    # ---- parity ------------------------------------------------------------------------------------------
    X = np.linspace(0.0, 100.0, Np)
    At = 0.05 * 511.0 / (Np - 1)
    u0 = 1.0 + 0.1 * np.sin(X / 9.0)
    parts = node_partition(Np, R)
    off, cnt = parts[rank]
    U, it = fom_burgers_partitioned(X, At, 12, u0[off:off + cnt], 4.75, 0.0, 0.02, ex, gl)
    blocks = [None] * R
    dist.all_gather_object(blocks, (U[0].cpu().numpy(), it[0].cpu().numpy()), group=gl)
    if rank == 0:
        from oracle import burgers_oracle as oracle
        Ufull = np.concatenate([b[0] for b in blocks], axis=1).T
        same_counts = all(np.array_equal(b[1], blocks[0][1]) for b in blocks)
        Uo, ito = oracle.fom_burgers(X, At, 12, u0, 4.75, 0.0, 0.02)
        rel = float(np.max(np.linalg.norm(Ufull - Uo, axis=0) / np.linalg.norm(Uo, axis=0)))
        fem = burgers_b200.FEMBurgers(X)
        Us, its = fem.fom_burgers_batch(At, 12, u0, [4.75], 0.0, [0.02])
        Us = Us[0].cpu().numpy().T
        rel1 = float(np.max(np.linalg.norm(Ufull - Us, axis=0) / np.linalg.norm(Us, axis=0)))
        ok = same_counts and np.array_equal(blocks[0][1], ito) and rel <= 1e-10 and rel1 <= 1e-10
        print(f"parity N={Np} on {R} ranks (blocks {parts}): max per-step rel L2 vs oracle {rel:.2e}, vs the single-GPU "
              f"kernel {rel1:.2e}, counts identical on all ranks and to the oracle: {same_counts and np.array_equal(blocks[0][1], ito)}"
              f" -> {'OK' if ok else 'FAIL'}", flush=True)

    # ---- timing ------------------------------------------------------------------------------------------
    X = np.linspace(0.0, 100.0, Nt)
    At = 0.05 * 511.0 / (Nt - 1)
    parts = node_partition(Nt, R)
    off, cnt = parts[rank]
    u0l = np.ones(cnt)
    out = torch.empty((1, steps + 1, cnt), dtype=torch.float64, device=dev)
    best = None
    for rep in range(3):
        U, it = fom_burgers_partitioned(X, At, steps, u0l, 4.75, 0.0, 0.02, ex, gl, out=out)
        t = torch.tensor([ex.last_kernel_ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=gl)
        best = t.item() if best is None else min(best, t.item())
    total = int(it.sum().item())
    finite = bool(torch.isfinite(U).all().item())
    if rank == 0:
        print(f"timing N={Nt} x {steps} steps on {R} ranks ({[-(-c // 8192) for _, c in parts]} CTAs): {best:.2f} ms "
              f"(kernel, max over ranks, best of 3), {total} Picard iterations, {1e3 * best / total:.2f} us per iteration, "
              f"finite={finite}", flush=True)
        if Nt <= 8192 * 148:
            fem = burgers_b200.FEMBurgers(X)
            o1 = torch.empty((1, steps + 1, Nt), dtype=torch.float64, device=dev)
            ts = []
            for rep in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                U1, it1 = fem.fom_burgers_batch(At, steps, np.ones(Nt), [4.75], 0.0, [0.02], out=o1)
                e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            t1 = int(it1.sum().item())
            print(f"single GPU, same mesh: {min(ts):.2f} ms (incl. the u0 upload), {t1} Picard iterations, "
                  f"{1e3 * min(ts) / t1:.2f} us per iteration", flush=True)
    dist.barrier(gl)
    ex.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
