"""Config 5 on N GPUs (weak scaling): every rank owns an n^3 slab of an (n N) x n x n box of Q1 hexahedra, SIMP
elasticity.  Launch with torchrun, one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29561 \\
        scripts/hex_bench_dist.py --cells 128 [--reps 5] [--krylov-iters 50] [--out FILE]

Assembly is owner-computes (no collective); the PCG iteration exchanges one halo plane per neighbour and all-reduces two
scalar pairs (NCCL inside libfemo_b200).  Times are CUDA-event times inside the library, maximum over the ranks; rank 0
prints one JSON line.  The N-GPU parity of the same path is tests/test_gpu_dist.py::test_multi_gpu_hexahedral_solid_matches_oracle
(green at 2 and 3 GPUs); this timing script itself has not had its hardware run (the round's GPU budget ended on a
command-line mistake: torchrun rejects `--n`)."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cells", "--n", dest="n", type=int, default=128,
                    help="cells per edge and rank (use --cells under torchrun: Python 3.12 argparse takes --n for an abbreviation of its own options)")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--krylov-iters", type=int, default=50)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from femo_alpha_b200.dist_solid import DistributedSolid
    from femo_alpha_b200.solid_device import box_hexes
    n = a.n
    nodes, cells = box_hexes(n * world, n, n, float(world), 1.0, 1.0)
    part = (np.arange(len(cells)) // (n * n * n)).astype(np.int32)          # box_hexes numbers cells x-major: slabs
    ds = DistributedSolid(nodes, cells, rank, world, part=part, device=local)
    ds.init_comm()
    rng = np.random.default_rng(20261104)
    f = np.zeros((len(nodes), 3)); f[:, 2] = -1.0
    ds.set_input("body_force", f)
    ds.set_strong_bc(lambda x: np.less(x[0], 1e-12))
    d = ds.dev
    ms = []
    for r in range(a.reps + 2):
        ds.set_input("density", rng.uniform(0.01, 1.0, len(cells)))        # same global field on every rank
        d.assemble_jacobian()
        ms.append(d.last_timings()["assemble_ms"])
    asm = float(np.mean(ms[2:]))
    pcg = d.pcg_fixed(a.krylov_iters) / a.krylov_iters
    t = torch.tensor([asm, pcg], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    asm, pcg = float(t[0]), float(t[1])
    ndof_total = 3 * len(nodes)
    if rank == 0:
        line = dict(metric="3-D elasticity Jacobian assembly throughput (config 5)", value=ndof_total / asm / 1e6, unit="GDOF/s",
                    n_gpus=world, scaling="weak", assemble_ms=asm, pcg_ms_per_iter=pcg,
                    config=dict(workload="box %dx%dx%d Q1 hexahedra, elasticity, SIMP rho^3, x-slabs" % (n * world, n, n),
                                dof_total=int(ndof_total), cells_per_gpu=int(ds.lp.n_owned_cells), local_cells=int(d.nc)),
                    dtype="f64", data="synthetic")
        s = json.dumps(line)
        print(s)
        if a.out:
            open(a.out, "w").write(s + "\n")
    ds.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
