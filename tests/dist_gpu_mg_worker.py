"""torchrun worker: V-cycle preconditioned state solve on N GPUs (femo_alpha_b200.dist_multigrid.DistributedHierarchy:
distributed finest level, replicated coarser levels) against the oracle's global solve.  One process per GPU."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from femo_alpha_b200.dist_multigrid import DistributedHierarchy
    from oracle.rm_shell_ref import RMShellOracle, plate_mesh
    nodes_c, quads_c = plate_mesh(8, 4, 10.0, 2.0)
    h = DistributedHierarchy(nodes_c, quads_c, 3, rank, world, device=local)
    h.init_comm()
    nodes_f, quads_f = h.meshes[0]
    nn = nodes_f.shape[0]
    x = nodes_f[:, 0]
    inputs = dict(thickness=0.1 * (1 + 0.3 * np.sin(2 * np.pi * x / 10)), E=np.full(nn, 1e8), nu=np.full(nn, 0.3),
                  density=np.full(nn, 10.0), F_solid=np.tile([0.0, 0.0, 5.0], nn))
    for k, v in inputs.items():
        h.set_input(k, v)
    clamp = lambda xx: np.less(xx[0], 3e-16)
    h.set_strong_bc(clamp)
    its, rr = h.solve_state(rtol=1e-11, max_it=200)
    w_loc = h.fine.dev.get_state().reshape(-1, 3)
    o = RMShellOracle(nodes_f, quads_f, False, False)
    o.set_inputs(**inputs)
    o.set_strong_bc(clamp)
    w = o.solve_state().reshape(-1, 3)
    t = o.topo
    gkey = np.empty(t.nb, dtype=np.int64)
    gkey[:t.nq2] = np.arange(t.nq2)
    gkey[t.nq2:] = t.nn + t.ne + t.nc + np.arange(t.nn)
    pos = np.searchsorted(gkey, h.fine.lp.block_gkey)
    err = np.abs(w_loc - w[pos]).max() / np.abs(w).max()
    hx, ar = h.fine.comm_counts()
    print("rank %d/%d distributed V-cycle PCG: %d iterations, relres %.1e | u err %.2e | %d halo exchanges, %d all-reduces"
          % (rank, world, its, rr, err, hx, ar), flush=True)
    ok = err <= 1e-8 and 0 < its <= 60 and rr <= 1e-11
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    h.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
