"""torchrun worker: V-cycle preconditioned state + adjoint solves on N GPUs (femo_alpha_b200.dist_multigrid.DistributedHierarchy)
against the oracle's global solve, and against the single-GPU hierarchy of the same problem for the iteration count
(the preconditioner is the same operator whatever the partition).  One process per GPU.

argv: [n_dist]  number of distributed levels (default 1: distributed finest level over replicated coarser ones)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


# Plate 10 x 2 with h ~ 0.3.  (With h ~ 0.1, kappa ~ 1e9, device and oracle differ by 0.8e-8 / 1.6e-8 in u / lambda whatever
# the solver does -- rtol 1e-11 or 1e-13, with or without refinement: the solution is that sensitive to last-bit
# differences between the two assembled matrices; tests/test_gpu_multigrid.py documents it.)  One refinement step on a
# double-double residual (femo_mg_set_refine) exercises the halo exchanges of that path.
RTOL = 1e-11


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    n_dist = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from femo_alpha_b200.dist_multigrid import DistributedHierarchy
    from femo_alpha_b200.multigrid import ShellHierarchy
    from oracle.rm_shell_ref import RMShellOracle, plate_mesh
    nodes_c, quads_c = plate_mesh(8, 4, 10.0, 2.0)
    levels = 4
    h = DistributedHierarchy(nodes_c, quads_c, levels, rank, world, device=local, n_dist=n_dist)
    h.init_comm()
    h.set_refine(1)
    nodes_f, quads_f = h.meshes[0]
    nn = nodes_f.shape[0]
    x = nodes_f[:, 0]
    inputs = dict(thickness=0.3 * (1 + 0.3 * np.sin(2 * np.pi * x / 10)), E=np.full(nn, 1e8), nu=np.full(nn, 0.3),
                  density=np.full(nn, 10.0), F_solid=np.tile([0.0, 0.0, 5.0], nn))
    for k, v in inputs.items():
        h.set_input(k, v)
    clamp = lambda xx: np.less(xx[0], 3e-16)
    h.set_strong_bc(clamp)
    its, rr = h.solve_state(rtol=RTOL, max_it=400)
    w_loc = h.fine.dev.get_state().reshape(-1, 3)
    # adjoint of the compliance: A lam = dC/du (local right-hand side; owned entries are complete)
    lam_loc = h.solve_adjoint(h.fine.dev.output_grad("compliance", "disp_solid"), rtol=RTOL, max_it=400).reshape(-1, 3)
    its_a = h.last_solve["iters"]
    o = RMShellOracle(nodes_f, quads_f, False, False)
    o.set_inputs(**inputs)
    o.set_strong_bc(clamp)
    w = o.solve_state()
    lam = o.solve_adjoint(o.dcompliance_du(w)).reshape(-1, 3)
    w = w.reshape(-1, 3)
    t = o.topo
    gkey = np.empty(t.nb, dtype=np.int64)
    gkey[:t.nq2] = np.arange(t.nq2)
    gkey[t.nq2:] = t.nn + t.ne + t.nc + np.arange(t.nn)
    pos = np.searchsorted(gkey, h.fine.lp.block_gkey)
    err = np.abs(w_loc - w[pos]).max() / np.abs(w).max()
    err_a = np.abs(lam_loc - lam[pos]).max() / np.abs(lam).max()
    # the same hierarchy on one GPU: iteration counts must agree
    s = ShellHierarchy(nodes_c, quads_c, levels, device=local)
    for k, v in inputs.items():
        s.set_input(k, v)
    s.set_strong_bc(clamp)
    s.set_refine(1)
    its1, _ = s.solve_state(rtol=RTOL, max_it=400)
    s.close()
    hx, ar = h.fine.comm_counts()
    print("rank %d/%d distributed V-cycle PCG (%d distributed levels of %d): state %d / adjoint %d iterations (single GPU: %d), "
          "relres %.1e | u err %.2e | lam err %.2e | %d halo exchanges, %d all-reduces on the finest level"
          % (rank, world, n_dist, levels, its, its_a, its1, rr, err, err_a, hx, ar), flush=True)
    ok = err <= 1e-8 and err_a <= 1e-8 and 0 < its <= its1 + 2 and rr <= 1e-8      # rr: true residual after refinement (2e-9: eps |A| |x| / |b|)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    h.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
