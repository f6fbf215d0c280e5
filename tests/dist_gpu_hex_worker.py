"""torchrun worker: hexahedral-solid forms on N GPUs (DistributedSolid: NCCL halo exchange + all-reduce inside
libfemo_b200) against the oracle's global solve.  One process per GPU."""
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
    from femo_alpha_b200.dist_solid import DistributedSolid
    from oracle.hex_ref import HexOracle
    from tests.dist_hex_worker import problem
    nodes, cells, rho, f = problem()
    ds = DistributedSolid(nodes, cells, rank, world, device=local, E0=4.0, nu=0.3)
    ds.init_comm()
    ds.set_input("density", rho); ds.set_input("body_force", f)
    ds.set_strong_bc(lambda x: np.less(x[0], 1e-12))
    it, rr = ds.solve_state(rtol=1e-13, max_it=50000)
    x = ds.dev.get_state().reshape(-1, 3)
    o = HexOracle(nodes, cells, E0=4.0, nu=0.3)
    o.set_inputs(density=rho, body_force=f)
    base = np.nonzero(nodes[:, 0] < 1e-12)[0]
    o.set_strong_bc((3 * base[:, None] + np.arange(3)).ravel())
    u = o.solve()
    err_u = np.abs(x - u.reshape(-1, 3)[ds.lp.vertex_global]).max() / np.abs(u).max()
    C, Cref = ds.output("compliance"), o.compliance(u)
    V, Vref = ds.output("volume"), o.volume()
    lam = ds.solve_adjoint(ds.dev.output_grad("compliance", "state"), rtol=1e-13, max_it=50000)
    idx, g_own = ds.owned_cell_values(-ds.dev.dRdf_T_apply("density", lam))
    g_ref = o.total_dcompliance_drho()
    err_g = np.abs(g_own - g_ref[idx]).max() / np.abs(g_ref).max()
    h, a = ds.comm_counts()
    print("rank %d/%d hex: PCG %d its relres %.1e | u err %.2e | C %.12g vs %.12g | volume %.12g vs %.12g | dC/drho err %.2e | "
          "%d halo exchanges, %d all-reduces" % (rank, world, it, rr, err_u, C, Cref, V, Vref, err_g, h, a), flush=True)
    ok = (err_u <= 1e-8 and abs(C - Cref) <= 1e-8 * abs(Cref) and abs(V - Vref) <= 1e-12 * abs(Vref) and err_g <= 1e-8
          and h > 0 and a > 0)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ds.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
