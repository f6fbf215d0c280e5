"""torchrun worker: the multi-GPU path (NCCL halo exchange + all-reduce inside libfemo_b200) against the
oracle's global solve.  One process per GPU."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main(name):
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from femo_alpha_b200.dist_shell import DistributedShell
    from tests.problems import make_problem, oracle_for
    p = make_problem(name)
    ds = DistributedShell(p["nodes"], p["quads"], rank, world, device=local, element_wise_material=p["ewm"],
                          elementwise_pressure=p["ewp"])
    ds.init_comm()
    for k in ("thickness", "E", "nu", "density", "F_solid"):
        ds.set_input(k, p[k])
    if p["uhat"] is not None:
        ds.set_input("uhat", p["uhat"])
    ds.set_strong_bc(p["bc"])
    it, rr = ds.solve_state(rtol=1e-13, max_it=50000)
    w_loc = ds.dev.get_state()
    o = oracle_for(p)
    w = o.solve_state()
    t = o.topo
    # map local blocks to the oracle's global blocks through the global keys
    gkey = np.empty(t.nb, dtype=np.int64)
    gkey[:t.nq2] = np.arange(t.nq2)
    gkey[t.nq2:] = t.nn + t.ne + t.nc + np.arange(t.nn)
    pos = np.searchsorted(gkey, ds.lp.block_gkey)
    assert np.array_equal(gkey[pos], ds.lp.block_gkey)
    err_w = np.abs(w_loc.reshape(-1, 3) - w.reshape(-1, 3)[pos]).max() / np.abs(w).max()
    C, Cref = ds.output("compliance"), o.compliance(w)
    M, Mref = ds.output("mass"), o.mass()
    # adjoint thickness gradient on the owned vertices / cells
    rhs = ds.dev.output_grad("compliance", "disp_solid")
    lam = ds.solve_adjoint(rhs, rtol=1e-13, max_it=50000)
    lam = ds.halo_exchange(lam)
    g = ds.dev.output_grad("compliance", "thickness") - ds.dev.dRdf_T_apply("thickness", lam)
    g_ref = o.total_dcompliance_dh()
    if p["ewm"]:
        idx = ds.lp.cell_global[:ds.lp.n_owned_cells]
        g_own = g[:ds.lp.n_owned_cells]
    else:
        own_v = ds.lp.own_block_mask[:ds.dev.nn]
        idx = ds.lp.vertex_global[own_v]
        g_own = g[own_v]
    err_g = np.abs(g_own - g_ref[idx]).max() / np.abs(g_ref).max()
    # the one-call residual evaluation on a rank: local arrays, bitwise the separate calls
    R_sep = ds.dev.assemble_residual(w_loc)
    R_one = ds.dev.evaluate_residual({}, w_loc)
    same_R = bool(np.array_equal(R_sep, R_one))
    h, a = ds.comm_counts()
    print("rank %d/%d %s: PCG %d its relres %.1e | u err %.2e | C %.12g vs %.12g | mass %.12g vs %.12g | dC/dh err %.2e | "
          "%d halo exchanges, %d all-reduces" % (rank, world, name, it, rr, err_w, C, Cref, M, Mref, err_g, h, a), flush=True)
    ok = err_w <= 1e-8 and abs(C - Cref) <= 1e-8 * abs(Cref) and abs(M - Mref) <= 1e-11 * abs(Mref) and err_g <= 1e-8 \
        and h > 0 and a > 0 and same_R
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main(sys.argv[1])
