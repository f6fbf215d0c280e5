"""torchrun worker: the time-stepped path on N GPUs (distributed PCG per step, recurrences rank-local) against the
oracle's global history, energies and time-adjoint thickness gradient.  One process per GPU."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main(name, sw):
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from femo_alpha_b200.dist_shell import DistributedShell
    from oracle.dynamic_ref import DynamicShellOracle
    from tests.problems import make_problem, oracle_for
    p = make_problem(name)
    n_steps, dt = 5, 0.01
    ds = DistributedShell(p["nodes"], p["quads"], rank, world, device=local, element_wise_material=p["ewm"],
                          elementwise_pressure=p["ewp"])
    ds.init_comm()
    for k in ("thickness", "E", "nu", "density", "F_solid"):
        ds.set_input(k, p[k])
    ds.set_strong_bc(p["bc"])
    dd = ds.dynamic(dt, n_steps, add_self_weight=sw, rtol=1e-13, max_it=50000)
    rng = np.random.default_rng(7)
    f_hist = np.stack([p["F_solid"] * (0.5 + 0.1 * i) + 0.05 * rng.standard_normal(p["F_solid"].size)
                       for i in range(n_steps)])
    its = dd.solve(ds.slice_force_history(f_hist))
    w_loc = dd.history()

    o = oracle_for(p)
    od = DynamicShellOracle(o, dt, n_steps, add_self_weight=sw)
    w = od.solve(f_hist)
    t = o.topo
    gkey = np.empty(t.nb, dtype=np.int64)
    gkey[:t.nq2] = np.arange(t.nq2)
    gkey[t.nq2:] = t.nn + t.ne + t.nc + np.arange(t.nn)
    pos = np.searchsorted(gkey, ds.lp.block_gkey)
    assert np.array_equal(gkey[pos], ds.lp.block_gkey)
    loc = lambda hist: hist.reshape(hist.shape[0], -1, 3)[:, pos, :].reshape(hist.shape[0], -1)
    err_w = np.abs(w_loc - loc(w)).max() / np.abs(w).max()
    err_e = np.abs(dd.energies() - od.energy).max() / np.abs(od.energy).max()
    J, Jref = dd.total_strain_energy(), od.total_strain_energy(w)
    # time adjoint of the total strain energy, thickness gradient on what this rank owns
    lam = dd.solve_adjoint(dd.total_strain_energy_grad("disp"))
    lam_ref = od.adjoint_recurrence(od.dtotal_strain_energy_dw(w))
    err_l = np.abs(lam - loc(lam_ref)).max() / np.abs(lam_ref).max()
    g = dd.total_strain_energy_grad("thickness") - dd.dRdh_T_apply(lam)
    g_ref = od.dtotal_strain_energy_dh(w) - od.dRdh_T(w, lam_ref)
    if p["ewm"]:
        idx = ds.lp.cell_global[:ds.lp.n_owned_cells]
        g_own = g[:ds.lp.n_owned_cells]
    else:
        own_v = ds.lp.own_block_mask[:ds.dev.nn]
        idx = ds.lp.vertex_global[own_v]
        g_own = g[own_v]
    err_g = np.abs(g_own - g_ref[idx]).max() / np.abs(g_ref).max()
    h, a = ds.comm_counts()
    print("rank %d/%d %s sw=%d: %d PCG its over %d steps (%s) | history err %.2e | energies err %.2e | J %.12g vs %.12g | "
          "lambda err %.2e | dJ/dh err %.2e | %d halo exchanges, %d all-reduces"
          % (rank, world, name, sw, its, n_steps, dd.info()["solver"], err_w, err_e, J, Jref, err_l, err_g, h, a), flush=True)
    ok = (err_w <= 1e-8 and err_e <= 1e-8 and abs(J - Jref) <= 1e-8 * abs(Jref) and err_l <= 1e-7 and err_g <= 1e-7
          and dd.info()["solver"] == "pcg" and h > 0 and a > 0)
    dd.close()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]))
