"""BASELINE config 4 (dynamic RM shell, 1-cos gust): seconds per time step (assemble + solve) and the cost of the
time-adjoint sweep on one B200, next to the CPU oracle on the box's host cores.  Writes one JSON object.

    python scripts/dyn_bench.py [--out gpurun_out/dyn.json] [--large]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def gust_history(nn, Nsteps):
    # examples/advanced_examples/simple_dynamic_shell_opt/ex_simple_dynamic_shell_opt.py:66-95
    V_inf, V_p, l_chord, GGLc, T0 = 50., 50., 1.2, 5, 0.02
    T1 = GGLc * l_chord / V_inf
    T = T0 + T1 + 2.86
    dt = T / Nsteps
    t = np.linspace(0, T, Nsteps + 1)
    fh = np.zeros((Nsteps + 1, nn * 3))
    for i, ti in enumerate(t):
        vg = V_p * (1 - np.cos(2 * np.pi * (ti - T0) / T1)) if T0 <= ti <= T0 + T1 else 0.0
        fh[i, 2::3] = 0.1 * vg
    return fh, dt


def run_gpu(mesh, Nsteps, solver, quad_deg=3, adjoint=True, dt_steps=None):
    """`dt_steps`: the step size is T / dt_steps (default Nsteps); only the first Nsteps steps are run."""
    from femo_alpha_b200.dynamic_rm_shell.plate_sim import PlateSim
    nn = mesh.geometry.x.shape[0]
    fh, dt = gust_history(nn, dt_steps or Nsteps)
    fh = fh[:Nsteps + 1]
    t0 = time.perf_counter()
    ps = PlateSim(mesh, 1e8, 0.3, 10., dt, Nsteps, quad_deg=quad_deg, solver=solver, rtol=1e-10)
    ps.update_t(np.full(nn, 0.1)); ps.update_f_history(fh)
    t_build = time.perf_counter() - t0
    dyn = ps._dyn
    t0 = time.perf_counter(); dyn.setup(); t_setup = time.perf_counter() - t0
    dyn.solve(fh)                                   # warm-up
    t0 = time.perf_counter(); its = dyn.solve(fh); t_fwd = time.perf_counter() - t0
    info = dyn.info()
    out = dict(n_dof=int(ps.fe_dofs), cells=int(ps.nel), n_steps=Nsteps, dt=dt, solver=info["solver"],
               build_s=t_build, setup_s=t_setup, forward_s=t_fwd, forward_device_ms=info["forward_ms"],
               s_per_step=t_fwd / Nsteps, solver_iters_total=int(its), tip_disp_max=float(np.abs(dyn.history()[:, 2::3]).max()),
               total_strain_energy=dyn.total_strain_energy())
    if adjoint:
        t0 = time.perf_counter()
        dEdw = dyn.total_strain_energy_grad("state")
        lam = dyn.solve_adjoint(dEdw)
        g = dyn.total_strain_energy_grad("thickness") - dyn.dRdh_T_apply(lam)
        out.update(adjoint_gradient_s=time.perf_counter() - t0, adjoint_device_ms=dyn.info()["adjoint_ms"],
                   grad_norm=float(np.linalg.norm(g)))
    ps._dyn.close()
    if ps.hierarchy is not None:
        ps.hierarchy.close()
    else:
        ps.device.close()
    return out


def run_cpu(nodes, cells, Nsteps, quad_deg=3, refactor_each_step=False, max_steps=None):
    import scipy.sparse.linalg as spla
    from oracle.dynamic_ref import DynamicShellOracle
    from oracle.rm_shell_ref import RMShellOracle
    nn = nodes.shape[0]
    fh, dt = gust_history(nn, Nsteps)
    o = RMShellOracle(nodes, cells, quadrature_degree=quad_deg)
    o.set_inputs(thickness=np.full(nn, 0.1), E=np.full(nn, 1e8), nu=np.full(nn, 0.3), density=np.full(nn, 10.), F_solid=fh[0])
    o.set_strong_bc(lambda x: np.isclose(x[0], 0., atol=1e-6))
    steps = Nsteps if max_steps is None else min(Nsteps, max_steps)
    d = DynamicShellOracle(o, dt, steps)
    t0 = time.perf_counter()
    if not refactor_each_step:
        d.solve(fh)
    else:
        # the reference's loop: one Newton iteration = assemble R, assemble J, LU factor, solve -- every step
        # (plate_sim.py:322, nonlinear_utils.py:210-233)
        d.operators()
        w = np.zeros(o.topo.ndof); v = np.zeros_like(w)
        for i in range(1, steps + 1):
            d.operators()
            F = d.load(fh[i])
            rhs = F - d.Bold @ w - d.Cdot @ v; rhs[o.bc_dofs] = 0
            wn = spla.splu(d.A.tocsc()).solve(rhs)
            v = 2 / dt * (wn - w) - v; w = wn
    el = time.perf_counter() - t0
    return dict(n_dof=int(o.topo.ndof), n_steps=steps, s_per_step=el / steps, total_s=el,
                kind="oracle (NumPy/SciPy SuperLU)" + (", assembled + factorised every step like the reference" if refactor_each_step else ", factorised once"),
                cores=1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--large", action="store_true")
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--only", default="", help="run a single small mesh, e.g. 40x8 (profiling)")
    ap.add_argument("--cfg3", type=int, default=0, metavar="STEPS",
                    help="time STEPS steps on the 816x816 plate of BASELINE config 3 (10.0 M dofs, 5-level V-cycle PCG)")
    a = ap.parse_args()
    from femo_alpha_b200.mesh import createRectangleMesh, refine_uniform
    res = dict(config="BASELINE config 4: plate 2x10, quad_deg=3, E=1e8, nu=0.3, rho=10, h=0.1, 1-cos gust, f_z = 0.1 V_g(t)",
               gpu=[], cpu=[], large=[])
    if a.cfg3:
        mesh = refine_uniform(createRectangleMesh((0., 0.), (10., 10.), 51, 51), 4)
        t0 = time.perf_counter()
        r = run_gpu(mesh, a.cfg3, "multigrid", quad_deg=6, adjoint=True, dt_steps=a.steps)
        r.update(mesh="816x816", quad_deg=6, wall_s_total=time.perf_counter() - t0,
                 workload="dynamic RM shell on the config-3 plate [0,10]^2, h = 0.1, dt = %.4g s, 1-cos gust" % r["dt"])
        print(json.dumps(r), flush=True)
        if a.out:
            json.dump(r, open(a.out, "w"), indent=1)
        return
    if a.only:
        nx, ny = [int(v) for v in a.only.split("x")]
        print(json.dumps(run_gpu(createRectangleMesh((0., 0.), (10., 2.), nx, ny), a.steps, "auto")), flush=True)
        return
    for nx, ny in ((10, 2), (20, 4), (40, 8)):
        mesh = createRectangleMesh((0., 0.), (10., 2.), nx, ny)
        r = run_gpu(mesh, a.steps, "auto"); r["mesh"] = "%dx%d" % (nx, ny); res["gpu"].append(r)
        print(json.dumps(r), flush=True)
    mesh = createRectangleMesh((0., 0.), (10., 2.), 40, 8)
    for refac, ms in ((False, None), (True, 10)):
        r = run_cpu(mesh.geometry.x, mesh.cells, a.steps, refactor_each_step=refac, max_steps=ms); r["mesh"] = "40x8"
        res["cpu"].append(r); print(json.dumps(r), flush=True)
    if a.large:
        base = createRectangleMesh((0., 0.), (10., 2.), 40, 8)
        for lv, steps in ((3, 20), (4, 10)):
            mesh = refine_uniform(base, lv)
            # quad_deg 3 (2x2 points, the reference's setting) leaves hourglass modes in K that only the mass term
            # controls; quad_deg 6 (4x4) is the fully integrated operator the multigrid study of the static path used
            for solver, deg in (("multigrid", 3), ("multigrid", 6), ("pcg", 3)):
                if solver == "pcg" and lv > 3:
                    continue
                try:
                    r = run_gpu(mesh, steps, solver, quad_deg=deg, adjoint=(solver == "multigrid"), dt_steps=a.steps)
                    r["mesh"] = "%dx%d" % (40 * 2 ** lv, 8 * 2 ** lv)
                except Exception as e:          # report, do not hide
                    r = dict(mesh="%dx%d" % (40 * 2 ** lv, 8 * 2 ** lv), solver=solver, error=str(e))
                r["quad_deg"] = deg
                res["large"].append(r); print(json.dumps(r), flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
