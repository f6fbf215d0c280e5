"""Generates tests/golden/*.npz from the CPU oracle (oracle/rm_shell_ref.py).

The reference itself (dolfinx 0.5.1 / PETSc) cannot be imported in this container, so these
fixtures pin the ORACLE, not the reference ("parity unpinned", see oracle header): they guard the
oracle and the CUDA path against regressions and travel to the GPU box.
Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import pickle
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests.problems import make_problem, oracle_for, PROBLEMS  # noqa: E402
from oracle.rm_shell_ref import RMShellOracle  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def summarize(o, name):
    t = o.topo
    rng = np.random.default_rng(7)
    v = rng.standard_normal(t.ndof)
    K = o.assemble_dRdu()
    w = o.solve_state()
    o.w = w
    lam = o.solve_adjoint(o.dcompliance_du(w))
    d = dict(ndof=t.ndof, nnz=K.nnz, probe=v, K_probe=K @ v, K_diag=K.diagonal(), K_fro=np.sqrt((K.data ** 2).sum()),
             load=o.assemble_load(), w=w, compliance=o.compliance(w), mass=o.mass(), elastic_energy=o.elastic_energy(w),
             dCdu=o.dcompliance_du(w), lam=lam, dCdh_total=o.total_dcompliance_dh(),
             pnorm=o.pnorm_stress(w, m=1e-4, rho=6.0), stress=o.stress_field(w))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
    print(name, t.ndof, K.nnz, d["compliance"])


DYN_CASES = [("plate_varh", 6, False), ("plate_dg0", 3, True), ("warped", 6, True)]      # (problem, quad_deg, self weight)


def dynamic_and_shape(name, deg, sw):
    """Fixtures of the time-stepped path (oracle/dynamic_ref.py; config-4 style: 6 implicit midpoint steps, dt = 0.01)
    and of the shape partials (oracle/shape_ref.py, complex step)."""
    from oracle.dynamic_ref import DynamicShellOracle
    from oracle import shape_ref
    p = make_problem(name)
    o = oracle_for(p, quadrature_degree=deg)
    rng = np.random.default_rng(11)
    n_steps, dt = 6, 0.01
    f_hist = np.stack([p["F_solid"] * (0.4 + 0.1 * i) for i in range(n_steps + 1)])
    od = DynamicShellOracle(o, dt, n_steps, add_self_weight=sw)
    w = od.solve(f_hist)
    dE = od.dtotal_strain_energy_dw(w)
    lam = od.adjoint_recurrence(dE)
    x = rng.standard_normal(o.topo.ndof)
    d = dict(quad_deg=deg, self_weight=int(sw), dt=dt, n_steps=n_steps, f_hist=f_hist, w_hist=w, energies=od.energy,
             total_strain_energy=od.total_strain_energy(w), lam_hist=lam,
             grad_total=od.dtotal_strain_energy_dh(w) - od.dRdh_T(w, lam), probe=x, M_probe=od.mass_matrix() @ x,
             volume=od.volume())
    o.set_inputs(F_solid=p["F_solid"])
    ws = rng.standard_normal(o.topo.ndof); ls = rng.standard_normal(o.topo.ndof); du = rng.standard_normal(3 * o.topo.nn)
    d.update(shape_w=ws, shape_lam=ls, shape_du=du, dRduhat_T=shape_ref.dRduhat_T_apply(o, ws, ls),
             dRduhat=shape_ref.dRduhat_apply(o, ws, du), dcompliance_duhat=shape_ref.doutput_duhat(o, "compliance", ws),
             dmass_duhat=shape_ref.doutput_duhat(o, "mass", ws), dpnorm_duhat=shape_ref.dpnorm_duhat(o, ws, 1e-6, 6.0))
    np.savez_compressed(os.path.join(OUT, "dyn_shape_" + name + ".npz"), **d)
    print("dyn_shape", name, d["total_strain_energy"])


if __name__ == "__main__":
    if "--dynamic-only" not in sys.argv:
        for name in PROBLEMS:
            summarize(oracle_for(make_problem(name)), name)
    for case in DYN_CASES:
        dynamic_and_shape(*case)
    if "--dynamic-only" in sys.argv:
        sys.exit(0)
    # wing box of the reference tree (the only real mesh in the checkout), clamp y <= 1e-3, Al, -z pressure
    wb = "/root/reference/examples/advanced_examples/lpc_shell_opt/data_files/mesh_bay_data.pickle"
    if os.path.exists(wb):
        d = pickle.load(open(wb, "rb"))
        nodes = np.asarray(d["nodes"], dtype=np.float64)
        quads = np.asarray(d["connectivity"]).astype(np.int64)
        np.savez_compressed(os.path.join(OUT, "wingbox_mesh.npz"), nodes=nodes, quads=quads.astype(np.int32))
        o = RMShellOracle(nodes, quads, element_wise_material=True, elementwise_pressure=False)
        h = 0.05 * 0.0254
        o.set_inputs(thickness=h, E=73.1e9, nu=0.33, density=2780.0,
                     F_solid=np.tile([0.0, 0.0, -9.81 * 2780.0 * h], nodes.shape[0]))
        o.set_strong_bc(lambda x: np.less_equal(x[1], 1e-3))
        summarize(o, "wingbox")
