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


if __name__ == "__main__":
    for name in PROBLEMS:
        summarize(oracle_for(make_problem(name)), name)
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
