"""Generates tests/golden/hex_*.npz from oracle/hex_ref.py (generic 3-D forms on hexahedra).

The reference has no form and no vector for this configuration ("parity unpinned", oracle header): the fixtures pin the
ORACLE, guard it and the CUDA path against regressions, and travel to the GPU box (nothing but NumPy is needed there).
Run from the repo root:  python tests/golden/make_golden_hex.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.hex_ref import HexOracle, box_mesh  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
CASES = [("elasticity", (4, 3, 3), 0.2, 5.0, 0.3), ("vector_poisson", (3, 4, 2), 0.25, 2.0, 0.0)]


def problem(form, shape, distort, E0, nu):
    nodes, cells = box_mesh(*shape, distort=distort, seed=5)
    rng = np.random.default_rng(20261104)
    rho = rng.uniform(0.05, 1.0, len(cells)); f = rng.standard_normal((len(nodes), 3))
    o = HexOracle(nodes, cells, form=form, E0=E0, nu=nu, simp_p=3.0)
    o.set_inputs(density=rho, body_force=f)
    base = np.nonzero(nodes[:, 0] < 1e-12)[0]
    o.set_strong_bc((3 * base[:, None] + np.arange(3)).ravel())
    return o, rng


def generate(form, shape, distort, E0, nu):
    o, rng = problem(form, shape, distort, E0, nu)
    K = o.stiffness(); K.sort_indices()
    u = o.solve()
    lam = rng.standard_normal(o.ndof); drho = rng.standard_normal(o.nc)
    return dict(form=form, E0=E0, nu=nu, simp_p=o.p, nodes=o.x, cells=o.cells.astype(np.int32), rho=o.rho, f=o.f,
                bc_dofs=o.bc_dofs, K_indptr=K.indptr, K_indices=K.indices, K_data=K.data, load=o.load(), u=u,
                compliance=o.compliance(u), volume=o.volume(), probe_lam=lam, probe_drho=drho,
                dRdrho_T=o.dRdrho_T(u, lam), dRdrho=o.dRdrho(u, drho), grad_total=o.total_dcompliance_drho())


if __name__ == "__main__":
    for case in CASES:
        d = generate(*case)
        np.savez_compressed(os.path.join(OUT, "hex_" + case[0] + ".npz"), **d)
        print("hex", case[0], d["compliance"], d["volume"])
