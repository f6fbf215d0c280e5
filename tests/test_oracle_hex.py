"""Pins of oracle/hex_ref.py (generic 3-D forms on hexahedra).  The reference holds no form and no
golden vector for this configuration; the pins are the patch test, closed-form energies / elongations, the convergence
rate on a manufactured solution and finite differences of the partials."""
import numpy as np
import pytest

from oracle.hex_ref import HexOracle, box_mesh


def _boundary_dofs(nodes, L=(1.0, 1.0, 1.0)):
    on = np.zeros(len(nodes), bool)
    for d in range(3):
        on |= (nodes[:, d] < 1e-12) | (nodes[:, d] > L[d] - 1e-12)
    return (3 * np.nonzero(on)[0][:, None] + np.arange(3)).ravel()


@pytest.mark.parametrize("form", ["elasticity", "vector_poisson"])
def test_patch_test_on_a_distorted_mesh(form):
    nodes, cells = box_mesh(3, 4, 2, distort=0.25, seed=1)
    o = HexOracle(nodes, cells, form=form, E0=7.0, nu=0.3)
    rng = np.random.default_rng(0)
    A = rng.standard_normal((3, 3)); b = rng.standard_normal(3)
    u = (nodes @ A.T + b).ravel()                       # linear field: zero body force -> zero interior residual
    r = o.stiffness() @ u
    interior = np.setdiff1d(np.arange(o.ndof), _boundary_dofs(nodes))
    assert np.abs(r[interior]).max() < 1e-12 * np.abs(r).max()
    # strain energy of the uniform strain state
    vol = 1.0
    if form == "elasticity":
        eps = 0.5 * (A + A.T)
        w = vol * (o.c1 * np.trace(eps) ** 2 + 2 * o.c2 * (eps * eps).sum())
    else:
        w = vol * o.c1 * (A * A).sum()
    assert np.isclose(u @ (o.stiffness() @ u), w, rtol=1e-12)
    assert np.isclose(o.vol.sum(), 1.0, rtol=1e-13)
    ones = np.ones(o.ndof)
    assert np.isclose(ones @ (o.mass_matrix() @ ones), 3.0, rtol=1e-13)


def test_uniaxial_bar():
    # bar [0,4]x[0,1]x[0,1], nu = 0: x = 0 face held in x (rigid modes removed), uniform traction-equivalent body force
    # replaced by a body force along x: u_x(L) = f L^2 / (2 E)
    E, fx, L = 50.0, 0.3, 4.0
    nodes, cells = box_mesh(16, 2, 2, L, 1.0, 1.0)
    o = HexOracle(nodes, cells, E0=E, nu=0.0)
    f = np.zeros((o.nn, 3)); f[:, 0] = fx
    o.set_inputs(body_force=f)
    base = np.nonzero(nodes[:, 0] < 1e-12)[0]
    o.set_strong_bc((3 * base[:, None] + np.arange(3)).ravel())
    u = o.solve().reshape(-1, 3)
    tip = nodes[:, 0] > L - 1e-12
    # Q1 elements reproduce nodal values of the 1-D quadratic solution exactly for a consistent load
    assert np.allclose(u[tip, 0], fx * L ** 2 / (2 * E), rtol=1e-10)
    assert np.isclose(o.compliance(u.ravel()), fx ** 2 * L ** 3 / (3 * E), rtol=2e-3)


def test_manufactured_poisson_converges_at_second_order():
    errs = []
    for n in (4, 8):
        nodes, cells = box_mesh(n, n, n)
        o = HexOracle(nodes, cells, form="vector_poisson", E0=1.0)
        ex = np.sin(np.pi * nodes[:, 0]) * np.sin(np.pi * nodes[:, 1]) * np.sin(np.pi * nodes[:, 2])
        f = np.zeros((o.nn, 3)); f[:, 1] = 3 * np.pi ** 2 * ex
        o.set_inputs(body_force=f)
        o.set_strong_bc(_boundary_dofs(nodes))
        u = o.solve().reshape(-1, 3)
        assert np.abs(u[:, 0]).max() < 1e-13 and np.abs(u[:, 2]).max() < 1e-13
        errs.append(np.abs(u[:, 1] - ex).max())
    assert errs[0] / errs[1] > 3.0


def test_density_partials_against_finite_differences():
    nodes, cells = box_mesh(3, 2, 2, distort=0.2, seed=3)
    o = HexOracle(nodes, cells, E0=3.0, nu=0.25, simp_p=3.0)
    rng = np.random.default_rng(2)
    rho = rng.uniform(0.2, 1.0, o.nc)
    o.set_inputs(density=rho, body_force=rng.standard_normal((o.nn, 3)))
    base = np.nonzero(nodes[:, 0] < 1e-12)[0]
    o.set_strong_bc((3 * base[:, None] + np.arange(3)).ravel())
    u = rng.standard_normal(o.ndof); lam = rng.standard_normal(o.ndof); d = rng.standard_normal(o.nc) * 0.1
    eps = 1e-6
    o.set_inputs(density=rho + eps * d); Rp = o.residual(u)
    o.set_inputs(density=rho - eps * d); Rm = o.residual(u)
    o.set_inputs(density=rho)
    fd = (Rp - Rm) / (2 * eps)
    assert np.abs(o.dRdrho(u, d) - fd).max() < 1e-8 * np.abs(fd).max()
    assert np.isclose(lam @ fd, o.dRdrho_T(u, lam) @ d, rtol=1e-8)
    g = o.total_dcompliance_drho()
    o.set_inputs(density=rho + eps * d); Cp = o.compliance(o.solve())
    o.set_inputs(density=rho - eps * d); Cm = o.compliance(o.solve())
    o.set_inputs(density=rho)
    assert np.isclose(g @ d, (Cp - Cm) / (2 * eps), rtol=1e-6)


@pytest.mark.parametrize("form", ["elasticity", "vector_poisson"])
def test_golden_fixtures_are_reproduced(form):
    """tests/golden/hex_*.npz (the fixtures the GPU box checks against) are what the oracle produces today."""
    import os
    from tests.golden.make_golden_hex import CASES, generate
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hex_" + form + ".npz"))
    d = generate(*[c for c in CASES if c[0] == form][0])
    for k in ("K_data", "load", "u", "dRdrho_T", "dRdrho", "grad_total"):
        assert np.abs(d[k] - g[k]).max() <= 1e-12 * np.abs(g[k]).max(), k
    assert np.array_equal(d["K_indices"], g["K_indices"]) and np.array_equal(d["cells"], g["cells"])
    assert np.isclose(d["compliance"], float(g["compliance"]), rtol=1e-12) and np.isclose(d["volume"], float(g["volume"]), rtol=1e-14)


@pytest.mark.parametrize("form", ["elasticity", "vector_poisson"])
def test_c_port_matches_the_numpy_oracle(form):
    """oracle/hex_c.c (the timed CPU baseline of scripts/hex_bench.py) against oracle/hex_ref.py"""
    from oracle.c_port import HexCsrAssembler
    nodes, cells = box_mesh(4, 3, 5, distort=0.25, seed=6)
    o = HexOracle(nodes, cells, form=form, E0=3.0, nu=0.25)
    o.set_inputs(density=np.random.default_rng(0).uniform(0.05, 1.0, o.nc))
    K = o.stiffness()
    Kc = HexCsrAssembler(o).matrix()
    assert abs(Kc - K).max() < 1e-13 * abs(K).max()
