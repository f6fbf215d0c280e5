"""Pins of the dynamic (time-stepped) oracle, oracle/dynamic_ref.py  (BASELINE config 4).

The reference holds no golden vector for this path; the pins are physical (first bending frequency of the
clamped plate strip, the weak pin of the same kind as the Euler-Bernoulli deflection of the static path),
algebraic (the O(N) adjoint recurrence against the literal restatement of the reference's O(N^2) sweep,
state_operation_dynamic.py:590-640) and finite differences of the total strain energy.
"""
import numpy as np
import pytest

from oracle.dynamic_ref import DynamicShellOracle
from oracle.rm_shell_ref import RMShellOracle, plate_mesh
from tests.problems import make_problem, oracle_for


def small_dynamic(name="plate_varh", dt=0.01, n_steps=5, quadrature_degree=6, self_weight=False):
    p = make_problem(name)
    o = oracle_for(p, quadrature_degree=quadrature_degree)
    d = DynamicShellOracle(o, dt, n_steps, add_self_weight=self_weight)
    rng = np.random.default_rng(7)
    f_hist = np.stack([p["F_solid"] * (0.5 + 0.1 * i) + 0.05 * rng.standard_normal(p["F_solid"].size)
                       for i in range(n_steps + 1)])
    return p, o, d, f_hist


def test_mass_matrix_total_mass():
    # sum of the u_z-u_z block = int rho h dx = the `mass` output (rm_shell_pde.py:101-102)
    p, o, d, _ = small_dynamic()
    M = d.mass_matrix()
    t = o.topo
    ez = np.zeros(t.ndof); ez[2:3 * t.nq2:3] = 1.0
    assert np.isclose(ez @ (M @ ez), o.mass(), rtol=1e-13)
    # rotational inertia rho h hE^2 on the theta blocks
    er = np.zeros(t.ndof); er[3 * t.nq2::3] = 1.0
    hE2 = np.unique(np.round(o.hE, 12))
    if hE2.size == 1:
        assert np.isclose(er @ (M @ er), o.mass() * hE2[0] ** 2, rtol=1e-13)
    assert abs(M - M.T).max() < 1e-18 + 1e-14 * abs(M).max()


def test_first_bending_frequency_of_the_clamped_strip():
    # cantilever strip, step load: the tip oscillates about the static deflection with the first
    # Euler-Bernoulli period T1 = 2 pi / (1.8751^2 sqrt(EI / (rho A L^4)))
    E, nu, h, rho, L, b = 1e8, 0.0, 0.1, 10.0, 10.0, 2.0
    nodes, quads = plate_mesh(20, 4, L, b)
    nn = nodes.shape[0]
    o = RMShellOracle(nodes, quads, quadrature_degree=4)
    o.set_inputs(thickness=np.full(nn, h), E=np.full(nn, E), nu=np.full(nn, nu), density=np.full(nn, rho),
                 F_solid=np.tile([0, 0, 1.0], nn))
    o.set_strong_bc(lambda x: np.less(x[0], 3e-16))
    T1 = 2 * np.pi / (1.8751 ** 2 * np.sqrt(E * b * h ** 3 / 12 / (rho * b * h * L ** 4)))
    dt = T1 / 80
    d = DynamicShellOracle(o, dt, 100)
    w = d.solve(np.tile([0, 0, 1.0], nn)[None, :])
    tip = np.nonzero(np.isclose(nodes[:, 0], L) & np.isclose(nodes[:, 1], 1.0))[0][0]
    uz = w[:, 3 * tip + 2]
    w_static = o.solve_state()[3 * tip + 2]
    i_peak = int(np.argmax(uz))
    assert abs(i_peak * dt - 0.5 * T1) < 0.04 * T1          # peak at half a period
    assert 1.9 < uz[i_peak] / w_static < 2.1                # dynamic overshoot factor ~2 of a step load
    # the implicit midpoint rule conserves  kinetic + strain energy - work  for a constant load
    v = np.zeros_like(w[0]); en = []
    F = o.assemble_load(); F[o.bc_dofs] = 0
    for i in range(1, d.levels):
        v = 2 / dt * (w[i] - w[i - 1]) - v
        en.append(0.5 * v @ (d.M @ v) + 0.5 * w[i] @ (d.K @ w[i]) - F @ w[i])
    assert np.ptp(en) < 1e-8 * np.abs(w_static * F.sum())


@pytest.mark.parametrize("name,sw", [("plate_varh", False), ("plate_dg0", True)])
def test_adjoint_recurrence_matches_reference_sweep(name, sw):
    p, o, d, f_hist = small_dynamic(name, n_steps=6, self_weight=sw)
    w = d.solve(f_hist)
    R = d.residual_history(w, f_hist)
    assert np.abs(R).max() < 1e-9 * np.abs(d.Acur @ w[-1]).max()
    rng = np.random.default_rng(3)
    dout = rng.standard_normal(w.shape)
    lam_ref = d.adjoint_reference(dout)
    lam = d.adjoint_recurrence(dout)
    assert np.abs(lam - lam_ref).max() < 1e-10 * np.abs(lam_ref).max()
    # (dR/dw)^T lam = dout on the free dofs of levels >= 1
    back = d.dRdw_apply(lam, transpose=True)
    free = d.keep > 0
    assert np.abs(back[1:, free] - dout[1:, free]).max() < 1e-9 * np.abs(dout).max()
    # forward solve inverts the product
    x = rng.standard_normal(w.shape)
    x[:, o.bc_dofs] = 0
    y = d.dRdw_apply(x)
    x2 = d.forward_recurrence(y)
    assert np.abs(x2 - x).max() < 1e-9 * np.abs(x).max()


@pytest.mark.parametrize("name,sw", [("plate_varh", True), ("plate_dg0", False)])
def test_total_strain_energy_gradient_fd(name, sw):
    p, o, d, f_hist = small_dynamic(name, n_steps=4, self_weight=sw)
    w = d.solve(f_hist)
    J0 = d.total_strain_energy(w)
    lam = d.adjoint_recurrence(d.dtotal_strain_energy_dw(w))
    grad = d.dtotal_strain_energy_dh(w) - d.dRdh_T(w, lam)
    rng = np.random.default_rng(5)
    dh = rng.standard_normal(o.n_mat) * p["thickness"]
    eps = 1e-6
    h0 = o.h.copy()
    o.set_inputs(thickness=h0 + eps * dh); Jp = d.total_strain_energy(d.solve(f_hist))
    o.set_inputs(thickness=h0 - eps * dh); Jm = d.total_strain_energy(d.solve(f_hist))
    o.set_inputs(thickness=h0)
    fd = (Jp - Jm) / (2 * eps)
    assert np.isclose(grad @ dh, fd, rtol=2e-6), (grad @ dh, fd, J0)
    # dR/dh . dh against a finite difference of the residual history
    d.operators()
    Rlin = d.dRdh(w, dh)
    o.set_inputs(thickness=h0 + eps * dh); d.operators(); Rp = d.residual_history(w, f_hist)
    o.set_inputs(thickness=h0 - eps * dh); d.operators(); Rm = d.residual_history(w, f_hist)
    o.set_inputs(thickness=h0); d.operators()
    fdR = (Rp - Rm) / (2 * eps)
    assert np.abs(Rlin - fdR).max() < 1e-6 * np.abs(fdR).max()


def test_volume():
    p, o, d, _ = small_dynamic("plate_varh")
    g = d.dvolume_dh()
    assert np.isclose(g @ o.h, d.volume(), rtol=1e-13)     # volume is linear in the nodal thickness


@pytest.mark.parametrize("name", ["plate_varh", "plate_dg0", "warped"])
def test_golden_fixtures_are_reproduced(name):
    """tests/golden/dyn_shape_*.npz (the fixtures the GPU box checks against) are what the oracles produce today."""
    import os
    from oracle import shape_ref
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dyn_shape_" + name + ".npz"))
    p = make_problem(name)
    o = oracle_for(p, quadrature_degree=int(g["quad_deg"]))
    d = DynamicShellOracle(o, float(g["dt"]), int(g["n_steps"]), add_self_weight=bool(g["self_weight"]))
    w = d.solve(g["f_hist"])
    assert np.abs(w - g["w_hist"]).max() <= 1e-12 * np.abs(g["w_hist"]).max()
    o.set_inputs(F_solid=p["F_solid"])
    got = shape_ref.dRduhat_T_apply(o, g["shape_w"], g["shape_lam"])
    assert np.abs(got - g["dRduhat_T"]).max() <= 1e-12 * np.abs(g["dRduhat_T"]).max()
