"""Sparse direct solver (csrc/coarse.cu: block-tridiagonal factorisation on level sets, hand-written Cholesky / GEMM
kernels) against the oracle's SuperLU solves.  The reference puts MUMPS behind every solve
(femo_alpha/fea/utils_dolfinx.py:465-466, 514-531); here the direct solver serves the coarsest multigrid level, the
dense time stepper and meshes without a hierarchy (the reference's wing box)."""
import os

import numpy as np
import pytest

from tests.problems import PROBLEMS, make_problem, oracle_for, device_for

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("name", PROBLEMS)
def test_direct_solver_small_problems(built_lib, name):
    """n <= 8192: one product with the explicit inverse built from the factor."""
    p = make_problem(name)
    o, d = oracle_for(p), device_for(p)
    info = d.enable_direct(refine=2)
    assert info["explicit_inverse"] and info["level_sets"] >= 3
    w = o.solve_state()
    d.solve_state()
    # refinement on a double-double residual: the reported residual is the true one
    print("%s: %s | relres %.2e | rel. error %.2e" % (name, info, d.last_solve["relres"], rel(d.get_state(), w)))
    assert d.last_solve["relres"] <= 1e-7
    assert rel(d.get_state(), w) <= 1e-8
    rhs = o.dcompliance_du(w)
    lam = d.solve_adjoint(rhs)
    assert rel(lam, o.solve_adjoint(rhs)) <= 1e-8             # plate_4x20 (kappa ~ 1e10) reaches 2e-9, the others 1e-11
    assert np.array_equal(lam, d.solve_adjoint(rhs))           # bitwise repeatable
    # new inputs -> the factor follows (femo_direct_setup re-factorises on demand)
    d.set_input("thickness", 1.3 * np.asarray(p["thickness"]))
    o.set_inputs(thickness=1.3 * np.asarray(p["thickness"]))
    d.solve_state()
    assert rel(d.get_state(), o.solve_state()) <= 1e-9
    d.close()


def test_direct_solver_sweeps(built_lib):
    """n > 8192: forward / backward sweeps over the level sets (one small kernel per level and direction)."""
    from oracle.rm_shell_ref import RMShellOracle, plate_mesh
    from femo_alpha_b200.shell_device import ShellDevice
    nodes, quads = plate_mesh(40, 30, 10.0, 7.5)
    nn = nodes.shape[0]
    x, y = nodes[:, 0], nodes[:, 1]
    inp = dict(thickness=0.05 * (1 + 0.4 * np.sin(x) * np.cos(y)), E=np.full(nn, 1e8), nu=np.full(nn, 0.3),
               density=np.full(nn, 10.0), F_solid=np.tile([0.0, 0.0, 5.0], nn))
    clamp = lambda xx: np.less(xx[0], 3e-16)
    o = RMShellOracle(nodes, quads, False, False)
    o.set_inputs(**inp); o.set_strong_bc(clamp)
    d = ShellDevice(nodes, quads)
    for k, v in inp.items():
        d.set_input(k, v)
    blocks = np.nonzero(clamp(d.block_coords().T))[0]
    d.set_strong_bc(np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()))
    info = d.enable_direct(refine=2)
    print("direct solver: %d dofs, %s" % (d.ndof, info))
    assert d.ndof > 8192 and not info["explicit_inverse"]
    w = o.solve_state()
    d.solve_state()
    print("state: relres %.2e, rel. error %.2e" % (d.last_solve["relres"], rel(d.get_state(), w)))
    assert d.last_solve["relres"] <= 1e-7 and rel(d.get_state(), w) <= 1e-8
    rng = np.random.default_rng(5)
    rhs = rng.standard_normal(d.ndof)
    lam = d.solve_adjoint(rhs)
    assert rel(lam, o.solve_adjoint(rhs)) <= 1e-8
    assert np.array_equal(lam, d.solve_adjoint(rhs))
    d.close()


def test_direct_solver_wingbox(built_lib):
    """BASELINE config 2 as shipped (no mesh hierarchy): the direct solver gives the state, the adjoint and the thickness
    gradient of the wing box to the north-star tolerance (block-Jacobi PCG needs ~1e5 iterations and stalls at 1e-7)."""
    from femo_alpha_b200.shell_device import ShellDevice
    m = np.load(os.path.join(GOLD, "wingbox_mesh.npz"))
    g = np.load(os.path.join(GOLD, "wingbox.npz"))
    nodes, quads = m["nodes"], m["quads"]
    d = ShellDevice(nodes, quads, True, False)
    nc = quads.shape[0]
    h = 0.05 * 0.0254
    d.set_input("thickness", np.full(nc, h)); d.set_input("E", np.full(nc, 73.1e9)); d.set_input("nu", np.full(nc, 0.33))
    d.set_input("density", np.full(nc, 2780.0))
    d.set_input("F_solid", np.tile([0.0, 0.0, -9.81 * 2780.0 * h], nodes.shape[0]))
    blocks = np.nonzero(d.block_coords()[:, 1] <= 1e-3)[0]
    d.set_strong_bc(np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()))
    info = d.enable_direct(refine=3)
    print("wing box direct solver: %s" % info)
    d.solve_state()
    e_w = rel(d.get_state(), g["w"])
    d.set_state(g["w"])
    lam = d.solve_adjoint(d.output_grad("compliance", "disp_solid"))
    e_l = rel(lam, g["lam"])
    total = d.output_grad("compliance", "thickness") - d.dRdf_T_apply("thickness", lam)
    e_g = rel(total, g["dCdh_total"])
    print("wing box: relres %.2e | u %.2e | lam %.2e | dC/dh %.2e (relative to the oracle's SuperLU solution)"
          % (d.last_solve["relres"], e_w, e_l, e_g))
    # u and lam agree with the oracle to 1e-10.  The total gradient  dC/dh - (dR/dh)^T lam  cancels ~2 digits, so the
    # oracle's own 1e-10 uncertainty in lam shows up as ~1e-8 there (unchanged whether or not the device refines further):
    # 1e-8 on u / lam is the measurable statement, the gradient is checked at 1e-7 like in test_gpu_golden.py
    assert e_w <= 1e-8 and e_l <= 1e-8 and e_g <= 1e-7
    d.close()


def test_multigrid_with_swept_coarse_level(built_lib):
    """V-cycle whose coarsest level (9 078 dofs > 8192) is solved by the sweeps instead of the explicit inverse."""
    from oracle.rm_shell_ref import RMShellOracle, plate_mesh
    from femo_alpha_b200.multigrid import ShellHierarchy
    nodes_c, quads_c = plate_mesh(24, 24, 10.0, 10.0)
    hier = ShellHierarchy(nodes_c, quads_c, 2)
    nodes, quads = hier.meshes[0]
    nn = nodes.shape[0]
    inp = dict(thickness=0.1 * (1 + 0.5 * np.sin(2 * np.pi * nodes[:, 0] / 10) * np.cos(2 * np.pi * nodes[:, 1] / 10)),
               E=np.full(nn, 1e8), nu=np.full(nn, 0.3), density=np.full(nn, 10.0), F_solid=np.tile([0.0, 0.0, 5.0], nn))
    for k, v in inp.items():
        hier.set_input(k, v)
    clamp = lambda xx: np.less(xx[0], 3e-16)
    hier.set_strong_bc(clamp)
    it, rr = hier.solve_state(rtol=1e-11)
    o = RMShellOracle(nodes, quads, False, False)
    o.set_inputs(**inp); o.set_strong_bc(clamp)
    e = rel(hier.fine.get_state(), o.solve_state())
    print("2-level V-cycle PCG over a swept coarse solve: %d iterations, relres %.1e, rel. error %.2e" % (it, rr, e))
    assert it <= 60 and e <= 1e-8
    hier.close()
