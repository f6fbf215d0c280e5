"""Multigrid-preconditioned state / adjoint solves (csrc/mg.cu) against the oracle's direct solve."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300)


@pytest.mark.parametrize("levels,deg", [(2, 3), (3, 3), (3, 2)])
def test_mg_solves_match_direct(built_lib, levels, deg):
    from femo_alpha_b200.multigrid import ShellHierarchy, refine_quads
    from oracle.rm_shell_ref import RMShellOracle, plate_mesh
    rng = np.random.default_rng(11)
    nodes, quads = plate_mesh(6, 4, 3.0, 2.0)
    nodes[:, :2] += 0.03 * rng.standard_normal((nodes.shape[0], 2)) * (nodes[:, :1] > 1e-9)
    thick = lambda x: 0.05 * (1.0 + 0.3 * np.sin(2.0 * x[0]) * np.cos(1.5 * x[1]))
    clamp = lambda x: np.less(x[0], 3e-16)
    h = ShellHierarchy(nodes, quads, levels, cheb_degree=deg)
    nf, qf = h.meshes[0]
    nn = nf.shape[0]
    F = np.tile([0.0, 0.0, 5.0], nn) + 0.1 * rng.standard_normal(3 * nn)
    h.set_input("thickness", thick)
    h.set_input("E", 1e6); h.set_input("nu", 0.3); h.set_input("density", 2.0)
    h.set_input("F_solid", F)
    h.set_strong_bc(clamp)
    o = RMShellOracle(nf, qf, False, False, 6)
    o.set_inputs(thickness=thick(nf.T), E=np.full(nn, 1e6), nu=np.full(nn, 0.3), density=np.full(nn, 2.0), F_solid=F)
    o.set_strong_bc(clamp)
    w = o.solve_state()
    it, rr = h.solve_state(rtol=1e-11)
    assert rr <= 1e-11 and it <= 150
    assert rel(h.fine.get_state(), w) <= 1e-8
    # block-Jacobi PCG on the same level needs far more iterations
    h.fine._mg = None
    itj, _ = h.fine.solve_state(rtol=1e-11)
    h.fine._mg = h
    assert itj > 5 * it
    rhs = rng.standard_normal(o.topo.ndof)
    lam = h.solve_adjoint(rhs, rtol=1e-11)
    assert rel(lam, o.solve_adjoint(rhs)) <= 1e-8
    # inputs change -> the preconditioner is rebuilt on the next solve
    h.set_input("thickness", lambda x: 1.3 * thick(x))
    o.set_inputs(thickness=1.3 * thick(nf.T))
    w2 = o.solve_state()
    h.solve_state(rtol=1e-11)
    assert rel(h.fine.get_state(), w2) <= 1e-8
    h.close()


def test_mg_junction_mesh(built_lib):
    """Non-manifold hierarchy: plate with a wall along one edge line, clamped at x = 0."""
    from femo_alpha_b200.multigrid import ShellHierarchy
    from oracle.rm_shell_ref import RMShellOracle, plate_mesh
    base, quads = plate_mesh(4, 3, 2.0, 1.5)
    n2, q2 = plate_mesh(4, 2, 2.0, 1.0)
    wall = np.stack([n2[:, 0], np.zeros(len(n2)), n2[:, 1]], 1)
    on_line = np.isclose(wall[:, 2], 0.0)
    remap = np.full(len(wall), -1)
    for i in np.nonzero(on_line)[0]:
        remap[i] = int(np.nonzero(np.isclose(base[:, 0], wall[i, 0]) & np.isclose(base[:, 1], 0.0))[0][0])
    new = np.nonzero(~on_line)[0]
    remap[new] = len(base) + np.arange(new.size)
    nodes = np.concatenate([base, wall[new]])
    quads = np.concatenate([quads, remap[q2]])
    clamp = lambda x: np.less(x[0], 3e-16)
    h = ShellHierarchy(nodes, quads, 3)
    nf, qf = h.meshes[0]
    nn = nf.shape[0]
    for k, v in (("thickness", 0.04), ("E", 2e6), ("nu", 0.25), ("density", 1.0)):
        h.set_input(k, v)
    F = np.tile([0.3, -0.2, 1.0], nn)
    h.set_input("F_solid", F)
    h.set_strong_bc(clamp)
    o = RMShellOracle(nf, qf, False, False, 6)
    o.set_inputs(thickness=np.full(nn, 0.04), E=np.full(nn, 2e6), nu=np.full(nn, 0.25), density=np.full(nn, 1.0), F_solid=F)
    o.set_strong_bc(clamp)
    w = o.solve_state()
    it, rr = h.solve_state(rtol=1e-11)
    assert rr <= 1e-11 and it <= 300
    assert rel(h.fine.get_state(), w) <= 1e-8
    h.close()


def test_operator_surface_uses_multigrid(built_lib):
    """RMShellPDE on a mesh from refine_uniform attaches the hierarchy: FEA-level solves go through the V-cycle PCG and
    give the same state / adjoint as the block-Jacobi path."""
    from femo_alpha_b200.mesh import createRectangleMesh, refine_uniform
    from femo_alpha_b200.rm_shell.rm_shell_pde import RMShellPDE
    coarse = createRectangleMesh((0.0, 0.0), (10.0, 2.0), 10, 2)
    mesh = refine_uniform(coarse, 2)
    assert len(mesh.mg_meshes) == 3 and mesh.cells.shape[0] == 320
    pde = RMShellPDE(mesh)
    d = pde.device
    assert pde.hierarchy is not None and d._mg is pde.hierarchy
    nn = d.nn
    for k, v in (("thickness", np.full(nn, 0.1)), ("E", np.full(nn, 1e8)), ("nu", np.full(nn, 0.3)),
                 ("density", np.full(nn, 10.0)), ("F_solid", np.tile([0.0, 0.0, 5.0], nn))):
        d.set_input(k, v)
    on = d.block_coords()[:, 0] < 3e-16
    blocks = np.nonzero(on)[0]
    d.set_strong_bc(np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()))
    it, rr = d.solve_state(rtol=1e-11)
    u_mg = d.get_state().copy()
    rhs = d.output_grad("compliance", "disp_solid")
    lam_mg = d.solve_adjoint(rhs, rtol=1e-11)
    assert it <= 150 and d.last_solve["iters"] <= 150
    d._mg = None
    itj, _ = d.solve_state(rtol=1e-11)
    assert itj > 5 * it
    assert rel(u_mg, d.get_state()) <= 1e-8
    assert rel(lam_mg, d.solve_adjoint(rhs, rtol=1e-11)) <= 1e-8
    tip = np.abs(u_mg.reshape(-1, 3)[:nn, 2]).max()
    assert abs(tip - 0.75) < 0.02                     # Euler-Bernoulli tip deflection, ex_simple_shell_opt.py:100-105
    d._mg = pde.hierarchy
    pde.hierarchy.close()


def test_mg_refined_wingbox(built_lib):
    """BASELINE config 2, synthetically refined: the reference's wing-box mesh (warped skins, rib / spar junctions,
    element-wise material) subdivided once; two-level V-cycle PCG against block-Jacobi PCG and the true residual."""
    import os
    from femo_alpha_b200.mesh import ShellMesh, refine_uniform
    from femo_alpha_b200.rm_shell.rm_shell_pde import RMShellPDE
    m = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "wingbox_mesh.npz"))
    fine = refine_uniform(ShellMesh(m["nodes"], m["quads"]), 1)
    pde = RMShellPDE(fine, element_wise_material=True)
    d = pde.device
    assert pde.hierarchy is not None and d.sizes.all_planar == 0 and d.nc == 4 * 2244
    h = 0.05 * 0.0254
    d.set_input("thickness", np.full(d.nc, h)); d.set_input("E", np.full(d.nc, 73.1e9)); d.set_input("nu", np.full(d.nc, 0.33))
    d.set_input("density", np.full(d.nc, 2780.0))
    d.set_input("F_solid", np.tile([0.0, 0.0, -9.81 * 2780.0 * h], d.nn))
    on = d.block_coords()[:, 1] <= 1e-3
    blocks = np.nonzero(on)[0]
    d.set_strong_bc(np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()))
    it, rr = d.solve_state(rtol=1e-12)
    u = d.get_state().copy()
    F = d.assemble_load()
    free = np.ones(d.ndof, bool); free[np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel())] = False
    res = (d.jacobian_apply(u, with_bc=True) - F)[free]
    print("wing box x4: %d dof, %d V-cycle PCG iterations, relres %.1e, true residual %.1e" %
          (d.ndof, it, rr, np.linalg.norm(res) / np.linalg.norm(F[free])))
    # kappa(K) ~ 1e12 on this thin-walled box: the attainable true residual in fp64 is ~1e-7 |F|
    assert it <= 1000 and np.linalg.norm(res) <= 1e-5 * np.linalg.norm(F[free])
    d._mg = None
    itj, rrj = d.solve_state(rtol=1e-12, max_it=400000)
    print("block-Jacobi PCG: %d iterations, relres %.1e" % (itj, rrj))
    assert rel(u, d.get_state()) <= 1e-5 and itj > 4 * it
    d._mg = pde.hierarchy
    pde.hierarchy.close()


@pytest.mark.parametrize("h0", [0.3, 0.1])
def test_mg_iterative_refinement(built_lib, h0):
    """Plate 10 x 2, 64 x 32 cells.  femo_mg_set_refine adds corrections x += PCG(b - A x) on a double-double residual and makes
    femo_mg_solve report the TRUE relative residual, which is bounded below by eps |A| |x| / |b| (x is stored in fp64).
    h ~ 0.3: everything agrees with the oracle's refined direct solve to 1e-9.  h ~ 0.1 (kappa ~ 1e9): the difference to the
    oracle stays at ~1e-8 with or without refinement -- it is the sensitivity of the solution to last-bit differences
    between the two assembled matrices (both within 1e-12 of each other), not solver error; asserted as such."""
    from femo_alpha_b200.multigrid import ShellHierarchy
    from oracle.rm_shell_ref import RMShellOracle, plate_mesh
    nodes_c, quads_c = plate_mesh(8, 4, 10.0, 2.0)
    h = ShellHierarchy(nodes_c, quads_c, 4)
    nodes, quads = h.meshes[0]
    nn = nodes.shape[0]
    inp = dict(thickness=h0 * (1 + 0.3 * np.sin(2 * np.pi * nodes[:, 0] / 10)), E=np.full(nn, 1e8), nu=np.full(nn, 0.3),
               density=np.full(nn, 10.0), F_solid=np.tile([0.0, 0.0, 5.0], nn))
    for k, v in inp.items():
        h.set_input(k, v)
    clamp = lambda xx: np.less(xx[0], 3e-16)
    h.set_strong_bc(clamp)
    o = RMShellOracle(nodes, quads, False, False)
    o.set_inputs(**inp); o.set_strong_bc(clamp)
    w = o.solve_state()
    emax = lambda a: np.abs(a - w).max() / np.abs(w).max()
    it0, rr0 = h.solve_state(rtol=1e-11)
    e0 = emax(h.fine.get_state())
    h.set_refine(1)
    it1, rr1 = h.solve_state(rtol=1e-11)
    e1 = emax(h.fine.get_state())
    print("h ~ %.1f | plain: %d iterations, recurrence relres %.1e, max error %.2e | one refinement step: %d iterations, "
          "true relres %.1e, max error %.2e" % (h0, it0, rr0, e0, it1, rr1, e1))
    tol = 1e-9 if h0 > 0.2 else 5e-8
    assert e0 <= tol and e1 <= tol and e1 <= 1.5 * e0 + 1e-12
    assert rr1 <= (1e-8 if h0 > 0.2 else 1e-7) and it0 < it1 <= 2 * it0       # measured 2.0e-9 / 1.5e-8: eps |A| |x| / |b|
    lam = h.solve_adjoint(o.dcompliance_du(w), rtol=1e-11)
    lam_o = o.solve_adjoint(o.dcompliance_du(w))
    assert np.abs(lam - lam_o).max() / np.abs(lam_o).max() <= tol
    h.close()
