"""GPU parity of the generic 3-D forms on hexahedra (csrc/hex.cu, SURVEY 8 f4 / BASELINE config 5) against
oracle/hex_ref.py.  Every call goes through the C ABI (femo_hex_* and the shared femo_* entry points); the oracle is only
the checker.  Tolerances: matrix / vectors 1e-12, state / compliance / gradients 1e-8 (north_star)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-300)


def _pair(built_lib, form, shape=(3, 4, 2), distort=0.25, E0=5.0, nu=0.3):
    from oracle.hex_ref import HexOracle, box_mesh
    from femo_alpha_b200.solid_device import SolidDevice
    nodes, cells = box_mesh(*shape, distort=distort, seed=4)
    o = HexOracle(nodes, cells, form=form, E0=E0, nu=nu, simp_p=3.0)
    d = SolidDevice(nodes, cells, form=form, E0=E0, nu=nu, simp_p=3.0)
    rng = np.random.default_rng(11)
    rho = rng.uniform(0.05, 1.0, o.nc); f = rng.standard_normal((o.nn, 3))
    o.set_inputs(density=rho, body_force=f)
    d.set_input("density", rho); d.set_input("body_force", f)
    base = np.nonzero(nodes[:, 0] < 1e-12)[0]
    bc = (3 * base[:, None] + np.arange(3)).ravel()
    o.set_strong_bc(bc); d.set_strong_bc(bc)
    return o, d, rng


@pytest.mark.parametrize("form", ["elasticity", "vector_poisson"])
def test_matrix_load_and_residual(built_lib, form):
    o, d, rng = _pair(built_lib, form)
    K = o.stiffness()
    assert d.ndof == o.ndof and d.nc == o.nc and d.nnzb * 9 >= K.nnz
    assert abs(d.get_csr() - K).max() < 1e-12 * abs(K).max()
    assert abs(d.get_csr(with_bc=True) - o.with_bc(K)).max() < 1e-12 * abs(K).max()
    x = rng.standard_normal(o.ndof)
    assert rel(d.jacobian_apply(x), K @ x) < 1e-12
    assert rel(d.jacobian_apply(x, with_bc=True), o.with_bc(K) @ x) < 1e-12
    assert rel(d.assemble_load(), o.load()) < 1e-12
    assert rel(d.assemble_residual(x), o.residual(x)) < 1e-11
    assert np.array_equal(d.get_input("density"), o.rho) and np.array_equal(d.get_input("body_force"), o.f.ravel())
    assert np.array_equal(d.cell_blocks(), o.cells) and np.array_equal(d.block_coords(), o.x)
    # assembly is bitwise repeatable (no atomics)
    v1 = d.get_bsr()[2].copy()
    d.set_input("density", o.rho); d.assemble_jacobian()
    assert np.array_equal(v1, d.get_bsr()[2])
    d.close()


@pytest.mark.parametrize("form", ["elasticity", "vector_poisson"])
def test_solve_outputs_partials_and_adjoint_gradient(built_lib, form):
    o, d, rng = _pair(built_lib, form, shape=(4, 3, 3))
    u = o.solve()
    it, rr = d.solve_state(1e-13, 20000)
    assert 0 < it < 20000 and rr <= 1e-12
    assert rel(d.get_state(), u) < 1e-8
    assert np.isclose(d.output("compliance"), o.compliance(u), rtol=1e-8)
    assert np.isclose(d.output("volume"), o.volume(), rtol=1e-13)
    # partials at the oracle's state
    d.set_state(u)
    lam = rng.standard_normal(o.ndof); drho = rng.standard_normal(o.nc); df = rng.standard_normal(o.ndof)
    M = o.mass_matrix()
    assert rel(d.dRdf_T_apply("density", lam), o.dRdrho_T(u, lam)) < 1e-12
    assert rel(d.dRdf_apply("density", drho), o.dRdrho(u, drho)) < 1e-12
    assert rel(d.dRdf_T_apply("body_force", lam), -(M @ lam)) < 1e-12
    assert rel(d.dRdf_apply("body_force", df), -(M @ df)) < 1e-12
    assert rel(d.output_grad("compliance", "state"), o.load()) < 1e-12
    assert rel(d.output_grad("compliance", "body_force"), M @ u) < 1e-12
    assert np.abs(d.output_grad("compliance", "density")).max() == 0.0
    assert rel(d.output_grad("volume", "density"), o.vol) < 1e-13
    assert np.abs(d.output_grad("volume", "state")).max() == 0.0
    # adjoint total gradient of the compliance w.r.t. the cell densities
    lam = d.solve_adjoint(d.output_grad("compliance", "state"), 1e-13, 20000)
    g = d.output_grad("compliance", "density") - d.dRdf_T_apply("density", lam)
    assert rel(g, o.total_dcompliance_drho()) < 1e-8
    # forward solve inverts the operator with BCs
    y = d.solve_forward(o.with_bc(o.stiffness()) @ df, 1e-13, 20000)
    assert rel(y, df) < 1e-8
    d.close()


def test_unstructured_valence_and_errors(built_lib):
    from oracle.hex_ref import HexOracle, box_mesh
    from femo_alpha_b200.solid_device import SolidDevice
    th = np.deg2rad([0, 120, 240])
    ring = np.stack([np.cos(th), np.sin(th)], 1)
    mid = 0.5 * (ring + np.roll(ring, -1, axis=0))
    pts2 = np.vstack([[0, 0], ring, mid])
    nodes = np.vstack([np.c_[pts2, np.zeros(7)], np.c_[pts2, np.ones(7)]])
    quads = [[0, 6, 1, 4], [0, 4, 2, 5], [0, 5, 3, 6]]
    cells = np.array([q + [v + 7 for v in q] for q in quads], dtype=np.int32)
    o = HexOracle(nodes, cells, E0=2.0, nu=0.2)
    o.set_inputs(density=np.array([0.3, 0.6, 1.0]))
    d = SolidDevice(nodes, cells, E0=2.0, nu=0.2)
    d.set_input("density", o.rho)
    K = o.stiffness()
    assert abs(d.get_csr() - K).max() < 1e-12 * abs(K).max()
    # ten cells around one edge: 80 work items on the axis rows -> two staging passes of k_hex_assemble
    from tests.test_hex_host import fan_mesh
    n2, c2 = fan_mesh(10)
    o2 = HexOracle(n2, c2, E0=2.0, nu=0.2)
    o2.set_inputs(density=np.linspace(0.2, 1.0, o2.nc))
    d2 = SolidDevice(n2, c2, E0=2.0, nu=0.2)
    d2.set_input("density", o2.rho)
    K2 = o2.stiffness()
    assert abs(d2.get_csr() - K2).max() < 1e-12 * abs(K2).max()
    d2.close()
    # RM-shell entry points refuse a solid context; wrong sizes and inverted cells are reported
    from femo_alpha_b200 import _lib as L
    import ctypes as C
    v = C.c_double(0.0)
    assert d.lib.femo_output(d._ctx, 0, C.byref(v)) != 0 and b"solid" in d.lib.femo_last_error()
    x = np.zeros(5)
    assert d.lib.femo_set_input(d._ctx, 0, L.ptr(x), 5, 0) != 0
    with pytest.raises(RuntimeError):
        d.set_input("density", np.ones(2))
    with pytest.raises(RuntimeError):
        d.set_input("body_force", np.ones(5))
    with pytest.raises(RuntimeError):
        SolidDevice(nodes, cells[:, [1, 0, 3, 2, 5, 4, 7, 6]].copy())      # mirrored cells: negative Jacobians
    with pytest.raises(RuntimeError):
        SolidDevice(np.vstack([nodes, [[9.0, 9.0, 9.0]]]), cells)      # a node that belongs to no cell
    with pytest.raises(ValueError):
        SolidDevice(nodes, cells, form="stokes")
    d.close()


def test_cube_properties_at_scale(built_lib):
    """24^3 elasticity cube (46 875 dofs): rigid translations are in the kernel of K, K is symmetric, and the clamped
    problem solves to the requested tolerance; residual of the solution through the matrix-free partial."""
    from femo_alpha_b200.solid_device import SolidDevice, box_hexes
    n = 24
    nodes, cells = box_hexes(n, n, n)
    d = SolidDevice(nodes, cells, E0=1.0, nu=0.3)
    rng = np.random.default_rng(20261104)
    rho = rng.uniform(0.01, 1.0, d.nc)
    d.set_input("density", rho)
    f = np.zeros((d.nb, 3)); f[:, 2] = -1.0
    d.set_input("body_force", f)
    t = np.tile([0.3, -1.0, 2.0], d.nb)
    x = rng.standard_normal(d.ndof); y = rng.standard_normal(d.ndof)
    Kx = d.jacobian_apply(x)
    assert np.abs(d.jacobian_apply(t)).max() < 1e-12 * np.abs(Kx).max()
    assert np.isclose(y @ Kx, x @ d.jacobian_apply(y), rtol=1e-11)
    base = np.nonzero(nodes[:, 0] < 1e-12)[0]
    d.set_strong_bc((3 * base[:, None] + np.arange(3)).ravel())
    it, rr = d.solve_state(1e-10, 50000)
    assert rr <= 1e-10
    u = d.get_state()
    r = d.assemble_residual(u)
    r[(3 * base[:, None] + np.arange(3)).ravel()] = 0
    assert np.abs(r).max() < 1e-6 * np.abs(d.assemble_load()).max()
    # K(rho) u = sum_e rho_e^3 K0_e u_e = (dR/drho)(rho / 3) by Euler's theorem for the degree-3 SIMP law
    assert rel(d.dRdf_apply("density", rho / 3.0), d.jacobian_apply(u)) < 1e-11
    assert d.output("compliance") > 0 and np.isclose(d.output("volume"), rho.mean(), rtol=1e-12)
    d.close()


def test_generic_fea_operations_on_the_device(built_lib):
    """FEA.add_state / add_output with a named form, StateOperation + OutputOperation through FEAModel, FD totals."""
    import femo_alpha_b200.fea.named_forms as NF
    from femo_alpha_b200 import csdl_shim as csdl
    from femo_alpha_b200.csdl_alpha_opt.fea_model import FEAModel
    from femo_alpha_b200.fea.utils_dolfinx import locate_dofs_geometrical
    mesh = NF.createBoxMesh((0, 0, 0), (2.0, 1.0, 1.0), 6, 3, 3)
    fea, fn = NF.build_fea("elasticity3d", mesh, E0=10.0, nu=0.3, simp_p=3.0)
    fea.backend_options.update(rtol=1e-13)
    fea.add_strong_bc(None, [locate_dofs_geometrical(fn["u"].function_space, lambda x: np.less(x[0], 1e-12))])
    rng = np.random.default_rng(3)
    rec = csdl.Recorder(inline=True); rec.start()
    g = csdl.VariableGroup()
    f = np.zeros((fea.nn, 3)); f[:, 2] = -1.0
    g.density = csdl.Variable(value=rng.uniform(0.3, 1.0, fea.nel), name="density")
    g.body_force = csdl.Variable(value=f.ravel(), name="body_force")
    out = FEAModel(fea=[fea]).evaluate(g)
    rec.stop()
    sim = csdl.experimental.PySimulator(rec)
    rep = sim.check_totals(ofs=[out.compliance, out.volume], wrts=[g.density, g.body_force], step_size=1e-6, verbose=False)
    assert rep[out.compliance, g.density]["relative_error"] < 1e-5
    assert rep[out.compliance, g.body_force]["relative_error"] < 1e-5
    assert rep[out.volume, g.density]["relative_error"] < 1e-7
    fea.device.close()


def test_gradient_records(built_lib):
    """FEMO_HEX_GRAD=1: assembly from per-point gradient records (k_hex_geom2 + k_hex_assemble_g) gives the same matrix.
    Runs in its own process (tests/hex_grad_worker.py): the switch is read from the environment at set-up.  First run on
    hardware: profiles/r01v_hex_grad_worker.log (4e-16 against the oracle, bitwise equal to the default kernel)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, FEMO_HEX_GRAD="1")
    r = subprocess.run([sys.executable, os.path.join(root, "tests", "hex_grad_worker.py")], cwd=root, env=env, capture_output=True,
                       text=True, timeout=600)
    sys.stdout.write(r.stdout[-3000:]); sys.stderr.write(r.stderr[-3000:])
    assert r.returncode == 0
