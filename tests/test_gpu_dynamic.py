"""GPU parity of the dynamic (time-stepped) RM-shell path against oracle/dynamic_ref.py  (SURVEY 8 f1).

Every call goes through the C ABI (femo_dyn_*); the oracle is only the checker.  Tolerances: the state history and
every derived quantity to 1e-8 relative (north_star: displacement / gradient relative error <= 1e-8); the block-scalar
mass operator to 1e-12.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from tests.problems import make_problem, oracle_for, device_for

CASES = [("plate_varh", 6, False), ("plate_dg0", 3, True), ("warped", 6, True), ("skew_planar", 4, False)]


def _pair(name, deg, sw, n_steps=5, dt=0.01, solver="auto"):
    from oracle.dynamic_ref import DynamicShellOracle
    from femo_alpha_b200.dynamic_device import DynamicShellDevice
    p = make_problem(name)
    o = oracle_for(p, quadrature_degree=deg)
    d = device_for(p, quadrature_degree=deg)
    od = DynamicShellOracle(o, dt, n_steps, add_self_weight=sw)
    dd = DynamicShellDevice(d, dt, n_steps, add_self_weight=sw, solver=solver, rtol=1e-13)
    rng = np.random.default_rng(7)
    f_hist = np.stack([p["F_solid"] * (0.5 + 0.1 * i) + 0.05 * rng.standard_normal(p["F_solid"].size)
                       for i in range(n_steps)])          # one row short: the last row repeats (plate_sim.py:314-319)
    return p, o, d, od, dd, f_hist


def rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-300)


@pytest.mark.parametrize("name,deg,sw", CASES)
def test_mass_operator(name, deg, sw):
    p, o, d, od, dd, f_hist = _pair(name, deg, sw)
    dd.setup()
    M = od.mass_matrix()
    rng = np.random.default_rng(1)
    x = rng.standard_normal(o.topo.ndof)
    assert rel(dd.mass_apply(x), M @ x) < 1e-12
    # the effective operator the solvers see: A = 2/dt^2 M + K/2 with BCs
    od.operators()
    assert rel(d.jacobian_apply(x, with_bc=True), od.A @ x) < 1e-11
    assert rel(d.jacobian_apply(x, with_bc=False), od.Acur @ x) < 1e-11


@pytest.mark.parametrize("name,deg,sw", CASES)
def test_history_energy_and_sweeps(name, deg, sw):
    p, o, d, od, dd, f_hist = _pair(name, deg, sw)
    w_ref = od.solve(f_hist)
    dd.solve(f_hist)
    w = dd.history()
    assert dd.info()["solver"] == "dense"
    assert rel(w, w_ref) < 1e-8
    assert rel(dd.energies(), od.energy) < 1e-8
    assert np.isclose(dd.total_strain_energy(), od.total_strain_energy(w_ref), rtol=1e-8)
    # evaluate everything else at the oracle's history so that the comparisons do not compound
    dd.set_history(w_ref)
    rng = np.random.default_rng(3)
    dout = rng.standard_normal(w_ref.shape)
    lam_ref = od.adjoint_recurrence(dout)
    lam = dd.solve_adjoint(dout)
    assert rel(lam, lam_ref) < 1e-8
    assert np.abs(lam[0]).max() == 0.0 and np.abs(lam[:, o.bc_dofs]).max() == 0.0
    x = rng.standard_normal(w_ref.shape)
    assert rel(dd.dRdw_apply(x), od.dRdw_apply(x)) < 1e-10
    assert rel(dd.dRdw_apply(x, transpose=True), od.dRdw_apply(x, transpose=True)) < 1e-10
    assert rel(dd.solve_forward(dout), od.forward_recurrence(dout)) < 1e-8
    assert rel(dd.dRdh_T_apply(lam_ref), od.dRdh_T(w_ref, lam_ref)) < 1e-9
    dh = rng.standard_normal(o.n_mat) * p["thickness"]
    assert rel(dd.dRdh_apply(dh), od.dRdh(w_ref, dh)) < 1e-9
    assert rel(dd.total_strain_energy_grad("thickness"), od.dtotal_strain_energy_dh(w_ref)) < 1e-9
    assert rel(dd.total_strain_energy_grad("state"), od.dtotal_strain_energy_dw(w_ref)) < 1e-10
    assert np.isclose(d.output("volume"), od.volume(), rtol=1e-12)
    assert rel(d.output_grad("volume", "thickness"), od.dvolume_dh()) < 1e-12


def test_pcg_stepping_matches_dense_and_is_reproducible():
    p, o, d, od, dd, f_hist = _pair("plate_varh", 6, False, n_steps=4, solver="pcg")
    w_ref = od.solve(f_hist)
    its = dd.solve(f_hist)
    assert dd.info()["solver"] == "pcg" and its > 0
    w1 = dd.history()
    assert rel(w1, w_ref) < 1e-8
    dd.solve(f_hist)
    assert np.array_equal(w1, dd.history())            # no atomics anywhere: bitwise repeatable
    lam = dd.solve_adjoint(od.dtotal_strain_energy_dw(w_ref))
    dd.set_history(w_ref)
    assert rel(lam, od.adjoint_recurrence(od.dtotal_strain_energy_dw(w_ref))) < 1e-7


def test_thickness_change_rebuilds_the_operator():
    p, o, d, od, dd, f_hist = _pair("plate_varh", 6, True, n_steps=3)
    dd.solve(f_hist)
    h2 = p["thickness"] * 1.3
    d.set_input("thickness", h2); o.set_inputs(thickness=h2)
    d.set_input("density", p["density"] * 0.7); o.set_inputs(density=p["density"] * 0.7)
    w_ref = od.solve(f_hist)
    dd.solve(f_hist)
    assert rel(dd.history(), w_ref) < 1e-8


def test_multigrid_stepping_on_a_refined_plate():
    from femo_alpha_b200.mesh import createRectangleMesh, refine_uniform
    from femo_alpha_b200.dynamic_rm_shell.plate_sim import PlateSim
    from oracle.dynamic_ref import DynamicShellOracle
    from oracle.rm_shell_ref import RMShellOracle
    mesh = refine_uniform(createRectangleMesh((0., 0.), (10., 2.), 10, 2), 2)       # 40 x 8 cells
    nn = mesh.geometry.x.shape[0]
    sims = {s: PlateSim(mesh, 1e8, 0.3, 10., 0.01, 3, quad_deg=6, solver=s, rtol=1e-12) for s in ("dense", "multigrid")}
    f = np.tile([0., 0., -5.], nn)[None, :]
    hist = {}
    for s, ps in sims.items():
        ps.update_t(np.full(nn, 0.1)); ps.update_f_history(f)
        hist[s] = ps.solve_dynamic_problem(ps.SVK_residual())
        assert ps._dyn.info()["solver"] == s
    assert rel(hist["multigrid"], hist["dense"]) < 1e-8
    assert sims["multigrid"]._dyn.last_iters <= 3 * 100      # V-cycle counts (13 per step at 3e5 dofs), not O(1e3..1e4) Jacobi iterations
    o = RMShellOracle(mesh.geometry.x, mesh.cells, quadrature_degree=6)
    o.set_inputs(thickness=np.full(nn, 0.1), E=np.full(nn, 1e8), nu=np.full(nn, 0.3), density=np.full(nn, 10.), F_solid=f[0])
    o.set_strong_bc(lambda x: np.isclose(x[0], 0., atol=1e-6))
    w_ref = DynamicShellOracle(o, 0.01, 3).solve(f)
    assert rel(hist["dense"].T, w_ref) < 1e-8


def test_operator_surface_check_totals():
    """The reference's usage (ex_simple_dynamic_shell_opt.py:97-140): StateOperation -> TotalStrainEnergyOperation,
    VolumeOperation; adjoint totals against finite differences through the CSDL (stand-in) simulator."""
    from femo_alpha_b200.csdl_compat import csdl
    from femo_alpha_b200.mesh import createRectangleMesh
    from femo_alpha_b200.dynamic_rm_shell.plate_sim import PlateSim
    from femo_alpha_b200.dynamic_rm_shell.state_operation_dynamic import StateOperation
    from femo_alpha_b200.dynamic_rm_shell.total_strain_energy_operation import TotalStrainEnergyOperation
    from femo_alpha_b200.dynamic_rm_shell.volume_operation import VolumeOperation
    mesh = createRectangleMesh((0., 0.), (10., 2.), 5, 2)
    nn = mesh.geometry.x.shape[0]
    Nsteps, dt = 6, 0.02
    ps = PlateSim(mesh, 1e8, 0.3, 10., dt, Nsteps, add_self_weight=True)
    rec = csdl.Recorder(inline=True); rec.start()
    rng = np.random.default_rng(11)
    thickness = csdl.Variable(value=0.1 * (1 + 0.2 * rng.random(nn)), name='thickness')
    fh = np.zeros((Nsteps + 1, nn * 3))
    for i in range(Nsteps + 1):
        fh[i, 2::3] = 0.1 * 50. * (1 - np.cos(2 * np.pi * i / Nsteps))            # 1 - cos gust (ex_...:66-95)
    force_history = csdl.Variable(value=fh, name='force_history')
    inputs = csdl.VariableGroup(); inputs.thickness = thickness; inputs.force_history = force_history
    disp_history = StateOperation(plate_sim=ps, record=False).evaluate(inputs)
    assert disp_history.value.shape == (ps.fe_dofs * (Nsteps + 1),)
    g2 = csdl.VariableGroup(); g2.thickness = thickness; g2.disp_history = disp_history
    tse = TotalStrainEnergyOperation(ps).evaluate(g2)
    g3 = csdl.VariableGroup(); g3.thickness = thickness
    vol = VolumeOperation(ps).evaluate(g3)
    rec.stop()
    assert tse.value[0] > 0 and np.isclose(vol.value[0], 20. * thickness.value.mean(), rtol=0.05)
    sim = csdl.experimental.PySimulator(rec)
    rep = sim.check_totals(ofs=[tse, vol], wrts=[thickness], step_size=1e-6, verbose=False)
    assert rep[tse, thickness]["relative_error"] < 1e-5
    assert rep[vol, thickness]["relative_error"] < 1e-6
    with pytest.raises(ValueError):
        StateOperation(plate_sim=ps, record=False).apply_inverse_jacobian(
            {'thickness': thickness.value}, {'disp_history': disp_history.value}, {}, {'disp_history': disp_history.value}, 'sideways')


def test_energy_balance_at_scale():
    """Size-independent property at a size the oracle cannot reach (320x64 cells, 3.1e5 dofs, V-cycle PCG): for a
    constant load the implicit midpoint rule conserves  kinetic + strain energy - work  exactly (to the solver
    tolerance), level by level."""
    from femo_alpha_b200.mesh import createRectangleMesh, refine_uniform
    from femo_alpha_b200.dynamic_rm_shell.plate_sim import PlateSim
    mesh = refine_uniform(createRectangleMesh((0., 0.), (10., 2.), 40, 8), 3)
    nn = mesh.geometry.x.shape[0]
    n_steps, dt = 8, 0.015
    ps = PlateSim(mesh, 1e8, 0.3, 10., dt, n_steps, quad_deg=6, solver="multigrid", rtol=1e-12)
    f = np.tile([0., 0., -5.], nn)[None, :]
    ps.update_t(np.full(nn, 0.1)); ps.update_f_history(f)
    w = ps.solve_dynamic_problem(ps.SVK_residual()).T                 # (levels, ndof)
    assert ps._dyn.info()["solver"] == "multigrid" and ps.fe_dofs == 310662
    strain = ps._dyn.energies()
    F = ps.device.assemble_load()
    F[ps.bc_dofs] = 0.0
    v = np.zeros(ps.fe_dofs)
    bal = []
    for i in range(1, n_steps + 1):
        v = 2.0 / dt * (w[i] - w[i - 1]) - v
        bal.append(0.5 * v @ ps._dyn.mass_apply(v) + strain[i] - F @ w[i])
    scale = abs(F @ w[-1])
    # w^T K w loses ~8 digits to the membrane / bending / shear scale separation at this element size (cf. test_gpu_golden)
    assert scale > 0 and np.ptp(bal) < 1e-5 * scale and abs(bal[0]) < 1e-5 * scale, (bal, scale)
    ps.hierarchy.close()


def test_dynamic_edge_cases():
    """Single step, over-long and all-zero load histories, history round trip, argument errors."""
    from oracle.dynamic_ref import DynamicShellOracle
    from femo_alpha_b200.dynamic_device import DynamicShellDevice
    p = make_problem("plate_dg0")
    o, d = oracle_for(p), device_for(p)
    # one step; the history has more rows than time levels: only rows 0..1 are used
    od, dd = DynamicShellOracle(o, 0.02, 1), DynamicShellDevice(d, 0.02, 1)
    f_long = np.stack([p["F_solid"] * (1.0 + i) for i in range(5)])
    w_ref = od.solve(f_long)
    assert dd.solve(f_long) == 1
    assert rel(dd.history(), w_ref) < 1e-8 and dd.energies()[0] == 0.0
    # zero loads: the plate stays at rest, and so do the adjoints of a zero right-hand side
    dd.solve(np.zeros((1, d.n_f)))
    assert np.abs(dd.history()).max() == 0.0 and dd.total_strain_energy() == 0.0
    assert np.abs(dd.solve_adjoint(np.zeros((2, d.ndof)))).max() == 0.0
    # history round trip through the device
    rng = np.random.default_rng(4)
    h = rng.standard_normal((2, d.ndof))
    dd.set_history(h)
    assert np.array_equal(dd.history(), h)
    with pytest.raises(ValueError):
        dd.solve(np.zeros((2, d.n_f + 1)))
    dd.close()
    with pytest.raises(RuntimeError, match="multigrid"):
        DynamicShellDevice(d, 0.02, 2, solver="multigrid").solve(f_long)
    with pytest.raises(RuntimeError):
        DynamicShellDevice(d, -1.0, 2)
    # the context is back in static mode once the handles are gone: K, not A
    import gc; gc.collect()
    x = rng.standard_normal(d.ndof)
    assert rel(d.jacobian_apply(x), o.assemble_dRdu() @ x) < 1e-12
