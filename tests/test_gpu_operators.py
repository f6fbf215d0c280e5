"""The reference's operator surface (RMShellModel / FEAModel / StateOperation / OutputOperation) on the
CUDA back-end: forward evaluation against the oracle, adjoint totals against finite differences
(the reference's own acceptance checks)."""
import numpy as np
import pytest

from tests.problems import make_problem, oracle_for

pytestmark = pytest.mark.gpu

DOLFIN_EPS = 3E-16


def ClampedBoundary(x):
    return np.less(x[0], 0.0 + DOLFIN_EPS)


def _run_model(nx, ny, ewm=False, penalty=True, E_val=1e8, nu_val=0.3, h_val=0.1, f_d=5.0, rho_val=10.0):
    from femo_alpha_b200.csdl_compat import csdl
    from femo_alpha_b200.mesh import createRectangleMesh
    from femo_alpha_b200.rm_shell.rm_shell_model import RMShellModel
    mesh = createRectangleMesh((0.0, 0.0), (10.0, 2.0), nx, ny)
    nel = mesh.topology.index_map(mesh.topology.dim).size_local
    nn = mesh.topology.index_map(0).size_local
    rec = csdl.Recorder(inline=True)
    rec.start()
    pressure = csdl.Variable(value=np.zeros((nn, 3)), name='force_vector')
    pressure.value[:, 2] = f_d
    node_disp = csdl.Variable(value=np.zeros((nn, 3)), name='node_disp')
    n = nel if ewm else nn
    thickness = csdl.Variable(value=h_val * np.ones(n), name='thickness')
    E = csdl.Variable(value=E_val * np.ones(n), name='E')
    nu = csdl.Variable(value=nu_val * np.ones(n), name='nu')
    density = csdl.Variable(value=rho_val * np.ones(n), name='density')
    model = RMShellModel(mesh, shell_bc_func=ClampedBoundary, element_wise_material=ewm, PENALTY_BC=penalty,
                         record=False)
    out = model.evaluate(pressure, thickness, E, nu, density, node_disp, debug_mode=False, is_pressure=True)
    rec.stop()
    return mesh, model, out, dict(thickness=thickness, E=E, nu=nu, density=density, pressure=pressure), rec, csdl


def test_simple_shell_forward(built_lib):
    """ex_simple_shell_opt.py:38-45,100-105: 2x10 plate, 4x20 quads; EB deflection 0.75 beside the FE value."""
    mesh, model, out, v, rec, csdl = _run_model(20, 4)
    p = make_problem("plate_4x20")
    o = oracle_for(p, penalty=True)
    w = o.solve_state()
    tip = np.abs(out.disp_extracted.value).max()
    assert abs(tip - np.abs(o.disp_extracted(w)).max()) <= 1e-7 * tip
    EB = 5.0 * 2.0 * 10.0 ** 4 / (8 * 1e8 * 2.0 * 0.1 ** 3 / 12)
    assert 0.9 * EB < tip < 1.0 * EB                      # plate (nu=0.3) is a little stiffer than the beam
    assert abs(out.mass.value[0] - 20.0) <= 1e-10
    assert abs(out.compliance.value[0] - o.compliance(w)) <= 1e-7 * o.compliance(w)
    assert out.disp_solid.value.shape == (o.topo.ndof,)
    assert out.stress.value.shape == (4 * 80,)
    assert np.isfinite(out.aggregated_stress.value).all() and out.aggregated_stress.value[0] > 0
    ref_agg = 1 / 1e-6 * o.pnorm_stress(w) ** (1 / 100)
    assert abs(out.aggregated_stress.value[0] - ref_agg) <= 1e-6 * ref_agg
    # post-processing access used by the example (ex_simple_shell_opt.py:142-144)
    wfn = model.fea.states_dict['disp_solid']['function']
    assert abs(np.abs(wfn.sub(0).collapse().x.array).max() - np.abs(w[:3 * o.topo.nq2]).max()) <= 1e-7 * tip
    assert (model.fea.nel, model.fea.nn) == (80, 105)


@pytest.mark.parametrize("ewm,penalty", [(False, False), (True, True)])
def test_check_totals(ewm, penalty, built_lib):
    """sim.check_totals([compliance],[thickness]) (ex_simple_shell_opt.py:109-112): adjoint vs FD."""
    mesh, model, out, v, rec, csdl = _run_model(6, 2, ewm=ewm, penalty=penalty, E_val=1e6)
    sim = csdl.experimental.PySimulator(rec)
    rep = sim.check_totals([out.compliance, out.mass, out.elastic_energy], [v["thickness"]], step_size=2e-5,
                           verbose=False)
    for key, r in rep.items():
        assert r["relative_error"] <= 5e-6, (key[0].name, r["relative_error"])
    rep = sim.check_totals([out.aggregated_stress], [v["thickness"]], step_size=2e-5, verbose=False)
    for key, r in rep.items():
        assert r["relative_error"] <= 5e-5, (key[0].name, key[1].name, r["relative_error"])


def test_operator_errors(built_lib):
    from femo_alpha_b200.mesh import createRectangleMesh
    from femo_alpha_b200.rm_shell.rm_shell_model import RMShellModel
    mesh = createRectangleMesh((0.0, 0.0), (1.0, 1.0), 2, 2)
    with pytest.raises(ValueError):
        RMShellModel(mesh, shell_bc_func=None)
    model = RMShellModel(mesh, shell_bc_func=ClampedBoundary, record=False)
    with pytest.raises(ValueError):
        model.fea.add_input('thickness', model.fea.inputs_dict['E']['function'])
    from femo_alpha_b200.csdl_alpha_opt.state_operation import StateOperation
    op = StateOperation(model.fea, ['thickness', 'F_solid', 'E', 'nu', 'uhat'], 'disp_solid')
    with pytest.raises(ValueError):
        op.apply_inverse_jacobian({}, {}, {}, {}, 'sideways')
    with pytest.raises(ValueError):
        op.compute_jacvec_product({}, {}, {}, {}, {}, 'sideways')
