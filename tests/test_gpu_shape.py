"""GPU parity of the shape partials dR/d(uhat), dOutput/d(uhat) against the complex-step oracle
(oracle/shape_ref.py), through the C ABI (femo_dRdf_T_apply / femo_dRdf_apply / femo_output_grad with FEMO_UHAT).
Tolerance 1e-9 relative (north_star: gradients within 1e-8)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import shape_ref as S
from tests.problems import make_problem, oracle_for, device_for


def rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / np.abs(np.asarray(b)).max()


@pytest.mark.parametrize("name", ["warped", "skew_planar", "plate_varh", "plate_dg0"])
def test_shape_partials(name):
    p = make_problem(name)
    o, d = oracle_for(p), device_for(p)
    rng = np.random.default_rng(0)
    n = o.topo.ndof
    w, lam, du = rng.standard_normal(n), rng.standard_normal(n), rng.standard_normal(3 * o.topo.nn)
    d.set_state(w)
    assert rel(d.dRdf_T_apply("uhat", lam), S.dRduhat_T_apply(o, w, lam)) < 1e-9
    assert rel(d.dRdf_apply("uhat", du), S.dRduhat_apply(o, w, du)) < 1e-9
    for out in ("compliance", "mass", "elastic_energy"):
        assert rel(d.output_grad(out, "uhat"), S.doutput_duhat(o, out, w)) < 1e-9
    assert np.abs(d.output_grad("volume", "uhat")).max() == 0.0
    d.set_stress_params(1e-6, 6.0)
    assert rel(d.output_grad("pnorm_stress", "uhat"), S.dpnorm_duhat(o, w, 1e-6, 6.0)) < 1e-9
    cells = np.arange(o.topo.nc)[::2]
    tags = np.full(o.topo.nc, 3, dtype=np.int32); tags[cells] = 7
    d.set_cell_tags(tags)
    assert rel(d.pnorm_stress_grad("uhat", tag=7), S.dpnorm_duhat(o, w, 1e-6, 6.0, cells=cells)) < 1e-9


def test_total_shape_derivative_of_the_compliance():
    """dC/d(uhat) by the adjoint (dC/duhat - (dR/duhat)^T lam, A lam = dC/du) against finite differences of the
    oracle's solve -- the chain CSDL runs when `node_disp` is a design variable (rm_shell_model.py:432-438)."""
    p = make_problem("skew_planar")
    o, d = oracle_for(p), device_for(p)
    d.solve_state(rtol=1e-13, max_it=50000)
    lam = d.solve_adjoint(d.output_grad("compliance", "disp_solid"), rtol=1e-13, max_it=50000)
    total = d.output_grad("compliance", "uhat") - d.dRdf_T_apply("uhat", lam)
    rng = np.random.default_rng(2)
    du = rng.standard_normal(3 * o.topo.nn)
    eps = 1e-6
    u0 = o.uhat.copy()
    vals = []
    for sgn in (+1, -1):
        o.set_inputs(uhat=u0 + sgn * eps * du)
        vals.append(o.compliance(o.solve_state()))
    o.set_inputs(uhat=u0)
    fd = (vals[0] - vals[1]) / (2 * eps)
    assert np.isclose(total @ du, fd, rtol=1e-5), (total @ du, fd)


def test_penalty_shape(built_lib):
    """Shape partial of a penalty-BC context, facet terms included (Nanson's factor; on by default), in its own process
    (tests/penalty_shape_worker.py)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tests", "penalty_shape_worker.py")], cwd=root, capture_output=True,
                       text=True, timeout=900)
    sys.stdout.write(r.stdout[-3000:]); sys.stderr.write(r.stderr[-3000:])
    assert r.returncode == 0
