"""Host logic of the dynamic operator mirror (femo_alpha_b200/dynamic_rm_shell/*) without a GPU: the two device
handles are replaced by oracle-backed stand-ins with the same method surface, so what is exercised here is the
reference-facing wiring -- history layout (ravel order 'F', dynamic_rm_shell/utils.py:9-16), argument names, the
apply_inverse_jacobian / compute_jacvec_product contract and the CSDL reverse sweep -- not the arithmetic (the GPU
tier checks that through the C ABI, tests/test_gpu_dynamic.py)."""
import numpy as np
import pytest

from oracle.dynamic_ref import DynamicShellOracle
from oracle.rm_shell_ref import RMShellOracle


class FakeShellDevice:
    def __init__(self, nodes, quads, ewm=False, ewp=False, quadrature_degree=6, device=0):
        self.o = RMShellOracle(nodes, quads, ewm, ewp, quadrature_degree)
        t = self.o.topo
        self.nn, self.nc, self.nq2, self.ndof = t.nn, t.nc, t.nq2, t.ndof
        self.n_mat, self.n_f = self.o.n_mat, self.o.n_f
        self._children, self._w = [], np.zeros(t.ndof)

    def set_input(self, name, v):
        self.o.set_inputs(**{name: np.asarray(v, dtype=np.float64)})

    def block_coords(self):
        return self.o.topo.block_coords()

    def set_strong_bc(self, dofs):
        self.o.bc_dofs = np.asarray(dofs, dtype=np.int64)

    def output(self, name):
        assert name == "volume"
        return DynamicShellOracle(self.o, 1.0, 1).volume()

    def output_grad(self, name, arg):
        assert (name, arg) == ("volume", "thickness")
        return DynamicShellOracle(self.o, 1.0, 1).dvolume_dh()

    def close(self):
        pass


class FakeDynamicShellDevice:
    def __init__(self, device, dt, n_steps, add_self_weight=False, g_factor=None, solver="auto", rtol=1e-12, max_it=20000):
        self.dev, self.levels = device, n_steps + 1
        self.od = DynamicShellOracle(device.o, dt, n_steps, add_self_weight, g_factor)
        self.w = np.zeros((self.levels, device.ndof))

    def _ops(self):
        self.od.operators()

    def solve(self, f):
        self.w = self.od.solve(f); return self.levels - 1

    def history(self):
        return self.w.copy()

    def set_history(self, w):
        self.w = np.array(w, dtype=np.float64).reshape(self.levels, -1); self._ops()

    def energies(self):
        return np.array([0.5 * x @ (self.od.K @ x) for x in self.w])

    def solve_adjoint(self, d):
        return self.od.adjoint_recurrence(np.asarray(d).reshape(self.levels, -1))

    def solve_forward(self, d):
        return self.od.forward_recurrence(np.asarray(d).reshape(self.levels, -1))

    def dRdw_apply(self, x, transpose=False):
        return self.od.dRdw_apply(np.asarray(x).reshape(self.levels, -1), transpose)

    def dRdh_T_apply(self, lam):
        return self.od.dRdh_T(self.w, np.asarray(lam).reshape(self.levels, -1))

    def dRdh_apply(self, dh):
        return self.od.dRdh(self.w, np.asarray(dh))

    def total_strain_energy(self):
        return self.od.total_strain_energy(self.w)

    def total_strain_energy_grad(self, arg):
        return self.od.dtotal_strain_energy_dh(self.w) if arg == "thickness" else self.od.dtotal_strain_energy_dw(self.w)

    def info(self):
        return dict(solver="oracle", forward_ms=0.0, adjoint_ms=0.0)

    def close(self):
        pass


@pytest.fixture
def plate(monkeypatch):
    from femo_alpha_b200.dynamic_rm_shell import plate_sim as PS
    from femo_alpha_b200.mesh import createRectangleMesh
    monkeypatch.setattr(PS, "ShellDevice", FakeShellDevice)
    monkeypatch.setattr(PS, "DynamicShellDevice", FakeDynamicShellDevice)
    mesh = createRectangleMesh((0., 0.), (10., 2.), 4, 2)
    return mesh, PS.PlateSim(mesh, 1e8, 0.3, 10., 0.02, 4, add_self_weight=True)


def _graph(mesh, ps):
    from femo_alpha_b200 import csdl_shim as csdl
    from femo_alpha_b200.dynamic_rm_shell.state_operation_dynamic import StateOperation
    from femo_alpha_b200.dynamic_rm_shell.total_strain_energy_operation import TotalStrainEnergyOperation
    from femo_alpha_b200.dynamic_rm_shell.volume_operation import VolumeOperation
    nn = mesh.geometry.x.shape[0]
    rec = csdl.Recorder(inline=True); rec.start()
    rng = np.random.default_rng(2)
    thickness = csdl.Variable(value=0.1 * (1 + 0.3 * rng.random(nn)), name='thickness')
    fh = np.zeros((ps.time_levels, 3 * nn)); fh[:, 2::3] = np.linspace(0, 5, ps.time_levels)[:, None]
    g = csdl.VariableGroup(); g.thickness = thickness; g.force_history = csdl.Variable(value=fh, name='force_history')
    op = StateOperation(plate_sim=ps, gradient_mode='petsc', record=False)
    g.disp_history = op.evaluate(g)
    tse = TotalStrainEnergyOperation(ps).evaluate(g)
    vol = VolumeOperation(ps).evaluate(g)
    rec.stop()
    return csdl, rec, op, thickness, g.disp_history, tse, vol, fh


def test_state_history_layout_and_outputs(plate):
    mesh, ps = plate
    csdl, rec, op, thickness, disp, tse, vol, fh = _graph(mesh, ps)
    L, n = ps.time_levels, ps.fe_dofs
    assert disp.value.shape == (n * L,)
    w = ps._dyn.history()
    assert np.array_equal(disp.value.reshape(L, n), w)                 # level-major = ravel('F') of (fe_dofs, levels)
    assert np.array_equal(disp.value, np.ravel(w.T, order='F'))
    assert np.isclose(tse.value[0], ps._dyn.total_strain_energy())
    assert np.isclose(vol.value[0], 20.0 * thickness.value.mean(), rtol=0.1)
    assert ps.tip_disp_history.shape == (L,) and ps.strain_energy_history.shape == (L,)
    assert np.array_equal(ps.bc_dofs, np.sort(ps.bc_dofs)) and ps.bc_dofs.size > 0
    with pytest.raises(ValueError):
        op.apply_inverse_jacobian({'thickness': thickness.value}, {'disp_history': disp.value}, {}, {}, 'sideways')
    with pytest.raises(ValueError):
        op.compute_jacvec_product({'thickness': thickness.value}, {'disp_history': disp.value}, {}, {}, {}, 'sideways')


def test_reverse_sweep_totals_and_forward_mode(plate):
    mesh, ps = plate
    csdl, rec, op, thickness, disp, tse, vol, fh = _graph(mesh, ps)
    sim = csdl.experimental.PySimulator(rec)
    rep = sim.check_totals(ofs=[tse, vol], wrts=[thickness], step_size=1e-6, verbose=False)
    assert rep[tse, thickness]["relative_error"] < 1e-5
    assert rep[vol, thickness]["relative_error"] < 1e-7
    # forward and reverse products of the operation are transposes of each other
    rng = np.random.default_rng(5)
    iv, ov = {'thickness': thickness.value, 'force_history': fh}, {'disp_history': disp.value}
    dw, dh, lam = rng.standard_normal(disp.value.size), rng.standard_normal(thickness.value.size), rng.standard_normal(disp.value.size)
    lam = (lam.reshape(ps.time_levels, -1) * ps._dyn.od.keep).ravel(); lam[:ps.fe_dofs] = 0
    d_res = {'disp_history': np.zeros_like(dw)}
    op.compute_jacvec_product(iv, ov, {'thickness': dh}, {'disp_history': dw}, d_res, 'fwd')
    d_in = {'thickness': np.zeros_like(dh)}
    op.compute_jacvec_product(iv, ov, d_in, {'disp_history': np.zeros_like(dw)}, {'disp_history': lam}, 'rev')
    back_w = ps._dyn.dRdw_apply(lam, transpose=True).ravel()
    assert np.isclose(lam @ d_res['disp_history'], d_in['thickness'] @ dh + back_w @ dw, rtol=1e-9)
    # apply_inverse_jacobian: fwd then the operator gives the right-hand side back
    d_out = {}
    op.apply_inverse_jacobian(iv, ov, d_out, {'disp_history': dw}, 'fwd')
    assert np.allclose(ps._dyn.dRdw_apply(d_out["disp_history"]).ravel(), dw, rtol=1e-6, atol=1e-6 * np.abs(dw).max())
