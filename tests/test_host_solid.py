"""Host logic of the generic `FEA` path with a named 3-D form (femo_alpha_b200/fea/named_forms.py): registry,
spaces, BC lookup, the StateOperation / OutputOperation protocol and a finite-difference check of the totals -- with an
oracle-backed stand-in for the device (no GPU here; the CUDA path is tests/test_gpu_hex.py)."""
import numpy as np
import pytest
import scipy.sparse.linalg as spla

from oracle.hex_ref import HexOracle


class FakeSolidDevice:
    """SolidDevice surface on top of oracle/hex_ref.py"""

    def __init__(self, nodes, hexes, form="elasticity", E0=1.0, nu=0.3, simp_p=3.0, device=0):
        self.o = HexOracle(nodes, hexes, form=form, E0=E0, nu=nu, simp_p=simp_p)
        o = self.o
        self.nn, self.nc, self.nb, self.ndof, self.nq2 = o.nn, o.nc, o.nn, o.ndof, o.nn
        self.n_mat, self.n_f = o.nc, o.ndof
        self.w = np.zeros(o.ndof)
        self.last_solve = dict(iters=1, relres=0.0)

    def block_coords(self):
        return self.o.x.copy()

    def set_strong_bc(self, dofs):
        self.o.set_strong_bc(dofs)

    def set_input(self, name, values):
        v = np.asarray(values, dtype=float).ravel()
        if name == "density":
            self.o.set_inputs(density=v if v.size > 1 else v[0])
        elif name == "body_force":
            self.o.set_inputs(body_force=v)
        else:
            raise KeyError(name)

    def set_state(self, w):
        self.w = np.array(w, dtype=float)

    def get_state(self, out=None):
        if out is None:
            return self.w.copy()
        out[:] = self.w
        return out

    def _A(self):
        return self.o.with_bc(self.o.stiffness()).tocsc()

    def solve_state(self, rtol=1e-12, max_it=0):
        self.w = self.o.solve()
        return 1, 0.0

    def _lin(self, rhs, zero_bc):
        x = spla.spsolve(self._A(), np.asarray(rhs, dtype=float))
        if zero_bc:
            x[self.o.bc_dofs] = 0
        return x

    def solve_adjoint(self, rhs, rtol=1e-12, max_it=0):
        return self._lin(rhs, True)

    def solve_forward(self, rhs, rtol=1e-12, max_it=0):
        return self._lin(rhs, False)

    def assemble_jacobian(self):
        pass

    def jacobian_apply(self, x, with_bc=False):
        K = self.o.stiffness()
        return (self.o.with_bc(K) if with_bc else K) @ x

    def get_csr(self, with_bc=False):
        K = self.o.stiffness()
        return self.o.with_bc(K) if with_bc else K

    def assemble_residual(self, w=None, out=None):
        return self.o.residual(self.w if w is None else np.asarray(w))

    def dRdf_apply(self, arg, df):
        if arg == "density":
            return self.o.dRdrho(self.w, np.asarray(df))
        return -(self.o.mass_matrix() @ np.asarray(df))

    def dRdf_T_apply(self, arg, lam):
        if arg == "density":
            return self.o.dRdrho_T(self.w, np.asarray(lam))
        return -(self.o.mass_matrix() @ np.asarray(lam))

    def output(self, name):
        return self.o.compliance(self.w) if name == "compliance" else self.o.volume()

    def output_grad(self, name, arg):
        n = self.nc if arg == "density" else self.ndof
        if name == "compliance" and arg in ("disp_solid", "state", "u"):
            return self.o.load()
        if name == "compliance" and arg == "body_force":
            return self.o.mass_matrix() @ self.w
        if name == "volume" and arg == "density":
            return self.o.vol.copy()
        return np.zeros(n)

    def close(self):
        pass


@pytest.fixture
def solid(monkeypatch):
    import femo_alpha_b200.fea.named_forms as NF
    monkeypatch.setattr(NF, "SolidDevice", FakeSolidDevice)
    mesh = NF.createBoxMesh((0, 0, 0), (2.0, 1.0, 1.0), 4, 2, 2)
    fea, fn = NF.build_fea("elasticity3d", mesh, E0=10.0, nu=0.3, simp_p=3.0)
    return NF, mesh, fea, fn


def test_registry_spaces_and_bc_lookup(solid):
    NF, mesh, fea, fn = solid
    from femo_alpha_b200.fea.utils_dolfinx import locate_dofs_geometrical, dirichletbc
    assert fea.nn == 5 * 3 * 3 and fea.nel == 16
    assert set(fea.inputs_dict) == {"density", "body_force"} and set(fea.states_dict) == {"u"}
    assert set(fea.outputs_dict) == {"compliance", "volume"}
    assert fea.inputs_dict["density"]["shape"] == 16 and fea.states_dict["u"]["shape"] == 3 * 45
    dofs = locate_dofs_geometrical(fn["u"].function_space, lambda x: np.less(x[0], 1e-12))
    assert dofs.size == 3 * 9 and np.array_equal(dofs, np.sort(dofs))
    x = fn["u"].function_space.tabulate_dof_coordinates()
    assert x.shape == (135, 3) and np.all(x[dofs, 0] < 1e-12)
    fea.add_strong_bc(None, [dofs])
    assert len(fea.bc) == 1 and isinstance(dirichletbc(None, dofs), type(fea.bc[0]))
    with pytest.raises(ValueError):
        NF.build_fea("stokes3d", mesh)
    with pytest.raises(ValueError):
        fea.add_input("density", fn["density"])
    with pytest.raises(ValueError):
        NF.SolidMesh(mesh.geometry.x, mesh.cells[:, :4])


def test_operations_and_totals(solid):
    NF, mesh, fea, fn = solid
    from femo_alpha_b200 import csdl_shim as csdl
    from femo_alpha_b200.csdl_alpha_opt.fea_model import FEAModel
    from femo_alpha_b200.fea.utils_dolfinx import locate_dofs_geometrical
    fea.add_strong_bc(None, [locate_dofs_geometrical(fn["u"].function_space, lambda x: np.less(x[0], 1e-12))])
    rng = np.random.default_rng(3)
    rec = csdl.Recorder(inline=True); rec.start()
    g = csdl.VariableGroup()
    f = np.zeros((fea.nn, 3)); f[:, 2] = -1.0
    g.density = csdl.Variable(value=rng.uniform(0.3, 1.0, fea.nel), name="density")
    g.body_force = csdl.Variable(value=f.ravel(), name="body_force")
    out = FEAModel(fea=[fea]).evaluate(g)
    rec.stop()
    o = fea.device.o
    assert np.allclose(out.u.value, o.solve(), rtol=1e-10, atol=1e-14)
    assert np.isclose(np.ravel(out.compliance.value)[0], o.compliance(o.solve()), rtol=1e-10)
    assert np.isclose(np.ravel(out.volume.value)[0], o.volume(), rtol=1e-13)
    sim = csdl.experimental.PySimulator(rec)
    rep = sim.check_totals(ofs=[out.compliance, out.volume], wrts=[g.density, g.body_force], step_size=1e-6, verbose=False)
    assert rep[out.compliance, g.density]["relative_error"] < 1e-5
    assert rep[out.compliance, g.body_force]["relative_error"] < 1e-5
    assert rep[out.volume, g.density]["relative_error"] < 1e-7


def test_field_output_operation_protocol(solid):
    """OutputFieldOperation (output_operation.py:72-128 of the reference): shape from the registry, projection on compute,
    recorder hook, no derivatives."""
    NF, mesh, fea, fn = solid
    from femo_alpha_b200 import csdl_shim as csdl
    from femo_alpha_b200.csdl_alpha_opt.output_operation import OutputFieldOperation, OutputOperation
    from femo_alpha_b200.fea.utils_dolfinx import FunctionSpace

    class CellField:
        kind = "field"

        def field_space(self, spec):
            return FunctionSpace(mesh, "DG0", fea.nel, fea.device, "cell_density")

        def project(self):
            return 2.0 * fea.device.o.rho

    fea.add_field_output("twice_density", CellField(), ["density"], record=True)
    written = []
    fea.outputs_field_dict["twice_density"]["recorder"] = type("R", (), {"write_function": lambda self, f, t: written.append(t)})()
    rec = csdl.Recorder(inline=True); rec.start()
    g = csdl.VariableGroup()
    rho = np.linspace(0.2, 1.0, fea.nel)
    g.density = csdl.Variable(value=rho, name="density")
    out = OutputFieldOperation(fea, ["density"], "twice_density").evaluate(g)
    rec.stop()
    assert out.value.shape == (fea.nel,) and np.allclose(out.value, 2.0 * rho) and len(written) == 1
    with pytest.raises(ValueError):
        OutputOperation(fea, ["density", "body_force"], "volume").evaluate(g)       # body_force missing from the group
    with pytest.raises(KeyError):
        OutputFieldOperation(fea, ["density"], "no_such_field")
