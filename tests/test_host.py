"""Host-side logic without a GPU: mesh stand-in, csdl stand-in (tape, custom operations, totals)."""
import numpy as np
import pytest

from femo_alpha_b200 import csdl_shim as csdl
from femo_alpha_b200.mesh import ShellMesh, createRectangleMesh, reconstructFEAMesh, refine_uniform


def test_mesh_standin_attributes():
    m = createRectangleMesh((0.0, 0.0), (10.0, 2.0), 20, 4)
    assert m.topology.dim == 2 and m.geometry.dim == 3
    assert m.topology.index_map(0).size_local == 105 and m.topology.index_map(2).size_local == 80
    assert np.array_equal(m.geometry.input_global_indices, np.arange(105))
    assert np.array_equal(m.topology.original_cell_index, np.arange(80))
    r = refine_uniform(m, 1)
    assert r.cells.shape == (320, 4) and r.geometry.x.shape == (369, 3)
    # positively oriented children, same total area
    def area(mm):
        x = mm.geometry.x[mm.cells]
        return 0.5 * np.abs(np.cross(x[:, 2] - x[:, 0], x[:, 3] - x[:, 1])[:, 2]).sum()
    assert abs(area(r) - 20.0) < 1e-12 and abs(area(m) - 20.0) < 1e-12
    with pytest.raises(ValueError):
        reconstructFEAMesh("x.xdmf", m.geometry.x, np.zeros((3, 5), dtype=int))
    with pytest.raises(NotImplementedError):
        ShellMesh(m.geometry.x, np.array([[0, 1, 2]]))


class _Square(csdl.CustomExplicitOperation):
    def evaluate(self, x):
        self.declare_input("x", x)
        y = self.create_output("y", (1,))
        return y

    def compute(self, iv, ov):
        ov["y"] = np.sum(iv["x"] ** 2)

    def compute_derivatives(self, iv, ov, d):
        d["y", "x"] = 2 * iv["x"]


class _LinSolve(csdl.experimental.CustomImplicitOperation):
    """R(a, y) = (A0 + diag(a)) y - b = 0"""
    A0 = np.array([[4.0, 1.0, 0.0], [1.0, 3.0, 1.0], [0.0, 1.0, 5.0]])
    b = np.array([1.0, 2.0, 3.0])

    def evaluate(self, a):
        self.declare_input("a", a)
        return self.create_output("y", (3,))

    def solve_residual_equations(self, iv, ov):
        ov["y"] = np.linalg.solve(self.A0 + np.diag(iv["a"]), self.b)

    def apply_inverse_jacobian(self, iv, ov, d_out, d_res, mode):
        A = self.A0 + np.diag(iv["a"])
        if mode == "rev":
            d_res["y"] = np.linalg.solve(A.T, d_out["y"])
        else:
            raise ValueError("mode")

    def compute_jacvec_product(self, iv, ov, d_in, d_out, d_res, mode):
        d_in["a"] += ov["y"] * d_res["y"]            # (dR/da)^T lam, dR_i/da_i = y_i


def test_shim_totals_explicit_and_implicit():
    rec = csdl.Recorder(inline=True); rec.start()
    a = csdl.Variable(value=np.array([0.5, 0.2, 0.1, 0.7]), name="a")
    y = _LinSolve().evaluate(a[[0, 1, 3]])
    z = _Square().evaluate(y * 2.0)
    f = 3.0 * z ** 0.5
    f.add_name("f")
    rec.stop()
    sim = csdl.experimental.PySimulator(rec)
    rep = sim.check_totals([f], [a], step_size=1e-6, verbose=False)
    r = rep[f, a]
    assert r["relative_error"] < 1e-7
    assert abs(r["analytical"][0, 2]) == 0.0       # a[2] is not used


def test_shim_reshape_transpose_matvec():
    import scipy.sparse as sp
    rec = csdl.Recorder(inline=True); rec.start()
    x = csdl.Variable(value=np.arange(6.0).reshape(2, 3), name="x")
    M = sp.csr_matrix(np.arange(12.0).reshape(2, 6))
    v = csdl.sparse.matvec(M, csdl.reshape(csdl.transpose(x), (-1,)))
    s = _Square().evaluate(v)
    rec.stop()
    sim = csdl.experimental.PySimulator(rec)
    assert sim.check_totals([s], [x], verbose=False)[s, x]["relative_error"] < 1e-7
    g = csdl.VariableGroup()
    assert g.missing is None
    g.thing = x
    assert g.thing is x
