"""Registry of named forms behind the generic `FEA.add_state` / `add_output` API (BASELINE config 5).

The reference's generic FEA takes arbitrary UFL forms (femo_alpha/fea/fea_dolfinx.py:71-116); there is no UFL here, so a
problem is built from a *named* form whose kernels live in libfemo_b200 (csrc/hex.cu):

    fea, fn = build_fea("elasticity3d", mesh, E0=1.0, nu=0.3, simp_p=3.0)
    fea.add_strong_bc(None, [locate_dofs_geometrical(fn["u"].function_space, lambda x: x[0] < 1e-12)])
    update(fn["density"], rho); update(fn["body_force"], f)
    fea.solve(fea.states_dict["u"]["residual_form"], fn["u"], fea.bc)

after which `StateOperation(fea, ["density", "body_force"], "u")` and `OutputOperation(fea, [...], "compliance")` of
csdl_alpha_opt/ work as for the shell.
"""
from __future__ import annotations

import types

import numpy as np

from .fea_dolfinx import FEA
from .utils_dolfinx import Function, FunctionSpace, OutputForm, ResidualForm
from ..mesh import _Geometry, _IndexMap
from ..solid_device import SolidDevice, box_hexes


class _Topology3:
    def __init__(self, nn, nc):
        self.dim = 3
        self._sizes = {0: nn, 3: nc}
        self.original_cell_index = np.arange(nc, dtype=np.int64)
        self.cell_name = "hexahedron"

    def index_map(self, d):
        if d not in self._sizes:
            raise ValueError("index_map(%d): only vertices (0) and cells (3) are numbered on the host" % d)
        return _IndexMap(self._sizes[d])

    def create_connectivity(self, d0, d1):
        return None


class SolidMesh:
    """Hexahedral mesh in R^3 (`cells` in VTK_HEXAHEDRON order): the dolfinx.mesh.Mesh stand-in of the solid forms."""

    def __init__(self, nodes, cells):
        x = np.ascontiguousarray(np.asarray(nodes, dtype=np.float64))
        c = np.ascontiguousarray(np.asarray(cells).astype(np.int32))
        if x.ndim != 2 or x.shape[1] != 3:
            raise ValueError("nodes must have shape (n, 3)")
        if c.ndim != 2 or c.shape[1] != 8:
            raise ValueError("Invalid cell shape--hexahedra (n, 8) expected")
        self.cells = c
        self.geometry = _Geometry(x)
        self.topology = _Topology3(x.shape[0], c.shape[0])
        self.comm = None

    def ufl_cell(self):
        return types.SimpleNamespace(cellname=lambda: "hexahedron")


def createBoxMesh(pt1, pt2, nx, ny, nz):
    nodes, cells = box_hexes(nx, ny, nz, pt2[0] - pt1[0], pt2[1] - pt1[1], pt2[2] - pt1[2])
    return SolidMesh(nodes + np.asarray(pt1, dtype=np.float64)[None, :], cells)


def createUnitCubeMesh(n):
    return createBoxMesh((0.0, 0.0, 0.0), (1.0, 1.0, 1.0), n, n, n)


# name -> (device form, inputs, outputs with their arguments)
REGISTRY = {
    "elasticity3d": dict(form="elasticity", inputs=("density", "body_force"),
                         outputs={"compliance": ["u", "body_force"], "volume": ["density"]}),
    "vector_poisson3d": dict(form="vector_poisson", inputs=("density", "body_force"),
                             outputs={"compliance": ["u", "body_force"], "volume": ["density"]}),
}


def build_fea(name, mesh, E0=1.0, nu=0.3, simp_p=3.0, state_name="u", device=0):
    """FEA with the inputs, the state and the outputs of the named form registered; returns (fea, {name: Function})."""
    if name not in REGISTRY:
        raise ValueError("unknown named form %r (registered: %s)" % (name, ", ".join(sorted(REGISTRY))))
    spec = REGISTRY[name]
    dev = SolidDevice(mesh.geometry.x, mesh.cells, spec["form"], E0, nu, simp_p, device)
    fea = FEA(mesh)
    fea.linear_problem = True
    fea.device = dev
    W = FunctionSpace(mesh, "W3", dev.ndof, dev, label=state_name)
    spaces = {"density": FunctionSpace(mesh, "DG0", dev.nc, dev, "density"),
              "body_force": FunctionSpace(mesh, "VF", dev.ndof, dev, "body_force")}
    fn = {state_name: Function(W, state_name)}
    for k in spec["inputs"]:
        fn[k] = Function(spaces[k], k)
        fea.add_input(k, fn[k], init_val=1.0 if k == "density" else 0.0)
        fn[k]._push()
    res = ResidualForm(dev, fn[state_name], [fn[k] for k in spec["inputs"]])
    fea.add_state(state_name, fn[state_name], res, arguments=list(spec["inputs"]))
    for out, args in spec["outputs"].items():
        fea.add_output(out, OutputForm(dev, out, fn[state_name]), [state_name if a == "u" else a for a in args])
    return fea, fn
