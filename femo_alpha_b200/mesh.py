"""`ShellMesh`: the light object that stands in for `dolfinx.mesh.Mesh` at the drop-in boundary.

Only the attributes the reference touches are provided (femo_alpha/fea/fea_dolfinx.py:50-51,
rm_shell/rm_shell_model.py:68-69,115,399-401,432): `topology.dim`,
`topology.index_map(d).size_local`, `topology.original_cell_index`, `geometry.x`,
`geometry.dim`, `geometry.input_global_indices`.  Our numbering IS the external
numbering, so both index maps are the identity.
"""
from __future__ import annotations

import pickle
import types

import numpy as np


class _IndexMap:
    def __init__(self, n):
        self.size_local = int(n)
        self.size_global = int(n)
        self.num_ghosts = 0


class _Topology:
    def __init__(self, nn, ne_unknown, nc):
        self.dim = 2
        self._sizes = {0: nn, 2: nc}
        self.original_cell_index = np.arange(nc, dtype=np.int64)
        self.cell_name = "quadrilateral"

    def index_map(self, d):
        if d not in self._sizes:
            raise ValueError("index_map(%d): only vertices (0) and cells (2) are numbered on the host" % d)
        return _IndexMap(self._sizes[d])

    def create_connectivity(self, d0, d1):
        return None


class _Geometry:
    def __init__(self, x):
        self.x = x
        self.dim = 3
        self.input_global_indices = np.arange(x.shape[0], dtype=np.int64)


class ShellMesh:
    """Quadrilateral surface mesh in R^3. `cells` in VTK/meshio cyclic vertex order."""

    def __init__(self, nodes, cells):
        x = np.ascontiguousarray(np.asarray(nodes, dtype=np.float64))
        if x.ndim != 2 or x.shape[1] not in (2, 3):
            raise ValueError("nodes must have shape (n, 2) or (n, 3)")
        if x.shape[1] == 2:
            x = np.concatenate([x, np.zeros((x.shape[0], 1))], 1)
        c = np.ascontiguousarray(np.asarray(cells).astype(np.int32))
        if c.ndim != 2 or c.shape[1] not in (3, 4):
            raise ValueError("Invalid cell shape--should be either triangular or quadrilateral")
        if c.shape[1] == 3:
            raise NotImplementedError("triangular (P2xP1) shells are not implemented in femo_alpha_b200 yet")
        self.cells = c
        self.geometry = _Geometry(x)
        self.topology = _Topology(x.shape[0], None, c.shape[0])
        self.comm = None

    def ufl_cell(self):
        return types.SimpleNamespace(cellname=lambda: "quadrilateral")


def createRectangleMesh(pt1, pt2, nx, ny):
    """Structured quads on [pt1, pt2] (fea/utils_dolfinx.py:164-169 createRectangleMesh)."""
    xs = np.linspace(pt1[0], pt2[0], nx + 1)
    ys = np.linspace(pt1[1], pt2[1], ny + 1)
    X, Y = np.meshgrid(xs, ys, indexing="ij")
    nodes = np.stack([X.ravel(), Y.ravel(), np.zeros(X.size)], 1)
    idx = np.arange((nx + 1) * (ny + 1), dtype=np.int64).reshape(nx + 1, ny + 1)
    cells = np.stack([idx[:-1, :-1].ravel(), idx[1:, :-1].ravel(), idx[1:, 1:].ravel(), idx[:-1, 1:].ravel()], 1)
    return ShellMesh(nodes, cells)


def createUnitSquareMesh(n):
    return createRectangleMesh((0.0, 0.0), (1.0, 1.0), n, n)


def reconstructFEAMesh(filename, nodes, connectivity):
    """fea/utils_dolfinx.py:653-668: build the FEA mesh from node coordinates + connectivity.
    `filename` (the XDMF scratch file of the reference) is accepted and ignored."""
    connectivity = np.asarray(connectivity)
    if connectivity.shape[1] not in (3, 4):
        raise ValueError('Invalid cell shape--should be either triangular or quadrilateral')
    return ShellMesh(nodes, connectivity)


def readPickleMesh(path):
    """Mesh dictionaries like examples/advanced_examples/lpc_shell_opt/data_files/mesh_bay_data.pickle
    ({'nodes': (nn,3), 'connectivity': (nc,4), ...})."""
    with open(path, "rb") as f:
        d = pickle.load(f)
    return ShellMesh(d["nodes"], np.asarray(d["connectivity"]).astype(np.int64))


def refine_uniform(mesh, levels=1):
    """Uniform 1->4 quad subdivision (keeps non-manifold junctions); used for synthetic refinements.
    The returned mesh remembers its parents (`mg_meshes`, coarse to fine): RMShellPDE then builds the nested
    Q2xQ1 hierarchy and the state / adjoint solves run multigrid-preconditioned (femo_alpha_b200/multigrid.py)."""
    from .multigrid import refine_quads
    chain = list(getattr(mesh, "mg_meshes", None) or [(mesh.geometry.x, mesh.cells)])
    for _ in range(levels):
        chain.append(refine_quads(*chain[-1]))
    fine = ShellMesh(*chain[-1])
    fine.mg_meshes = chain
    return fine
