"""CPU oracle of the generic 3-D forms on trilinear hexahedra (SURVEY 8 f4 / BASELINE config 5).  TEST INFRASTRUCTURE ONLY:
nothing under femo_alpha_b200/ may import this module.

PARITY UNPINNED: the reference has no in-tree form for config 5 -- its generic FEA accepts any UFL form
(femo_alpha/fea/fea_dolfinx.py:71-116) and dolfinx/UFL/FFCx are absent here.  What is restated is the
published algorithm those packages implement for the two forms SURVEY 8d names:

    linear elasticity   a(u, v) = int s_e (lam div u div v + 2 mu eps(u):eps(v)) dx,   s_e = rho_e^p  (SIMP)
    vector Poisson      a(u, v) = int s_e kappa grad u : grad v dx
    load                l(v)    = int f . v dx,  f a nodal Q1 field
    residual            R(u; rho, f) = K(rho) u - M f         (strong BCs the dolfinx way: rows + columns zeroed, diagonal 1)

on Q1 hexahedra (dolfinx `hexahedron` cells, tensor Gauss rule with 2 points per direction -- the rule basix picks for
degree 2..3 integrands).  Pins (tests/test_oracle_hex.py): patch test (linear fields are reproduced exactly on distorted
meshes), strain energy of a uniform strain state = |Omega| (lam tr^2 + 2 mu eps:eps), uniaxial bar elongation F L / (E A),
manufactured Poisson solution at the O(h^2) rate, finite differences of every partial.
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

_G = 0.57735026918962576451
# VTK_HEXAHEDRON corner signs
_SX = np.array([-1, 1, 1, -1, -1, 1, 1, -1], dtype=float)
_SY = np.array([-1, -1, 1, 1, -1, -1, 1, 1], dtype=float)
_SZ = np.array([-1, -1, -1, -1, 1, 1, 1, 1], dtype=float)


def box_mesh(nx, ny, nz, lx=1.0, ly=1.0, lz=1.0, distort=0.0, seed=0):
    """Structured box of hexahedra in VTK node order; `distort` moves interior nodes randomly (fraction of the cell size)."""
    xs, ys, zs = np.linspace(0, lx, nx + 1), np.linspace(0, ly, ny + 1), np.linspace(0, lz, nz + 1)
    X, Y, Z = np.meshgrid(xs, ys, zs, indexing="ij")
    nodes = np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1)
    nid = lambda i, j, k: (i * (ny + 1) + j) * (nz + 1) + k
    I, J, K = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    I, J, K = I.ravel(), J.ravel(), K.ravel()
    cells = np.stack([nid(I, J, K), nid(I + 1, J, K), nid(I + 1, J + 1, K), nid(I, J + 1, K),
                      nid(I, J, K + 1), nid(I + 1, J, K + 1), nid(I + 1, J + 1, K + 1), nid(I, J + 1, K + 1)], axis=1)
    if distort > 0:
        rng = np.random.default_rng(seed)
        inner = np.ones(len(nodes), bool)
        for d, L in enumerate((lx, ly, lz)):
            inner &= (nodes[:, d] > 1e-12) & (nodes[:, d] < L - 1e-12)
        h = np.array([lx / nx, ly / ny, lz / nz])
        nodes[inner] += distort * h * rng.uniform(-1, 1, (int(inner.sum()), 3))
    return nodes, cells.astype(np.int32)


class HexOracle:
    def __init__(self, nodes, cells, form="elasticity", E0=1.0, nu=0.3, simp_p=3.0):
        self.x = np.asarray(nodes, float)
        self.cells = np.asarray(cells, np.int64)
        self.nn, self.nc = len(self.x), len(self.cells)
        self.ndof = 3 * self.nn
        self.form = form
        if form == "elasticity":
            self.c1, self.c2 = E0 * nu / ((1 + nu) * (1 - 2 * nu)), E0 / (2 * (1 + nu))
        elif form == "vector_poisson":
            self.c1, self.c2 = E0, 0.0
        else:
            raise ValueError("unknown form")
        self.p = float(simp_p)
        self.rho = np.ones(self.nc)
        self.f = np.zeros((self.nn, 3))
        self.bc_dofs = np.zeros(0, np.int64)
        self._geometry()

    # ---- geometry: shape values, physical gradients and weights at the 8 Gauss points ---------------------------
    def _geometry(self):
        N = np.empty((8, 8)); dN = np.empty((8, 8, 3))                    # [q, a], [q, a, j]
        for q in range(8):
            xi = np.array([_G if q & 1 else -_G, _G if q & 2 else -_G, _G if q & 4 else -_G])
            fx, fy, fz = 1 + _SX * xi[0], 1 + _SY * xi[1], 1 + _SZ * xi[2]
            N[q] = 0.125 * fx * fy * fz
            dN[q] = 0.125 * np.stack([_SX * fy * fz, fx * _SY * fz, fx * fy * _SZ], axis=1)
        xe = self.x[self.cells]                                           # (nc, 8, 3)
        J = np.einsum("eai,qaj->eqij", xe, dN)                            # d x_i / d xi_j
        self.det = np.linalg.det(J)
        if not (self.det > 0).all():
            raise ValueError("non-positive Jacobian (VTK_HEXAHEDRON order expected)")
        Jinv = np.linalg.inv(J)                                           # [j, i] = d xi_j / d x_i
        self.N = N
        self.g = np.einsum("qaj,eqji->eqai", dN, Jinv)                    # physical gradients (nc, 8, 8, 3)
        self.vol = self.det.sum(axis=1)

    # ---- inputs ---------------------------------------------------------------------------------------------------
    def set_inputs(self, density=None, body_force=None):
        if density is not None:
            self.rho = np.broadcast_to(np.asarray(density, float), (self.nc,)).copy()
        if body_force is not None:
            self.f = np.asarray(body_force, float).reshape(self.nn, 3).copy()

    def set_strong_bc(self, dofs):
        self.bc_dofs = np.unique(np.asarray(dofs, np.int64))

    # ---- element matrices --------------------------------------------------------------------------------------------
    def element_matrices0(self):
        """(nc, 24, 24) unit-density element matrices, dof order 3 a + i."""
        g, w = self.g, self.det
        I3 = np.eye(3)
        dots = np.einsum("eqai,eqbi->eqab", g, g)
        if self.form == "elasticity":
            K = (self.c1 * np.einsum("eq,eqar,eqbc->earbc", w, g, g) + self.c2 * np.einsum("eq,eqbr,eqac->earbc", w, g, g)
                 + self.c2 * np.einsum("eq,eqab,rc->earbc", w, dots, I3))
        else:
            K = self.c1 * np.einsum("eq,eqab,rc->earbc", w, dots, I3)
        return K.reshape(self.nc, 24, 24)

    def _edofs(self):
        return (3 * self.cells[:, :, None] + np.arange(3)[None, None, :]).reshape(self.nc, 24)

    def _assemble(self, Ke):
        ed = self._edofs()
        rows = np.repeat(ed, 24, axis=1).ravel()
        cols = np.tile(ed, (1, 24)).ravel()
        return sp.csr_matrix((Ke.ravel(), (rows, cols)), shape=(self.ndof, self.ndof))

    def stiffness(self):
        return self._assemble(self.element_matrices0() * (self.rho ** self.p)[:, None, None])

    def mass_matrix(self):
        m = np.einsum("eq,qa,qb->eab", self.det, self.N, self.N)
        Me = np.einsum("eab,rc->earbc", m, np.eye(3)).reshape(self.nc, 24, 24)
        return self._assemble(Me)

    def with_bc(self, K):
        keep = np.ones(self.ndof); keep[self.bc_dofs] = 0
        D = sp.diags(keep)
        return (D @ K @ D + sp.diags(1 - keep)).tocsr()

    # ---- residual, solve, outputs, partials ---------------------------------------------------------------------------------
    def load(self):
        return self.mass_matrix() @ self.f.ravel()

    def residual(self, u):
        return self.stiffness() @ u - self.load()

    def solve(self):
        b = self.load(); b[self.bc_dofs] = 0
        return spla.spsolve(self.with_bc(self.stiffness()).tocsc(), b)

    def compliance(self, u):
        return float(self.load() @ u)

    def volume(self):
        return float(self.rho @ self.vol)

    def dRdrho_T(self, u, lam):
        ed = self._edofs()
        K0 = self.element_matrices0()
        return self.p * self.rho ** (self.p - 1) * np.einsum("ei,eij,ej->e", lam[ed], K0, u[ed])

    def dRdrho(self, u, drho):
        ed = self._edofs()
        ye = np.einsum("eij,ej->ei", self.element_matrices0(), u[ed]) * (self.p * self.rho ** (self.p - 1) * drho)[:, None]
        out = np.zeros(self.ndof)
        np.add.at(out, ed.ravel(), ye.ravel())
        return out

    def total_dcompliance_drho(self):
        """adjoint gradient of the compliance w.r.t. the cell densities (fixed load): -lam^T dR/drho, A lam = dC/du."""
        u = self.solve()
        rhs = self.load(); rhs[self.bc_dofs] = 0
        lam = spla.spsolve(self.with_bc(self.stiffness()).tocsc(), rhs)
        lam[self.bc_dofs] = 0
        return -self.dRdrho_T(u, lam)
