"""Mesh hierarchies by uniform quad subdivision + the multigrid-preconditioned state / adjoint solves.

The reference factorises dR/du with MUMPS (femo_alpha/fea/utils_dolfinx.py:438-468, 514-531).  On the GPU the
Krylov solver of csrc/solver.cu needs O(1e5) block-Jacobi iterations on thin shells at 1e7 DOF; with a hierarchy
of nested meshes a V-cycle preconditioner (csrc/mg.cu) brings that to ~1e2, mesh-independently.

A hierarchy is produced by subdividing every quad of a COARSE mesh into four (new vertices = edge mid-points and
cell centres of the bilinear parent), `levels - 1` times; the analysis runs on the finest mesh.  Because the children
are exact sub-squares of the parent's reference square, the Q2xQ1 spaces are nested and the prolongation is the
parent's basis evaluated at the children's nodes.  Works for any quad surface mesh, non-manifold junctions included.

    h = ShellHierarchy(coarse_nodes, coarse_quads, levels=4)
    h.set_input("thickness", lambda x: 0.1 + 0.0 * x[0])      # callable of the vertex coordinates, or finest-level array
    h.set_strong_bc(lambda x: x[0] < 1e-12)
    h.setup()                                                  # assemble every level, smoother bounds, coarse inverse
    iters, relres = h.solve_state()
    u = h.fine.get_state()
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L
from .shell_device import ShellDevice

# reference coordinates of the 13 local block nodes of a cell, in the order of femo_get_cell_blocks:
# 4 Q2 vertices (tensor order), 4 Q2 edge mid-points (edges {0,1},{0,2},{1,3},{2,3}), Q2 centre, 4 Q1 vertices
_REF2 = np.array([(0, 0), (1, 0), (0, 1), (1, 1), (.5, 0), (0, .5), (1, .5), (.5, 1), (.5, .5)], dtype=np.float64)
_REF1 = np.array([(0, 0), (1, 0), (0, 1), (1, 1)], dtype=np.float64)


def _lag2(x, node):
    """1-D quadratic Lagrange polynomial of the node at `node` (0, .5 or 1), evaluated at x."""
    if node == 0.0:
        return 2.0 * (x - 0.5) * (x - 1.0)
    if node == 0.5:
        return 4.0 * x * (1.0 - x)
    return 2.0 * x * (x - 0.5)


def _q2(X0, X1):
    return np.array([_lag2(X0, a) * _lag2(X1, b) for a, b in _REF2])


def _q1(X0, X1):
    return np.array([(1 - X0) * (1 - X1), X0 * (1 - X1), (1 - X0) * X1, X0 * X1])


def refine_quads(nodes, quads):
    """Uniform subdivision.  `quads` in VTK cyclic order (as reconstructFEAMesh takes them,
    femo_alpha/fea/utils_dolfinx.py:653-668).  Returns (nodes_f, quads_f); fine cell 4 c + k is child k = ci + 2 cj of
    cell c and covers [ci/2, (ci+1)/2] x [cj/2, (cj+1)/2] of the parent's (tensor-ordered) reference square.
    The parent's vertices keep their numbers; edge mid-points follow (sorted unique vertex pairs), then centres."""
    nodes = np.asarray(nodes, dtype=np.float64)
    q = np.asarray(quads, dtype=np.int64)
    nn, nc = nodes.shape[0], q.shape[0]
    a, b, c, d = q[:, 0], q[:, 1], q[:, 2], q[:, 3]          # tensor vertices: (0,0)=a (1,0)=b (0,1)=d (1,1)=c
    pairs = np.stack([np.stack([a, b], 1), np.stack([a, d], 1), np.stack([b, c], 1), np.stack([d, c], 1)], 1)   # nc,4,2
    key = np.sort(pairs, axis=2).reshape(-1, 2)
    ukey, inv = np.unique(key[:, 0] * nn + key[:, 1], return_inverse=True)
    ne = ukey.size
    mid = nn + inv.reshape(nc, 4)                             # m_ab, m_ad, m_bc, m_dc
    ctr = nn + ne + np.arange(nc)
    xm = 0.5 * (nodes[ukey // nn] + nodes[ukey % nn])
    xc = 0.25 * (nodes[a] + nodes[b] + nodes[c] + nodes[d])
    nodes_f = np.concatenate([nodes, xm, xc])
    m_ab, m_ad, m_bc, m_dc = mid[:, 0], mid[:, 1], mid[:, 2], mid[:, 3]
    # children as tensor vertex lists [t0 t1 t2 t3] -> VTK cyclic [t0 t1 t3 t2]
    ch = [(a, m_ab, m_ad, ctr), (m_ab, b, ctr, m_bc), (m_ad, ctr, d, m_dc), (ctr, m_bc, m_dc, c)]
    qf = np.empty((nc, 4, 4), dtype=np.int64)
    for k, (t0, t1, t2, t3) in enumerate(ch):
        qf[:, k, 0], qf[:, k, 1], qf[:, k, 2], qf[:, k, 3] = t0, t1, t3, t2
    return nodes_f, qf.reshape(4 * nc, 4).astype(np.int32)


def prolongation(fine_blocks, coarse_blocks, nb_fine, nb_coarse, parent=None, child=None, allow_missing_rows=False):
    """Block-scalar prolongation between a mesh and its `refine_quads` parent.
    fine_blocks (4 nc, 13), coarse_blocks (nc, 13): femo_get_cell_blocks of the two levels.
    Returns scipy CSR (nb_fine x nb_coarse).

    General form (one rank of a partitioned hierarchy, dist_multigrid.py): `parent[i]` = row of `coarse_blocks` that is the
    parent of fine cell i and `child[i]` = its child index (0..3); fine cells may then come in any number and order, and
    with `allow_missing_rows` fine block nodes that no listed cell holds get an empty row instead of an error."""
    fb = np.asarray(fine_blocks, dtype=np.int64)
    cb = np.asarray(coarse_blocks, dtype=np.int64)
    if parent is None:
        assert fb.shape[0] == 4 * cb.shape[0]
        parent = np.arange(fb.shape[0], dtype=np.int64) // 4          # refine_quads order: children of a cell are adjacent
        child = np.arange(fb.shape[0], dtype=np.int64) % 4
    return _prolongation_general(fb, cb, nb_fine, nb_coarse, parent, child, allow_missing_rows)


def _child_weights():
    """weight tables per (child, local fine node): 9 Q2 nodes -> 9 coarse Q2 weights, 4 Q1 nodes -> 4 coarse Q1 weights"""
    W2 = np.zeros((4, 9, 9)); W1 = np.zeros((4, 4, 4))
    for k in range(4):
        ci, cj = k % 2, k // 2
        for a in range(9):
            W2[k, a] = _q2((ci + _REF2[a, 0]) / 2.0, (cj + _REF2[a, 1]) / 2.0)
        for a in range(4):
            W1[k, a] = _q1((ci + _REF1[a, 0]) / 2.0, (cj + _REF1[a, 1]) / 2.0)
    return W2, W1


def _prolongation_general(fine_blocks, coarse_blocks, nb_fine, nb_coarse, parent, child, allow_missing_rows):
    import scipy.sparse as sp
    fb = np.asarray(fine_blocks, dtype=np.int64)
    cb = np.asarray(coarse_blocks, dtype=np.int64)
    parent = np.asarray(parent, dtype=np.int64); child = np.asarray(child, dtype=np.int64)
    nf = fb.shape[0]
    assert parent.shape == (nf,) and child.shape == (nf,)
    W2, W1 = _child_weights()
    # a fine node shared by several fine cells gets the same row from each of them (conforming spaces): the (fine cell,
    # local node) pair with the lowest index writes it.  Decided per pair (13 per cell), before the rows are expanded.
    flat = fb.ravel()
    gid = np.arange(flat.size, dtype=np.int64)
    first = np.full(nb_fine, -1, dtype=np.int64)
    first[flat[::-1]] = gid[::-1]                        # repeated indices: the last write (lowest pair index) stays
    if (first < 0).any() and not allow_missing_rows:
        raise RuntimeError("prolongation: a fine block node belongs to no fine cell")
    own = (first[flat] == gid).reshape(nf, 13)
    cbp = cb[parent]
    c2, a2 = np.nonzero(own[:, :9])
    c1, a1 = np.nonzero(own[:, 9:])
    rows = np.concatenate([np.repeat(fb[c2, a2], 9), np.repeat(fb[c1, 9 + a1], 4)])
    cols = np.concatenate([cbp[c2, :9].ravel(), cbp[c1, 9:].ravel()])
    vals = np.concatenate([W2[child[c2], a2].ravel(), W1[child[c1], a1].ravel()])
    keep = np.abs(vals) > 1e-14
    P = sp.csr_matrix((vals[keep], (rows[keep], cols[keep])), shape=(nb_fine, nb_coarse))
    P.sum_duplicates(); P.sort_indices()
    return P


class ShellHierarchy:
    """Levels 0 (finest, the analysis mesh) .. levels-1 (the mesh handed in), each a ShellDevice on one GPU."""

    def __init__(self, coarse_nodes, coarse_quads, levels, element_wise_material=False, elementwise_pressure=False,
                 quadrature_degree=6, device=0, cheb_degree=3, meshes=None, fine_device=None, coarse_refresh=1):
        self.ewm, self.ewp = bool(element_wise_material), bool(elementwise_pressure)
        if meshes is None:
            meshes = [(np.asarray(coarse_nodes, dtype=np.float64), np.asarray(coarse_quads, dtype=np.int32))]
            for _ in range(levels - 1):
                meshes.append(refine_quads(*meshes[-1]))
        meshes = [(np.asarray(n, dtype=np.float64), np.asarray(q, dtype=np.int32)) for n, q in meshes]
        levels = len(meshes)
        if levels < 2:
            raise ValueError("a hierarchy needs at least two levels")
        meshes = meshes[::-1]                                  # level 0 = finest
        self.meshes = meshes
        self.levels = [fine_device if (l == 0 and fine_device is not None) else
                       ShellDevice(n, q, self.ewm, self.ewp, quadrature_degree, device) for l, (n, q) in enumerate(meshes)]
        self.fine = self.levels[0]
        self.fine._mg = self
        self._closing = False
        for d in self.levels:           # whichever object is finalised first, the tear-down runs through close() below
            d._owner = self
        self.cheb_degree = int(cheb_degree)
        # the dense coarse inverse is recomputed at every `coarse_refresh`-th set-up (1 = always); in between the
        # previous one is reused (lagged coarse solve)
        self.coarse_refresh = max(1, int(coarse_refresh))
        self._n_setups = 0
        self.lib = L.load()
        arr = (C.c_void_p * levels)(*[d._ctx for d in self.levels])
        self._mg = C.c_void_p()
        L.check(self.lib.femo_mg_create(levels, arr, C.byref(self._mg)))
        self._keep = []
        for l in range(levels - 1):
            f, c = self.levels[l], self.levels[l + 1]
            P = prolongation(f.cell_blocks(), c.cell_blocks(), f.nb, c.nb)
            R = P.T.tocsr(); R.sort_indices()
            a = [np.ascontiguousarray(P.indptr, np.int32), np.ascontiguousarray(P.indices, np.int32),
                 np.ascontiguousarray(P.data, np.float64), np.ascontiguousarray(R.indptr, np.int32),
                 np.ascontiguousarray(R.indices, np.int32), np.ascontiguousarray(R.data, np.float64)]
            L.check(self.lib.femo_mg_set_transfer(self._mg, l, *[L.ptr(v) for v in a]))
        self._dirty = True
        self._children = []             # weak references to DynamicShellDevice handles that use this hierarchy
        self.last_solve = dict(iters=0, relres=0.0)

    @classmethod
    def for_device(cls, fine_device, mesh_chain, quadrature_degree=6, device=0, cheb_degree=3,
                   element_wise_material=False, elementwise_pressure=False):
        """Hierarchy under an existing finest-level device; `mesh_chain` = [(nodes, quads), ...] coarse to fine as
        recorded by femo_alpha_b200.mesh.refine_uniform."""
        return cls(None, None, len(mesh_chain), element_wise_material, elementwise_pressure, quadrature_degree, device,
                   cheb_degree, meshes=mesh_chain, fine_device=fine_device)

    def set_input(self, name, values):
        """`values`: callable of the vertex coordinates x (3, n) returning (n,) or (n, 3), a scalar, or the finest
        level's array.  The finest device forwards it to the coarser levels (ShellDevice.set_input)."""
        nodes, quads = self.meshes[0]
        cellwise = self._cellwise(name)
        n = quads.shape[0] if cellwise else nodes.shape[0]
        if callable(values):
            x = nodes[quads].mean(axis=1) if cellwise else nodes
            v = np.asarray(values(x.T), dtype=np.float64)
        else:
            v = np.asarray(values, dtype=np.float64)
            if v.size == 1:
                v = np.full(n, float(v))
        if name in ("F_solid", "uhat") and v.size == n:
            raise ValueError(name + " needs 3 components per entry")
        self.fine.set_input(name, np.ascontiguousarray(v).ravel())

    def _cellwise(self, name):
        return (self.ewm and name in ("thickness", "E", "nu", "density")) or (self.ewp and name == "F_solid")

    # the parent's vertices keep their numbers under refine_quads: a nodal array of the finest level restricts by slicing;
    # cell-wise arrays (children 4 c .. 4 c + 3 of cell c) restrict by averaging
    def _coarse_inputs(self, name, values):
        if hasattr(values, "detach"):
            values = values.detach().cpu().numpy()
        if self._cellwise(name):
            v = np.asarray(values, dtype=np.float64).reshape(self.meshes[0][1].shape[0], -1)
            for lvl in self.levels[1:]:
                v = v.reshape(v.shape[0] // 4, 4, -1).mean(axis=1)
                lvl.set_input(name, np.ascontiguousarray(v).ravel())
        else:
            v = np.asarray(values, dtype=np.float64).reshape(self.meshes[0][0].shape[0], -1)
            for lvl, (nodes, _) in zip(self.levels[1:], self.meshes[1:]):
                lvl.set_input(name, np.ascontiguousarray(v[:nodes.shape[0]]).ravel())
        self._dirty = True

    def set_strong_bc(self, predicate):
        """predicate(x) with x of shape (3, n): True on clamped block nodes (locate_dofs_geometrical analogue,
        femo_alpha/rm_shell/rm_shell_model.py:168-180)."""
        on = np.asarray(predicate(self.fine.block_coords().T)).astype(bool)
        blocks = np.nonzero(on)[0]
        self.fine.set_strong_bc(np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()))

    # coarse Q2 block j is the fine vertex block j, coarse Q1 block v the fine Q1 block v: clamped dofs map by index
    def _coarse_bcs(self, dofs):
        dofs = np.asarray(dofs, dtype=np.int64)
        for f, c in zip(self.levels[:-1], self.levels[1:]):
            blk, cmp = dofs // 3, dofs % 3
            q2 = blk < c.nq2
            q1 = (blk >= f.nq2) & (blk - f.nq2 < c.nn)
            dofs = np.sort(np.concatenate([3 * blk[q2] + cmp[q2], 3 * (c.nq2 + blk[q1] - f.nq2) + cmp[q1]]))
            c.set_strong_bc(dofs)
        self._dirty = True

    def setup(self, reuse_coarse=None):
        if reuse_coarse is None:
            reuse_coarse = self._n_setups % self.coarse_refresh != 0
        L.check(self.lib.femo_mg_setup(self._mg, self.cheb_degree, int(bool(reuse_coarse))))
        self._n_setups += 1
        self._dirty = False

    def set_refine(self, steps, inner_rtol=1e-6):
        """Iterative refinement around the V-cycle PCG (femo_mg_set_refine): `steps` corrections x += PCG(b - A x) with the
        residual accumulated in double-double; solve() then reports the TRUE relative residual (bounded below by
        eps |A| |x| / |b|, x being stored in fp64) instead of CG's recurrence residual.  Stands in for the direct solve
        behind the reference's KSP (femo_alpha/fea/utils_dolfinx.py:514-531)."""
        L.check(self.lib.femo_mg_set_refine(self._mg, int(steps), float(inner_rtol)))

    def lambda_max(self, level):
        v = C.c_double(0.0)
        L.check(self.lib.femo_mg_get_lambda(self._mg, int(level), C.byref(v)))
        return v.value

    def solve(self, rhs, rtol=1e-10, max_it=2000, zero_bc=True, out=None):
        """A x = rhs on the finest level (A = dR/du with BCs)."""
        if self._dirty:
            self.setup()
        if out is None:
            out = np.empty(self.fine.ndof) if isinstance(rhs, np.ndarray) else rhs.new_empty(rhs.shape)
        if isinstance(rhs, np.ndarray):
            rhs = np.ascontiguousarray(rhs, dtype=np.float64)
        it = C.c_int(0); rr = C.c_double(0.0)
        L.check(self.lib.femo_mg_solve(self._mg, L.ptr(rhs), L.ptr(out), float(rtol), int(max_it), C.byref(it), C.byref(rr),
                                       L.on_device(rhs), int(bool(zero_bc))))
        self.last_solve = dict(iters=it.value, relres=rr.value)
        return out

    def solve_state(self, rtol=1e-10, max_it=2000):
        """K u = F with homogeneous Dirichlet values (FEA.solve, fea_dolfinx.py:159-170); the state lands in `fine`."""
        F = self.fine.assemble_load()
        u = self.solve(F, rtol, max_it, zero_bc=True)
        self.fine.set_state(u)
        self.fine.last_solve = dict(self.last_solve)
        return self.last_solve["iters"], self.last_solve["relres"]

    def solve_adjoint(self, rhs, rtol=1e-10, max_it=2000):
        """A^T lam = rhs, lam[bc] = 0 (state_operation.py:188-220); A is symmetric."""
        return self.solve(rhs, rtol, max_it, zero_bc=True)

    def close(self):
        """Ordered tear-down: dynamic handles, then the multigrid object, then the level contexts (the C side keeps
        plain pointers between them)."""
        if getattr(self, "_closing", False):
            return
        self._closing = True
        for ref in getattr(self, "_children", []):
            child = ref()
            if child is not None:
                child.close()
        self._children = []
        if getattr(self, "_mg", None) is not None and self._mg.value:
            self.lib.femo_mg_destroy(self._mg)
            self._mg = C.c_void_p()
        for d in getattr(self, "levels", []):
            d._mg = None
            d._owner = None
            d.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
