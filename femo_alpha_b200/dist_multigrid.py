"""Host side of the N-rank V-cycle (open item 2): nested partitions of a `refine_quads` hierarchy and the
rank-local prolongation between two partitioned levels (NumPy), and `DistributedHierarchy`, the handle that drives the
device side (csrc/mg.cu: distributed smoothing with a halo exchange per operator application, restriction through
`femo_halo_accumulate`, replicated coarse levels below).  The algorithm was prototyped on these functions over gloo
(tests/dist_mg_worker.py).

Nested partition: a fine cell belongs to the part of its parent.  Then every fine block node owned by a rank lies in an
owned fine cell whose parent is an owned coarse cell, so its interpolation stencil (the block nodes of that parent) is
inside the rank's local coarse mesh and the local prolongation needs no communication."""
from __future__ import annotations

import numpy as np

from .multigrid import prolongation


def nested_parts(part_coarse, levels):
    """[part of level 0 (finest), ..., part of the coarsest level]: `refine_quads` numbers the children of cell c as
    4 c + k, so the part array of a level is its parent's repeated four times."""
    parts = [np.asarray(part_coarse, dtype=np.int32)]
    for _ in range(levels - 1):
        parts.append(np.repeat(parts[-1], 4))
    return parts[::-1]


def local_prolongation(lp_fine, lp_coarse, fine_cell_blocks, coarse_cell_blocks):
    """Prolongation from the local coarse blocks to the local fine blocks of one rank (scipy CSR, block-scalar).

    lp_fine, lp_coarse: partition.LocalPart of the two levels (nested partition); *_cell_blocks: the (n_local_cells, 13)
    block tables of the two local meshes (femo_get_cell_blocks / ShellTopology.cell_blocks).  Rows of fine block nodes that
    no local fine cell with a local parent holds (ghost nodes at the outer rim) are empty: their values come from the halo
    exchange that follows a prolongation."""
    fcb = np.asarray(fine_cell_blocks, dtype=np.int64)
    ccb = np.asarray(coarse_cell_blocks, dtype=np.int64)
    parent_global = lp_fine.cell_global // 4
    child = lp_fine.cell_global % 4
    to_local = np.full(lp_coarse.n_global_cells, -1, dtype=np.int64)
    to_local[lp_coarse.cell_global] = np.arange(lp_coarse.cell_global.size)
    parent = to_local[parent_global]
    have = parent >= 0
    nb_f, nb_c = int(fcb.max()) + 1, int(ccb.max()) + 1
    P = prolongation(fcb[have], ccb, nb_f, nb_c, parent=parent[have], child=child[have], allow_missing_rows=True)
    owned_rows = np.nonzero(lp_fine.own_block_mask)[0]
    if (np.diff(P.indptr)[owned_rows] == 0).any():
        raise RuntimeError("local_prolongation: an owned fine block node has no local parent cell (partition not nested?)")
    return P


def replicated_part(nodes, quads):
    """LocalPart of a level that lives entirely on every rank (world 1): local numbering = global numbering."""
    from .partition import build_local_part
    return build_local_part(nodes, quads, np.zeros(np.asarray(quads).shape[0], dtype=np.int32), 0, 1)


def transfer_arrays(P, row_weight=None):
    """The six arrays of femo_mg_set_transfer for a prolongation P (scipy CSR, block-scalar): P itself and R = P^T with
    the same number of stored entries.  `row_weight` (one value per fine block, the ownership mask of a distributed
    level) scales the rows of P inside R only: restriction then sums each fine node once over the ranks, while
    prolongation still fills every local row it can."""
    import scipy.sparse as sp
    P = P.tocsr(); P.sort_indices()
    coo = P.tocoo()
    w = coo.data if row_weight is None else coo.data * np.asarray(row_weight, dtype=np.float64)[coo.row]
    R = sp.csr_matrix((w, (coo.col, coo.row)), shape=(P.shape[1], P.shape[0]))      # explicit zeros are kept
    R.sort_indices()
    assert R.nnz == P.nnz
    return [np.ascontiguousarray(P.indptr, np.int32), np.ascontiguousarray(P.indices, np.int32),
            np.ascontiguousarray(P.data, np.float64), np.ascontiguousarray(R.indptr, np.int32),
            np.ascontiguousarray(R.indices, np.int32), np.ascontiguousarray(R.data, np.float64)]


class _AllCells:
    """Stand-in for the LocalPart of a replicated level (every rank holds every cell; local = global numbering)."""

    def __init__(self, n_cells):
        self.cell_global = np.arange(n_cells, dtype=np.int64)
        self.n_global_cells = int(n_cells)


class DistributedHierarchy:
    """V-cycle preconditioned solves on N GPUs, one process per GPU (csrc/mg.cu).

    The finest `n_dist` levels are DistributedShell contexts over NESTED partitions (a fine cell lives on the rank of its
    parent); the coarser levels are replicated on every rank and the coarsest is solved directly (csrc/coarse.cu).
    Communication per V-cycle: one halo exchange per operator application of a distributed level, a reverse halo
    exchange after a restriction between two distributed levels, one all-reduce of the first replicated residual.
    The preconditioner is the same linear operator as on one GPU: iteration counts do not depend on the rank count
    (tests/dist_gpu_mg_worker.py)."""

    def __init__(self, coarse_nodes, coarse_quads, levels, rank, world, device=0, quadrature_degree=6, cheb_degree=3,
                 part_coarse=None, n_dist=1):
        import ctypes as C
        from . import _lib as L
        from .dist_shell import DistributedShell
        from .multigrid import prolongation, refine_quads
        from .partition import partition_cells
        from .shell_device import ShellDevice
        if levels < 2:
            raise ValueError("a hierarchy needs at least two levels")
        n_dist = max(1, min(int(n_dist), levels - 1))          # the coarsest level is always replicated
        self.n_dist, self.rank, self.world = n_dist, int(rank), int(world)
        meshes = [(np.asarray(coarse_nodes, dtype=np.float64), np.asarray(coarse_quads, dtype=np.int32))]
        for _ in range(levels - 1):
            meshes.append(refine_quads(*meshes[-1]))
        self.meshes = meshes[::-1]                                 # level 0 = finest
        if part_coarse is None:
            n, q = self.meshes[-1]
            part_coarse = partition_cells(n, q, world)
        self.parts = nested_parts(part_coarse, levels)
        self.dist = [DistributedShell(self.meshes[l][0], self.meshes[l][1], rank, world, part=self.parts[l], device=device,
                                      quadrature_degree=quadrature_degree) for l in range(n_dist)]
        self.fine = self.dist[0]
        self.coarse = [ShellDevice(n, q, False, False, quadrature_degree, device) for n, q in self.meshes[n_dist:]]
        self.devs = [d.dev for d in self.dist] + self.coarse       # one device handle per level
        self.lib, self._L, self._C = L.load(), L, C
        self._mg = C.c_void_p()
        self._keep = []
        self._transfers = []
        for l in range(levels - 1):
            f, c = self.devs[l], self.devs[l + 1]
            if l < n_dist:
                lp_c = self.dist[l + 1].lp if l + 1 < n_dist else _AllCells(self.meshes[l + 1][1].shape[0])
                P = local_prolongation(self.dist[l].lp, lp_c, f.cell_blocks(), c.cell_blocks())
                self._transfers.append((P, self.dist[l].lp.own_block_mask.astype(np.float64)))
            else:
                self._transfers.append((prolongation(f.cell_blocks(), c.cell_blocks(), f.nb, c.nb), None))
        self.cheb_degree = int(cheb_degree)
        self._dirty = True
        self._comm = False
        self.last_solve = dict(iters=0, relres=0.0)

    def _create(self):
        L, C = self._L, self._C
        levels = len(self.devs)
        ctxs = [d._ctx for d in self.devs]
        L.check(self.lib.femo_mg_create(levels, (C.c_void_p * levels)(*ctxs), C.byref(self._mg)))
        for l, (P, w) in enumerate(self._transfers):
            a = transfer_arrays(P, w)
            self._keep.append(a)
            L.check(self.lib.femo_mg_set_transfer(self._mg, l, *[L.ptr(v) for v in a]))
        self._transfers = None
        if getattr(self, "_refine", None):
            L.check(self.lib.femo_mg_set_refine(self._mg, *self._refine))

    def set_refine(self, steps, inner_rtol=1e-6):
        """Iterative refinement around the distributed V-cycle PCG (see ShellHierarchy.set_refine); kept until close()."""
        self._refine = (int(steps), float(inner_rtol))
        if self._mg.value:
            self._L.check(self.lib.femo_mg_set_refine(self._mg, *self._refine))

    def init_comm(self, dist=None):
        """Collective: one NCCL communicator (finest level), shared by the other distributed levels."""
        self.fine.init_comm(dist)
        for d in self.dist[1:]:
            self._L.check(self.lib.femo_dist_share_comm(d.dev._ctx, self.fine.dev._ctx))
        self._comm = True

    def set_input(self, name, global_values):
        """global nodal arrays of the FINEST mesh (mesh order); parents keep their vertex numbers under refine_quads, so
        the coarser levels take the leading entries"""
        v = np.asarray(global_values, dtype=np.float64)
        for l, (nodes, _) in enumerate(self.meshes):
            if v.size == 1:
                vl = v.ravel()
            else:
                k = v.size // self.meshes[0][0].shape[0]
                vl = np.ascontiguousarray(v.reshape(-1, k)[:nodes.shape[0]]).ravel()
            (self.dist[l] if l < self.n_dist else self.coarse[l - self.n_dist]).set_input(name, vl)
        self._dirty = True

    def set_strong_bc(self, predicate):
        for d in self.dist:
            d.set_strong_bc(predicate)
        for d in self.coarse:
            blocks = np.nonzero(np.asarray(predicate(d.block_coords().T)).astype(bool))[0]
            d.set_strong_bc(np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()))
        self._dirty = True

    def prepare(self):
        """Create the device-side hierarchy (femo_mg_create + upload of the transfer operators) ahead of the first setup();
        setup() does it on demand otherwise.  One-time work, not part of the per-design set-up."""
        if not self._mg.value:
            self._create()

    def setup(self, reuse_coarse=False):
        if self.world > 1 and not self._comm:
            raise RuntimeError("DistributedHierarchy: call init_comm() first")
        if not self._mg.value:
            self._create()
        self._L.check(self.lib.femo_mg_setup(self._mg, self.cheb_degree, int(bool(reuse_coarse))))
        self._dirty = False

    def solve(self, rhs_local, rtol=1e-10, max_it=2000, zero_bc=True):
        """A x = rhs on the finest level; rhs and x in this rank's local block order (collective).  Owned entries of
        rhs must be complete (assemble_load / output_grad of an owner-computes context are); ghost entries are ignored."""
        if self._dirty:
            self.setup()
        C, L = self._C, self._L
        rhs = np.ascontiguousarray(rhs_local, dtype=np.float64)
        out = np.empty(self.fine.dev.ndof)
        it = C.c_int(0); rr = C.c_double(0.0)
        L.check(self.lib.femo_mg_solve(self._mg, L.ptr(rhs), L.ptr(out), float(rtol), int(max_it), C.byref(it), C.byref(rr), 0,
                                       int(bool(zero_bc))))
        self.last_solve = dict(iters=it.value, relres=rr.value, device_ms=self.fine.dev.last_timings()["solve_ms"])
        return out

    def solve_state(self, rtol=1e-10, max_it=2000):
        u = self.solve(self.fine.dev.assemble_load(), rtol, max_it)
        self.fine.dev.set_state(u)
        return self.last_solve["iters"], self.last_solve["relres"]

    def solve_adjoint(self, rhs_local, rtol=1e-10, max_it=2000):
        """A^T lam = rhs, lam[bc] = 0 (state_operation.py:188-220; A is symmetric)"""
        return self.solve(rhs_local, rtol, max_it, zero_bc=True)

    def close(self):
        if getattr(self, "_mg", None) is not None and self._mg.value:
            self.lib.femo_mg_destroy(self._mg)
            self._mg = self._C.c_void_p()
        for d in list(getattr(self, "dist", []))[::-1] + list(getattr(self, "coarse", [])):
            if d is not None:
                d.close()
        self.dist, self.coarse = [], []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
