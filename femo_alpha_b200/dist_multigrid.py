"""Host side of the N-rank V-cycle (open item 2): nested partitions of a `refine_quads` hierarchy and the
rank-local prolongation between two partitioned levels.  NumPy only; the device side (distributed smoothing with the
existing halo exchange, restriction through `femo_halo_accumulate`, replicated coarse levels) is the next round's work --
the algorithm is prototyped on these functions in tests/dist_mg_worker.py.

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


class DistributedHierarchy:
    """V-cycle preconditioned solves on N GPUs: the finest level is a DistributedShell (nested partition), the coarser
    levels are replicated on every rank (csrc/mg.cu: halo exchange before every operator application of the finest
    level, one all-reduce of the level-1 residual per V-cycle).  One process per GPU.

    NOT YET RUN ON HARDWARE (written after the round-1 GPU budget was spent); the algorithm is the one of
    tests/dist_mg_worker.py, the transfer arrays are checked on the CPU in tests/test_multigrid_host.py."""

    def __init__(self, coarse_nodes, coarse_quads, levels, rank, world, device=0, quadrature_degree=6, cheb_degree=3,
                 part_coarse=None):
        import ctypes as C
        from . import _lib as L
        from .dist_shell import DistributedShell
        from .multigrid import refine_quads
        from .partition import partition_cells_rcb
        from .shell_device import ShellDevice
        if levels < 2:
            raise ValueError("a hierarchy needs at least two levels")
        meshes = [(np.asarray(coarse_nodes, dtype=np.float64), np.asarray(coarse_quads, dtype=np.int32))]
        for _ in range(levels - 1):
            meshes.append(refine_quads(*meshes[-1]))
        self.meshes = meshes[::-1]                                 # level 0 = finest
        if part_coarse is None:
            n, q = self.meshes[-1]
            part_coarse = partition_cells_rcb(n[q].mean(axis=1), world)
        self.parts = nested_parts(part_coarse, levels)
        self.fine = DistributedShell(self.meshes[0][0], self.meshes[0][1], rank, world, part=self.parts[0], device=device,
                                     quadrature_degree=quadrature_degree)
        self.coarse = [ShellDevice(n, q, False, False, quadrature_degree, device) for n, q in self.meshes[1:]]
        self.lib, self._L, self._C = L.load(), L, C
        ctxs = [self.fine.dev._ctx] + [d._ctx for d in self.coarse]
        self._mg = C.c_void_p()
        L.check(self.lib.femo_mg_create(levels, (C.c_void_p * levels)(*ctxs), C.byref(self._mg)))
        self._keep = []
        lp_all = replicated_part(*self.meshes[1])
        P0 = local_prolongation(self.fine.lp, lp_all, self.fine.dev.cell_blocks(), self.coarse[0].cell_blocks())
        self._set_transfer(0, P0, self.fine.lp.own_block_mask.astype(np.float64))
        from .multigrid import prolongation as _prolongation
        for l in range(1, levels - 1):
            f, c = self.coarse[l - 1], self.coarse[l]
            self._set_transfer(l, _prolongation(f.cell_blocks(), c.cell_blocks(), f.nb, c.nb), None)
        self.cheb_degree = int(cheb_degree)
        self._dirty = True
        self.last_solve = dict(iters=0, relres=0.0)

    def _set_transfer(self, level, P, row_weight):
        a = transfer_arrays(P, row_weight)
        self._keep.append(a)
        self._L.check(self.lib.femo_mg_set_transfer(self._mg, level, *[self._L.ptr(v) for v in a]))

    def init_comm(self, dist=None):
        self.fine.init_comm(dist)

    def set_input(self, name, global_values):
        """global nodal arrays of the FINEST mesh (mesh order); parents keep their vertex numbers under refine_quads, so
        the coarser levels take the leading entries"""
        v = np.asarray(global_values, dtype=np.float64)
        self.fine.set_input(name, v)
        for d, (nodes, _) in zip(self.coarse, self.meshes[1:]):
            if v.size == 1:
                d.set_input(name, v.ravel())
            else:
                k = v.size // self.meshes[0][0].shape[0]
                d.set_input(name, np.ascontiguousarray(v.reshape(-1, k)[:nodes.shape[0]]).ravel())
        self._dirty = True

    def set_strong_bc(self, predicate):
        self.fine.set_strong_bc(predicate)
        for d in self.coarse:
            blocks = np.nonzero(np.asarray(predicate(d.block_coords().T)).astype(bool))[0]
            d.set_strong_bc(np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()))
        self._dirty = True

    def setup(self):
        self._L.check(self.lib.femo_mg_setup(self._mg, self.cheb_degree, 0))
        self._dirty = False

    def solve(self, rhs_local, rtol=1e-10, max_it=2000, zero_bc=True):
        """A x = rhs on the finest level; rhs and x in this rank's local block order (collective)."""
        if self._dirty:
            self.setup()
        C, L = self._C, self._L
        rhs = np.ascontiguousarray(rhs_local, dtype=np.float64)
        out = np.empty(self.fine.dev.ndof)
        it = C.c_int(0); rr = C.c_double(0.0)
        L.check(self.lib.femo_mg_solve(self._mg, L.ptr(rhs), L.ptr(out), float(rtol), int(max_it), C.byref(it), C.byref(rr), 0,
                                       int(bool(zero_bc))))
        self.last_solve = dict(iters=it.value, relres=rr.value)
        return out

    def solve_state(self, rtol=1e-10, max_it=2000):
        u = self.solve(self.fine.dev.assemble_load(), rtol, max_it)
        self.fine.dev.set_state(u)
        return self.last_solve["iters"], self.last_solve["relres"]

    def close(self):
        if getattr(self, "_mg", None) is not None and self._mg.value:
            self.lib.femo_mg_destroy(self._mg)
            self._mg = self._C.c_void_p()
        for d in [getattr(self, "fine", None)] + list(getattr(self, "coarse", [])):
            if d is not None:
                d.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
