"""Mesh partitioning for the multi-GPU RM-shell path: host-side NumPy only.

One process per GPU.  Cells are split into ``world`` parts by METIS k-way partitioning of the cell-adjacency graph
(curved / non-manifold meshes such as wing boxes; `partition_cells_graph`) or by recursive coordinate bisection of the
cell centroids (flat plates; `partition_cells_rcb`); `partition_cells` chooses.  A Q2/Q1 block node (vertex, edge mid-point, cell centre, rotation vertex) is
*owned* by the lowest part among the cells that contain it.  A rank keeps every cell that contains
a node it owns -- its own cells plus one layer of ghost cells -- so that the rows of its owned
nodes are assembled completely from local data (owner-computes: no collective in assembly, and the
result is bitwise independent of the number of ranks' scheduling).  The only data that ever crosses
ranks are the ghost entries of a vector before an operator application (halo exchange) and the
scalars of dot products.

Every rank evaluates this module on the *global* mesh (cheap: integer work on the connectivity),
so no set-up communication is needed: both sides of a halo pair derive the same list, ordered by a
global node key.

The reference has no counterpart (its wrapper is serial); dolfinx's own ghosting
(`GhostMode.shared_facet`) follows the same owner/ghost idea.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

_LOCAL_EDGES = np.array([(0, 1), (0, 2), (1, 3), (2, 3)])   # tensor vertex order, as csrc/topology.cu
_VTK_TO_TENSOR = np.array([0, 1, 3, 2])


def partition_cells_rcb(centroids: np.ndarray, world: int) -> np.ndarray:
    """Recursive coordinate bisection; parts are numbered so that part ids increase along the cut axis.
    Works for any ``world`` >= 1 (uneven splits are proportional)."""
    n = centroids.shape[0]
    part = np.zeros(n, dtype=np.int32)

    def rec(idx, first, count):
        if count == 1:
            part[idx] = first
            return
        left = count // 2
        pts = centroids[idx]
        axis = int(np.argmax(pts.max(0) - pts.min(0)))
        k = int(round(len(idx) * left / count))
        order = np.argsort(pts[:, axis], kind="stable")
        rec(idx[order[:k]], first, left)
        rec(idx[order[k:]], first + left, count - left)

    rec(np.arange(n), 0, int(world))
    return part


def cell_adjacency(quads: np.ndarray):
    """Dual graph of a quad surface mesh as CSR (xadj, adjncy), int64: two cells are adjacent when they share an edge.
    Non-manifold edges (wing-box junctions: 3 or 4 cells on one edge) connect every pair of their cells."""
    q = np.asarray(quads, dtype=np.int64)
    nc = q.shape[0]
    nn = int(q.max()) + 1
    ev = np.stack([q, np.roll(q, -1, axis=1)], axis=2)                     # (nc, 4, 2): VTK cyclic order -> the 4 edges
    key = (np.minimum(ev[..., 0], ev[..., 1]) * nn + np.maximum(ev[..., 0], ev[..., 1])).ravel()
    cell = np.repeat(np.arange(nc, dtype=np.int64), 4)
    order = np.argsort(key, kind="stable")
    key, cell = key[order], cell[order]
    start = np.r_[0, np.nonzero(key[1:] != key[:-1])[0] + 1, key.size]
    cnt = np.diff(start)
    rows, cols = [], []
    for k in np.unique(cnt[cnt > 1]):                                         # all pairs inside every group of k cells
        g = start[:-1][cnt == k]
        members = cell[g[:, None] + np.arange(k)[None, :]]                    # (ng, k)
        i, j = np.nonzero(~np.eye(int(k), dtype=bool))
        rows.append(members[:, i].ravel()); cols.append(members[:, j].ravel())
    if not rows:
        return np.zeros(nc + 1, dtype=np.int64), np.zeros(0, dtype=np.int64)
    pair = np.unique(np.concatenate(rows) * nc + np.concatenate(cols))       # two cells sharing two edges: one arc
    r, c = pair // nc, pair % nc
    xadj = np.zeros(nc + 1, dtype=np.int64)
    np.add.at(xadj, r + 1, 1)
    return np.cumsum(xadj), np.ascontiguousarray(c)


_METIS = None


def _metis_lib():
    """libfemo_metis.so (csrc/metis_shim.c linked against CUDA's libmetis_static.a by build.py), or None."""
    global _METIS
    if _METIS is None:
        import ctypes as C
        import os
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libfemo_metis.so")
        if not os.path.exists(path):
            _METIS = False
        else:
            lib = C.CDLL(path)
            lib.femo_metis_kway.restype = C.c_int
            lib.femo_metis_kway.argtypes = [C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
            _METIS = lib
    return _METIS or None


def partition_cells_graph(quads: np.ndarray, world: int, centroids=None, seed: int = 20261104) -> np.ndarray:
    """k-way partition of the cell-adjacency graph (METIS, fixed seed: every rank computes the same partition).
    Parts are renumbered by the position of their centroid along the longest axis of the mesh, so that neighbouring
    part ids tend to be neighbouring ranks.  Falls back to recursive coordinate bisection (with a warning) when
    libfemo_metis.so was not built; `centroids` (nc, 3) is then required."""
    import ctypes as C
    q = np.asarray(quads)
    nc = q.shape[0]
    if world <= 1:
        return np.zeros(nc, dtype=np.int32)
    lib = _metis_lib()
    if lib is None:
        import warnings
        warnings.warn("femo_alpha_b200: libfemo_metis.so not built (no libmetis_static.a?) -- falling back to RCB")
        if centroids is None:
            raise RuntimeError("partition_cells_graph: RCB fallback needs the cell centroids")
        return partition_cells_rcb(np.asarray(centroids), world)
    xadj, adjncy = cell_adjacency(q)
    part = np.empty(nc, dtype=np.int64)
    cut = C.c_int64(0)
    r = lib.femo_metis_kway(nc, xadj.ctypes.data_as(C.c_void_p), adjncy.ctypes.data_as(C.c_void_p), int(world), 0, int(seed),
                            part.ctypes.data_as(C.c_void_p), C.byref(cut))
    if r != 0:
        raise RuntimeError("METIS_PartGraphKway failed (status %d)" % r)
    if np.unique(part).size != world:
        raise RuntimeError("METIS returned %d of %d parts" % (np.unique(part).size, world))
    if centroids is not None:                                                    # order the parts along the longest axis
        cen = np.asarray(centroids, dtype=np.float64)
        axis = int(np.argmax(cen.max(0) - cen.min(0)))
        pos = np.array([cen[part == p, axis].mean() for p in range(world)])
        rank_of = np.empty(world, dtype=np.int64)
        rank_of[np.argsort(pos, kind="stable")] = np.arange(world)
        part = rank_of[part]
    partition_cells_graph.last_edge_cut = int(cut.value)
    return part.astype(np.int32)


def partition_cells(nodes: np.ndarray, quads: np.ndarray, world: int, method: str = "auto") -> np.ndarray:
    """'rcb', 'graph' (METIS k-way on the dual graph) or 'auto': graph partitioning for meshes with non-manifold
    junctions or curvature (wing boxes), coordinate bisection for flat single-sheet meshes (plates), where RCB gives
    straight cuts."""
    nodes = np.asarray(nodes, dtype=np.float64)
    q = np.asarray(quads)
    cen = nodes[q].mean(axis=1)
    if method == "rcb":
        return partition_cells_rcb(cen, world)
    if method == "graph":
        return partition_cells_graph(q, world, cen)
    if method != "auto":
        raise ValueError("partition method must be 'auto', 'rcb' or 'graph'")
    ext = np.sort(nodes.max(0) - nodes.min(0))
    flat = ext[0] <= 1e-9 * max(ext[2], 1e-300)
    return partition_cells_rcb(cen, world) if (flat or _metis_lib() is None) else partition_cells_graph(q, world, cen)


@dataclass
class LocalPart:
    """What one rank needs: its local mesh (owned cells first, then ghost cells) and the maps back."""
    rank: int
    world: int
    nodes: np.ndarray            # (nn_loc, 3) local vertex coordinates
    quads: np.ndarray            # (nc_loc, 4) local connectivity, VTK cyclic order (as the caller's)
    vertex_global: np.ndarray    # (nn_loc,) global vertex id of each local vertex (ascending)
    cell_global: np.ndarray      # (nc_loc,) global cell id; the first n_owned_cells are owned
    n_owned_cells: int
    # global ownership tables restricted to the local cells, tensor order: (nc_loc, 9) owners of the 9 Q2 nodes
    own9: np.ndarray
    gkey9: np.ndarray            # (nc_loc, 9) global keys of the 9 Q2 nodes
    n_global_vertices: int = 0
    n_global_cells: int = 0
    n_global_edges: int = 0
    # filled by attach_blocks()
    block_owner: np.ndarray = field(default=None)     # (nb_loc,) owning rank of every local block node
    block_gkey: np.ndarray = field(default=None)      # (nb_loc,) global key (unique per global block node)
    neighbors: list = field(default_factory=list)     # sorted neighbour ranks
    send_blocks: dict = field(default_factory=dict)   # q -> local block ids (owned here, ghost on q), key-sorted
    recv_blocks: dict = field(default_factory=dict)   # q -> local block ids (ghost here, owned by q), key-sorted

    @property
    def own_block_mask(self):
        return self.block_owner == self.rank


def _global_tables(quads_vtk: np.ndarray, n_vertices: int, part: np.ndarray, world: int):
    cv = quads_vtk[:, _VTK_TO_TENSOR].astype(np.int64)
    nc = cv.shape[0]
    v_owner = np.full(n_vertices, world, dtype=np.int32)
    np.minimum.at(v_owner, cv.ravel(), np.repeat(part, 4))
    a = cv[:, _LOCAL_EDGES[:, 0]]
    b = cv[:, _LOCAL_EDGES[:, 1]]
    ekey = np.minimum(a, b) * n_vertices + np.maximum(a, b)            # (nc,4)
    uniq, inv = np.unique(ekey.ravel(), return_inverse=True)
    inv = inv.reshape(nc, 4)
    e_owner = np.full(uniq.size, world, dtype=np.int32)
    np.minimum.at(e_owner, inv.ravel(), np.repeat(part, 4))
    own9 = np.concatenate([v_owner[cv], e_owner[inv], part[:, None].astype(np.int32)], axis=1)   # (nc,9)
    # global keys: vertices [0,NV) | edges NV + edge id | centres NV + NE + cell
    ne = uniq.size
    gkey9 = np.concatenate([cv, n_vertices + inv, (n_vertices + ne + np.arange(nc))[:, None]], axis=1).astype(np.int64)
    return cv, own9, gkey9, ne


def build_local_part(nodes: np.ndarray, quads_vtk: np.ndarray, part: np.ndarray, rank: int, world: int) -> LocalPart:
    """Local mesh of ``rank``: every cell containing a block node owned by ``rank``."""
    nodes = np.asarray(nodes, dtype=np.float64)
    quads_vtk = np.asarray(quads_vtk)
    part = np.asarray(part, dtype=np.int32)
    nv = nodes.shape[0]
    cv, own9, gkey9, ne = _global_tables(quads_vtk, nv, part, world)
    needed = (own9 == rank).any(axis=1)
    owned = part == rank
    assert not (owned & ~needed).any()
    loc_cells = np.concatenate([np.nonzero(owned)[0], np.nonzero(needed & ~owned)[0]])
    vglob = np.unique(quads_vtk[loc_cells].ravel())
    remap = np.full(nv, -1, dtype=np.int64)
    remap[vglob] = np.arange(vglob.size)
    lp = LocalPart(rank=rank, world=world, nodes=np.ascontiguousarray(nodes[vglob]),
                   quads=np.ascontiguousarray(remap[quads_vtk[loc_cells]].astype(np.int32)),
                   vertex_global=vglob, cell_global=loc_cells, n_owned_cells=int(owned.sum()),
                   own9=own9[loc_cells], gkey9=gkey9[loc_cells],
                   n_global_vertices=nv, n_global_cells=quads_vtk.shape[0], n_global_edges=ne)
    # halo pairs need to know which *other* ranks hold my owned nodes: cells needed by q that contain nodes of mine
    lp._global_own9 = own9
    lp._global_gkey9 = gkey9
    return lp


def attach_blocks(lp: LocalPart, cell_blocks: np.ndarray) -> LocalPart:
    """Map ownership onto the library's local block numbering.  ``cell_blocks`` is (nc_loc, 13): the
    9 Q2 + 4 Q1 local block ids of every local cell (femo_get_cell_blocks / ShellTopology.cell_blocks)."""
    cb = np.asarray(cell_blocks, dtype=np.int64)
    nb = int(cb.max()) + 1
    nq1_off = lp.n_global_vertices + lp.n_global_edges + lp.n_global_cells       # keys of rotation nodes
    owner = np.full(nb, -1, dtype=np.int32)
    gkey = np.full(nb, -1, dtype=np.int64)
    owner[cb[:, :9].ravel()] = lp.own9.ravel()
    gkey[cb[:, :9].ravel()] = lp.gkey9.ravel()
    owner[cb[:, 9:].ravel()] = lp.own9[:, :4].ravel()
    gkey[cb[:, 9:].ravel()] = nq1_off + lp.gkey9[:, :4].ravel()
    assert (owner >= 0).all() and (gkey >= 0).all()
    lp.block_owner, lp.block_gkey = owner, gkey
    order = np.argsort(gkey)
    sorted_keys = gkey[order]

    def to_local(keys):
        pos = np.searchsorted(sorted_keys, keys)
        assert np.array_equal(sorted_keys[pos], keys)
        return order[pos].astype(np.int32)

    # receive side: my ghost blocks grouped by owner, key-sorted
    recv = {}
    ghosts = np.nonzero(owner != lp.rank)[0]
    for q in np.unique(owner[ghosts]):
        ids = ghosts[owner[ghosts] == q]
        recv[int(q)] = ids[np.argsort(gkey[ids])].astype(np.int32)
    # send side: nodes I own inside cells that rank q keeps (q owns another node of the same cell)
    g_own9, g_key9 = lp._global_own9, lp._global_gkey9
    mine = (g_own9 == lp.rank)
    cells_with_mine = mine.any(axis=1)
    send = {}
    others = np.unique(g_own9[cells_with_mine])
    for q in others:
        if q == lp.rank:
            continue
        cells_q = cells_with_mine & (g_own9 == q).any(axis=1)
        if not cells_q.any():
            continue
        m = mine[cells_q]
        k9 = g_key9[cells_q]
        keys_u = np.unique(k9[m])                                   # displacement nodes I own in those cells
        mv = m[:, :4]
        keys_t = np.unique(nq1_off + k9[:, :4][mv])                 # rotation nodes follow their vertex
        send[int(q)] = to_local(np.concatenate([keys_u, keys_t]))
    lp.send_blocks, lp.recv_blocks = send, recv
    lp.neighbors = sorted(set(send) | set(recv))
    for q in lp.neighbors:
        send.setdefault(q, np.zeros(0, dtype=np.int32))
        recv.setdefault(q, np.zeros(0, dtype=np.int32))
    return lp


def halo_arrays(lp: LocalPart):
    """Flat arrays for femo_dist_setup: neighbours, block offsets and block lists (send, recv)."""
    nbr = np.asarray(lp.neighbors, dtype=np.int32)
    s_off, r_off = [0], [0]
    for q in lp.neighbors:
        s_off.append(s_off[-1] + lp.send_blocks[q].size)
        r_off.append(r_off[-1] + lp.recv_blocks[q].size)
    s = np.concatenate([lp.send_blocks[q] for q in lp.neighbors]) if lp.neighbors else np.zeros(0, np.int32)
    r = np.concatenate([lp.recv_blocks[q] for q in lp.neighbors]) if lp.neighbors else np.zeros(0, np.int32)
    return (nbr, np.asarray(s_off, dtype=np.int32), np.ascontiguousarray(s, dtype=np.int32),
            np.asarray(r_off, dtype=np.int32), np.ascontiguousarray(r, dtype=np.int32))


# ---- hexahedral solids (csrc/hex.cu, SURVEY 8 f4 on N GPUs): block nodes = mesh nodes ------------------------------------
def build_local_part_hex(nodes: np.ndarray, hexes: np.ndarray, part: np.ndarray, rank: int, world: int) -> LocalPart:
    """Local mesh of ``rank`` for a hexahedral solid, halo lists included (the library numbers the block nodes of a solid
    context like the mesh nodes, so no attach_blocks step is needed).  Same rules as for the shell: a node is owned by
    the lowest part among its cells; a rank keeps every cell that holds a node it owns (owned cells first)."""
    nodes = np.asarray(nodes, dtype=np.float64)
    hexes = np.asarray(hexes).astype(np.int64)
    part = np.asarray(part, dtype=np.int32)
    nv, npc = nodes.shape[0], hexes.shape[1]
    v_owner = np.full(nv, world, dtype=np.int32)
    for r in range(world - 1, -1, -1):                              # lowest part written last = minimum over the cells
        v_owner[hexes[part == r].ravel()] = r
    own = v_owner[hexes]                                            # (nc, 8) owner of every cell node
    needed = (own == rank).any(axis=1)
    owned = part == rank
    assert not (owned & ~needed).any()
    loc_cells = np.concatenate([np.nonzero(owned)[0], np.nonzero(needed & ~owned)[0]])
    vglob = np.unique(hexes[loc_cells].ravel())
    remap = np.full(nv, -1, dtype=np.int64)
    remap[vglob] = np.arange(vglob.size)
    lp = LocalPart(rank=rank, world=world, nodes=np.ascontiguousarray(nodes[vglob]),
                   quads=np.ascontiguousarray(remap[hexes[loc_cells]].astype(np.int32)),
                   vertex_global=vglob, cell_global=loc_cells, n_owned_cells=int(owned.sum()),
                   own9=own[loc_cells], gkey9=hexes[loc_cells],
                   n_global_vertices=nv, n_global_cells=hexes.shape[0], n_global_edges=0)
    lp.block_owner = v_owner[vglob].astype(np.int32)
    lp.block_gkey = vglob.astype(np.int64)
    # receive side: my ghost nodes grouped by owner; vglob is ascending, so local order = key order
    recv = {}
    ghosts = np.nonzero(lp.block_owner != rank)[0]
    for q in np.unique(lp.block_owner[ghosts]):
        recv[int(q)] = ghosts[lp.block_owner[ghosts] == q].astype(np.int32)
    # send side: nodes I own inside cells that rank q keeps (q owns another node of the same cell)
    mine = own == rank
    cells_with_mine = mine.any(axis=1)
    send = {}
    for q in np.unique(own[cells_with_mine]):
        if q == rank:
            continue
        cells_q = cells_with_mine & (own == q).any(axis=1)
        keys = np.unique(hexes[cells_q][mine[cells_q]])
        send[int(q)] = remap[keys].astype(np.int32)
    lp.send_blocks, lp.recv_blocks = send, recv
    lp.neighbors = sorted(set(send) | set(recv))
    for q in lp.neighbors:
        send.setdefault(q, np.zeros(0, dtype=np.int32))
        recv.setdefault(q, np.zeros(0, dtype=np.int32))
    return lp
