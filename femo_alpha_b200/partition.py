"""Mesh partitioning for the multi-GPU RM-shell path: host-side NumPy only.

One process per GPU.  Cells are split into ``world`` parts by recursive coordinate bisection of
the cell centroids.  A Q2/Q1 block node (vertex, edge mid-point, cell centre, rotation vertex) is
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
