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
