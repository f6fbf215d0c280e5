"""Graph partitioner of the multi-GPU path (femo_alpha_b200/partition.py: METIS k-way on the cell-adjacency graph through
csrc/metis_shim.c) -- CPU only.  SURVEY 8e; the reference is serial and has no counterpart."""
import os

import numpy as np
import pytest

from femo_alpha_b200 import build
from femo_alpha_b200.partition import (attach_blocks, build_local_part, cell_adjacency, partition_cells,
                                       partition_cells_graph, partition_cells_rcb)

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def metis():
    if build.build_metis() is None:
        pytest.skip("libmetis_static.a not present on this machine")


def _wingbox():
    m = np.load(os.path.join(GOLD, "wingbox_mesh.npz"))
    return m["nodes"], m["quads"]


def test_cell_adjacency_handles_non_manifold_edges():
    # three quads around one common edge (a T-junction of a rib and a skin) + one quad hanging on the first
    nodes = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [1, 0, 1], [1, 1, 1], [2, 0, 0], [2, 1, 0], [-1, 0, 0], [-1, 1, 0]], float)
    quads = np.array([[0, 1, 2, 3], [1, 4, 5, 2], [1, 6, 7, 2], [8, 0, 3, 9]])
    xadj, adj = cell_adjacency(quads)
    nb = [sorted(adj[xadj[i]:xadj[i + 1]].tolist()) for i in range(4)]
    assert nb == [[1, 2, 3], [0, 2], [0, 1], [0]]
    assert nodes.shape[0] == 10


def test_metis_partition_of_the_wing_box(metis):
    nodes, quads = _wingbox()
    xadj, adj = cell_adjacency(quads)
    rows = np.repeat(np.arange(quads.shape[0]), np.diff(xadj))
    for world in (2, 3, 8):
        part = partition_cells(nodes, quads, world)                      # curved + non-manifold -> graph partitioning
        assert np.array_equal(part, partition_cells_graph(quads, world, nodes[quads].mean(axis=1)))
        assert np.array_equal(part, partition_cells(nodes, quads, world))   # fixed seed: every rank computes the same
        sizes = np.bincount(part, minlength=world)
        assert sizes.min() > 0 and sizes.max() <= 1.05 * quads.shape[0] / world + 1
        cut = int((part[rows] != part[adj]).sum() // 2)
        rcb = partition_cells_rcb(nodes[quads].mean(axis=1), world)
        assert cut <= int((rcb[rows] != rcb[adj]).sum() // 2)             # never worse than coordinate bisection here


def test_flat_plates_keep_coordinate_bisection():
    from oracle.rm_shell_ref import plate_mesh
    nodes, quads = plate_mesh(12, 6, 4.0, 2.0)
    assert np.array_equal(partition_cells(nodes, quads, 3), partition_cells_rcb(nodes[quads].mean(axis=1), 3))


def test_local_parts_of_a_graph_partition_are_consistent(metis):
    """ownership covers every block node exactly once and the halo lists of neighbouring ranks mirror each other"""
    from oracle.rm_shell_ref import ShellTopology
    nodes, quads = _wingbox()
    world = 4
    part = partition_cells(nodes, quads, world)
    lps = []
    for r in range(world):
        lp = build_local_part(nodes, quads, part, r, world)
        attach_blocks(lp, ShellTopology(lp.nodes, lp.quads).cell_blocks)
        lps.append(lp)
    owned = np.concatenate([lp.block_gkey[lp.own_block_mask] for lp in lps])
    assert np.unique(owned).size == owned.size                            # nobody owns a node twice
    t = ShellTopology(nodes, quads)
    assert owned.size == t.nb                                             # ... and every node has an owner
    for lp in lps:
        for q in lp.neighbors:
            other = lps[q]
            assert lp.rank in other.neighbors
            assert np.array_equal(lp.block_gkey[lp.send_blocks[q]], other.block_gkey[other.recv_blocks[lp.rank]])
            assert lp.own_block_mask[lp.send_blocks[q]].all() and not lp.own_block_mask[lp.recv_blocks[q]].any()
