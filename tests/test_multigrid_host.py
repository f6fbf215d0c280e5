"""Host side of the multigrid hierarchy (femo_alpha_b200/multigrid.py): subdivision and the nested Q2xQ1 prolongation,
checked on the CPU with the oracle's dofmap (the device numbers block nodes the same way, tests/test_gpu_parity.py)."""
import numpy as np

from femo_alpha_b200.multigrid import prolongation, refine_quads
from oracle.rm_shell_ref import ShellTopology, plate_mesh


def _levels(nodes, quads):
    nf, qf = refine_quads(nodes, quads)
    return ShellTopology(nodes, quads), ShellTopology(nf, qf), nf, qf


def test_refine_counts_and_geometry():
    nodes, quads = plate_mesh(3, 2, 3.0, 2.0)
    tc, tf, nf, qf = _levels(nodes, quads)
    assert qf.shape == (4 * 6, 4) and nf.shape[0] == tc.nn + tc.ne + tc.nc
    assert np.allclose(nf[:tc.nn], nodes)                       # parent vertices keep their numbers
    # children tile the parent: areas add up, orientation (z normal) preserved
    def area(n, q):
        a, b, c, d = n[q[:, 0]], n[q[:, 1]], n[q[:, 2]], n[q[:, 3]]
        return 0.5 * np.cross(c - a, d - b)[:, 2]
    assert np.all(area(nf, qf) > 0) and np.isclose(area(nf, qf).sum(), area(nodes, quads).sum())
    assert np.allclose(area(nf, qf).reshape(6, 4).sum(1), area(nodes, quads))


def test_prolongation_reproduces_polynomials_on_affine_mesh():
    nodes, quads = plate_mesh(4, 3, 2.0, 1.5)
    S = np.array([[1.0, 0.3, 0.0], [0.1, 0.9, 0.0], [0.2, -0.1, 1.0]])   # affine map: parallelograms out of plane
    nodes = nodes @ S.T
    tc, tf, nf, qf = _levels(nodes, quads)
    P = prolongation(tf.cell_blocks, tc.cell_blocks, tf.nb, tc.nb)
    xc, xf = tc.block_coords(), tf.block_coords()
    quad = lambda x: 1.0 + 2.0 * x[:, 0] - x[:, 1] + 0.5 * x[:, 0] * x[:, 1] + 0.3 * x[:, 0] ** 2 - 0.7 * x[:, 1] ** 2
    lin = lambda x: 0.5 - x[:, 0] + 2.0 * x[:, 1] + 0.4 * x[:, 2]
    f_c = np.where(np.arange(tc.nb) < tc.nq2, quad(xc), lin(xc))          # Q2 block nodes: quadratic, Q1: linear
    f_f = np.where(np.arange(tf.nb) < tf.nq2, quad(xf), lin(xf))
    assert np.abs(P @ f_c - f_f).max() <= 1e-12
    # Q2 rows only touch coarse Q2 nodes, Q1 rows only coarse Q1 nodes
    Pc = P.tocoo()
    assert np.all((Pc.row < tf.nq2) == (Pc.col < tc.nq2))


def test_prolongation_partition_of_unity_on_skew_and_junction_meshes():
    rng = np.random.default_rng(3)
    nodes, quads = plate_mesh(5, 4, 2.0, 1.5)
    nodes[:, :2] += 0.05 * rng.standard_normal((nodes.shape[0], 2))
    # T-junction: a wall attached along y = 0 (three cells share each edge of that line)
    n2, q2 = plate_mesh(5, 2, 2.0, 1.0)
    base = plate_mesh(5, 4, 2.0, 1.5)[0]
    wall = np.stack([n2[:, 0], np.zeros(len(n2)), n2[:, 1]], 1)
    on_line = np.isclose(wall[:, 2], 0.0)
    remap = np.full(len(wall), -1)
    for i in np.nonzero(on_line)[0]:
        remap[i] = int(np.nonzero(np.isclose(base[:, 0], wall[i, 0]) & np.isclose(base[:, 1], 0.0))[0][0])
    new = np.nonzero(~on_line)[0]
    remap[new] = len(nodes) + np.arange(new.size)
    nodes = np.concatenate([nodes, wall[new]])
    quads = np.concatenate([quads, remap[q2]])
    tc, tf, nf, qf = _levels(nodes, quads)
    P = prolongation(tf.cell_blocks, tc.cell_blocks, tf.nb, tc.nb)
    assert np.abs(np.asarray(P.sum(axis=1)).ravel() - 1.0).max() <= 1e-13
    # coarse vertices are fine vertices: injection rows
    sub = P[:tc.nn].tocsr()
    assert sub.nnz == tc.nn and np.allclose(sub.diagonal(), 1.0)
    # nested geometry: coordinates are reproduced at the fine Q1 / vertex nodes (bilinear map)
    xc, xf = tc.block_coords(), tf.block_coords()
    assert np.abs((P @ xc)[tf.nq2:] - xf[tf.nq2:]).max() <= 1e-13


def test_nested_partition_and_local_prolongation_three_levels():
    """dist_multigrid.py (host side of the N-rank V-cycle): on every rank and between every pair of consecutive levels the
    rank-local prolongation reproduces the rows of the global one for the fine block nodes the rank owns."""
    import scipy.sparse as sp
    from femo_alpha_b200.dist_multigrid import local_prolongation, nested_parts
    from femo_alpha_b200.partition import attach_blocks, build_local_part, partition_cells_rcb
    rng = np.random.default_rng(5)
    nodes, quads = plate_mesh(5, 3, 2.0, 1.5)
    nodes[:, :2] += 0.04 * rng.standard_normal((nodes.shape[0], 2))
    meshes = [(nodes, quads)]
    for _ in range(2):
        meshes.append(refine_quads(*meshes[-1]))
    meshes = meshes[::-1]                                      # finest first
    world = 3
    parts = nested_parts(partition_cells_rcb(meshes[-1][0][meshes[-1][1]].mean(axis=1), world), 3)
    assert [p.size for p in parts] == [m[1].shape[0] for m in meshes]
    gtopos = [ShellTopology(*m) for m in meshes]

    def gkeys(t):
        k = np.empty(t.nb, dtype=np.int64); k[:t.nq2] = np.arange(t.nq2); k[t.nq2:] = t.nn + t.ne + t.nc + np.arange(t.nn)
        return k

    for lf in range(2):
        tf, tc = gtopos[lf], gtopos[lf + 1]
        Pg = prolongation(tf.cell_blocks, tc.cell_blocks, tf.nb, tc.nb).tocsr()
        kf, kc = gkeys(tf), gkeys(tc)
        of, oc = np.argsort(kf), np.argsort(kc)
        for rank in range(world):
            lps, topos = [], []
            for lvl in (lf, lf + 1):
                lp = build_local_part(meshes[lvl][0], meshes[lvl][1], parts[lvl], rank, world)
                t = ShellTopology(lp.nodes, lp.quads)
                attach_blocks(lp, t.cell_blocks)
                lps.append(lp); topos.append(t)
            Pl = local_prolongation(lps[0], lps[1], topos[0].cell_blocks, topos[1].cell_blocks)
            f2g = of[np.searchsorted(kf[of], lps[0].block_gkey)]
            c2g = oc[np.searchsorted(kc[oc], lps[1].block_gkey)]
            own = np.nonzero(lps[0].own_block_mask)[0]
            # scatter the local rows to global columns and compare with the global rows
            Pl = Pl.tocoo()
            G = sp.csr_matrix((Pl.data, (Pl.row, c2g[Pl.col])), shape=(topos[0].nb, tc.nb))
            assert abs(G[own] - Pg[f2g[own]]).max() < 1e-15
            assert np.allclose(np.asarray(G[own].sum(axis=1)).ravel(), 1.0)       # partition of unity on the owned rows


def test_transfer_arrays_of_a_distributed_finest_level():
    """dist_multigrid.transfer_arrays / replicated_part: P maps the replicated level-1 vector to the local fine blocks, and
    the ownership-weighted R restricts so that the sum over the ranks is the global restriction."""
    import scipy.sparse as sp
    from femo_alpha_b200.dist_multigrid import local_prolongation, nested_parts, replicated_part, transfer_arrays
    from femo_alpha_b200.partition import attach_blocks, build_local_part, partition_cells_rcb
    nodes_c, quads_c = plate_mesh(6, 3, 3.0, 1.5)
    nodes_f, quads_f = refine_quads(nodes_c, quads_c)
    tf, tc = ShellTopology(nodes_f, quads_f), ShellTopology(nodes_c, quads_c)
    Pg = prolongation(tf.cell_blocks, tc.cell_blocks, tf.nb, tc.nb).tocsr()
    rng = np.random.default_rng(0)
    r_f, e_c = rng.standard_normal((tf.nb, 3)), rng.standard_normal((tc.nb, 3))
    kf = np.empty(tf.nb, dtype=np.int64); kf[:tf.nq2] = np.arange(tf.nq2); kf[tf.nq2:] = tf.nn + tf.ne + tf.nc + np.arange(tf.nn)
    of = np.argsort(kf)
    world = 3
    parts = nested_parts(partition_cells_rcb(nodes_c[quads_c].mean(axis=1), world), 2)
    lp_all = replicated_part(nodes_c, quads_c)
    assert np.array_equal(lp_all.cell_global, np.arange(quads_c.shape[0])) and np.array_equal(lp_all.vertex_global, np.arange(len(nodes_c)))
    total = np.zeros((tc.nb, 3))
    for rank in range(world):
        lp = build_local_part(nodes_f, quads_f, parts[0], rank, world)
        t = ShellTopology(lp.nodes, lp.quads)
        attach_blocks(lp, t.cell_blocks)
        P = local_prolongation(lp, lp_all, t.cell_blocks, tc.cell_blocks)
        Pp, Pc, Pw, Rp, Rc, Rw = transfer_arrays(P, lp.own_block_mask.astype(np.float64))
        assert Pp[-1] == Rp[-1] and Pp.dtype == np.int32 and Rw.dtype == np.float64
        f2g = of[np.searchsorted(kf[of], lp.block_gkey)]
        own = lp.own_block_mask
        # prolongation of the replicated coarse vector on the owned rows
        Pm = sp.csr_matrix((Pw, Pc, Pp), shape=(t.nb, tc.nb))
        assert np.abs((Pm @ e_c)[own] - (Pg @ e_c)[f2g[own]]).max() < 1e-14
        # this rank's share of the restriction of a (ghost-consistent) fine vector
        Rm = sp.csr_matrix((Rw, Rc, Rp), shape=(tc.nb, t.nb))
        total += Rm @ r_f[f2g]
    assert np.abs(total - Pg.T @ r_f).max() < 1e-13
