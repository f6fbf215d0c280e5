"""Worker for the gloo test of the N-rank V-cycle (open item 2): a NumPy prototype of the distributed two-grid
preconditioner on top of femo_alpha_b200.partition, with the local operators of the CPU oracle.

What it establishes for the CUDA implementation of the next round:
  * nested partition: a fine cell goes to the part of its parent, so the stencil of every OWNED fine block node lies in the
    rank's local coarse mesh -- the local prolongation is a row/column selection of the global one, no communication;
  * restriction is P_loc^T (own_f * r_f) on the local coarse blocks followed by a REVERSE halo exchange (ghost -> owner,
    accumulate) over the same neighbour lists as the forward exchange -- the one new primitive the library needs;
  * Chebyshev-Jacobi smoothing and the owner-masked dot products are the distributed operations csrc/dist.cu already has;
  * the coarse level is replicated: one all-reduce of the coarse residual, every rank solves, no scatter needed.
The outer PCG must take the same number of iterations for every world size (the preconditioner is the same operator)."""
import os
import sys

import numpy as np
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tests.dist_worker import halo_exchange            # noqa: E402


def reverse_halo_add(dist, torch, lp, x):
    """owners += the values their ghosts hold on other ranks (blocks of 3 dofs); ghost entries are left as they are."""
    reqs, bufs = [], []
    for q in lp.neighbors:
        rb = lp.recv_blocks[q]
        if rb.size:
            reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(x.reshape(-1, 3)[rb])), q))
        sb = lp.send_blocks[q]
        if sb.size:
            r = torch.empty((sb.size, 3), dtype=torch.float64)
            reqs.append(dist.irecv(r, q))
            bufs.append((sb, r))
    for r in reqs:
        r.wait()
    for sb, r in bufs:                                   # fixed neighbour order: deterministic
        x.reshape(-1, 3)[sb] += r.numpy()


def problem():
    from femo_alpha_b200.multigrid import refine_quads
    from oracle.rm_shell_ref import plate_mesh
    nodes_c, quads_c = plate_mesh(8, 4, 10.0, 2.0)
    nodes_f, quads_f = refine_quads(nodes_c, quads_c)
    return nodes_c, quads_c, nodes_f, quads_f


def global_keys(topo):
    """the global key convention of femo_alpha_b200.partition on an un-partitioned ShellTopology"""
    key = np.empty(topo.nb, dtype=np.int64)
    key[:topo.nq2] = np.arange(topo.nq2)
    key[topo.nq2:] = topo.nn + topo.ne + topo.nc + np.arange(topo.nn)
    return key


class Level:
    def __init__(self, nodes, quads, part, rank, world):
        from femo_alpha_b200.partition import attach_blocks, build_local_part
        from oracle.rm_shell_ref import RMShellOracle, ShellTopology
        self.lp = build_local_part(nodes, quads, part, rank, world)
        self.topo = ShellTopology(self.lp.nodes, self.lp.quads)
        attach_blocks(self.lp, self.topo.cell_blocks)
        o = RMShellOracle(self.lp.nodes, self.lp.quads, False, False)
        nn = self.lp.nodes.shape[0]
        o.set_inputs(thickness=np.full(nn, 0.1), E=np.full(nn, 1e8), nu=np.full(nn, 0.3), density=np.full(nn, 10.0),
                     F_solid=np.tile([0.0, 0.0, 5.0], nn))
        o.set_strong_bc(lambda x: np.less(x[0], 3e-16))
        self.o = o
        self.A = o.apply_bc_matrix(o.assemble_dRdu()).tocsr()
        self.own = np.repeat(self.lp.own_block_mask, 3).astype(np.float64)
        self.free = np.ones(self.topo.ndof); self.free[o.bc_dofs] = 0.0
        self.dinv = 1.0 / self.A.diagonal()
        gtopo = ShellTopology(nodes, quads)
        gk = global_keys(gtopo)
        order = np.argsort(gk)
        self.to_global = order[np.searchsorted(gk[order], self.lp.block_gkey)]      # local block -> global block
        self.gtopo = gtopo


def main(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from femo_alpha_b200.multigrid import prolongation
    from femo_alpha_b200.partition import partition_cells_rcb
    nodes_c, quads_c, nodes_f, quads_f = problem()
    part_c = partition_cells_rcb(nodes_c[quads_c].mean(axis=1), world)
    part_f = np.repeat(part_c, 4)                                   # refine_quads: fine cell 4 c + k is a child of c
    Lc, Lf = Level(nodes_c, quads_c, part_c, rank, world), Level(nodes_f, quads_f, part_f, rank, world)
    # local prolongation = selection of the global one; complete on the owned fine rows
    Pg = prolongation(Lf.gtopo.cell_blocks, Lc.gtopo.cell_blocks, Lf.gtopo.nb, Lc.gtopo.nb).tocsr()
    inv_c = np.full(Lc.gtopo.nb, -1, dtype=np.int64); inv_c[Lc.to_global] = np.arange(Lc.topo.nb)
    rows = Pg[Lf.to_global].tocoo()
    owned_row = Lf.lp.own_block_mask[rows.row]
    assert (inv_c[rows.col[owned_row]] >= 0).all(), "stencil of an owned fine node left the local coarse mesh"
    keep = inv_c[rows.col] >= 0
    Pb = sp.csr_matrix((rows.data[keep], (rows.row[keep], inv_c[rows.col[keep]])), shape=(Lf.topo.nb, Lc.topo.nb))
    # the product's rank-local builder (femo_alpha_b200/dist_multigrid.py) gives the same rows without the global matrix
    from femo_alpha_b200.dist_multigrid import local_prolongation, nested_parts
    assert np.array_equal(nested_parts(part_c, 2)[0], part_f) and np.array_equal(nested_parts(part_c, 2)[1], part_c)
    Pl = local_prolongation(Lf.lp, Lc.lp, Lf.topo.cell_blocks, Lc.topo.cell_blocks)
    own_rows = np.nonzero(Lf.lp.own_block_mask)[0]
    assert abs(Pl[own_rows] - Pb[own_rows]).max() < 1e-15
    filled = np.nonzero(np.diff(Pl.indptr) > 0)[0]
    assert abs(Pl[filled] - Pb[filled]).max() < 1e-15              # rows of ghost nodes, where present, agree as well
    P = sp.kron(Pl, sp.identity(3)).tocsr()

    def gdot(L, u, v):
        t = torch.tensor([float(np.dot(L.own * u, v))], dtype=torch.float64)
        dist.all_reduce(t)
        return float(t.item())

    def spmv(L, x):
        halo_exchange(dist, torch, L.lp, x)
        return L.own * (L.A @ x)

    def lam_max(L):
        # start vector defined through the global keys: the same global vector for every world size
        v = np.cos(1.0 + 3.0 * L.lp.block_gkey.repeat(3) + np.tile(np.arange(3), L.topo.nb)) * L.free
        lam = 1.0
        for _ in range(12):
            w = L.dinv * spmv(L, v)
            lam = np.sqrt(gdot(L, w, w) / gdot(L, v, v))
            v = w / lam
        return 1.1 * lam

    def chebyshev(L, b, x, lmax, degree=3):
        """x <- x + Chebyshev-Jacobi correction for A x = b on [lmax / 4, lmax]"""
        theta, delta = 0.625 * lmax, 0.375 * lmax
        sigma = theta / delta
        rho = 1.0 / sigma
        r = b - spmv(L, x)
        d = L.dinv * r / theta
        for _ in range(degree):
            x += d
            r = r - spmv(L, d)
            rho_new = 1.0 / (2.0 * sigma - rho)
            d = rho_new * rho * d + 2.0 * rho_new / delta * (L.dinv * r)
            rho = rho_new
        return x

    def pcg(L, b, precond, rtol, max_it):
        x = np.zeros(L.topo.ndof)
        r = L.own * b
        z = precond(r)
        p = z.copy()
        rz = gdot(L, r, z)
        bb = gdot(L, b, b)
        it = 0
        if bb == 0.0:
            return x, 0
        for it in range(1, max_it + 1):
            Ap = spmv(L, p)
            alpha = rz / gdot(L, p, Ap)
            x += alpha * p
            r -= alpha * Ap
            if np.sqrt(gdot(L, r, r) / bb) < rtol:
                break
            z = precond(r)
            rz_new = gdot(L, r, z)
            p = z + (rz_new / rz) * p
            rz = rz_new
        halo_exchange(dist, torch, L.lp, x)
        return x, it

    lmax_f = lam_max(Lf)
    # replicated coarse operator: the oracle on the global coarse mesh, in the partition's global block numbering
    import scipy.sparse.linalg as spla
    from oracle.rm_shell_ref import RMShellOracle
    oc = RMShellOracle(nodes_c, quads_c, False, False)
    nnc = nodes_c.shape[0]
    oc.set_inputs(thickness=np.full(nnc, 0.1), E=np.full(nnc, 1e8), nu=np.full(nnc, 0.3), density=np.full(nnc, 10.0),
                  F_solid=np.tile([0.0, 0.0, 5.0], nnc))
    oc.set_strong_bc(lambda x: np.less(x[0], 3e-16))
    coarse_lu = spla.splu(oc.apply_bc_matrix(oc.assemble_dRdu()).tocsc())
    coarse_dofs = (3 * Lc.to_global[:, None] + np.arange(3)[None, :]).ravel()

    def vcycle(r):
        z = chebyshev(Lf, r, np.zeros(Lf.topo.ndof), lmax_f)                       # pre-smoothing from zero
        res = Lf.own * (r - spmv(Lf, z))
        rc = P.T @ res                                                              # local coarse blocks, ghosts included
        reverse_halo_add(dist, torch, Lc.lp, rc)
        rc *= Lc.own * Lc.free
        # coarse solve on the REPLICATED coarse level (the plan of DESIGN 15: small levels live on every rank): owned
        # entries are summed into the global vector, every rank solves, local entries (ghosts included) are read back
        g = torch.zeros(3 * Lc.gtopo.nb, dtype=torch.float64)
        g[coarse_dofs] = torch.from_numpy(rc)
        dist.all_reduce(g)
        ec = coarse_lu.solve(g.numpy())[coarse_dofs]
        z = z + Lf.own * (P @ ec)
        z = chebyshev(Lf, r, z, lmax_f)                                             # post-smoothing
        return Lf.own * z

    b = Lf.o.assemble_load(); b[Lf.o.bc_dofs] = 0.0
    x, its = pcg(Lf, b, vcycle, 1e-11, 200)
    its_j = -1
    if world == 1:                                                                  # plain Jacobi PCG, for the iteration ratio
        _, its_j = pcg(Lf, b, lambda v: Lf.dinv * v, 1e-11, 20000)
    np.savez(os.path.join(out_dir, "mg_rank%d.npz" % rank), x=x, gkey=Lf.lp.block_gkey, its=its, its_jacobi=its_j, lmax=lmax_f)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4])
