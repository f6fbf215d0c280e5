"""Worker for the gloo tests of the multi-rank hexahedral-solid host logic (femo_alpha_b200.partition.build_local_part_hex):
a NumPy distributed PCG whose local operators come from oracle/hex_ref.py on the rank's local mesh (owned cells + one
ghost layer), compared by the test with the oracle's global direct solve.  Mirrors csrc/dist.cu (halo exchange before
every operator application, owner-masked dot products, all-reduce of the scalars) and the owner-masked output sums of
femo_hex_output."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tests.dist_worker import halo_exchange            # noqa: E402


def problem():
    from oracle.hex_ref import box_mesh
    nodes, cells = box_mesh(6, 4, 3, 3.0, 2.0, 1.5, distort=0.2, seed=2)
    rng = np.random.default_rng(9)
    return nodes, cells, rng.uniform(0.05, 1.0, len(cells)), rng.standard_normal((len(nodes), 3))


def main(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from femo_alpha_b200.partition import build_local_part_hex, partition_cells_rcb
    from oracle.hex_ref import HexOracle
    nodes, cells, rho, f = problem()
    part = partition_cells_rcb(nodes[cells].mean(axis=1), world)
    lp = build_local_part_hex(nodes, cells, part, rank, world)
    o = HexOracle(lp.nodes, lp.quads, E0=4.0, nu=0.3)
    o.set_inputs(density=rho[lp.cell_global], body_force=f[lp.vertex_global])
    base = np.nonzero(lp.nodes[:, 0] < 1e-12)[0]
    o.set_strong_bc((3 * base[:, None] + np.arange(3)).ravel())
    A = o.with_bc(o.stiffness())
    b = o.load(); b[o.bc_dofs] = 0.0
    own = np.repeat(lp.own_block_mask, 3).astype(np.float64)
    dinv = 1.0 / A.diagonal()

    def gdot(u, v):
        t = torch.tensor([float(np.dot(own * u, v))], dtype=torch.float64)
        dist.all_reduce(t)
        return float(t.item())

    x = np.zeros(o.ndof)
    r = own * b
    z = dinv * r
    pvec = z.copy()
    rz = gdot(r, z)
    bb = gdot(b, b)
    for it in range(20000):
        halo_exchange(dist, torch, lp, pvec)
        Ap = own * (A @ pvec)
        alpha = rz / gdot(pvec, Ap)
        x += alpha * pvec
        r -= alpha * Ap
        z = dinv * r
        rz_new = gdot(r, z)
        if np.sqrt(gdot(r, r) / bb) < 1e-13:
            break
        pvec = z + (rz_new / rz) * pvec
        rz = rz_new
    halo_exchange(dist, torch, lp, x)
    compliance = gdot(o.load(), x)                                   # owner-masked, summed over ranks
    n_own = lp.n_owned_cells
    t = torch.tensor([float(o.rho[:n_own] @ o.vol[:n_own])], dtype=torch.float64)
    dist.all_reduce(t)
    # density gradient of the compliance on the owned cells (self-adjoint: lam = u)
    g_own = -o.dRdrho_T(x, x * (1 - np.isin(np.arange(o.ndof), o.bc_dofs)))[:n_own]
    np.savez(os.path.join(out_dir, "hex_rank%d.npz" % rank), x=x, gkey=lp.block_gkey, owner=lp.block_owner, iters=it,
             n_owned_cells=n_own, cell_global=lp.cell_global, compliance=compliance, volume=float(t.item()), g_own=g_own,
             send=sum(lp.send_blocks[q].size for q in lp.neighbors), recv=sum(lp.recv_blocks[q].size for q in lp.neighbors))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4])
