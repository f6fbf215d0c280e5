"""Worker for the world_size-2/3 gloo tests: the multi-rank host logic (femo_alpha_b200.partition) drives a
NumPy distributed PCG whose local operators come from the CPU oracle; the result is compared with the
oracle's global direct solve.  Mirrors what csrc/dist.cu does on the GPUs (halo exchange before every
operator application, owner-masked dot products, all-reduce of the scalars)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def halo_exchange(dist, torch, lp, x):
    """ghost entries of x <- owners' values (blocks of 3 dofs), neighbour by neighbour."""
    reqs, bufs = [], []
    for q in lp.neighbors:
        sb = lp.send_blocks[q]
        if sb.size:
            t = torch.from_numpy(np.ascontiguousarray(x.reshape(-1, 3)[sb]))
            reqs.append(dist.isend(t, q))
        rb = lp.recv_blocks[q]
        if rb.size:
            r = torch.empty((rb.size, 3), dtype=torch.float64)
            reqs.append(dist.irecv(r, q))
            bufs.append((rb, r))
    for r in reqs:
        r.wait()
    for rb, r in bufs:
        x.reshape(-1, 3)[rb] = r.numpy()


def main(rank, world, port, name, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from femo_alpha_b200.partition import attach_blocks, build_local_part, partition_cells_rcb
    from oracle.rm_shell_ref import RMShellOracle, ShellTopology
    from tests.problems import make_problem
    p = make_problem(name)
    nodes, quads = p["nodes"], p["quads"]
    cent = nodes[quads].mean(axis=1)
    part = partition_cells_rcb(cent, world)
    lp = build_local_part(nodes, quads, part, rank, world)
    topo = ShellTopology(lp.nodes, lp.quads)
    attach_blocks(lp, topo.cell_blocks)

    def nodal(a, k=1):
        a = np.asarray(a)
        return a.reshape(-1, k)[lp.vertex_global].ravel() if k > 1 else a[lp.vertex_global]
    cellw = lambda a, k=1: np.asarray(a).reshape(-1, k)[lp.cell_global].ravel() if k > 1 else np.asarray(a)[lp.cell_global]
    mat = cellw if p["ewm"] else nodal
    o = RMShellOracle(lp.nodes, lp.quads, p["ewm"], p["ewp"])
    o.set_inputs(thickness=mat(p["thickness"]), E=mat(p["E"]), nu=mat(p["nu"]), density=mat(p["density"]),
                 F_solid=(cellw if p["ewp"] else nodal)(p["F_solid"], 3),
                 uhat=None if p["uhat"] is None else nodal(p["uhat"], 3))
    o.set_strong_bc(p["bc"])
    A = o.apply_bc_matrix(o.assemble_dRdu())
    b = o.assemble_load(); b[o.bc_dofs] = 0.0
    own = np.repeat(lp.own_block_mask, 3).astype(np.float64)
    # block-Jacobi on the owned diagonal (point Jacobi is enough for the test)
    dinv = 1.0 / A.diagonal()

    def gdot(u, v):
        t = torch.tensor([float(np.dot(own * u, v))], dtype=torch.float64)
        dist.all_reduce(t)
        return float(t.item())

    x = np.zeros(topo.ndof)
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
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), x=x, gkey=lp.block_gkey, owner=lp.block_owner, iters=it,
             n_owned_cells=lp.n_owned_cells, cell_global=lp.cell_global,
             send=np.concatenate([lp.send_blocks[q] for q in lp.neighbors] or [np.zeros(0, np.int32)]).size,
             recv=np.concatenate([lp.recv_blocks[q] for q in lp.neighbors] or [np.zeros(0, np.int32)]).size)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5])
