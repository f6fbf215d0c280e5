"""One rank of a multi-GPU RM-shell problem: partition on the host, everything else in the library.

Usage (one process per GPU, after torch.distributed.init_process_group):

    ds = DistributedShell(nodes, quads, rank, world, device=local_rank)
    ds.init_comm()                       # NCCL id from rank 0, broadcast with torch.distributed
    ds.set_input("thickness", h_global)  # global (external-order) arrays; each rank keeps its slice
    ds.set_strong_bc(lambda x: x[0] < 3e-16)
    iters, relres = ds.solve_state()
    C = ds.output("compliance")          # identical on every rank

Assembly is owner-computes (no collective).  Solves exchange the halo of the search direction and all-reduce two
scalars pairs per iteration inside libfemo_b200 (NCCL on the context's stream)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L
from .partition import attach_blocks, build_local_part, halo_arrays, partition_cells
from .shell_device import ShellDevice


class DistributedShell:
    def __init__(self, nodes, quads, rank, world, part=None, device=0, element_wise_material=False,
                 elementwise_pressure=False, quadrature_degree=6):
        nodes = np.asarray(nodes, dtype=np.float64)
        quads = np.asarray(quads)
        self.rank, self.world = int(rank), int(world)
        if part is None:            # METIS k-way on the dual graph for curved / non-manifold meshes, RCB for flat plates
            part = partition_cells(nodes, quads, world)
        self.part = np.asarray(part, dtype=np.int32)
        lp = build_local_part(nodes, quads, self.part, self.rank, self.world)
        self.dev = ShellDevice(lp.nodes, lp.quads, element_wise_material, elementwise_pressure, quadrature_degree, device)
        attach_blocks(lp, self.dev.cell_blocks())
        self.lp = lp
        self.ewm, self.ewp = bool(element_wise_material), bool(elementwise_pressure)
        nbr, s_off, s_blk, r_off, r_blk = halo_arrays(lp)
        own = np.ascontiguousarray(lp.own_block_mask.astype(np.uint8))
        self._keep = (nbr, s_off, s_blk, r_off, r_blk, own)
        L.check(self.dev.lib.femo_dist_setup(self.dev._ctx, self.rank, self.world, int(nbr.size), L.ptr(nbr), L.ptr(s_off),
                                             L.ptr(s_blk), L.ptr(r_off), L.ptr(r_blk), L.ptr(own), int(lp.n_owned_cells)))
        self.ndof_owned = 3 * int(own.sum())

    # ---- communicator ------------------------------------------------------------------------
    def init_comm(self, dist=None):
        """Collective.  The 128-byte NCCL id travels through torch.distributed (any backend)."""
        if self.world == 1:
            return
        import torch
        if dist is None:
            import torch.distributed as dist
        buf = np.zeros(128, dtype=np.uint8)
        if self.rank == 0:
            L.check(self.dev.lib.femo_nccl_unique_id(L.ptr(buf)))
        t = torch.from_numpy(buf)
        if dist.get_backend() == "nccl":
            t = t.cuda()
        dist.broadcast(t, src=0)
        buf = t.cpu().numpy().copy()
        L.check(self.dev.lib.femo_dist_init(self.dev._ctx, L.ptr(buf)))
        # NCCL establishes its peer-to-peer connections lazily, at the first send / receive and the first all-reduce of
        # a communicator (~1.5 s at 4 GPUs): pay that here, once, not inside the first solver set-up
        self.halo_exchange(np.zeros(self.dev.ndof))
        L.check(self.dev.lib.femo_dist_allreduce(self.dev._ctx, (C.c_double * 1)(0.0), 1))

    # ---- inputs: global arrays in external (mesh) order ------------------------------------------
    def _slice(self, name, values):
        v = np.asarray(values, dtype=np.float64)
        if v.size == 1:
            return v.ravel()
        if name in ("F_solid",):
            idx = self.lp.cell_global if self.ewp else self.lp.vertex_global
            return v.reshape(-1, 3)[idx].ravel()
        if name == "uhat":
            return v.reshape(-1, 3)[self.lp.vertex_global].ravel()
        idx = self.lp.cell_global if self.ewm else self.lp.vertex_global
        return v.ravel()[idx]

    def set_input(self, name, global_values):
        self.dev.set_input(name, self._slice(name, global_values))

    def set_strong_bc(self, predicate):
        on = np.asarray(predicate(self.dev.block_coords().T)).astype(bool)
        blocks = np.nonzero(on)[0]
        self.dev.set_strong_bc(np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()))

    # ---- pass-through (collective where the library communicates) -----------------------------------
    def assemble_jacobian(self):
        self.dev.assemble_jacobian()

    def solve_state(self, rtol=1e-12, max_it=200000):
        return self.dev.solve_state(rtol, max_it)

    def solve_adjoint(self, rhs_local, rtol=1e-12, max_it=200000):
        return self.dev.solve_adjoint(rhs_local, rtol, max_it)

    def output(self, name):
        return self.dev.output(name)

    # ---- time-stepped path (SURVEY 8 f1 on N GPUs) ----------------------------------------------------
    def dynamic(self, dt, n_steps, add_self_weight=False, g_factor=None, rtol=1e-12, max_it=20000):
        """DynamicShellDevice on this rank's context.  Every rank steps its own rows; the effective operator
        cM + K/2 is solved by the distributed block-Jacobi PCG (halo exchange + all-reduce per iteration), the
        recurrences in between need no communication (csrc/dynamic.cu, dyn_enable)."""
        from .dynamic_device import DynamicShellDevice
        return DynamicShellDevice(self.dev, dt, n_steps, add_self_weight, g_factor, "pcg", rtol, max_it)

    def slice_force_history(self, f_history_global):
        """(rows, 3*n_global) force history in mesh order -> this rank's (rows, n_f) columns."""
        f = np.atleast_2d(np.asarray(f_history_global, dtype=np.float64))
        idx = self.lp.cell_global if self.ewp else self.lp.vertex_global
        return np.ascontiguousarray(f.reshape(f.shape[0], -1, 3)[:, idx, :].reshape(f.shape[0], -1))

    def halo_exchange(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        L.check(self.dev.lib.femo_halo_exchange(self.dev._ctx, L.ptr(x), 0))
        return x

    def halo_accumulate(self, x):
        """owned entries += the values of their ghost copies on the other ranks (collective); returns x"""
        x = np.ascontiguousarray(x, dtype=np.float64)
        L.check(self.dev.lib.femo_halo_accumulate(self.dev._ctx, L.ptr(x), 0))
        return x

    def overlap_info(self):
        """(halo done, interior rows done, boundary rows done) in ms after the start of the last operator application;
        needs FEMO_TRACE_OVERLAP=1 in the environment when the object was built"""
        v = (C.c_double * 3)()
        L.check(self.dev.lib.femo_dist_overlap_info(self.dev._ctx, v))
        return float(v[0]), float(v[1]), float(v[2])

    def comm_counts(self):
        h = C.c_int64(0); a = C.c_int64(0)
        L.check(self.dev.lib.femo_dist_counts(self.dev._ctx, C.byref(h), C.byref(a)))
        return h.value, a.value

    # ---- owned pieces back in global numbering -----------------------------------------------------------
    def owned_vertex_values(self, local_state):
        """(global vertex ids, (n,3) displacements) of the vertices this rank owns: disp_extracted slice."""
        w = np.asarray(local_state).reshape(-1, 3)
        nn = self.dev.nn
        own_v = self.lp.own_block_mask[:nn]
        return self.lp.vertex_global[own_v], w[:nn][own_v]

    def close(self):
        self.dev.close()
