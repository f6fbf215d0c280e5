"""Thin Python handle on one hexahedral-solid context of libfemo_b200 (generic 3-D forms, csrc/hex.cu; SURVEY 8 f4 /
BASELINE config 5).  Same method surface as ShellDevice, so the form handles of fea/utils_dolfinx.py (ResidualForm,
OutputForm, JacobianMat, PartialMat, KSP) and the CSDL operations of csdl_alpha_opt/ work on it unchanged.

Inputs:  "density" (one value per cell, SIMP factor rho^p) and "body_force" ((n_nodes, 3) nodal field, flattened).
State:   3 dofs per node, dof = 3 * node + component (own numbering = mesh numbering).
Outputs: "compliance" = int f . u dx, "volume" = sum rho_e |e|.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L
from .shell_device import ShellDevice

FORMS = {"vector_poisson": 0, "poisson": 0, "elasticity": 1}
INPUTS = {"density": 0, "body_force": 1}
OUTPUTS = {"compliance": 0, "volume": 1}


class SolidDevice(ShellDevice):
    def __init__(self, nodes, hexes, form="elasticity", E0=1.0, nu=0.3, simp_p=3.0, device=0):
        self.lib = L.load()
        self._ctx = C.c_void_p()
        self._children, self._owner, self._mg = [], None, None
        if form not in FORMS:
            raise ValueError("unknown form %r (registered: %s)" % (form, ", ".join(sorted(FORMS))))
        xyz = np.ascontiguousarray(nodes, dtype=np.float64)
        conn = np.ascontiguousarray(hexes, dtype=np.int32)
        if xyz.ndim != 2 or xyz.shape[1] != 3:
            raise ValueError("nodes must have shape (n_nodes, 3)")
        if conn.ndim != 2 or conn.shape[1] != 8:
            raise ValueError("connectivity must have shape (n_cells, 8) (VTK_HEXAHEDRON order)")
        L.check(self.lib.femo_ctx_create(int(device), C.byref(self._ctx)))
        L.check(self.lib.femo_hex_setup(self._ctx, xyz.shape[0], L.ptr(xyz), conn.shape[0], L.ptr(conn), FORMS[form], float(E0),
                                        float(nu), float(simp_p)))
        s = L.Sizes()
        L.check(self.lib.femo_get_sizes(self._ctx, C.byref(s)))
        self.sizes = s
        self.form = form
        self.nn, self.ne, self.nc = xyz.shape[0], 0, s.n_cells
        self.nb, self.ndof, self.nnzb = s.n_blocks, s.n_dof, s.nnz_blocks
        self.nq2 = self.nb
        self.n_mat, self.n_f = s.n_material, s.n_force
        self.last_solve = dict(iters=0, relres=0.0)

    # ---- helpers ---------------------------------------------------------------------------------------------------
    def _arg_size(self, arg):
        return {0: self.nc, 1: self.ndof, L.ARG_STATE: self.ndof}[arg]

    @staticmethod
    def _arg_id(arg):
        if isinstance(arg, str):
            if arg in ("disp_solid", "state", "u"):
                return L.ARG_STATE
            return INPUTS[arg]
        return int(arg)

    def block_coords(self):
        out = np.empty((self.nb, 3))
        L.check(self.lib.femo_hex_get_mesh(self._ctx, L.ptr(out), None))
        return out

    def cell_blocks(self):
        out = np.empty((self.nc, 8), dtype=np.int32)
        L.check(self.lib.femo_hex_get_mesh(self._ctx, None, L.ptr(out)))
        return out

    def set_penalty_bc(self, vertex_on):
        raise NotImplementedError("solid forms take strong Dirichlet BCs (set_strong_bc)")

    # ---- inputs ---------------------------------------------------------------------------------------------------------
    def set_input(self, name, values):
        which = self._arg_id(name)
        if isinstance(values, np.ndarray) or not hasattr(values, "data_ptr"):
            values = np.ascontiguousarray(np.asarray(values, dtype=np.float64).ravel())
            n = values.size
        else:
            n = values.numel()
        L.check(self.lib.femo_hex_set_input(self._ctx, which, L.ptr(values), n, L.on_device(values)))

    def get_input(self, name):
        which = self._arg_id(name)
        out = np.empty(self._arg_size(which))
        L.check(self.lib.femo_hex_get_input(self._ctx, which, L.ptr(out), out.size, 0))
        return out

    # ---- partials ----------------------------------------------------------------------------------------------------------
    def dRdf_T_apply(self, arg, lam):
        a = self._arg_id(arg)
        lam = np.ascontiguousarray(lam, dtype=np.float64)
        out = np.empty(self._arg_size(a))
        L.check(self.lib.femo_hex_dRdf_T_apply(self._ctx, a, L.ptr(lam), L.ptr(out), 0))
        return out

    def dRdf_apply(self, arg, df):
        a = self._arg_id(arg)
        df = np.ascontiguousarray(df, dtype=np.float64)
        if df.size != self._arg_size(a):
            raise ValueError("dRdf_apply: %d values expected, got %d" % (self._arg_size(a), df.size))
        out = np.empty(self.ndof)
        L.check(self.lib.femo_hex_dRdf_apply(self._ctx, a, L.ptr(df), L.ptr(out), 0))
        return out

    # ---- outputs --------------------------------------------------------------------------------------------------------------
    def output(self, name):
        v = C.c_double(0.0)
        L.check(self.lib.femo_hex_output(self._ctx, OUTPUTS[name] if isinstance(name, str) else int(name), C.byref(v)))
        return v.value

    def output_grad(self, name, arg):
        a = self._arg_id(arg)
        out = np.empty(self._arg_size(a))
        L.check(self.lib.femo_hex_output_grad(self._ctx, OUTPUTS[name] if isinstance(name, str) else int(name), a, L.ptr(out), 0))
        return out

    # ---- bench helpers ------------------------------------------------------------------------------------------------------------
    def last_timings(self):
        ms = np.zeros(4)
        L.check(self.lib.femo_last_timings(self._ctx, ms.ctypes.data_as(C.POINTER(C.c_double))))
        return dict(assemble_ms=ms[1], solve_ms=ms[2], iter_ms=ms[3])

    def pcg_fixed(self, n_iter):
        t = C.c_double(0.0)
        L.check(self.lib.femo_pcg_fixed(self._ctx, int(n_iter), C.byref(t)))
        return t.value

    def launch_count(self):
        n = C.c_int64(0)
        L.check(self.lib.femo_launch_count(self._ctx, C.byref(n)))
        return n.value


def _shell_only(name):
    def f(self, *a, **k):
        raise NotImplementedError("%s is an RM-shell functional; not defined for solid forms" % name)
    f.__name__ = name
    return f


for _n in ("set_stress_params", "set_cell_tags", "pnorm_stress", "pnorm_stress_grad", "stress_field", "_cell_vertices",
           "_q1_cell_mass"):
    setattr(SolidDevice, _n, _shell_only(_n))


def box_hexes(nx, ny, nz, lx=1.0, ly=1.0, lz=1.0):
    """Structured box mesh of hexahedra (VTK_HEXAHEDRON order): (nodes (n, 3), cells (nx ny nz, 8) int32)."""
    xs, ys, zs = np.linspace(0, lx, nx + 1), np.linspace(0, ly, ny + 1), np.linspace(0, lz, nz + 1)
    X, Y, Z = np.meshgrid(xs, ys, zs, indexing="ij")
    nodes = np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1)
    I, J, K = np.meshgrid(np.arange(nx, dtype=np.int64), np.arange(ny, dtype=np.int64), np.arange(nz, dtype=np.int64),
                          indexing="ij")
    I, J, K = I.ravel(), J.ravel(), K.ravel()
    nid = lambda i, j, k: (i * (ny + 1) + j) * (nz + 1) + k
    cells = np.stack([nid(I, J, K), nid(I + 1, J, K), nid(I + 1, J + 1, K), nid(I, J + 1, K),
                      nid(I, J, K + 1), nid(I + 1, J, K + 1), nid(I + 1, J + 1, K + 1), nid(I, J + 1, K + 1)], axis=1)
    return nodes, cells.astype(np.int32)
