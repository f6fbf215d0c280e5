"""Thin Python handle on one libfemo_b200 context (one RM-shell problem resident on one B200).

All arithmetic happens in the CUDA library; this class only marshals NumPy
arrays (or torch CUDA tensors) across the C ABI.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L


class ShellDevice:
    def __init__(self, nodes, quads, element_wise_material=False, elementwise_pressure=False,
                 quadrature_degree=6, device=0):
        self.lib = L.load()
        self._ctx = C.c_void_p()
        L.check(self.lib.femo_ctx_create(int(device), C.byref(self._ctx)))
        xyz = np.ascontiguousarray(nodes, dtype=np.float64)
        conn = np.ascontiguousarray(quads, dtype=np.int32)
        if xyz.ndim != 2 or xyz.shape[1] != 3:
            raise ValueError("nodes must have shape (n_nodes, 3)")
        if conn.ndim != 2:
            raise ValueError("connectivity must have shape (n_cells, nodes_per_cell)")
        L.check(self.lib.femo_mesh_set(self._ctx, xyz.shape[0], L.ptr(xyz), conn.shape[0], conn.shape[1], L.ptr(conn)))
        opts = L.ShellOpts(int(bool(element_wise_material)), int(bool(elementwise_pressure)), int(quadrature_degree), 0)
        L.check(self.lib.femo_shell_setup(self._ctx, C.byref(opts)))
        s = L.Sizes()
        L.check(self.lib.femo_get_sizes(self._ctx, C.byref(s)))
        self.sizes = s
        self.nn, self.ne, self.nc = s.n_nodes, s.n_edges, s.n_cells
        self.nb, self.ndof, self.nnzb = s.n_blocks, s.n_dof, s.nnz_blocks
        self.nq2 = self.nn + self.ne + self.nc
        self.n_mat, self.n_f = s.n_material, s.n_force
        self.last_solve = dict(iters=0, relres=0.0)
        self._mg = None                 # ShellHierarchy this device is the finest level of (multigrid.py)
        self._direct = None             # dict(refine=..) when solves go through the sparse direct solver (csrc/coarse.cu)
        self._children = []             # weak references to handles that borrow this context (DynamicShellDevice)
        self._owner = None              # ShellHierarchy this device is a level of: it tears down in order (handles, mg, levels)

    def close(self):
        owner = getattr(self, "_owner", None)
        if owner is not None and not owner._closing:
            owner.close()               # closes every level, this one included
            return
        for ref in getattr(self, "_children", []):
            child = ref()
            if child is not None:
                child.close()
        self._children = []
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            self.lib.femo_ctx_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- helpers --------------------------------------------------------
    def _arg_size(self, arg):
        return {L.THICKNESS: self.n_mat, L.E: self.n_mat, L.NU: self.n_mat, L.DENSITY: self.n_mat,
                L.F_SOLID: self.n_f, L.UHAT: 3 * self.nn, L.ARG_STATE: self.ndof}[arg]

    @staticmethod
    def _arg_id(arg):
        if isinstance(arg, str):
            if arg in ("disp_solid", "state"):
                return L.ARG_STATE
            return L.INPUT_IDS[arg]
        return int(arg)

    def set_stream(self, stream_ptr):
        L.check(self.lib.femo_set_stream(self._ctx, C.c_void_p(stream_ptr)))

    # ---- mesh-derived data -------------------------------------------------
    def block_coords(self):
        out = np.empty((self.nb, 3))
        L.check(self.lib.femo_get_block_coords(self._ctx, L.ptr(out)))
        return out

    def cell_blocks(self):
        out = np.empty((self.nc, 13), dtype=np.int32)
        L.check(self.lib.femo_get_cell_blocks(self._ctx, L.ptr(out)))
        return out

    # ---- boundary conditions --------------------------------------------------
    def set_strong_bc(self, dofs):
        d = np.ascontiguousarray(dofs, dtype=np.int32)
        L.check(self.lib.femo_set_strong_bc(self._ctx, d.size, L.ptr(d)))
        if self._mg is not None:
            self._mg._coarse_bcs(d)

    def set_penalty_bc(self, vertex_on):
        v = np.ascontiguousarray(vertex_on, dtype=np.uint8)
        L.check(self.lib.femo_set_penalty_bc(self._ctx, v.size, L.ptr(v)))
        if self._mg is not None and v.any():
            import warnings
            warnings.warn("femo_alpha_b200: penalty BCs are not carried to the coarse levels; multigrid is switched off")
            self._mg = None

    # ---- inputs / state ----------------------------------------------------------
    def set_input(self, name, values):
        which = self._arg_id(name)
        if isinstance(values, np.ndarray) or not hasattr(values, "data_ptr"):
            values = np.ascontiguousarray(np.asarray(values, dtype=np.float64).ravel())
            n = values.size
        else:
            n = values.numel()
        L.check(self.lib.femo_set_input(self._ctx, which, L.ptr(values), n, L.on_device(values)))
        if self._mg is not None:
            self._mg._coarse_inputs(name, values)

    def get_input(self, name):
        which = self._arg_id(name)
        out = np.empty(self._arg_size(which))
        L.check(self.lib.femo_get_input(self._ctx, which, L.ptr(out), out.size, 0))
        return out

    def set_state(self, w):
        w = np.ascontiguousarray(w, dtype=np.float64) if isinstance(w, np.ndarray) else w
        L.check(self.lib.femo_set_state(self._ctx, L.ptr(w), L.on_device(w)))

    def get_state(self, out=None):
        if out is None:
            out = np.empty(self.ndof)
        L.check(self.lib.femo_get_state(self._ctx, L.ptr(out), L.on_device(out)))
        return out

    # ---- assembly --------------------------------------------------------------------
    def assemble_jacobian(self):
        L.check(self.lib.femo_assemble_jacobian(self._ctx))

    def assemble_residual(self, w=None, out=None):
        if out is None:
            out = np.empty(self.ndof)
        if w is not None and isinstance(w, np.ndarray):
            w = np.ascontiguousarray(w, dtype=np.float64)
        L.check(self.lib.femo_assemble_residual(self._ctx, L.ptr(w), L.ptr(out), L.on_device(out)))
        return out

    def evaluate_residual(self, inputs, w=None, out=None):
        """R(w; inputs) through host buffers in one call with the copies overlapped (femo_evaluate_residual): `inputs`
        maps input names to host arrays (pinned memory makes the overlap real), `w` the state or None for the resident
        one.  Equivalent to set_input for every entry followed by assemble_residual(w, out)."""
        if self._mg is not None:                    # the coarser levels of a hierarchy follow through set_input
            for k, v in inputs.items():
                self.set_input(k, v)
            return self.assemble_residual(w, out)
        if out is None:
            out = np.empty(self.ndof)
        names = list(inputs)
        arrs = [np.ascontiguousarray(np.asarray(inputs[k], dtype=np.float64).ravel()) for k in names]
        ids = (C.c_int * len(names))(*[self._arg_id(k) for k in names])
        ptrs = (C.c_void_p * len(names))(*[a.ctypes.data for a in arrs])
        sizes = (C.c_int64 * len(names))(*[a.size for a in arrs])
        if w is not None:
            w = np.ascontiguousarray(w, dtype=np.float64)
        L.check(self.lib.femo_evaluate_residual(self._ctx, len(names), ids, ptrs, sizes, L.ptr(w), L.ptr(out)))
        return out

    def assemble_load(self):
        out = np.empty(self.ndof)
        L.check(self.lib.femo_assemble_load(self._ctx, L.ptr(out), 0))
        return out

    def jacobian_apply(self, x, with_bc=False):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty(self.ndof)
        L.check(self.lib.femo_jacobian_apply(self._ctx, L.ptr(x), L.ptr(y), int(with_bc), 0))
        return y

    def get_bsr(self, with_bc=False, values=True):
        rowptr = np.empty(self.nb + 1, dtype=np.int32)
        colidx = np.empty(self.nnzb, dtype=np.int32)
        vals = np.empty((self.nnzb, 3, 3)) if values else None
        L.check(self.lib.femo_get_bsr(self._ctx, L.ptr(rowptr), L.ptr(colidx), L.ptr(vals), int(with_bc)))
        return rowptr, colidx, vals

    def get_csr(self, with_bc=False):
        """dR/du as scipy CSR (diagnostics / tests; the reference's getValuesCSR analogue)."""
        import scipy.sparse as sp
        rowptr, colidx, vals = self.get_bsr(with_bc)
        return sp.bsr_matrix((vals, colidx, rowptr), shape=(self.ndof, self.ndof)).tocsr()

    # ---- solves -------------------------------------------------------------------------
    def enable_direct(self, refine=2, lazy=False):
        """Route solve_state / solve_adjoint / solve_forward through the sparse direct solver (csrc/coarse.cu: what MUMPS
        is to the reference, fea/utils_dolfinx.py:465-466,514-531).  By default the symbolic analysis and a first
        factorisation run now and raise if the factor would not fit (FEMO_DIRECT_MAX_GB); `refine` steps of iterative
        refinement per solve.  lazy=True (what RMShellPDE uses for meshes without a hierarchy) defers that to the first
        solve -- BCs and inputs are not known yet -- and falls back to the Krylov solver, with a warning, if it fails."""
        self._direct = dict(refine=int(refine), auto=bool(lazy))
        if lazy:
            return None
        try:
            L.check(self.lib.femo_direct_setup(self._ctx))
        except Exception:
            self._direct = None
            raise
        return self.direct_info()

    def _direct_or_none(self, rhs, zero_bc):
        """direct solve, or None after switching an auto-enabled direct solver off because it cannot serve this problem"""
        try:
            return self.solve_direct(rhs, self._direct["refine"], zero_bc=zero_bc)
        except RuntimeError as e:
            if not self._direct.get("auto"):
                raise
            import warnings
            warnings.warn("femo_alpha_b200: direct solver unavailable for this problem (%s); using the Krylov solver" % e)
            self._direct = None
            return None

    def disable_direct(self):
        self._direct = None

    def direct_info(self):
        v = (C.c_double * 4)()
        L.check(self.lib.femo_direct_info(self._ctx, v))
        return dict(level_sets=int(v[0]), max_front_dofs=int(v[1]), factor_bytes=float(v[2]), explicit_inverse=bool(v[3]))

    def solve_direct(self, rhs, refine=2, zero_bc=True):
        """A x = rhs by the direct solver (factorises first if the inputs or BCs changed)."""
        rhs = np.ascontiguousarray(rhs, dtype=np.float64)
        L.check(self.lib.femo_direct_setup(self._ctx))
        out = np.empty(self.ndof)
        rr = C.c_double(0.0)
        L.check(self.lib.femo_direct_solve(self._ctx, L.ptr(rhs), L.ptr(out), int(refine), C.byref(rr), 0, int(bool(zero_bc))))
        self.last_solve = dict(iters=1 + int(refine), relres=rr.value, solver="direct")
        return out

    def solve_state(self, rtol=1e-12, max_it=200000):
        if self._mg is not None:          # nested mesh hierarchy available: V-cycle preconditioned PCG (csrc/mg.cu)
            return self._mg.solve_state(rtol, min(max_it, 5000))
        if getattr(self, "_direct", None) is not None:
            u = self._direct_or_none(self.assemble_load(), True)
            if u is not None:
                self.set_state(u)
                return self.last_solve["iters"], self.last_solve["relres"]
        it = C.c_int(0); rr = C.c_double(0.0)
        L.check(self.lib.femo_solve_state(self._ctx, float(rtol), int(max_it), C.byref(it), C.byref(rr)))
        self.last_solve = dict(iters=it.value, relres=rr.value)
        return it.value, rr.value

    def _solve(self, fn, rhs, rtol, max_it):
        rhs = np.ascontiguousarray(rhs, dtype=np.float64)
        if self._mg is not None:
            out = self._mg.solve(rhs, rtol, min(max_it, 5000), zero_bc=(fn is self.lib.femo_solve_adjoint))
            self.last_solve = dict(self._mg.last_solve)
            return out
        if getattr(self, "_direct", None) is not None:
            out = self._direct_or_none(rhs, fn is self.lib.femo_solve_adjoint)
            if out is not None:
                return out
        out = np.empty(self.ndof)
        it = C.c_int(0); rr = C.c_double(0.0)
        L.check(fn(self._ctx, L.ptr(rhs), L.ptr(out), float(rtol), int(max_it), C.byref(it), C.byref(rr), 0))
        self.last_solve = dict(iters=it.value, relres=rr.value)
        return out

    def solve_adjoint(self, rhs, rtol=1e-12, max_it=200000):
        return self._solve(self.lib.femo_solve_adjoint, rhs, rtol, max_it)

    def solve_forward(self, rhs, rtol=1e-12, max_it=200000):
        return self._solve(self.lib.femo_solve_forward, rhs, rtol, max_it)

    # ---- partials ---------------------------------------------------------------------------
    def dRdf_T_apply(self, arg, lam):
        a = self._arg_id(arg)
        lam = np.ascontiguousarray(lam, dtype=np.float64)
        out = np.empty(self._arg_size(a))
        L.check(self.lib.femo_dRdf_T_apply(self._ctx, a, L.ptr(lam), L.ptr(out), 0))
        return out

    def dRdf_apply(self, arg, df):
        a = self._arg_id(arg)
        df = np.ascontiguousarray(df, dtype=np.float64)
        out = np.empty(self.ndof)
        L.check(self.lib.femo_dRdf_apply(self._ctx, a, L.ptr(df), L.ptr(out), 0))
        return out

    # ---- outputs ---------------------------------------------------------------------------------
    def output(self, name):
        v = C.c_double(0.0)
        L.check(self.lib.femo_output(self._ctx, L.OUTPUT_IDS[name] if isinstance(name, str) else int(name), C.byref(v)))
        return v.value

    def output_grad(self, name, arg):
        a = self._arg_id(arg)
        out = np.empty(self._arg_size(a))
        L.check(self.lib.femo_output_grad(self._ctx, L.OUTPUT_IDS[name] if isinstance(name, str) else int(name), a,
                                          L.ptr(out), 0))
        return out

    # ---- stress functionals ----------------------------------------------------------------------
    def set_stress_params(self, m, rho):
        L.check(self.lib.femo_set_stress_params(self._ctx, float(m), float(rho)))

    def set_cell_tags(self, tags):
        t = np.ascontiguousarray(tags, dtype=np.int32)
        L.check(self.lib.femo_set_cell_tags(self._ctx, t.size, L.ptr(t)))

    def pnorm_stress(self, tag=-1):
        v = C.c_double(0.0)
        L.check(self.lib.femo_pnorm_stress(self._ctx, int(tag), C.byref(v)))
        return v.value

    def pnorm_stress_grad(self, arg, tag=-1):
        a = self._arg_id(arg)
        out = np.empty(self._arg_size(a))
        L.check(self.lib.femo_pnorm_stress_grad(self._ctx, int(tag), a, L.ptr(out), 0))
        return out

    def stress_field(self, surface="Top"):
        out = np.empty(4 * self.nc)
        L.check(self.lib.femo_stress_field(self._ctx, {"Top": 0, "Mid": 1, "Bot": 2}[surface], L.ptr(out), 0))
        return out

    # ---- small host-side pre-processing maps (not on the hot path) -------------------------------
    def _cell_vertices(self):
        return self.cell_blocks()[:, 9:13] - self.nq2          # tensor-ordered vertex ids

    def _q1_cell_mass(self):
        """int N_i N_j dx per cell, 2x2 Gauss (exact for the bilinear measure on planar cells)."""
        cv = self._cell_vertices()
        x = self.block_coords()[:self.nn][cv]                   # (nc,4,3)
        g = 0.5 + np.array([-1.0, 1.0]) * 0.5 / np.sqrt(3.0)
        M = np.zeros((self.nc, 4, 4)); area = np.zeros(self.nc); b = np.zeros((self.nc, 4))
        for X0 in g:
            for X1 in g:
                N = np.array([(1 - X0) * (1 - X1), X0 * (1 - X1), (1 - X0) * X1, X0 * X1])
                dN = np.array([[-(1 - X1), -(1 - X0)], [(1 - X1), -X0], [-X1, (1 - X0)], [X1, X0]])
                J = np.einsum("cvi,va->cia", x, dN)
                dJ = np.linalg.norm(np.cross(J[:, :, 0], J[:, :, 1]), axis=1) * 0.25
                M += dJ[:, None, None] * np.outer(N, N)[None]
                b += dJ[:, None] * N[None]
                area += dJ
        return cv, M, b, area

    def support_sizes(self):
        """assemble_vector(v*dx) of the scalar load space (rm_shell_pde.py:46-47)."""
        cv, M, b, area = self._q1_cell_mass()
        if self.n_f == 3 * self.nc:
            return area
        out = np.zeros(self.nn)
        np.add.at(out, cv.ravel(), b.ravel())
        return out

    def load_mass_matrix(self):
        """int Pv.w dx on the vector load space as SciPy CSR (rm_shell_pde.py:194-209)."""
        import scipy.sparse as sp
        cv, M, b, area = self._q1_cell_mass()
        if self.n_f == 3 * self.nc:
            return sp.diags(np.repeat(area, 3)).tocsr()
        rows, cols, vals = [], [], []
        for i in range(3):
            rows.append(np.repeat(3 * cv + i, 4, axis=1).ravel())
            cols.append(np.tile(3 * cv + i, (1, 4)).ravel())
            vals.append(M.ravel())
        A = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))),
                          shape=(3 * self.nn, 3 * self.nn)).tocsr()
        A.eliminate_zeros()
        return A

    # ---- instrumentation ------------------------------------------------------------------------------
    def launch_count(self):
        n = C.c_int64(0)
        L.check(self.lib.femo_launch_count(self._ctx, C.byref(n)))
        return n.value

    def last_timings(self):
        t = (C.c_double * 4)()
        L.check(self.lib.femo_last_timings(self._ctx, t))
        return dict(records_ms=t[0], gather_ms=t[1], solve_ms=t[2], iter_ms=t[3])

    def pcg_fixed(self, n_iter):
        ms = C.c_double(0.0)
        L.check(self.lib.femo_pcg_fixed(self._ctx, int(n_iter), C.byref(ms)))
        return ms.value

    def measure_peaks(self):
        tf = C.c_double(0.0); gb = C.c_double(0.0)
        L.check(self.lib.femo_measure_peaks(self._ctx, C.byref(tf), C.byref(gb)))
        return tf.value, gb.value

    def dmma_tflops(self):
        tf = C.c_double(0.0)
        L.check(self.lib.femo_last_dmma_tflops(self._ctx, C.byref(tf)))
        return tf.value
