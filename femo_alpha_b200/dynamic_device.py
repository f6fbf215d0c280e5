"""Thin Python handle on one `femo_dyn` (time-stepped RM shell resident on one B200); see csrc/dynamic.cu.

All arithmetic happens in the CUDA library; histories are (levels, n_dof) row-major = the level-major vectors of
the reference (femo_alpha/dynamic_rm_shell/utils.py:9-16, ravel order 'F' of (n_dof, levels)).
"""
from __future__ import annotations

import ctypes as C
import weakref

import numpy as np

from . import _lib as L

SOLVERS = {"auto": 0, "dense": 1, "pcg": 2, "multigrid": 3}


class DynamicShellDevice:
    def __init__(self, device, dt, n_steps, add_self_weight=False, g_factor=None, solver="auto", rtol=1e-12,
                 max_it=20000):
        self.dev = device
        self.lib = device.lib
        self.dt, self.n_steps, self.levels = float(dt), int(n_steps), int(n_steps) + 1
        self.rtol, self.max_it = float(rtol), int(max_it)
        self.solver = SOLVERS[solver] if isinstance(solver, str) else int(solver)
        self._h = C.c_void_p()
        gf = -1.0 if g_factor is None else float(g_factor)           # plate_sim.py:205-209
        L.check(self.lib.femo_dyn_create(device._ctx, self.dt, self.n_steps, int(bool(add_self_weight)), gf,
                                         C.byref(self._h)))
        self.last_iters = 0
        self._configured = False        # femo_dyn_setup(solver) runs on first use; the library keeps the choice afterwards
        self._hier = None
        device._children.append(weakref.ref(self))     # the context must outlive this handle (ShellDevice.close)

    def close(self):
        """Destroy the handle.  Must run before the context (and an attached hierarchy) is destroyed; ShellDevice.close
        and ShellHierarchy.close call it for the handles registered with them."""
        if getattr(self, "_h", None) is not None and self._h.value:
            hier = getattr(self, "_hier", None)
            ctx_alive = getattr(self.dev, "_ctx", None) is not None and self.dev._ctx.value
            hier_alive = hier is None or (getattr(hier, "_mg", None) is not None and hier._mg.value)
            if ctx_alive and hier_alive:
                self.lib.femo_dyn_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _hist(self, a=None):
        if a is None:
            return np.empty((self.levels, self.dev.ndof))
        a = np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(self.levels, self.dev.ndof))
        return a

    def attach_multigrid(self, hierarchy):
        L.check(self.lib.femo_dyn_set_mg(self._h, hierarchy._mg))
        self._hier = hierarchy
        hierarchy._children.append(weakref.ref(self))
        self._configured = False

    def setup(self):
        L.check(self.lib.femo_dyn_setup(self._h, self.solver))
        self._configured = True

    def _ensure(self):
        if not self._configured:
            self.setup()

    def solve(self, f_history):
        self._ensure()
        f = np.ascontiguousarray(np.atleast_2d(np.asarray(f_history, dtype=np.float64)))
        if f.shape[1] != self.dev.n_f:
            raise ValueError("force history must have %d columns, got %d" % (self.dev.n_f, f.shape[1]))
        it = C.c_int64(0)
        L.check(self.lib.femo_dyn_solve(self._h, L.ptr(f), f.shape[0], self.rtol, self.max_it, C.byref(it), 0))
        self.last_iters = it.value
        return it.value

    def history(self, out=None):
        out = self._hist() if out is None else out
        L.check(self.lib.femo_dyn_get_history(self._h, L.ptr(out), L.on_device(out)))
        return out

    def set_history(self, w_hist):
        w = w_hist if L.on_device(w_hist) else self._hist(w_hist)
        L.check(self.lib.femo_dyn_set_history(self._h, L.ptr(w), L.on_device(w)))

    def energies(self):
        self._ensure()
        e = np.empty(self.levels)
        L.check(self.lib.femo_dyn_get_energies(self._h, L.ptr(e)))
        return e

    def _sweep(self, fn, rhs):
        self._ensure()
        rhs = self._hist(rhs)
        out = self._hist()
        it = C.c_int64(0)
        L.check(fn(self._h, L.ptr(rhs), L.ptr(out), self.rtol, self.max_it, C.byref(it), 0))
        self.last_iters = it.value
        return out

    def solve_adjoint(self, d_out_hist):
        return self._sweep(self.lib.femo_dyn_solve_adjoint, d_out_hist)

    def solve_forward(self, d_res_hist):
        return self._sweep(self.lib.femo_dyn_solve_forward, d_res_hist)

    def dRdw_apply(self, x_hist, transpose=False):
        self._ensure()
        x = self._hist(x_hist)
        y = self._hist()
        L.check(self.lib.femo_dyn_dRdw_apply(self._h, L.ptr(x), L.ptr(y), int(bool(transpose)), 0))
        return y

    def dRdh_T_apply(self, lam_hist):
        self._ensure()
        lam = self._hist(lam_hist)
        out = np.empty(self.dev.n_mat)
        L.check(self.lib.femo_dyn_dRdh_T_apply(self._h, L.ptr(lam), L.ptr(out), 0))
        return out

    def dRdh_apply(self, dh):
        self._ensure()
        dh = np.ascontiguousarray(dh, dtype=np.float64)
        out = self._hist()
        L.check(self.lib.femo_dyn_dRdh_apply(self._h, L.ptr(dh), L.ptr(out), 0))
        return out

    def total_strain_energy(self):
        self._ensure()
        v = C.c_double(0.0)
        L.check(self.lib.femo_dyn_total_strain_energy(self._h, C.byref(v)))
        return v.value

    def total_strain_energy_grad(self, arg):
        self._ensure()
        if arg in ("thickness", L.THICKNESS):
            out = np.empty(self.dev.n_mat)
            L.check(self.lib.femo_dyn_total_strain_energy_grad(self._h, L.THICKNESS, L.ptr(out), 0))
            return out
        out = self._hist()
        L.check(self.lib.femo_dyn_total_strain_energy_grad(self._h, L.ARG_STATE, L.ptr(out), 0))
        return out

    def mass_apply(self, x):
        self._ensure()
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty(self.dev.ndof)
        L.check(self.lib.femo_dyn_mass_apply(self._h, L.ptr(x), L.ptr(y), 0))
        return y

    def info(self):
        a = C.c_double(0.0); b = C.c_double(0.0); s = C.c_int(0)
        L.check(self.lib.femo_dyn_info(self._h, C.byref(a), C.byref(b), C.byref(s)))
        return dict(forward_ms=a.value, adjoint_ms=b.value, solver={1: "dense", 2: "pcg", 3: "multigrid"}.get(s.value, "?"))
