"""PlateSim: time-stepped Reissner-Mindlin plate/shell -- drop-in for
femo_alpha/dynamic_rm_shell/plate_sim.py:61-361 on the CUDA back-end (csrc/dynamic.cu).

Same constructor arguments, attribute names and method names as the reference class; the UFL residual object
returned by `SVK_residual()` becomes a light handle, and `solve_dynamic_problem(res)` runs the whole implicit
midpoint loop on the GPU (one assembly of K and M per design, one solve per step).  The element classes of the
reference come from the un-vendored `shell_analysis_fenicsx`; the forms used here are its in-tree analogue
(rm_shell/linear_shell_fenicsx/linear_shell_model.py:260-348), see oracle/dynamic_ref.py.
"""
from __future__ import annotations

from time import perf_counter

import numpy as np

from ..dynamic_device import DynamicShellDevice
from ..shell_device import ShellDevice


class _Vec:
    """`function.vector` / `.x.array` stand-in: a host array."""
    def __init__(self, n):
        self.array = np.zeros(n)
        self.size = n

    def __getitem__(self, k):
        return self.array[k]


class _Func:
    def __init__(self, n):
        self.vector = _Vec(n)
        self.x = self.vector


class BC_list(object):
    """Clamped dofs (plate_sim.py:23-59): default x[0] ~ 0 (atol 1e-6), or `custom_bc_func`."""

    def __init__(self, device, custom_bc_func=None):
        self.device = device
        self.custom_bc_func = custom_bc_func

    def _on(self):
        xc = self.device.block_coords()
        f = self.custom_bc_func if self.custom_bc_func is not None else (lambda x: np.isclose(x[0], 0., atol=1e-6))
        return np.asarray(f(xc.T)).astype(bool)

    def return_affected_dof_tuple(self):
        on = self._on()
        nq2 = self.device.nq2
        bu = np.nonzero(on[:nq2])[0]
        bt = nq2 + np.nonzero(on[nq2:])[0]
        dofs = lambda b: np.sort((3 * b[:, None] + np.arange(3)[None, :]).ravel())
        return (dofs(bu), None), (dofs(bt), None)


class DynamicResidual:
    """Handle for M wddot + K w_mid - F - W (plate_sim.py:190-215)."""
    def __init__(self, sim):
        self.sim = sim


class PlateSim(object):
    def __init__(self, mesh, E, nu, rho, dt, Nsteps, element_wise_thickness=False,
                 custom_bc_func=None, add_self_weight=False, g_factor=None,
                 quad_deg=3, comm=None, device=0, solver="auto", rtol=1e-12):
        self.mesh = mesh
        self.E, self.nu, self.rho = E, nu, rho
        self.dt = dt
        self.Nsteps = Nsteps
        self.time_levels = Nsteps + 1
        self.comm = comm
        self.nn = mesh.topology.index_map(0).size_local
        self.nel = mesh.topology.index_map(mesh.topology.dim).size_local
        self.ALPHA = 1
        self.quad_deg = quad_deg
        self.element_wise_thickness = element_wise_thickness
        self.add_self_weight = add_self_weight
        self.g_factor = g_factor
        self._solver, self._rtol = solver, rtol
        self.device = dev = ShellDevice(mesh.geometry.x, mesh.cells, element_wise_thickness, False, quad_deg, device)
        # a mesh from refine_uniform carries its parents: V-cycle PCG for the step solves of large plates
        self.hierarchy = None
        chain = getattr(mesh, "mg_meshes", None)
        if chain and len(chain) >= 2 and (solver == "multigrid" or (solver == "auto" and dev.ndof > 46000)):
            from ..multigrid import ShellHierarchy
            self.hierarchy = ShellHierarchy.for_device(dev, chain, quad_deg, device, element_wise_material=element_wise_thickness)
        for name, val in (("E", E), ("nu", nu), ("density", rho)):
            dev.set_input(name, np.full(dev.n_mat, float(val)))
        self.bc_obj = BC_list(dev, custom_bc_func=custom_bc_func)
        (bu, _), (bt, _) = self.bc_obj.return_affected_dof_tuple()
        self.bc_dofs = np.union1d(bu, bt)
        dev.set_strong_bc(self.bc_dofs)
        self.bc_list = [self.bc_dofs]
        self.fe_dofs = dev.ndof
        self.num_var = dev.n_mat
        self.t = _Func(self.num_var)
        self.f = _Func(dev.n_f)
        self.w, self.w_old, self.wdot_old = _Func(self.fe_dofs), _Func(self.fe_dofs), _Func(self.fe_dofs)
        self.tip_disp_history = np.zeros((self.time_levels,))
        self.strain_energy_history = np.zeros((self.time_levels,))
        self.opt_iter = 0
        self.x_tip = None
        self.cell_tip = None
        self.f_history = None
        self._dyn = None
        self._make_dyn()

    def _make_dyn(self):
        if self._dyn is not None:
            self._dyn.close()
        self._dyn = DynamicShellDevice(self.device, self.dt, self.Nsteps, self.add_self_weight, self.g_factor,
                                       self._solver, self._rtol)
        if self.hierarchy is not None:
            self._dyn.attach_multigrid(self.hierarchy)

    # ---- reference method surface (plate_sim.py:153-279) -----------------------------------------------
    def set_up_tip_dofs(self, x_tip, cell_tip):
        self.x_tip = x_tip
        self.cell_tip = cell_tip

    def update_t(self, t_array):
        t_array = np.asarray(t_array, dtype=np.float64).ravel()
        if not np.array_equal(self.t.vector.array, t_array):
            self.t.vector.array = t_array.copy()
            self.device.set_input("thickness", t_array)

    def update_f(self, f_array):
        self.f.vector.array = np.asarray(f_array, dtype=np.float64).ravel().copy()

    def update_f_history(self, f_history_array):
        self.f_history = np.asarray(f_history_array, dtype=np.float64)

    def update_materialmodel(self):
        self.material_model = ("CLT", self.E, self.nu)

    @staticmethod
    def set_solution_vector(func_obj, arr):
        func_obj.vector.array = np.asarray(arr, dtype=np.float64).copy()

    def set_solution_vectors_for_adjoint(self, w, w_old, wdot_old):
        self.set_solution_vector(self.w, w)
        self.set_solution_vector(self.w_old, w_old)
        self.set_solution_vector(self.wdot_old, wdot_old)

    def reset_solution_vectors(self):
        z = np.zeros(self.fe_dofs)
        self.set_solution_vectors_for_adjoint(z, z, z)

    def SVK_residual(self):
        self.update_materialmodel()
        return DynamicResidual(self)

    def update_nsteps(self, Nsteps):
        self.Nsteps = Nsteps
        self.time_levels = Nsteps + 1
        self.tip_disp_history = np.zeros((self.time_levels,))
        self.strain_energy_history = np.zeros((self.time_levels,))
        self._make_dyn()

    def _tip_vertex(self):
        x = self.mesh.geometry.x
        xt = np.asarray(self.x_tip if self.x_tip is not None else [10., 0., 0.], dtype=np.float64)   # plate_sim.py:20
        return int(np.argmin(np.linalg.norm(x - xt[None, :], axis=1)))

    def solve_dynamic_problem(self, residual, saving_outputs=False, PATH=None, POD_matrix=None, POD_mean=None, timing=False):
        """Implicit midpoint loop from zero initial conditions (plate_sim.py:281-361); returns (fe_dofs, time_levels).
        `saving_outputs` (XDMF/npy records) is accepted and ignored."""
        if self.f_history is None:
            raise ValueError("update_f_history must be called before solve_dynamic_problem")
        t0 = perf_counter()
        f = self.f_history.reshape(self.f_history.shape[0], -1) if self.f_history.ndim > 1 else self.f_history[None, :]
        self._dyn.solve(f)
        w_hist = self._dyn.history()
        self.strain_energy_history = self._dyn.energies()
        self.tip_disp_history = w_hist[:, 3 * self._tip_vertex() + 2].copy()
        self.tip_disp_history[0] = 0.0
        self.set_solution_vector(self.w, w_hist[-1])
        self.set_solution_vector(self.w_old, w_hist[-1])
        if timing:
            print("Dynamic simulation wall time: {}".format(perf_counter() - t0))
        return np.ascontiguousarray(w_hist.T)

    def assembleStrainEnergy(self, w):
        arr = w.vector.array if hasattr(w, "vector") else np.asarray(w)
        self.device.set_state(np.ascontiguousarray(arr))
        return self.device.output("elastic_energy")

    # ---- device-side history used by the operations --------------------------------------------------------
    def load_history(self, disp_arr):
        """(fe_dofs, time_levels) array of the CSDL state -> resident history of the device."""
        self._dyn.set_history(np.ascontiguousarray(np.asarray(disp_arr).T))
