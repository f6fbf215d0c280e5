"""`FEA`: registry of inputs / states / outputs of one finite-element problem, with the solve
entry points -- same public surface as femo_alpha/fea/fea_dolfinx.py:22-221, backed by one
libfemo_b200 context instead of dolfinx/PETSc objects.
"""
from __future__ import annotations

import os

import numpy as np

from .utils_dolfinx import *  # noqa: F401,F403  (the reference re-exports its utils the same way, fea_dolfinx.py:5)
from .utils_dolfinx import (Function, FunctionSpace, DirichletBC, KSP, JacobianMat, derivative, getFuncArray,
                            setFuncArray)


class _NpyRecorder:
    """`record=True` writes XDMF/VTX in the reference (fea_dolfinx.py:209-222); no HDF5/ADIOS2
    here, so snapshots go to `<recorder_path>/record_<name>_<iter>.npy`."""

    def __init__(self, path, name):
        self.path, self.name = path, name

    def write_mesh(self, mesh):
        pass

    def write_function(self, u, t=0.0):
        os.makedirs(self.path, exist_ok=True)
        np.save(os.path.join(self.path, "record_%s_%s.npy" % (self.name, str(t))), np.asarray(u.x.array))


class FEA(object):
    def __init__(self, mesh):
        self.mesh = mesh
        self.inputs_dict = dict()
        self.states_dict = dict()
        self.outputs_dict = dict()
        self.outputs_field_dict = dict()
        self.bc = []

        self.PDE_SOLVER = 'Newton'
        self.REPORT = False

        self.ubc = None
        self.custom_solve = None

        self.opt_iter = 0
        self.initial_solve = True
        self.initialize = False
        self.record = False
        self.recorder_path = 'records'
        self.linear_problem = False
        # B200 back-end knobs (SURVEY 5 "config"): Krylov tolerance / cap of the state and adjoint solves
        self.backend_options = dict(rtol=1e-12, max_it=500000)
        self.solve_log = []

        self.nel = mesh.topology.index_map(mesh.topology.dim).size_local
        self.nn = mesh.topology.index_map(0).size_local
        self._bc_signature = None

    # ---- registry (fea_dolfinx.py:53-157) ---------------------------------------------------
    def add_input(self, name: str, function: Function, init_val=1.0, record=False):
        if name in self.inputs_dict:
            raise ValueError('name has already been used for an input')
        function._role = ("input", name)
        function.name = name
        function.x.array[:] = init_val
        self.inputs_dict[name] = dict(
            function=function,
            function_space=function.function_space,
            shape=len(getFuncArray(function)),
            recorder=self.createRecorder(name, record),
            record=record)

    def add_state(self, name: str, function: Function, residual_form, arguments: list, dR_du=None, dR_df_list=None,
                  record=False):
        function._role = ("state",)
        function.name = name
        if dR_du is None:
            dR_du = derivative(residual_form, function)
        self.states_dict[name] = dict(
            function=function,
            residual_form=residual_form,
            function_space=function.function_space,
            shape=len(getFuncArray(function)),
            d_residual=Function(function.function_space),
            d_state=Function(function.function_space),
            dR_du=dR_du,
            dR_df_list=dR_df_list,
            arguments=arguments,
            recorder=self.createRecorder(name, record),
            record=record)

    def add_output(self, name: str, form, arguments: list):
        partials = []
        for argument in arguments:
            if argument in self.inputs_dict:
                partials.append(derivative(form, self.inputs_dict[argument]['function']))
            elif argument in self.states_dict:
                partials.append(derivative(form, self.states_dict[argument]['function']))
        self.outputs_dict[name] = dict(form=form, shape=1, arguments=arguments, partials=partials)

    def add_field_output(self, name, form, arguments, function_space=('CG', 1), record=False, vtk=False):
        V = form.field_space(function_space)
        output_func = Function(V)
        self.outputs_field_dict[name] = dict(
            form=form, function=output_func, shape=len(getFuncArray(output_func)), arguments=arguments, partials=[],
            recorder=self.createRecorder(name, record, vtk=vtk), record=record)

    def add_strong_bc(self, ubc, locate_BC_list, function_space=None):
        for locate_BC in locate_BC_list:
            self.bc.append(DirichletBC(locate_BC))

    # ---- solves (fea_dolfinx.py:159-203) -----------------------------------------------------
    def _sync_bcs(self, device, bc):
        dofs = np.unique(np.concatenate([b.dof_indices()[0] for b in bc])) if len(bc) else np.zeros(0, np.int32)
        sig = (dofs.size, int(dofs.sum()) if dofs.size else 0)
        if sig != self._bc_signature:
            device.set_strong_bc(dofs)
            self._bc_signature = sig

    def solve(self, res, func, bc):
        """Newton on a linear problem == one linear solve; warm-started from `func`."""
        if self.custom_solve is not None and self.initial_solve is True:
            self.custom_solve(res, func, bc, self.REPORT)
            return
        device = res.device
        self._sync_bcs(device, bc)
        if self.initialize:
            func.x.array[:] = 0.1          # utils_dolfinx.py:451-453
        device.set_state(func.x.array)
        its, relres = device.solve_state(self.backend_options['rtol'], self.backend_options['max_it'])
        self.solve_log.append(dict(kind='state', iters=its, relres=relres))
        if relres > 10 * self.backend_options['rtol']:
            import warnings
            warnings.warn("femo_alpha_b200: state solve stopped at relres %.2e after %d PCG iterations" % (relres, its))
        device.get_state(func.x.array)
        if self.REPORT:
            print('Solve finished: %d PCG iterations, relative residual %.3e' % (its, relres))

    def _linear(self, A, rhs_fn, sol_fn, rhs_array, ksp):
        setFuncArray(rhs_fn, rhs_array)
        sol_fn.x.array[:] = 0.0
        if ksp is None:
            ksp = KSP(A)
        ksp.rtol, ksp.max_it = self.backend_options['rtol'], self.backend_options['max_it']
        ksp.solve(rhs_fn, sol_fn)
        self.solve_log.append(dict(kind='linear', iters=ksp.its, relres=ksp.relres))
        return sol_fn.x.array

    def solveLinearFwd(self, du, A, dR, dR_array, ksp=None):
        """solve A du = dR (the reference's fwd path passes its vectors in the wrong order)"""
        return self._linear(A, dR, du, dR_array, ksp)

    def solveLinearBwd(self, dR, A, du, du_array, ksp=None):
        """solve A^T dR = du; A is symmetric (fea_dolfinx.py:189-203)"""
        return self._linear(A, du, dR, du_array, ksp)

    def projectFieldOutput(self, form, func):
        func.x.array[:] = form.project()

    def createRecorder(self, name, record=False, vtk=False):
        if record or self.record:
            return _NpyRecorder(self.recorder_path, name)
        return None
