"""Implicit CSDL operation of one FEA state -- drop-in for
femo_alpha/csdl_alpha_opt/state_operation.py:8-295 (same ctor, methods, dict protocol).

Differences by design: derivatives are refreshed at EVERY evaluation
(B1: the reference only assembles them at opt_iter == 1 for linear problems, leaving a stale
adjoint afterwards; set `reference_stale_derivatives=True` to emulate), the partials dR/df are
matrix-free device operators (including the shape partial dR/d(uhat), csrc/shape_partials.cuh), and forward mode
works (B5).
"""
import numpy as np

from ..csdl_compat import csdl
from ..fea.fea_dolfinx import (FEA, update, getFuncArray, computePartials, assembleMatrix, computeMatVecProductFwd,
                               computeMatVecProductBwd, createFunction, assembleSystem, setUpKSP_MUMPS)


class StateOperation(csdl.experimental.CustomImplicitOperation):
    reference_stale_derivatives = False

    def __init__(self, fea, args_name_list, state_name, debug_mode=False):
        super().__init__()
        csdl.check_parameter(fea, 'fea', types=FEA)
        csdl.check_parameter(args_name_list, 'args_name_list', types=list)
        csdl.check_parameter(state_name, 'state_name', types=str)
        self.fea = fea
        self.args_dict = {a: fea.inputs_dict[a] for a in args_name_list}
        self.state_name = state_name
        self.debug_mode = debug_mode
        self.fea_state = fea.states_dict[state_name]
        self.fea_dR = self.fea_state['d_residual']
        self.fea_du = self.fea_state['d_state']
        self.set_up_fea_derivatives()

    def _banner(self, what):
        if self.debug_mode is True:
            print('=' * 15 + str(self.state_name) + '=' * 15)
            print('CSDL: Running %s()...' % what)
            print('=' * 40)

    def evaluate(self, inputs):
        self._banner('evaluate')
        for arg_name in self.args_dict:
            if getattr(inputs, arg_name) is not None:
                self.declare_input(arg_name, getattr(inputs, arg_name))
            else:
                raise ValueError(f'Variable {arg_name} not found in the FEA model.')
        state = self.create_output(self.state_name, shape=(self.fea_state['shape'],))
        state.add_name(self.state_name)
        self.declare_derivative_parameters(self.state_name, '*', dependent=True)
        return state

    def solve_residual_equations(self, input_vals, output_vals):
        self._banner('solve_residual_equations')
        self.fea.opt_iter += 1
        for arg_name in input_vals:
            arg = self.args_dict[arg_name]
            update(arg['function'], input_vals[arg_name])
            if arg['record'] and arg['recorder'] is not None:
                arg['recorder'].write_function(arg['function'], self.fea.opt_iter)
        self.fea.solve(self.fea_state['residual_form'], self.fea_state['function'], self.fea.bc)
        output_vals[self.state_name] = getFuncArray(self.fea_state['function']).copy()
        if self.fea.record and self.fea_state['recorder'] is not None:
            self.fea_state['recorder'].write_function(self.fea_state['function'], self.fea.opt_iter)
        if (not self.reference_stale_derivatives) or self.fea.linear_problem is False or self.fea.opt_iter == 1:
            self.assemble_derivatives(input_vals, output_vals)

    def compute_jacvec_product(self, input_vals, output_vals, d_inputs, d_outputs, d_residuals, mode):
        self._banner('compute_jacvec_product')
        state_name = self.state_name
        if mode == 'fwd':
            if state_name in d_residuals:
                if state_name in d_outputs:
                    update(self.fea_du, d_outputs[state_name])
                    d_residuals[state_name] += computeMatVecProductFwd(self.dRdu, self.fea_du)
                for arg_name in self.dR_df_dict:
                    if arg_name in d_inputs:
                        update(self.dR_df_dict[arg_name]['fea_df'], d_inputs[arg_name])
                        d_residuals[state_name] += computeMatVecProductFwd(
                            self.dR_df_dict[arg_name]['dRdf'], self.dR_df_dict[arg_name]['fea_df'])
        elif mode == 'rev':
            if state_name in d_residuals:
                update(self.fea_dR, d_residuals[state_name])
                for arg_name in self.dR_df_dict:
                    if arg_name in d_inputs:
                        d_inputs[arg_name] += computeMatVecProductBwd(self.dR_df_dict[arg_name]['dRdf'], self.fea_dR)
        else:
            raise ValueError("mode must be either 'fwd' or 'rev'.")

    def apply_inverse_jacobian(self, input_vals, output_vals, d_outputs, d_residuals, mode):
        self._banner('apply_inverse_jacobian')
        state_name = self.state_name
        if mode == 'fwd':
            d_outputs[state_name] = self.fea.solveLinearFwd(self.fea_du, self.A, self.fea_dR, d_residuals[state_name],
                                                            self.ksp).copy()
        elif mode == 'rev':
            d_residuals[state_name] = self.fea.solveLinearBwd(self.fea_dR, self.A, self.fea_du, d_outputs[state_name],
                                                              self.ksp).copy()
            for bc in self.fea.bc:                       # state_operation.py:217-218
                d_residuals[state_name][bc.dof_indices()[0]] = 0.0
        else:
            raise ValueError("mode must be either 'fwd' or 'rev'.")

    def set_up_fea_derivatives(self):
        self._banner('set_up_fea_derivatives')
        if self.fea_state['dR_du'] is None:
            self.fea_state['dR_du'] = computePartials(self.fea_state['residual_form'], self.fea_state['function'])
        dR_df_dict = dict()
        dR_df_list = self.fea_state['dR_df_list']
        for i, arg_name in enumerate(self.fea_state['arguments']):
            fn = self.args_dict[arg_name]['function']
            dR_df = computePartials(self.fea_state['residual_form'], fn) if dR_df_list is None else dR_df_list[i]
            dR_df_dict[arg_name] = dict(dR_df=dR_df, fea_df=createFunction(fn))
        self.dR_df_dict = dR_df_dict

    def assemble_derivatives(self, input_vals, output_vals):
        self._banner('assemble_derivatives')
        for arg_name in input_vals:
            update(self.args_dict[arg_name]['function'], input_vals[arg_name])
        update(self.fea_state['function'], output_vals[self.state_name])
        for arg_name in self.dR_df_dict:
            self.dR_df_dict[arg_name]['dRdf'] = assembleMatrix(self.dR_df_dict[arg_name]['dR_df'])
        self.dRdu = assembleMatrix(self.fea_state['dR_du'])
        self.A, _ = assembleSystem(self.fea_state['dR_du'], self.fea_state['residual_form'], bcs=self.fea.bc)
        self.ksp = setUpKSP_MUMPS(self.A)
