"""Implicit CSDL operation of the dynamic plate -- drop-in for
femo_alpha/dynamic_rm_shell/state_operation_dynamic.py:19-702 on the CUDA back-end.

Same constructor, argument names ('thickness', 'force_history'), state name ('disp_history'), and the
solve_residual_equations / compute_jacvec_product / apply_inverse_jacobian contract; `gradient_mode`
('numpy' | 'petsc') is accepted and ignored -- both are one recurrence on the GPU (csrc/dynamic.cu), without the
reference's dense matrices or its N^2 work vectors.  The forward-mode chain-rule sign slip of the reference
(:253-257,531-536) is not reproduced (see oracle/dynamic_ref.py).
"""
import numpy as np

from ..csdl_compat import csdl
from .plate_sim import PlateSim
from .utils import stack_array_into_vector, reshape_vector_into_array


class StateOperation(csdl.experimental.CustomImplicitOperation):
    def __init__(self, plate_sim, gradient_mode='numpy', debug_mode=False, record=True, path="./"):
        super().__init__()
        csdl.check_parameter(plate_sim, 'plate_sim')
        self.plate_sim = plate_sim
        self.gradient_mode = gradient_mode
        self.debug_mode = debug_mode
        self.state_name = 'disp_history'
        self.path = path
        self.args_dict = ['thickness', 'force_history']
        self.input_name = 'thickness'
        self.record = record          # XDMF thickness records are not written
        self.eval_iter = 0
        self.bc_dofs = plate_sim.bc_dofs

    def _debug(self, what):
        if self.debug_mode is True:
            print('=' * 15 + str(self.state_name) + '=' * 15)
            print('CSDL: Running %s()...' % what)
            print('=' * 40)

    def evaluate(self, inputs):
        self._debug('evaluate')
        for arg_name in self.args_dict:
            if getattr(inputs, arg_name) is not None:
                self.declare_input(arg_name, getattr(inputs, arg_name))
            else:
                raise ValueError(f'Variable {arg_name} not found in the FEA model.')
        state = self.create_output(self.state_name, shape=(self.plate_sim.fe_dofs * self.plate_sim.time_levels,))
        state.add_name(self.state_name)
        self.declare_derivative_parameters(self.state_name, '*', dependent=True)
        return state

    def solve_residual_equations(self, input_vals, output_vals):
        self._debug('solve_residual_equations')
        ps = self.plate_sim
        ps.update_t(input_vals['thickness'])
        fh = np.asarray(input_vals['force_history'], dtype=np.float64)
        ps.update_f_history(fh.reshape(-1, ps.device.n_f))
        ps.update_materialmodel()
        ps.reset_solution_vectors()
        svk_res = ps.SVK_residual()
        self.eval_iter += 1
        sol_output = ps.solve_dynamic_problem(svk_res, saving_outputs=False, PATH=self.path, timing=False)
        output_vals[self.state_name] = stack_array_into_vector(sol_output)

    def _prepare(self, input_vals, output_vals):
        ps = self.plate_sim
        ps.update_t(input_vals[self.input_name])
        ps.update_materialmodel()
        ps.load_history(reshape_vector_into_array(np.asarray(output_vals[self.state_name]), ps.time_levels))
        return ps

    def compute_jacvec_product(self, input_vals, output_vals, d_inputs, d_outputs, d_residuals, mode):
        self._debug('compute_jacvec_product')
        ps = self._prepare(input_vals, output_vals)
        dyn, L = ps._dyn, ps.time_levels
        name = self.state_name
        if mode == 'fwd':
            if name in d_residuals:
                res = np.zeros((L, ps.fe_dofs))
                if name in d_outputs:
                    res += dyn.dRdw_apply(np.asarray(d_outputs[name]).reshape(L, -1))
                if 'thickness' in d_inputs:
                    res += dyn.dRdh_apply(np.asarray(d_inputs['thickness']))
                d_residuals[name] = res.ravel()
        elif mode == 'rev':
            if name in d_residuals:
                lam = np.asarray(d_residuals[name]).reshape(L, -1)
                if 'thickness' in d_inputs:
                    d_inputs['thickness'] += dyn.dRdh_T_apply(lam)
        else:
            raise ValueError("mode must be either 'fwd' or 'rev'.")

    def apply_inverse_jacobian(self, input_vals, output_vals, d_outputs, d_residuals, mode):
        self._debug('apply_inverse_jacobian')
        ps = self._prepare(input_vals, output_vals)
        dyn, L = ps._dyn, ps.time_levels
        name = self.state_name
        if mode == 'fwd':
            d_outputs[name] = dyn.solve_forward(np.asarray(d_residuals[name]).reshape(L, -1)).ravel()
        elif mode == 'rev':
            d_residuals[name] = dyn.solve_adjoint(np.asarray(d_outputs[name]).reshape(L, -1)).ravel()
        else:
            raise ValueError("mode must be either 'fwd' or 'rev'.")
