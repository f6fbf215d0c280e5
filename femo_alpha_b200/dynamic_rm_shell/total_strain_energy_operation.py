"""Total strain energy over all time levels -- drop-in for
femo_alpha/dynamic_rm_shell/total_strain_energy_operation.py:21-139 on the CUDA back-end."""
import numpy as np

from ..csdl_compat import csdl
from .utils import reshape_vector_into_array


class TotalStrainEnergyOperation(csdl.CustomExplicitOperation):
    def __init__(self, plate_sim):
        super().__init__()
        csdl.check_parameter(plate_sim, "plate_sim")
        self.plate_sim = plate_sim
        self.output_dim = 0
        self.args_dict = ['thickness', 'disp_history']
        self.output_name = 'total_strain_energy'
        self.regularization = False

    def evaluate(self, inputs):
        for arg_name in self.args_dict:
            if getattr(inputs, arg_name) is not None:
                self.declare_input(arg_name, getattr(inputs, arg_name))
            else:
                raise ValueError(f"Variable {arg_name} not found in the FEA model.")
        output = self.create_output(self.output_name, (1,))
        output.add_name(self.output_name)
        self.declare_derivative_parameters(self.output_name, '*', dependent=True)
        return output

    def _prepare(self, input_vals):
        ps = self.plate_sim
        ps.update_t(input_vals['thickness'])
        ps.update_materialmodel()
        ps.load_history(reshape_vector_into_array(np.asarray(input_vals['disp_history']), ps.time_levels))
        return ps

    def _reg(self, grad=False):
        # total_strain_energy_operation.py:127-134: alpha = 1e-2 * time_levels, (grad t)^2 (CG1) or t^2 (DG0):
        # 2 alpha / 1e-2 times the regulariser of the static compliance form (0.5 * 1e-2 * ...)
        ps = self.plate_sim
        scale = 2.0 * ps.time_levels
        dev = ps.device
        if grad:
            return scale * dev.output_grad("compliance", "thickness")
        w = dev.get_state()
        dev.set_state(np.zeros(dev.ndof))
        val = scale * dev.output("compliance")
        dev.set_state(w)
        return val

    def compute(self, input_vals, output_vals):
        ps = self._prepare(input_vals)
        total = ps._dyn.total_strain_energy()
        output_vals['total_strain_energy'] = total + (self._reg() if self.regularization else 0.)

    def compute_derivatives(self, input_vals, output_vals, derivatives):
        ps = self._prepare(input_vals)
        dEdt = ps._dyn.total_strain_energy_grad("thickness")
        if self.regularization:
            dEdt = dEdt + self._reg(grad=True)
        derivatives[self.output_name, 'thickness'] = dEdt
        derivatives[self.output_name, 'disp_history'] = ps._dyn.total_strain_energy_grad("state").ravel()
