"""Plate volume int t dx -- drop-in for femo_alpha/dynamic_rm_shell/volume_operation.py:19-70 on the CUDA back-end."""
import numpy as np

from ..csdl_compat import csdl


class VolumeOperation(csdl.CustomExplicitOperation):
    def __init__(self, plate_sim):
        super().__init__()
        csdl.check_parameter(plate_sim, "plate_sim")
        self.plate_sim = plate_sim
        self.output_dim = 0
        self.args_dict = ['thickness']
        self.output_name = 'volume'

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

    def compute(self, input_vals, output_vals):
        self.plate_sim.update_t(input_vals['thickness'])
        output_vals['volume'] = self.plate_sim.device.output("volume")

    def compute_derivatives(self, input_vals, output_vals, derivatives):
        self.plate_sim.update_t(input_vals['thickness'])
        derivatives['volume', 'thickness'] = self.plate_sim.device.output_grad("volume", "thickness")
