"""Explicit CSDL operations for scalar and field outputs -- drop-in for
femo_alpha/csdl_alpha_opt/output_operation.py:6-128."""

import numpy as np

from ..csdl_compat import csdl
from ..fea.fea_dolfinx import FEA, update, assemble, computePartials, getFuncArray


def _collect_args(fea, args_name_list):
    args = dict()
    for arg_name in args_name_list:
        if arg_name in fea.inputs_dict:
            args[arg_name] = fea.inputs_dict[arg_name]
        elif arg_name in fea.states_dict:
            args[arg_name] = fea.states_dict[arg_name]
    return args


class OutputOperation(csdl.CustomExplicitOperation):
    def __init__(self, fea, args_name_list, output_name):
        super().__init__()
        csdl.check_parameter(fea, "fea", types=FEA)
        csdl.check_parameter(args_name_list, "args_name_list", types=list)
        csdl.check_parameter(output_name, "output_name", types=str)
        self.fea = fea
        self.args_dict = _collect_args(fea, args_name_list)
        self.output_name = output_name
        self.fea_output = fea.outputs_dict[output_name]
        self.output_dim = 0

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
        for arg_name in input_vals:
            update(self.args_dict[arg_name]["function"], input_vals[arg_name])
        output_vals[self.output_name] = assemble(self.fea_output["form"])

    def compute_derivatives(self, input_vals, output_vals, derivatives):
        for arg_name in input_vals:
            update(self.args_dict[arg_name]["function"], input_vals[arg_name])
        for arg_name in input_vals:
            derivatives[self.output_name, arg_name] = assemble(
                computePartials(self.fea_output["form"], self.args_dict[arg_name]["function"]),
                dim=self.output_dim + 1)


class OutputFieldOperation(csdl.CustomExplicitOperation):
    def __init__(self, fea, args_name_list, output_name):
        super().__init__()
        csdl.check_parameter(fea, "fea", types=FEA)
        csdl.check_parameter(args_name_list, "args_name_list", types=list)
        csdl.check_parameter(output_name, "output_name", types=str)
        self.fea = fea
        self.args_dict = _collect_args(fea, args_name_list)
        self.output_name = output_name
        self.fea_output = fea.outputs_field_dict[output_name]
        self.output_dim = 1

    def evaluate(self, inputs):
        for arg_name in self.args_dict:
            if getattr(inputs, arg_name) is not None:
                self.declare_input(arg_name, getattr(inputs, arg_name))
            else:
                raise ValueError(f"Variable {arg_name} not found in the FEA model.")
        output = self.create_output(self.output_name, (self.fea_output['shape'],))
        output.add_name(self.output_name)
        self.declare_derivative_parameters(self.output_name, '*', dependent=True)
        return output

    def compute(self, input_vals, output_vals):
        for arg_name in input_vals:
            update(self.args_dict[arg_name]['function'], input_vals[arg_name])
        self.fea.projectFieldOutput(self.fea_output['form'], self.fea_output['function'])
        output_vals[self.output_name] = getFuncArray(self.fea_output['function']).copy()
        if self.fea_output['record'] and self.fea_output['recorder'] is not None:
            self.fea_output['recorder'].write_function(self.fea_output['function'], self.fea.opt_iter)

    def compute_derivatives(self, input_vals, output_vals, derivatives):
        pass                                   # the reference provides none (output_operation.py:72-128)
