"""Explicit CSDL operations for scalar and field outputs -- drop-in for
femo_alpha/csdl_alpha_opt/output_operation.py:6-128 (same constructors, `evaluate`, `compute`, `compute_derivatives`
and error behaviour).  Scalar outputs and their partials are device reductions behind `assemble(form)`; field outputs
are the per-cell projections of `FEA.projectFieldOutput`.
"""

import numpy as np  # noqa: F401  (kept for users who import it from here, as with the reference module)

from ..csdl_compat import csdl
from ..fea.fea_dolfinx import FEA, update, assemble, computePartials, getFuncArray


class _FEAOutputBase(csdl.CustomExplicitOperation):
    """What the two operations share: argument registry lookup, input declaration, pushing input values to the device."""

    registry = None        # name of the FEA dictionary that holds the output
    output_dim = 0         # 0 scalar, 1 field (rank of the derivative = output_dim + 1)

    def __init__(self, fea, args_name_list, output_name):
        super().__init__()
        for value, name, kind in ((fea, "fea", FEA), (args_name_list, "args_name_list", list), (output_name, "output_name", str)):
            csdl.check_parameter(value, name, types=kind)
        self.fea, self.output_name = fea, output_name
        known = {**fea.states_dict, **fea.inputs_dict}              # inputs win over states of the same name
        self.args_dict = {a: known[a] for a in args_name_list if a in known}
        self.fea_output = getattr(fea, self.registry)[output_name]

    def _output_shape(self):
        return (1,)

    def evaluate(self, inputs):
        for arg_name in self.args_dict:
            variable = getattr(inputs, arg_name)
            if variable is None:
                raise ValueError(f"Variable {arg_name} not found in the FEA model.")
            self.declare_input(arg_name, variable)
        output = self.create_output(self.output_name, self._output_shape())
        output.add_name(self.output_name)
        self.declare_derivative_parameters(self.output_name, '*', dependent=True)
        return output

    def _push(self, input_vals):
        for arg_name, value in input_vals.items():
            update(self.args_dict[arg_name]["function"], value)


class OutputOperation(_FEAOutputBase):
    registry = "outputs_dict"

    def compute(self, input_vals, output_vals):
        self._push(input_vals)
        output_vals[self.output_name] = assemble(self.fea_output["form"])

    def compute_derivatives(self, input_vals, output_vals, derivatives):
        self._push(input_vals)
        form = self.fea_output["form"]
        for arg_name in input_vals:
            partial = computePartials(form, self.args_dict[arg_name]["function"])
            derivatives[self.output_name, arg_name] = assemble(partial, dim=self.output_dim + 1)


class OutputFieldOperation(_FEAOutputBase):
    registry = "outputs_field_dict"
    output_dim = 1

    def _output_shape(self):
        return (self.fea_output["shape"],)

    def compute(self, input_vals, output_vals):
        self._push(input_vals)
        out = self.fea_output
        self.fea.projectFieldOutput(out["form"], out["function"])
        output_vals[self.output_name] = getFuncArray(out["function"]).copy()
        if out["record"] and out["recorder"] is not None:
            out["recorder"].write_function(out["function"], self.fea.opt_iter)

    def compute_derivatives(self, input_vals, output_vals, derivatives):
        pass                                   # the reference provides none (output_operation.py:72-128)
