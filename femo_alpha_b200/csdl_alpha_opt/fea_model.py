"""FEAModel: one StateOperation per state, one OutputOperation per output, one
OutputFieldOperation per field output -- femo_alpha/csdl_alpha_opt/fea_model.py:6-64."""
from .state_operation import StateOperation
from .output_operation import OutputOperation, OutputFieldOperation


class FEAModel():
    def __init__(self, fea, fea_name='fea'):
        self.parameters = {'fea': fea, 'fea_name': fea_name}

    def evaluate(self, inputs, debug_mode=False):
        self.fea_list = self.parameters['fea']
        variables = inputs
        for fea in self.fea_list:
            for state_name, st in fea.states_dict.items():
                op = StateOperation(fea=fea, state_name=state_name, args_name_list=st['arguments'], debug_mode=debug_mode)
                setattr(variables, state_name, op.evaluate(variables))
            for output_name, out in fea.outputs_dict.items():
                op = OutputOperation(fea=fea, output_name=output_name, args_name_list=out['arguments'])
                setattr(variables, output_name, op.evaluate(variables))
            for output_name, out in fea.outputs_field_dict.items():
                op = OutputFieldOperation(fea=fea, output_name=output_name, args_name_list=out['arguments'])
                setattr(variables, output_name, op.evaluate(variables))
        return variables
