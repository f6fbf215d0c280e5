"""RMShellModel: the RM-shell entry point for aircraft thickness optimisation -- drop-in for
femo_alpha/rm_shell/rm_shell_model.py:14-544 (same constructor and `evaluate` signature, same
attribute names on the returned variable group).
"""
from __future__ import annotations

import numpy as np

from ..csdl_compat import csdl
from ..csdl_alpha_opt.fea_model import FEAModel
from ..fea.fea_dolfinx import FEA
from ..fea.utils_dolfinx import Function, createCustomMeasure, locate_dofs_geometrical, dirichletbc
from .rm_shell_pde import RMShellPDE


class RMShellModel:
    def __init__(self, mesh, shell_bc_func=None, element_wise_material=False, rho=100, PENALTY_BC=True,
                 additional_outputs=None, mesh_tags=None, record=True, elementwise_pressure=False,
                 association_table=None, quadrature_degree=6, device=0):
        self.mesh = mesh
        self.mesh_tags = mesh_tags
        self.additional_outputs = additional_outputs
        self.shell_bc_func = shell_bc_func
        self.element_wise_material = element_wise_material
        self.record = record
        self.m, self.rho = 1e-6, rho
        self.PENALTY_BC = PENALTY_BC
        self.quadrature_degree = quadrature_degree
        self._device_index = device
        self.nel = mesh.topology.index_map(mesh.topology.dim).size_local
        self.nn = mesh.topology.index_map(0).size_local
        self.elementwise_pressure = elementwise_pressure
        if mesh_tags is not None:
            self.set_up_subdomains(mesh_tags, association_table)
        if shell_bc_func is not None:
            self.set_up_bcs(shell_bc_func, PENALTY_BC)
        else:
            raise ValueError('Please provide the shell bc location function.\n \
                             Example:\n \
                             def ClampedBoundary(x):\n \
                                return np.less(x[1], 0.0)')
        self.set_up_fea()

    def set_up_bcs(self, bc_locs_func, PENALTY_BC):
        if PENALTY_BC:
            fdim = self.mesh.topology.dim - 1
            self.dss = createCustomMeasure(self.mesh, fdim, bc_locs_func, measure='ds', tag=100)(100)
            self.dSS = createCustomMeasure(self.mesh, fdim, bc_locs_func, measure='dS', tag=100)(100)
        else:
            self.dss = None
            self.dSS = None

    def set_up_subdomains(self, mesh_tags, association_table=None):
        """{tag: [cell ids]} (external cell numbering == ours); rm_shell_model.py:101-133."""
        if not isinstance(mesh_tags, dict):
            raise TypeError("mesh_tags must be a dictionary {tag: [cell indices]}")
        vals = -np.ones(self.nel, dtype=np.int32)
        for i, inds in enumerate(mesh_tags.values()):
            vals[np.asarray(inds, dtype=np.int64)] = i
        self.cell_tag_values = vals
        self.association_table = association_table or {key: i for i, key in enumerate(mesh_tags.keys())}

    def set_up_fea(self):
        print('-' * 40)
        print('Setting up the FEA model for RM shell analysis ...')
        mesh = self.mesh
        shell_pde = self.shell_pde = RMShellPDE(mesh, element_wise_material=self.element_wise_material,
                                                elementwise_pressure=self.elementwise_pressure,
                                                quadrature_degree=self.quadrature_degree, device=self._device_index)
        fea = FEA(mesh)
        fea.PDE_SOLVER = 'Newton'
        fea.REPORT = False
        fea.record = self.record
        fea.linear_problem = True
        h, f = Function(shell_pde.VT), Function(shell_pde.VF)
        E, nu, density = Function(shell_pde.VT), Function(shell_pde.VT), Function(shell_pde.VT)
        uhat = Function(shell_pde.VU)
        w = Function(shell_pde.W)
        if not self.PENALTY_BC:
            W = shell_pde.W
            locate_BC1 = locate_dofs_geometrical((W.sub(0), W.sub(0).collapse()[0]), self.shell_bc_func)
            locate_BC2 = locate_dofs_geometrical((W.sub(1), W.sub(1).collapse()[0]), self.shell_bc_func)
            ubc = Function(W)
            fea.bc = [dirichletbc(ubc, locate_BC1, W.sub(0)), dirichletbc(ubc, locate_BC2, W.sub(1))]
        g = Function(shell_pde.W)
        residual_form = shell_pde.pdeRes(h=h, w=w, uhat=uhat, f=f, E=E, nu=nu, penalty=self.PENALTY_BC,
                                         dss=self.dss, dSS=self.dSS, g=g)
        u_mid, theta = w.sub(0), w.sub(1)
        compliance_form = shell_pde.compliance(u_mid, uhat, h, f)
        mass_form = shell_pde.mass(uhat, h, density)
        elastic_energy_form = shell_pde.elastic_energy(w, uhat, h, E)
        pnorm_stress_form = shell_pde.pnorm_stress(w, uhat, h, E, nu, None, m=self.m, rho=self.rho)
        stress_form = shell_pde.von_Mises_stress(w, uhat, h, E, nu, surface='Top')
        shell_pde.device.set_stress_params(self.m, self.rho)

        fea.add_input('thickness', h, init_val=0.001, record=self.record)
        fea.add_input('F_solid', f, init_val=1., record=self.record)
        fea.add_input('E', E, init_val=1., record=self.record)
        fea.add_input('nu', nu, init_val=1., record=self.record)
        fea.add_input('density', density, init_val=1., record=self.record)
        fea.add_input('uhat', uhat, init_val=0., record=self.record)
        fea.add_state(name='disp_solid', function=w, residual_form=residual_form,
                      arguments=['thickness', 'F_solid', 'E', 'nu', 'uhat'])
        fea.add_output(name='compliance', form=compliance_form, arguments=['disp_solid', 'F_solid', 'thickness', 'uhat'])
        fea.add_output(name='mass', form=mass_form, arguments=['thickness', 'density', 'uhat'])
        fea.add_output(name='elastic_energy', form=elastic_energy_form, arguments=['thickness', 'disp_solid', 'E', 'uhat'])
        fea.add_output(name='pnorm_stress', form=pnorm_stress_form,
                       arguments=['thickness', 'disp_solid', 'E', 'nu', 'uhat'])
        fea.add_field_output(name='stress', form=stress_form, arguments=['thickness', 'disp_solid', 'E', 'nu', 'uhat'],
                             function_space=('DG', 1), record=self.record, vtk=True)
        if self.mesh_tags is not None:
            shell_pde.device.set_cell_tags(self.cell_tag_values)
            for tag, i in self.association_table.items():
                if i == -1:
                    continue
                from ..fea.utils_dolfinx import OutputForm
                form_i = OutputForm(shell_pde.device, "pnorm_stress")
                form_i.tag = int(i)
                fea.add_output(name='pnorm_stress_' + str(tag), form=form_i,
                               arguments=['thickness', 'disp_solid', 'E', 'nu', 'uhat'])
        self.fea = fea

    def evaluate(self, force_vector, thickness, E, nu, density, node_disp=None, debug_mode=False, is_pressure=True):
        shell_inputs = csdl.VariableGroup()
        mesh = self.shell_pde.mesh
        if self.element_wise_material:
            material_mesh_indices = mesh.topology.original_cell_index.tolist()
        else:
            material_mesh_indices = mesh.geometry.input_global_indices.tolist()
        shell_inputs.thickness = thickness[material_mesh_indices]
        shell_inputs.E = E[material_mesh_indices]
        shell_inputs.nu = nu[material_mesh_indices]
        shell_inputs.density = density[material_mesh_indices]
        if self.elementwise_pressure:
            pressure_mesh_indices = mesh.topology.original_cell_index.tolist()
        else:
            pressure_mesh_indices = mesh.geometry.input_global_indices.tolist()
        reshaped_force = csdl.reshape(force_vector[pressure_mesh_indices], (-1,))
        if is_pressure:
            shell_inputs.F_solid = reshaped_force
        else:
            print('Converting forces to pressures ...')
            A = self.shell_pde.construct_force_to_pressure_map()
            shell_inputs.F_solid = csdl.solve_linear(A.toarray(), reshaped_force)   # dense, as the reference (B7)
        shell_inputs.F_solid.add_name('F_solid')
        deformation_mesh_indices = mesh.geometry.input_global_indices.tolist()
        if node_disp is None:
            node_disp = csdl.Variable(value=0.0, shape=(len(deformation_mesh_indices), 3), name='node_disp')
        reshaped_node_disp = node_disp[deformation_mesh_indices].reshape((-1,))
        reshaped_node_disp.add_name('uhat')
        shell_inputs.uhat = reshaped_node_disp

        print('Evaluating the RM shell model ...')
        solid_model = FEAModel(fea=[self.fea], fea_name='rm_shell')
        shell_outputs = solid_model.evaluate(shell_inputs, debug_mode=debug_mode)

        disp_extracted = DisplacementExtractionModel(shell_pde=self.shell_pde).evaluate(shell_outputs.disp_solid)
        disp_extracted.add_name('disp_extracted')
        shell_outputs.disp_extracted = disp_extracted
        aggregated_stress = AggregatedStressModel(m=self.m, rho=self.rho).evaluate(shell_outputs.pnorm_stress)
        aggregated_stress.add_name('aggregated_stress')
        shell_outputs.aggregated_stress = aggregated_stress
        print('RM shell model evaluation completed.')
        print('-' * 40)
        return shell_outputs


class AggregatedStressModel:
    """aggregated_stress = 1/m * pnorm^(1/rho)  (rm_shell_model.py:493-503)"""

    def __init__(self, m: float, rho: int):
        self.m = m
        self.rho = rho

    def evaluate(self, pnorm_stress):
        return 1 / self.m * pnorm_stress ** (1 / self.rho)


class DisplacementExtractionModel:
    """(nn,3) vertex displacements in external node order (rm_shell_model.py:505-527)"""

    def __init__(self, shell_pde: RMShellPDE):
        self.shell_pde = shell_pde

    def evaluate(self, disp_vec):
        shell_pde = self.shell_pde
        disp_extraction_mats = shell_pde.construct_nodal_disp_map()
        shape = shell_pde.mesh.geometry.x.shape
        nodal_disp_vec = csdl.sparse.matvec(disp_extraction_mats, disp_vec)
        nodal_disp_mat = csdl.transpose(csdl.reshape(nodal_disp_vec, shape=(shape[1], shape[0])))
        reverse = np.argsort(shell_pde.mesh.geometry.input_global_indices).tolist()
        return nodal_disp_mat[reverse, :]


class ForceReshapingModel:
    """(nn,3) nodal force matrix -> flat load vector in FEA order (rm_shell_model.py:529-544)"""

    def __init__(self, shell_pde: RMShellPDE):
        self.shell_pde = shell_pde

    def evaluate(self, nodal_force_mat):
        idx = self.shell_pde.mesh.geometry.input_global_indices.tolist()
        return csdl.reshape(nodal_force_mat[idx, :], shape=(self.shell_pde.VF.size,))
