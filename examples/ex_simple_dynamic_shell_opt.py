"""Time-stepped cantilever plate under a 1-cos gust: total strain energy + time-adjoint thickness gradient
(BASELINE config 4).

The flow of the reference's examples/advanced_examples/simple_dynamic_shell_opt/ex_simple_dynamic_shell_opt.py with
`femo_alpha_b200` in place of `femo_alpha` (run on a B200).  The mesh is generated instead of read from the git-LFS
pointer files of the reference checkout.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from femo_alpha_b200.csdl_compat import csdl
from femo_alpha_b200.mesh import createRectangleMesh
from femo_alpha_b200.dynamic_rm_shell.plate_sim import PlateSim
from femo_alpha_b200.dynamic_rm_shell.state_operation_dynamic import StateOperation
from femo_alpha_b200.dynamic_rm_shell.total_strain_energy_operation import TotalStrainEnergyOperation
from femo_alpha_b200.dynamic_rm_shell.volume_operation import VolumeOperation

mesh = createRectangleMesh((0.0, 0.0), (10.0, 2.0), 10, 2)           # plate_2_10_quad_2_10
nn = mesh.topology.index_map(0).size_local
E_val, nu_val, h_val, rho_val = 1e8, 0.3, 0.1, 10.0
Nsteps = int(os.environ.get("NSTEPS", "40"))

# gust loads (1-cosine)
V_inf, V_p, l_chord, GGLc, T0 = 50., 50., 1.2, 5, 0.02
T1 = GGLc * l_chord / V_inf
T = T0 + T1 + 2.86
dt = T / Nsteps


def V_g(t):
    return V_p * (1 - np.cos(2 * np.pi * (t - T0) / T1)) if T0 <= t <= T0 + T1 else 0.


force_history = np.zeros((Nsteps + 1, nn * 3))
for i, ti in enumerate(np.linspace(0, T, Nsteps + 1)):
    force_history[i, 2::3] = V_g(ti) * 0.1

plate_sim = PlateSim(mesh, E_val, nu_val, rho_val, dt, Nsteps)

recorder = csdl.Recorder(inline=True)
recorder.start()
force_history = csdl.Variable(value=force_history, name='force_history')
thickness = csdl.Variable(value=h_val * np.ones(nn), name='thickness')
input_vars = csdl.VariableGroup()
input_vars.thickness = thickness
input_vars.force_history = force_history
disp_history = StateOperation(plate_sim=plate_sim, gradient_mode='petsc', record=False).evaluate(input_vars)
disp_history.add_name('disp_history')
input_vars.disp_history = disp_history
total_strain_energy = TotalStrainEnergyOperation(plate_sim=plate_sim).evaluate(input_vars)
total_strain_energy.add_name('total_strain_energy')
volume = VolumeOperation(plate_sim=plate_sim).evaluate(input_vars)
volume.add_name('volume')
recorder.stop()

print("total strain energy:", total_strain_energy.value, " mass:", volume.value * rho_val)
print("max |tip displacement|:", np.abs(plate_sim.tip_disp_history).max(), " solver:", plate_sim._dyn.info())
sim = csdl.experimental.PySimulator(recorder)
sim.check_totals([total_strain_energy], [thickness], step_size=2e-5)
