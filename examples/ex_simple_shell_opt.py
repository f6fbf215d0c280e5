"""Clamped plate under uniform pressure: compliance + adjoint thickness gradient (BASELINE config 1).

The flow of the reference's examples/advanced_examples/simple_shell_opt/ex_simple_shell_opt.py with
`femo_alpha_b200` in place of `femo_alpha` (run on a B200: `python examples/ex_simple_shell_opt.py`).
The plate mesh is generated (createRectangleMesh) instead of read from plate_meshes/*.xdmf, which are git-LFS
pointers in the reference checkout.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from femo_alpha_b200.csdl_compat import csdl                                    # csdl_alpha if installed
from femo_alpha_b200.mesh import createRectangleMesh
from femo_alpha_b200.rm_shell.rm_shell_model import RMShellModel

DOLFIN_EPS = 3E-16
E_val, nu_val, h_val, rho_val, f_d = 1e8, 0.3, 0.1, 10.0, 5.0
width, length = 2.0, 10.0


def ClampedBoundary(x):
    return np.less(x[0], 0.0 + DOLFIN_EPS)


mesh = createRectangleMesh((0.0, 0.0), (length, width), 20, 4)
nn = mesh.topology.index_map(0).size_local

recorder = csdl.Recorder(inline=True)
recorder.start()
pressure = csdl.Variable(value=np.zeros((nn, 3)), name='force_vector')
pressure.value[:, 2] = f_d
node_disp = csdl.Variable(value=np.zeros((nn, 3)), name='node_disp')
thickness = csdl.Variable(value=h_val * np.ones(nn), name='thickness')
E = csdl.Variable(value=E_val * np.ones(nn), name='E')
nu = csdl.Variable(value=nu_val * np.ones(nn), name='nu')
density = csdl.Variable(value=rho_val * np.ones(nn), name='density')

shell_model = RMShellModel(mesh, shell_bc_func=ClampedBoundary, record=False)
out = shell_model.evaluate(pressure, thickness, E, nu, density, node_disp, debug_mode=False, is_pressure=True)
recorder.stop()

Ix = width * h_val ** 3 / 12
print("Euler-Bernoulli beam theory deflection:", f_d * width * length ** 4 / (8 * E_val * Ix))
print("Reissner-Mindlin FE deflection:", np.abs(out.disp_extracted.value).max())
print("Compliance:", out.compliance.value, " Mass:", out.mass.value, " Aggregated stress:", out.aggregated_stress.value)

sim = csdl.experimental.PySimulator(recorder)
rep = sim.check_totals([out.compliance], [thickness], step_size=2e-5)
