"""Cantilevered block with a cell-wise SIMP density: compliance, volume and their density gradients through the
generic `FEA` / `FEAModel` path with a named 3-D form (BASELINE config 5 at a small size).

The reference's generic FEA takes a UFL form here (femo_alpha/fea/fea_dolfinx.py:71-116); with `femo_alpha_b200` the
form is picked from the registry of `fea/named_forms.py` and its kernels run in libfemo_b200
(run on a B200: `python examples/ex_solid_simp_compliance.py [n]`).
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from femo_alpha_b200.csdl_compat import csdl                                    # csdl_alpha if installed
from femo_alpha_b200.csdl_alpha_opt.fea_model import FEAModel
from femo_alpha_b200.fea.named_forms import build_fea, createBoxMesh
from femo_alpha_b200.fea.utils_dolfinx import locate_dofs_geometrical

DOLFIN_EPS = 3E-16
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
mesh = createBoxMesh((0.0, 0.0, 0.0), (2.0, 1.0, 1.0), 2 * n, n, n)
fea, fn = build_fea("elasticity3d", mesh, E0=1.0, nu=0.3, simp_p=3.0)
fea.backend_options.update(rtol=1e-12)
fea.add_strong_bc(None, [locate_dofs_geometrical(fn["u"].function_space, lambda x: np.less(x[0], DOLFIN_EPS))])

recorder = csdl.Recorder(inline=True)
recorder.start()
inputs = csdl.VariableGroup()
rng = np.random.default_rng(20261104)
f = np.zeros((fea.nn, 3)); f[:, 2] = -1.0
inputs.density = csdl.Variable(value=rng.uniform(0.3, 1.0, fea.nel), name="density")
inputs.body_force = csdl.Variable(value=f.ravel(), name="body_force")
out = FEAModel(fea=[fea]).evaluate(inputs)
recorder.stop()

u = out.u.value.reshape(-1, 3)
print("cells %d, dofs %d, PCG iterations %d" % (fea.nel, 3 * fea.nn, fea.solve_log[-1]["iters"]))
print("compliance  %.10e" % np.ravel(out.compliance.value)[0])
print("volume      %.10e" % np.ravel(out.volume.value)[0])
print("max |u_z|   %.10e" % np.abs(u[:, 2]).max())

sim = csdl.experimental.PySimulator(recorder)
rep = sim.check_totals(ofs=[out.compliance, out.volume], wrts=[inputs.density], step_size=2e-5, verbose=False)
print("d compliance / d density: relative error against finite differences %.2e" % rep[out.compliance, inputs.density]["relative_error"])
print("d volume / d density:     relative error against finite differences %.2e" % rep[out.volume, inputs.density]["relative_error"])
fea.device.close()
