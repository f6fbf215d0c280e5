"""State + adjoint solve at BASELINE config 3 scale with the multigrid preconditioner (development / profiles)."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from femo_alpha_b200.multigrid import ShellHierarchy

base = int(sys.argv[1]) if len(sys.argv) > 1 else 51
levels = int(sys.argv[2]) if len(sys.argv) > 2 else 5
deg = int(sys.argv[3]) if len(sys.argv) > 3 else 3
rtol = float(sys.argv[4]) if len(sys.argv) > 4 else 1e-10
xs = np.linspace(0.0, 10.0, base + 1)
X, Y = np.meshgrid(xs, xs, indexing="ij")
nodes = np.stack([X.ravel(), Y.ravel(), np.zeros(X.size)], 1)
idx = np.arange((base + 1) ** 2).reshape(base + 1, base + 1)
quads = np.stack([idx[:-1, :-1].ravel(), idx[1:, :-1].ravel(), idx[1:, 1:].ravel(), idx[:-1, 1:].ravel()], 1).astype(np.int32)
t0 = time.time()
h = ShellHierarchy(nodes, quads, levels, cheb_degree=deg)
t_build = time.time() - t0
h.set_input("thickness", lambda x: 0.1 * (1 + 0.5 * np.sin(2 * np.pi * x[0] / 10) * np.cos(2 * np.pi * x[1] / 10)))
h.set_input("E", 1e8); h.set_input("nu", 0.3); h.set_input("density", 10.0)
h.set_input("F_solid", lambda x: np.tile([0.0, 0.0, 5.0], (x.shape[1], 1)))
h.set_strong_bc(lambda x: x[0] < 3e-16)
t0 = time.time(); h.setup(); t_setup = time.time() - t0
t0 = time.time(); it, rr = h.solve_state(rtol=rtol); t_state = time.time() - t0
tm = h.fine.last_timings()
u = h.fine.get_state()
n = h.meshes[0][0].shape[0]
tip = float(np.abs(u.reshape(-1, 3)[:n, 2]).max())
dC = h.fine.output_grad("compliance", "disp_solid") if hasattr(h.fine, "output_grad") else None
out = dict(base=base, levels=levels, cheb_degree=deg, rtol=rtol, n_dof=int(h.fine.ndof), t_hierarchy_s=t_build, t_setup_s=t_setup,
           state_iters=int(it), state_relres=float(rr), t_state_s=t_state, state_ms_device=tm.get("solve_ms"),
           lambda_max=[h.lambda_max(l) for l in range(levels - 1)], max_deflection=tip)
if dC is not None:
    t0 = time.time(); lam = h.solve_adjoint(dC, rtol=rtol); t_adj = time.time() - t0
    out.update(adjoint_iters=int(h.last_solve["iters"]), adjoint_relres=float(h.last_solve["relres"]), t_adjoint_s=t_adj)
print(json.dumps(out))
