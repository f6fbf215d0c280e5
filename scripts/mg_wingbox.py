"""Two-level V-cycle PCG on the refined wing-box mesh (config 2 x4): iterations against the
Chebyshev degree, and the block-Jacobi PCG baseline.  Usage: python scripts/mg_wingbox.py"""
import os, sys, time, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from femo_alpha_b200.mesh import ShellMesh, refine_uniform
from femo_alpha_b200.rm_shell.rm_shell_pde import RMShellPDE
from femo_alpha_b200.multigrid import ShellHierarchy

m = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "wingbox_mesh.npz"))
out = {}
for levels_extra in (1, 2):
    fine = refine_uniform(ShellMesh(m["nodes"], m["quads"]), levels_extra)
    for deg in (3, 5, 8):
        os.environ["FEMO_MG"] = "0"
        pde = RMShellPDE(fine, element_wise_material=True)
        d = pde.device
        hier = ShellHierarchy.for_device(d, fine.mg_meshes, cheb_degree=deg, element_wise_material=True)
        h = 0.05 * 0.0254
        for k, v in (("thickness", h), ("E", 73.1e9), ("nu", 0.33), ("density", 2780.0)):
            d.set_input(k, np.full(d.nc, v))
        d.set_input("F_solid", np.tile([0.0, 0.0, -9.81 * 2780.0 * h], d.nn))
        blocks = np.nonzero(d.block_coords()[:, 1] <= 1e-3)[0]
        d.set_strong_bc(np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()))
        t0 = time.time(); it, rr = d.solve_state(rtol=1e-10); t1 = time.time()
        key = "x%d_cheb%d" % (4 ** levels_extra, deg)
        out[key] = {"ndof": int(d.ndof), "iters": int(it), "relres": float(rr), "wall_s": t1 - t0}
        print(key, out[key], flush=True)
        if deg == 3:
            d._mg = None
            t0 = time.time(); itj, rrj = d.solve_state(rtol=1e-10, max_it=400000); t1 = time.time()
            out["x%d_jacobi" % (4 ** levels_extra)] = {"iters": int(itj), "relres": float(rrj), "wall_s": t1 - t0}
            print("jacobi", itj, rrj, t1 - t0, flush=True)
            d._mg = hier
        hier.close()
json.dump(out, open("gpurun_out/mg_wingbox.json", "w"), indent=1)
