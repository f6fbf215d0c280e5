"""Jacobian assembly on the reference's wing-box mesh (BASELINE config 2: tests/golden/wingbox_mesh.npz, every cell warped,
T-junctions) and its uniform subdivisions: cell-centric warped kernels against the round-1 contribution-gather kernel.
    python scripts/wingbox_asm_bench.py > gpurun_out/wingbox_asm.json"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from femo_alpha_b200.shell_device import ShellDevice
    from femo_alpha_b200.multigrid import refine_quads
    m = np.load(os.path.join(ROOT, "tests", "golden", "wingbox_mesh.npz"))
    mesh = (m["nodes"], m["quads"].astype(np.int32))
    out = {}
    for level in range(4):
        nodes, quads = mesh
        row = {}
        ref = None
        for tag, env in (("warped_kernels", None), ("round1_gather", "0")):
            if env is None:
                os.environ.pop("FEMO_WARPED", None)
            else:
                os.environ["FEMO_WARPED"] = env
            d = ShellDevice(nodes, quads, True, False)
            nc = quads.shape[0]
            h = 0.05 * 0.0254
            d.set_input("thickness", np.full(nc, h)); d.set_input("E", np.full(nc, 73.1e9)); d.set_input("nu", np.full(nc, 0.33))
            ts = []
            for i in range(8):
                d.set_input("thickness", np.full(nc, h * (1 + 1e-3 * (i % 2))))
                d.assemble_jacobian()
                t = d.last_timings()
                ts.append(t["records_ms"] + t["gather_ms"])
            v = np.random.default_rng(1).standard_normal(d.ndof)
            Kv = d.jacobian_apply(v)
            ref = Kv if ref is None else ref
            row[tag] = dict(ms=float(np.mean(ts[3:])), Kv_rel_diff=float(np.abs(Kv - ref).max() / np.abs(ref).max()))
            row["n_dof"], row["n_cells"], row["all_planar"] = int(d.ndof), int(nc), int(d.sizes.all_planar)
            d.close()
        row["speedup"] = row["round1_gather"]["ms"] / row["warped_kernels"]["ms"]
        out["x%d" % 4 ** level] = row
        print("wing box x%d: %d cells, %d dofs | %.3f ms vs %.3f ms (%.1fx)" % (4 ** level, row["n_cells"], row["n_dof"],
              row["warped_kernels"]["ms"], row["round1_gather"]["ms"], row["speedup"]), file=sys.stderr, flush=True)
        mesh = refine_quads(nodes, quads)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
