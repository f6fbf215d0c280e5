"""Worker of tests/test_gpu_hex.py::test_gradient_records: with FEMO_HEX_GRAD=1 in the environment the solid contexts
assemble from gradient records (k_hex_geom2 + k_hex_assemble_g); matrices against oracle/hex_ref.py on a distorted box,
on the ten-cell fan (two chunks of sub-passes) and, for the timing line, on a 48^3 cube against the default kernel."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    assert os.environ.get("FEMO_HEX_GRAD") == "1"
    from oracle.hex_ref import HexOracle, box_mesh
    from femo_alpha_b200.solid_device import SolidDevice, box_hexes
    from tests.test_hex_host import fan_mesh
    worst = 0.0
    for form in ("elasticity", "vector_poisson"):
        for nodes, cells in (box_mesh(3, 4, 2, distort=0.25, seed=4), fan_mesh(10)):
            o = HexOracle(nodes, cells, form=form, E0=2.0, nu=0.2)
            o.set_inputs(density=np.linspace(0.2, 1.0, o.nc))
            d = SolidDevice(nodes, cells, form=form, E0=2.0, nu=0.2)
            d.set_input("density", o.rho)
            K = o.stiffness()
            worst = max(worst, abs(d.get_csr() - K).max() / abs(K).max())
            d.close()
    # same matrix as the default kernel at a size with full interior rows, and the two assembly times
    nodes, cells = box_hexes(48, 48, 48)
    rho = np.random.default_rng(1).uniform(0.01, 1.0, len(cells))
    vals, ms = {}, {}
    for flag in ("1", "0"):
        os.environ["FEMO_HEX_GRAD"] = flag
        d = SolidDevice(nodes, cells)
        d.set_input("density", rho)
        d.assemble_jacobian(); d.set_input("density", rho); d.assemble_jacobian()
        ms[flag] = d.last_timings()["assemble_ms"]
        vals[flag] = d.get_bsr()[2]
        d.close()
    diff = np.abs(vals["1"] - vals["0"]).max() / np.abs(vals["0"]).max()
    print("gradient records: worst matrix rel.err %.2e | 48^3: rel. difference to the default kernel %.2e | assembly %.3f ms "
          "(records) vs %.3f ms (default)" % (worst, diff, ms["1"], ms["0"]), flush=True)
    sys.exit(0 if worst < 1e-12 and diff < 1e-12 else 1)


if __name__ == "__main__":
    main()
