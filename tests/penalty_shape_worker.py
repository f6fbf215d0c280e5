"""Worker of tests/test_gpu_shape.py::test_penalty_shape: shape partial dR/duhat of a penalty-BC context (cell terms +
the facet terms carrying Nanson's factor, csrc/penalty_shape.cu) against the complex-step oracles."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    assert os.environ.get("FEMO_PENALTY_SHAPE", "1") != "0"      # the facet term is on by default
    from oracle import shape_ref as S
    from tests.problems import make_problem, oracle_for, device_for
    worst = 0.0
    for name in ("plate_4x20", "warped", "skew_planar"):
        p = make_problem(name)
        rng = np.random.default_rng(2)
        if p["uhat"] is None:
            p["uhat"] = 0.05 * rng.standard_normal(3 * p["nodes"].shape[0])
        o, d = oracle_for(p, penalty=True), device_for(p, penalty=True)
        o.set_inputs(uhat=p["uhat"]); d.set_input("uhat", p["uhat"])
        n = o.topo.ndof
        w, lam, du = rng.standard_normal(n), rng.standard_normal(n), rng.standard_normal(3 * o.topo.nn)
        d.set_state(w)
        pen_T, pen_F = S.dRpen_duhat_T_apply(o, w, lam), S.dRpen_duhat_apply(o, w, du)
        ref_T, ref_F = S.dRduhat_T_apply(o, w, lam) + pen_T, S.dRduhat_apply(o, w, du) + pen_F
        got_T, got_F = d.dRdf_T_apply("uhat", lam), d.dRdf_apply("uhat", du)
        e_T = np.abs(got_T - ref_T).max() / np.abs(ref_T).max()
        e_F = np.abs(got_F - ref_F).max() / np.abs(ref_F).max()
        print("%s: (dR/duhat)^T lam rel.err %.2e | (dR/duhat) duhat rel.err %.2e | facet share %.1e" %
              (name, e_T, e_F, np.abs(pen_T).max() / np.abs(ref_T).max()), flush=True)
        worst = max(worst, e_T, e_F)
        d.close()
    sys.exit(0 if worst < 1e-10 else 1)


if __name__ == "__main__":
    main()
