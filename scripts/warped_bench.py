"""Warped-cell Jacobian assembly at config-3 size (doubly curved plate, 816 x 816 cells): the cell-centric kernels
(k_warped_elem<NB> + k_warped_gather, NB = 1, 2, 3 node pairs per task) against the round-1 contribution-gather kernel
(FEMO_WARPED=0).  Times are CUDA events inside the library (records kernel / assembly kernels).
    python scripts/warped_bench.py [n] > gpurun_out/warped_bench.json"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import plate, inputs_for, CELL_BYTES, measured_peaks   # noqa: E402


def main():
    import torch
    from femo_alpha_b200.shell_device import ShellDevice
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 816
    peak, _ = measured_peaks()
    nodes, quads = plate(n, n, 0.0, 10.0, 10.0)
    nodes[:, 2] = 0.05 * np.sin(np.pi * nodes[:, 0] / 10.0) * np.sin(np.pi * nodes[:, 1] / 10.0)
    inp = inputs_for(nodes)
    out = {"workload": "doubly curved plate %dx%d, Jacobian assembly" % (n, n)}
    ref = None
    for tag, env in (("nb1", {"FEMO_WARPED_NB": "1"}), ("nb2", {"FEMO_WARPED_NB": "2"}), ("nb3", {"FEMO_WARPED_NB": "3"}),
                     ("legacy", {"FEMO_WARPED": "0"})):
        for k in ("FEMO_WARPED", "FEMO_WARPED_NB"):
            os.environ.pop(k, None)
        os.environ.update(env)
        d = ShellDevice(nodes, quads)
        for k in ("thickness", "E", "nu"):
            d.set_input(k, inp[k])
        h_dev = [torch.from_numpy(inp["thickness"] * (1.0 + 1e-3 * k)).cuda() for k in range(2)]
        rec, ker = [], []
        for i in range(8):
            d.set_input("thickness", h_dev[i % 2])
            d.assemble_jacobian()
            t = d.last_timings()
            if i >= 3:
                rec.append(t["records_ms"]); ker.append(t["gather_ms"])
        rng = np.random.default_rng(3)
        v = rng.standard_normal(d.ndof)
        Kv = d.jacobian_apply(v)
        if ref is None:
            ref = Kv
        t_ms = float(np.mean(rec) + np.mean(ker))
        out[tag] = dict(n_dof=int(d.ndof), records_ms=float(np.mean(rec)), kernels_ms=float(np.mean(ker)), ms=t_ms,
                        gdof_s=d.ndof / (t_ms * 1e-3) / 1e9, roofline_frac=CELL_BYTES * d.nc / (t_ms * 1e-3) / 1e9 / peak,
                        Kv_rel_diff_to_nb1=float(np.abs(Kv - ref).max() / np.abs(ref).max()))
        print(tag, out[tag], file=sys.stderr, flush=True)
        d.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
