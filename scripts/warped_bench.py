"""Warped-cell Jacobian assembly at config-3 size (doubly curved plate, 816 x 816 cells): the cell-centric kernels
(k_warped_elem<NB> + k_warped_gather, NB = 1 .. 4 displacement-node pairs per task) against the round-1 contribution-gather kernel
(FEMO_WARPED=0).  Times are CUDA events inside the library (element-matrix kernel / gather kernel; records + gather for the legacy path).
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
    only = sys.argv[2] if len(sys.argv) > 2 else None          # e.g. 'default': one variant (ncu captures)
    for tag, env in (("default", {}), ("nb2", {"FEMO_WARPED_NB": "2"}), ("nb3_t4", {"FEMO_WARPED_NB": "3", "FEMO_WARPED_NBT": "4"}),
                     ("nb3_t1", {"FEMO_WARPED_NB": "3", "FEMO_WARPED_NBT": "1"}), ("nb4", {"FEMO_WARPED_NB": "4"}),
                     ("legacy", {"FEMO_WARPED": "0"})):
        if only and tag != only:
            continue
        for k in ("FEMO_WARPED", "FEMO_WARPED_NB", "FEMO_WARPED_NBT", "FEMO_WARPED_ORDER"):
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
        out[tag] = dict(n_dof=int(d.ndof), first_kernel_ms=float(np.mean(rec)), second_kernel_ms=float(np.mean(ker)), ms=t_ms,
                        gdof_s=d.ndof / (t_ms * 1e-3) / 1e9, roofline_frac=CELL_BYTES * d.nc / (t_ms * 1e-3) / 1e9 / peak,
                        Kv_rel_diff_to_first=float(np.abs(Kv - ref).max() / np.abs(ref).max()))
        print(tag, out[tag], file=sys.stderr, flush=True)
        d.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
