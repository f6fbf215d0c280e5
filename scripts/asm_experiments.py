"""Timing experiments on the tiled assembly kernel at config 3 (816 x 816): FEMO_TILE_DEBUG masks switch parts of the
kernel off (results are then wrong; only the time matters) to tell which side bounds it.
    python scripts/asm_experiments.py [--n 816] [--masks 0,1,2,4,8,...]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=816)
    ap.add_argument("--masks", default="0,1,2,4,8,9,6,3,12,15")
    a = ap.parse_args()
    import torch
    from bench import plate, inputs_for
    from femo_alpha_b200.shell_device import ShellDevice
    nodes, quads = plate(a.n, a.n, 0.0, 10.0, 10.0)
    d = ShellDevice(nodes, quads)
    inp = inputs_for(nodes)
    for k in ("thickness", "E", "nu"):
        d.set_input(k, inp[k])
    h = [torch.from_numpy(inp["thickness"] * (1 + 1e-3 * k)).cuda() for k in range(2)]
    names = {1: "no stores", 2: "no tile arithmetic", 4: "no halo items", 8: "no gather/rotate"}
    for m in [int(x) for x in a.masks.split(",")]:
        os.environ["FEMO_TILE_DEBUG"] = str(m)
        ts = []
        for i in range(6):
            d.set_input("thickness", h[i % 2])
            d.assemble_jacobian()
            ts.append(d.last_timings()["gather_ms"])
        what = " + ".join(v for k, v in names.items() if m & k) or "full kernel"
        print("mask %2d  %-60s %.3f ms" % (m, what, float(np.mean(ts[2:]))), flush=True)
    os.environ["FEMO_TILE_DEBUG"] = "0"
    d.close()


if __name__ == "__main__":
    main()
