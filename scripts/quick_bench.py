"""Ad-hoc timing of the assembly / SpMV / PCG kernels on a structured plate (development aid)."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from femo_alpha_b200.shell_device import ShellDevice


def plate(nx, ny, lx, ly):
    xs = np.linspace(0.0, lx, nx + 1); ys = np.linspace(0.0, ly, ny + 1)
    X, Y = np.meshgrid(xs, ys, indexing="ij")
    nodes = np.stack([X.ravel(), Y.ravel(), np.zeros(X.size)], 1)
    idx = np.arange((nx + 1) * (ny + 1)).reshape(nx + 1, ny + 1)
    quads = np.stack([idx[:-1, :-1].ravel(), idx[1:, :-1].ravel(), idx[1:, 1:].ravel(), idx[:-1, 1:].ravel()], 1)
    return nodes, quads.astype(np.int32)


n = int(sys.argv[1]) if len(sys.argv) > 1 else 816
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 100
nodes, quads = plate(n, n, 10.0, 10.0)
t0 = time.time()
d = ShellDevice(nodes, quads)
print("setup s", time.time() - t0, "ndof", d.ndof, "nnzb", d.nnzb, "planar", d.sizes.all_planar, flush=True)
x, y = nodes[:, 0], nodes[:, 1]
d.set_input("thickness", 0.1 * (1 + 0.5 * np.sin(2 * np.pi * x / 10) * np.cos(2 * np.pi * y / 10)))
d.set_input("E", np.full(d.nn, 1e8)); d.set_input("nu", np.full(d.nn, 0.3)); d.set_input("density", np.full(d.nn, 10.0))
d.set_input("F_solid", np.tile([0, 0, 5.0], d.nn))
on = d.block_coords()[:, 0] < 3e-16
blocks = np.nonzero(on)[0]
d.set_strong_bc((3 * blocks[:, None] + np.arange(3)[None, :]).ravel())
print("peaks (fp64 TFLOP/s, copy GB/s):", d.measure_peaks(), flush=True)
for rep in range(4):
    d.set_input("thickness", d.get_input("thickness"))
    d.assemble_jacobian()
    t = d.last_timings()
    tot = t["records_ms"] + t["gather_ms"]
    print("assemble: records %.3f ms gather %.3f ms -> %.3f GDOF/s, %.1f GB/s (9.1KB/cell)" % (
        t["records_ms"], t["gather_ms"], d.ndof / tot / 1e6, d.nc * 9100 / tot / 1e6), flush=True)
ms = d.pcg_fixed(iters)
bytes_it = 12 * d.nnzb * 9 + 188 * d.ndof
print("pcg %d its: %.3f ms/iter, %.1f GB/s (12B/nnz+188B/dof)" % (iters, ms / iters, bytes_it / (ms / iters) / 1e6), flush=True)
print("launches", d.launch_count())
