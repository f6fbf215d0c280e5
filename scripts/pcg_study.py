"""PCG (block-Jacobi) iteration counts vs mesh size on the cfg-3 plate; prints one JSON line per size."""
import json, sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import plate, inputs_for
from femo_alpha_b200.shell_device import ShellDevice

sizes = [int(a) for a in sys.argv[1:]] or [50, 100, 200, 400]
for n in sizes:
    nodes, quads = plate(n, n, 0.0, 10.0, 10.0)
    d = ShellDevice(nodes, quads)
    inp = inputs_for(nodes)
    for k in ("thickness", "E", "nu", "density", "F_solid"):
        d.set_input(k, inp[k])
    on = d.block_coords()[:, 0] < 3e-16
    blocks = np.nonzero(on)[0]
    d.set_strong_bc((3 * blocks[:, None] + np.arange(3)[None, :]).ravel())
    t0 = time.perf_counter()
    it, rr = d.solve_state(rtol=1e-10, max_it=int(os.environ.get("MAXIT", "150000")))
    dt = time.perf_counter() - t0
    w = d.get_state()
    print(json.dumps(dict(n=n, ndof=int(d.ndof), iters=it, relres=rr, solve_s=dt, ms_per_iter=d.last_timings()["iter_ms"],
                          tip=float(np.abs(w[:3 * d.nn]).max()))), flush=True)
    if n == sizes[0]:
        tf, gb = d.measure_peaks()
        print(json.dumps(dict(fp64_fma_tflops=tf, copy_gbs=gb, dmma_tflops=d.dmma_tflops())), flush=True)
    d.close()
