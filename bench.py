#!/usr/bin/env python
"""bench.py -- RM-shell Jacobian assembly GDOF/s (+ state/adjoint Krylov s/iter) on B200.

Workload (BASELINE.json configs[2], SURVEY 8d cfg 3): synthetic refined RM-shell plate,
816x816 Q2xQ1 quads on [0,10]^2 -> 665 856 cells, 10 002 534 DOF, ~737 M nnz, thickness
h(x,y) = 0.1 (1 + 0.5 sin(2 pi x/10) cos(2 pi y/10)), E = 1e8, nu = 0.3, f = (0,0,5), clamp x = 0.

A "step" = one pass of the assembly hot path: recompute the BSR(3x3) values of dR/du from the
inputs resident in HBM (quadrature records + tiled deterministic assembly kernel).  `value` = n_dof / step time.
`e2e` = the same through the plugin API with HOST buffers (H2D of thickness, F_solid, E, nu, uhat;
assemble; D2H of the residual vector).  `solve` = PCG (block-Jacobi) iteration time on A = dR/du.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--n 816]
N > 1: launched under torch.distributed.run; every rank owns one x-strip of the plate with the
same number of cells (weak scaling); assembly is owner-computes and needs no collective.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CELL_BYTES = 9100.0          # algorithmic bytes per Q2xQ1 cell of the Jacobian assembly


def plate(nx, ny, x0, lx, ly):
    xs = x0 + np.linspace(0.0, lx, nx + 1)
    ys = np.linspace(0.0, ly, ny + 1)
    X, Y = np.meshgrid(xs, ys, indexing="ij")
    nodes = np.stack([X.ravel(), Y.ravel(), np.zeros(X.size)], 1)
    idx = np.arange((nx + 1) * (ny + 1), dtype=np.int64).reshape(nx + 1, ny + 1)
    quads = np.stack([idx[:-1, :-1].ravel(), idx[1:, :-1].ravel(), idx[1:, 1:].ravel(), idx[:-1, 1:].ravel()], 1)
    return nodes, quads.astype(np.int32)


def inputs_for(nodes):
    x, y = nodes[:, 0], nodes[:, 1]
    nn = nodes.shape[0]
    return dict(thickness=0.1 * (1 + 0.5 * np.sin(2 * np.pi * x / 10) * np.cos(2 * np.pi * y / 10)),
                E=np.full(nn, 1e8), nu=np.full(nn, 0.3), density=np.full(nn, 10.0),
                F_solid=np.tile([0.0, 0.0, 5.0], nn), uhat=np.zeros(3 * nn))


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def cpu_baseline(n, budget_s=12.0, rows=None):
    """The CPU restatement (oracle/rm_shell_c.c, C + OpenMP, all host cores) timed on a bounded
    sample of the same workload: an n x rows strip of the n x n plate.  Checker / baseline only."""
    from oracle import c_port
    from oracle.rm_shell_ref import ShellTopology
    c_port.use_all_cores()
    rows = rows or max(8, min(n, 24))
    nodes, quads = plate(n, rows, 0.0, 10.0, 10.0 * rows / n)
    topo = ShellTopology(nodes, quads)
    inp = inputs_for(nodes)
    asm = c_port.CsrAssembler(topo)                      # pattern + element->CSR map: not timed (as on the GPU)
    threads = asm.assemble(inp["thickness"], inp["E"], inp["nu"])   # warm-up
    times = []
    t_end = time.time() + budget_s
    while time.time() < t_end or len(times) < 2:
        t0 = time.perf_counter()
        asm.assemble(inp["thickness"], inp["E"], inp["nu"])
        times.append(time.perf_counter() - t0)
        if len(times) >= 50:
            break
    t = float(np.median(times))
    return dict(value=topo.ndof / t / 1e9, unit="GDOF/s", cores=int(threads), kind="port",
                sample="%dx%d-cell strip of the %dx%d plate (%d cells, %d DOF), %d timed assemblies, median %.3f s"
                       % (n, rows, n, n, topo.nc, topo.ndof, len(times), t)), t, topo.ndof


def run_reference(args, rank):
    """--impl reference: the reference's CPU implementation of the path.  dolfinx/PETSc cannot be
    installed here (no network), so this arm times the oracle's C/OpenMP port with all
    host cores; each step = one Jacobian assembly of the bounded sample."""
    if rank != 0:
        return
    from oracle import c_port
    from oracle.rm_shell_ref import ShellTopology
    c_port.use_all_cores()           # torch.distributed.run exports OMP_NUM_THREADS=1
    n = args.n
    rows = max(8, min(n, 24))
    nodes, quads = plate(n, rows, 0.0, 10.0, 10.0 * rows / n)
    topo = ShellTopology(nodes, quads)
    inp = inputs_for(nodes)
    asm = c_port.CsrAssembler(topo)
    threads = 1
    for _ in range(max(args.warmup, 1)):
        threads = asm.assemble(inp["thickness"], inp["E"], inp["nu"])
    t0 = time.perf_counter()
    for _ in range(args.steps):
        asm.assemble(inp["thickness"], inp["E"], inp["nu"])
    dt = (time.perf_counter() - t0) / args.steps
    val = topo.ndof / dt / 1e9
    sample = "%dx%d-cell strip of the %dx%d plate (%d DOF) per step" % (n, rows, n, n, topo.ndof)
    print(json.dumps({
        "impl": "reference", "metric": "rm_shell_jacobian_assembly_throughput", "value": val, "unit": "GDOF/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "rm_shell_plate_%dx%d_q2q1_cfg3" % (n, n), "sample": sample},
        "cpu_baseline": {"value": val, "unit": "GDOF/s", "cores": int(threads), "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "GDOF/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "dolfinx 0.5.1/PETSc/MUMPS are not installable here; this is the oracle's C/OpenMP restatement"}))


def mg_solve_metric(n, device):
    """State + adjoint solve to relres 1e-10 on the same n x n plate, built as a nested hierarchy (coarse plate
    subdivided uniformly) so that the V-cycle preconditioner of csrc/mg.cu applies (SURVEY 8d metric 2)."""
    from femo_alpha_b200.multigrid import ShellHierarchy
    levels = 1
    while n % (2 ** levels) == 0 and n // (2 ** levels) >= 8:
        levels += 1
    base = n // 2 ** (levels - 1)
    if levels < 2:
        return None
    nodes, quads = plate(base, base, 0.0, 10.0, 10.0)
    t0 = time.perf_counter()
    h = ShellHierarchy(nodes, quads, levels, device=device)
    t_build = time.perf_counter() - t0
    h.set_input("thickness", lambda x: 0.1 * (1 + 0.5 * np.sin(2 * np.pi * x[0] / 10) * np.cos(2 * np.pi * x[1] / 10)))
    h.set_input("E", 1e8); h.set_input("nu", 0.3); h.set_input("density", 10.0)
    h.set_input("F_solid", lambda x: np.tile([0.0, 0.0, 5.0], (x.shape[1], 1)))
    h.set_strong_bc(lambda x: x[0] < 3e-16)
    t0 = time.perf_counter(); h.setup(); t_setup = time.perf_counter() - t0
    t0 = time.perf_counter(); it_s, rr_s = h.solve_state(rtol=1e-10); t_state = time.perf_counter() - t0
    dev_state_ms = h.fine.last_timings()["solve_ms"]
    rhs = h.fine.output_grad("compliance", "disp_solid")
    t0 = time.perf_counter(); h.solve_adjoint(rhs, rtol=1e-10); t_adj = time.perf_counter() - t0
    adj = dict(h.last_solve)
    # next optimisation iterate: thickness changed by 5 %, coarse inverse kept from the previous set-up (lagged)
    h.set_input("thickness", lambda x: 0.105 * (1 + 0.45 * np.sin(2 * np.pi * x[0] / 10) * np.cos(2 * np.pi * x[1] / 10)))
    t0 = time.perf_counter(); h.setup(reuse_coarse=True); t_setup2 = time.perf_counter() - t0
    t0 = time.perf_counter(); it_l, rr_l = h.solve_state(rtol=1e-10); t_state2 = time.perf_counter() - t0
    out = dict(levels=levels, coarse_cells="%dx%d" % (base, base), n_dof=int(h.fine.ndof), smoother="Chebyshev(3) of block-Jacobi",
               coarsest="direct (block-tridiagonal LDL^T on level sets, in-tree kernels), %d dof" % h.levels[-1].ndof, rtol=1e-10,
               state_iters=int(it_s), state_relres=float(rr_s), state_s=t_state, state_device_ms=dev_state_ms,
               adjoint_iters=int(adj["iters"]), adjoint_relres=float(adj["relres"]), adjoint_s=t_adj,
               setup_s=t_setup, hierarchy_build_s=t_build,
               state_plus_adjoint_s_per_opt_iteration=t_setup + t_state + t_adj,
               lagged_coarse=dict(setup_s=t_setup2, state_iters=int(it_l), state_relres=float(rr_l), state_s=t_state2,
                                  note="thickness changed by 5 %, coarse factor of the previous iterate reused"),
               note="setup = assemble all levels + smoother bounds + coarse factorisation, redone when the inputs change")
    h.close()
    return out


def mg_solve_metric_dist(n, rank, world, local_rank, dist, strong):
    """Metric 2 on N GPUs: V-cycle PCG state + adjoint solve to relres 1e-10 with the finest levels distributed over
    nested x-strip partitions and the coarsest level (51 cells across) replicated and solved directly
    (femo_alpha_b200/dist_multigrid.py, csrc/mg.cu, csrc/coarse.cu).  Weak: (n*world) x n plate; strong: the n x n plate."""
    import torch
    from femo_alpha_b200.dist_multigrid import DistributedHierarchy
    levels = 1
    while n % (2 ** levels) == 0 and n // (2 ** levels) >= 8:
        levels += 1
    base = n // 2 ** (levels - 1)
    if levels < 2:
        return None
    bx = base if strong else base * world
    nodes, quads = plate(bx, base, 0.0, 10.0 * (1 if strong else world), 10.0)
    ix = np.arange(quads.shape[0]) // base                      # plate() numbers cells x-major
    part_c = ((ix * world) // bx).astype(np.int32)               # x-strips of whole coarse columns

    def timed(fn):
        dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = fn()
        torch.cuda.synchronize()
        t = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return out, float(t.item())

    # distribute a level only while a rank keeps >= 250 k dofs of it: below that the halo exchanges (NCCL latency, ~0.15 ms
    # each, 8 per level and V-cycle) cost more than smoothing the whole level redundantly on every rank
    n_dist = 0
    for l in range(levels - 1):
        cells_l = (bx * base) * 4 ** (levels - 1 - l)
        if 15.0 * cells_l / world >= 2.5e5:
            n_dist += 1
    n_dist = max(1, n_dist)
    h, t_build = timed(lambda: DistributedHierarchy(nodes, quads, levels, rank, world, device=local_rank, part_coarse=part_c,
                                                    n_dist=n_dist))
    h.init_comm(dist)
    _, t_prep = timed(h.prepare)             # one-time: device-side hierarchy + transfer operators
    t_build += t_prep
    xf = h.meshes[0][0]
    nn = xf.shape[0]
    lx = 10.0
    h.set_input("thickness", 0.1 * (1 + 0.5 * np.sin(2 * np.pi * xf[:, 0] / lx) * np.cos(2 * np.pi * xf[:, 1] / 10)))
    h.set_input("E", np.full(nn, 1e8)); h.set_input("nu", np.full(nn, 0.3)); h.set_input("density", np.full(nn, 10.0))
    h.set_input("F_solid", np.tile([0.0, 0.0, 5.0], nn))
    h.set_strong_bc(lambda x: x[0] < 3e-16)
    _, t_setup = timed(h.setup)
    hx0, ar0 = h.fine.comm_counts()
    (it_s, rr_s), t_state = timed(lambda: h.solve_state(rtol=1e-10))
    dev_state_ms = h.last_solve["device_ms"]
    hx1, ar1 = h.fine.comm_counts()
    rhs = h.fine.dev.output_grad("compliance", "disp_solid")
    _, t_adj = timed(lambda: h.solve_adjoint(rhs, rtol=1e-10))
    adj = dict(h.last_solve)
    C = h.fine.output("compliance")
    tt = torch.tensor([h.fine.ndof_owned], device="cuda", dtype=torch.float64)
    dist.all_reduce(tt)
    out = dict(levels=levels, distributed_levels=h.n_dist, coarse_cells="%dx%d" % (bx, base), n_dof=int(tt.item()),
               smoother="Chebyshev(3) of block-Jacobi", coarsest="direct (block-tridiagonal LDL^T on level sets), %d dof, replicated"
               % h.coarse[-1].ndof, rtol=1e-10, state_iters=int(it_s), state_relres=float(rr_s), state_s=t_state,
               state_device_ms=dev_state_ms, adjoint_iters=int(adj["iters"]), adjoint_relres=float(adj["relres"]),
               adjoint_s=t_adj, setup_s=t_setup, hierarchy_build_s=t_build, compliance=float(C),
               state_plus_adjoint_s_per_opt_iteration=t_setup + t_state + t_adj,
               ms_per_vcycle_iteration=dev_state_ms / max(int(it_s), 1),
               finest_level_comm_per_iteration=dict(halo_exchanges=(hx1 - hx0) / max(int(it_s), 1),
                                                    allreduces=(ar1 - ar0) / max(int(it_s), 1)),
               partition="nested x-strips; times = max over ranks (host clock around synchronised calls)")
    h.close()
    return out


def warped_metric(n, device, steps=5):
    """Non-flat variants of config 3, Jacobian assembly timed by CUDA events inside the library, each with its
    own roofline line (same 9.1 KB of algorithmic bytes per cell):
      ruled          z = 0.05 sin(pi x / 10), SURVEY's variant: a ruled surface -- every cell is planar but has its own
                     frame, so the tiled kernel runs with per-contribution rotations;
      doubly_curved  z = 0.05 sin(pi x / 10) sin(pi y / 10): every cell is genuinely warped (per-point frames and the
                     dE2/dX bending terms of kinematics.py:54-82) -- the cell-centric warped kernels;
      wingbox_x16 / wingbox_x64   the reference's wing-box mesh of config 2 (tests/golden/wingbox_mesh.npz: warped cells,
                     T-junctions) subdivided twice / three times, element-wise material."""
    import torch
    from femo_alpha_b200.shell_device import ShellDevice
    peak, _ = measured_peaks()
    out = {}
    for name, zf in (("ruled", lambda x, y: 0.05 * np.sin(np.pi * x / 10.0)),
                     ("doubly_curved", lambda x, y: 0.05 * np.sin(np.pi * x / 10.0) * np.sin(np.pi * y / 10.0))):
        try:
            nodes, quads = plate(n, n, 0.0, 10.0, 10.0)
            nodes[:, 2] = zf(nodes[:, 0], nodes[:, 1])
            d = ShellDevice(nodes, quads, device=device)
            inp = inputs_for(nodes)
            for k in ("thickness", "E", "nu", "density", "F_solid"):
                d.set_input(k, inp[k])
            h_dev = [torch.from_numpy(inp["thickness"] * (1.0 + 1e-3 * k)).cuda() for k in range(2)]
            ms, rec_ms, ker_ms = [], [], []
            for i in range(3 + steps):
                d.set_input("thickness", h_dev[i % 2])
                d.assemble_jacobian()
                t = d.last_timings()
                if i >= 3:
                    ms.append(t["records_ms"] + t["gather_ms"]); rec_ms.append(t["records_ms"]); ker_ms.append(t["gather_ms"])
            t_ms = float(np.mean(ms))
            ach = CELL_BYTES * d.nc / (t_ms * 1e-3) / 1e9
            out[name] = dict(all_planar=int(d.sizes.all_planar), kernel="k_asm_tiles_ws" if d.sizes.all_planar else ("k_gather<false>" if os.environ.get("FEMO_WARPED") == "0" else "k_warped_elem<%s> + k_warped_gather" % os.environ.get("FEMO_WARPED_NB", "3")),
                             first_kernel_ms=float(np.mean(rec_ms)), second_kernel_ms=float(np.mean(ker_ms)),
                             n_dof=int(d.ndof), ms_per_assembly=t_ms, value=d.ndof / (t_ms * 1e-3) / 1e9, unit="GDOF/s",
                             roofline=dict(bound="hbm", achieved=ach, peak=peak, unit="GB/s", frac=ach / peak,
                                           algorithmic_bytes_per_launch=CELL_BYTES * d.nc))
            d.close()
        except Exception as e:                      # reported, never hidden
            out[name] = {"error": "%s: %s" % (type(e).__name__, e)}
    # the reference's own wing-box mesh (config 2: every cell warped, T-junctions), subdivided x16 and x64
    try:
        from femo_alpha_b200.multigrid import refine_quads
        m = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests", "golden", "wingbox_mesh.npz"))
        mesh = (m["nodes"], m["quads"].astype(np.int32))
        for level in (1, 2, 3):
            mesh = refine_quads(*mesh)
            if level < 2:
                continue
            nodes, quads = mesh
            nc = quads.shape[0]
            d = ShellDevice(nodes, quads, True, False, device=device)
            hh = 0.05 * 0.0254
            d.set_input("E", np.full(nc, 73.1e9)); d.set_input("nu", np.full(nc, 0.33))
            ts = []
            for i in range(3 + steps):
                d.set_input("thickness", np.full(nc, hh * (1.0 + 1e-3 * (i % 2))))
                d.assemble_jacobian()
                t = d.last_timings()
                if i >= 3:
                    ts.append(t["records_ms"] + t["gather_ms"])
            t_ms = float(np.mean(ts))
            ach = CELL_BYTES * d.nc / (t_ms * 1e-3) / 1e9
            out["wingbox_x%d" % 4 ** level] = dict(
                all_planar=int(d.sizes.all_planar), kernel="k_warped_elem + k_warped_gather", n_cells=int(nc), n_dof=int(d.ndof),
                ms_per_assembly=t_ms, value=d.ndof / (t_ms * 1e-3) / 1e9, unit="GDOF/s",
                roofline=dict(bound="hbm", achieved=ach, peak=peak, unit="GB/s", frac=ach / peak,
                              algorithmic_bytes_per_launch=CELL_BYTES * d.nc))
            d.close()
    except Exception as e:
        out["wingbox"] = {"error": "%s: %s" % (type(e).__name__, e)}
    return out


def dynamic_metric():
    """BASELINE config 4 (dynamic_rm_shell, 1-cos gust, plate 2x10 with 40x8 quads, quad_deg 3, 200 steps): seconds per
    time step (assemble once + solve per step) and the time-adjoint gradient of the total strain energy."""
    try:
        from scripts.dyn_bench import run_gpu
        from femo_alpha_b200.mesh import createRectangleMesh
        r = run_gpu(createRectangleMesh((0., 0.), (10., 2.), 40, 8), 200, "auto")
        r["workload"] = "dynamic_rm_shell plate 2x10, 40x8 quads, quad_deg 3, 200 implicit-midpoint steps (config 4)"
        return r
    except Exception as e:                      # reported, never hidden
        return {"error": "%s: %s" % (type(e).__name__, e)}


def solid_metric():
    """BASELINE config 5 (generic 3-D form): unit cube, 160^3 Q1 hexahedra, SIMP elasticity -- Jacobian assembly and
    PCG iteration times (scripts/hex_bench.py)."""
    try:
        from scripts.hex_bench import run
        return run(160, 3, 20)
    except Exception as e:                      # reported, never hidden
        return {"error": "%s: %s" % (type(e).__name__, e)}


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from femo_alpha_b200.shell_device import ShellDevice
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        # torchrun exports OMP_NUM_THREADS=1: give the library's host-side plan construction this rank's share of the cores
        os.environ.setdefault("FEMO_HOST_THREADS", str(max(1, (os.cpu_count() or 1) // world)))
    n = args.n
    strong = args.scaling == "strong" and world > 1
    if world > 1:
        # weak (default): one global (n*world) x n plate on [0, 10 world] x [0, 10] in x-strips of n x n cells per GPU;
        # strong: the ONE n x n plate of config 3 cut into `world` x-strips.  Owner-computes assembly with one
        # ghost-cell layer, NCCL halo + all-reduce in the solves
        from femo_alpha_b200.dist_shell import DistributedShell
        nx = n if strong else n * world
        nodes, quads = plate(nx, n, 0.0, 10.0 * (nx // n), 10.0)
        part = (((np.arange(quads.shape[0]) // n) * world) // nx).astype(np.int32)     # plate() numbers cells x-major
        ds = DistributedShell(nodes, quads, rank, world, part=part, device=local_rank)
        ds.init_comm(dist)
        d = ds.dev
        d.set_stream(torch.cuda.current_stream().cuda_stream)
        ginp = inputs_for(nodes)
        inp = {k: ds._slice(k, ginp[k]) for k in ("thickness", "E", "nu", "density", "F_solid", "uhat")}
        for k in ("thickness", "E", "nu", "density", "F_solid"):
            d.set_input(k, inp[k])
        ds.set_strong_bc(lambda x: x[0] < 3e-16)
        del nodes, quads, ginp
    else:
        nodes, quads = plate(n, n, 0.0, 10.0, 10.0)
        d = ShellDevice(nodes, quads, device=local_rank)
        d.set_stream(torch.cuda.current_stream().cuda_stream)
        inp = inputs_for(nodes)
        for k in ("thickness", "E", "nu", "density", "F_solid"):
            d.set_input(k, inp[k])
        on = d.block_coords()[:, 0] < 3e-16
        blocks = np.nonzero(on)[0]
        d.set_strong_bc((3 * blocks[:, None] + np.arange(3)[None, :]).ravel())
    ndof_total = d.ndof
    if world > 1:
        tt = torch.tensor([ds.ndof_owned], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt)
        ndof_total = int(tt.item())                   # owned DOFs summed over ranks (ghosts not counted)
    # device-resident thickness variants so that every step really recomputes from changed inputs
    h_dev = [torch.from_numpy(inp["thickness"] * (1.0 + 1e-3 * k)).cuda() for k in range(2)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(i):
        d.set_input("thickness", h_dev[i % 2])          # D2D, 8 B/vertex: invalidates the matrix
        d.assemble_jacobian()

    for i in range(args.warmup):
        step(i)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = d.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    gather_ms, rec_ms = [], []
    ev0.record()
    for i in range(args.steps):
        step(i)
        t = d.last_timings()
        gather_ms.append(t["gather_ms"]); rec_ms.append(t["records_ms"])
    ev1.record()
    barrier()
    launches = d.launch_count() - l0
    ms = ev0.elapsed_time(ev1) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        tt = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())

    # ---- end to end through the plugin API with host buffers -------------------------------
    host = {k: np.ascontiguousarray(inp[k]) for k in ("thickness", "F_solid", "E", "nu", "uhat")}
    try:
        pinned = {k: torch.from_numpy(v).pin_memory().numpy() for k, v in host.items()}
    except Exception:
        pinned = host
    R = torch.empty(d.ndof, dtype=torch.float64).pin_memory().numpy()
    d.set_state(np.zeros(d.ndof))

    def e2e_step_separate(i):
        pinned["thickness"][0] = inp["thickness"][0] * (1.0 + 1e-6 * i)
        for k in ("thickness", "F_solid", "E", "nu", "uhat"):
            d.set_input(k, pinned[k])                   # H2D
        d.assemble_residual(None, R)                    # assembles dR/du (inputs changed) and R; D2H of R

    def e2e_step_fused(i):                              # the same work as ONE call with the copies overlapped
        pinned["thickness"][0] = inp["thickness"][0] * (1.0 + 1e-6 * i)
        d.evaluate_residual(pinned, None, R)

    e_steps = max(2, min(args.steps, 5))

    def time_e2e(fn):
        for i in range(2):
            fn(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(e_steps):
            fn(i)
        barrier()
        return (time.perf_counter() - t0) / e_steps * 1e3

    e_sep_ms = time_e2e(e2e_step_separate)
    e_ms = time_e2e(e2e_step_fused)
    if world > 1:
        tt = torch.tensor([e_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e_ms = float(tt.item())
    h2d = int(sum(v.nbytes for v in host.values())); d2h = int(R.nbytes)

    # the same through the API for the Jacobian alone: the design variable comes from the host, dR/du stays on the device
    # (what StateOperation does between two optimiser iterates when only the thickness changed)
    def e2e_jac_step(i):
        pinned["thickness"][0] = inp["thickness"][0] * (1.0 + 1e-6 * (i + 7))
        d.set_input("thickness", pinned["thickness"])   # H2D
        d.assemble_jacobian()
    for i in range(2):
        e2e_jac_step(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(e_steps):
        e2e_jac_step(i)
    barrier()
    ej_ms = (time.perf_counter() - t0) / e_steps * 1e3
    if world > 1:
        tt = torch.tensor([ej_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ej_ms = float(tt.item())

    # ---- Krylov iteration (state / adjoint solves share it): fixed number of PCG iterations ---
    d.set_input("thickness", h_dev[0])
    kiters = args.krylov_iters
    d.pcg_fixed(10)
    pcg_ms = d.pcg_fixed(kiters) / kiters
    barrier()
    overlap = None
    if world > 1 and os.environ.get("FEMO_TRACE_OVERLAP") == "1":
        tr = ds.overlap_info()                        # CUDA-event trace of the last operator application of the loop
        tt = torch.tensor(tr, device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        overlap = dict(halo_done_ms=float(tt[0]), interior_rows_done_ms=float(tt[1]), boundary_rows_done_ms=float(tt[2]),
                       hidden=bool(tt[0] <= tt[1]), note="max over ranks, ms after the start of the last overlapped SpMV of the timed PCG loop")

    mg = None
    dyn = solid = warped = None
    if not args.no_mg:
        if world == 1:
            mg = mg_solve_metric(n, local_rank)
            warped = warped_metric(n, local_rank)
            dyn = dynamic_metric()
            solid = solid_metric()
        else:
            try:
                mg = mg_solve_metric_dist(n, rank, world, local_rank, dist, strong)
            except Exception as e:                  # reported, never hidden (every rank raises or none does: collective calls)
                mg = {"error": "%s: %s" % (type(e).__name__, e)}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = measured_peaks()
    g_ms = float(np.mean(gather_ms)); r_ms = float(np.mean(rec_ms))
    # fp64 denominators measured live (FMA chain and DMMA m8n8k4 microbenchmarks of the library, csrc/solver.cu)
    try:
        fma_tf, copy_gbs = d.measure_peaks()
        dmma_tf = d.dmma_tflops()
    except Exception:
        fma_tf, copy_gbs, dmma_tf = None, None, None
    achieved = CELL_BYTES * d.nc / (g_ms * 1e-3) / 1e9
    pcg_bytes = 12.0 * d.nnzb * 9 + 188.0 * d.ndof      # CSR-equivalent figure
    # what the BSR layout has to move: 72 B values + 4 B column word per block, row pointers, and per DOF the
    # vectors of SpMV / update / direction kernels (~96 B) plus the block-Jacobi inverses (~34 B)
    bsr_bytes = 76.0 * d.nnzb + 4.0 * d.ndof / 3 + 130.0 * d.ndof
    cpu, _, _ = cpu_baseline(n) if world == 1 else (None, None, None)
    traffic = None
    try:                                            # DRAM bytes of one launch from the committed ncu --set full capture
        tj = json.load(open(os.path.join(ROOT, "profiles", "asm_traffic.json")))
        if tj.get("workload") == "rm_shell_plate_%dx%d_q2q1_cfg3" % (n, n) and world == 1:
            traffic = float(tj["dram_bytes_per_launch"])
    except Exception:
        pass
    out = {
        "metric": "rm_shell_jacobian_assembly_throughput", "value": ndof_total / (ms * 1e-3) / 1e9, "unit": "GDOF/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "rm_shell_plate_%dx%d_q2q1_cfg3" % (n, n), "cells_per_gpu": int(d.nc),
                   "dof_per_gpu": int(d.ndof), "dof_total": int(ndof_total), "nnz_per_gpu": int(d.nnzb * 9), "quadrature": "4x4 Gauss",
                   "l2_policy": "working set (%.1f GB written per step) >> 126 MB L2" % (d.nnzb * 72 / 1e9),
                   "partition": "x-strips (RCB), owner-computes + 1 ghost-cell layer, no collective in assembly"},
        "e2e": {"value": ndof_total / (e_ms * 1e-3) / 1e9, "unit": "GDOF/s", "ms_per_step": e_ms,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "api": "ShellDevice.evaluate_residual(host inputs x5, host out): one call, copies overlapped with the kernels",
                "separate_calls": {"value": ndof_total / (e_sep_ms * 1e-3) / 1e9, "ms_per_step": e_sep_ms,
                                   "api": "ShellDevice.set_input x5 (host) + assemble_residual(host out)"},
                "jacobian_only": {"value": ndof_total / (ej_ms * 1e-3) / 1e9, "unit": "GDOF/s", "ms_per_step": ej_ms,
                                  "h2d_bytes_per_step": int(host["thickness"].nbytes), "d2h_bytes_per_step": 0,
                                  "api": "ShellDevice.set_input('thickness', host) + assemble_jacobian()"}},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"kernel": "k_asm_tiles_ws (register-tiled moments, warp-specialised compute/writer roles, deterministic)", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": CELL_BYTES * d.nc, "avg_launch_ms": g_ms,
                     "records_kernel_ms": r_ms,
                     # "report both fractions".  dense_equivalent = the 0.44 MFLOP/cell of a dense B^T D B
                     # (what the reference's generated kernel evaluates); moment_form = the ~20 k fp64 FMA per cell
                     # this kernel actually issues (local-frame moments)
                     "fp64": None if not fma_tf else {
                         "peak_tflops_fma": fma_tf, "peak_tflops_dmma": dmma_tf, "library_copy_gbs": copy_gbs,
                         "dense_equivalent_tflops": 0.44e6 * d.nc / (g_ms * 1e-3) / 1e12,
                         "dense_equivalent_frac": 0.44e6 * d.nc / (g_ms * 1e-3) / 1e12 / fma_tf,
                         "moment_form_tflops": 2 * 20e3 * d.nc / (g_ms * 1e-3) / 1e12,
                         "fp64_frac": 2 * 20e3 * d.nc / (g_ms * 1e-3) / 1e12 / fma_tf}},
        "solve": {"krylov": "PCG + block-Jacobi (6x6 vertex / 3x3 edge, centre)", "iters_timed": kiters,
                  "s_per_krylov_iter": pcg_ms * 1e-3, "algorithmic_bytes_per_iter": pcg_bytes,
                  "csr_equivalent_gbs": pcg_bytes / (pcg_ms * 1e-3) / 1e9, "bsr_bytes_per_iter": bsr_bytes,
                  "achieved_gbs": bsr_bytes / (pcg_ms * 1e-3) / 1e9, "frac_of_hbm_peak": bsr_bytes / (pcg_ms * 1e-3) / 1e9 / peak,
                  "converged": False, "halo_overlap_trace": overlap,
                  "note": "fixed iteration count; N>1: one coupled system, halo exchange overlapped with the interior rows, one all-reduce per iteration (Chronopoulos-Gear)",
                  "multigrid": mg},
        "warped": warped,
        "dynamic": dyn,
        "solid": solid,
        "cpu_baseline": cpu,
    }
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", "--cells", dest="n", type=int, default=816,
                    help="cells per side of the (per-GPU) plate (--cells: the spelling torch.distributed.run passes through)")
    ap.add_argument("--no-mg", action="store_true", help="skip the multigrid state+adjoint solve metric")
    ap.add_argument("--krylov-iters", type=int, default=100)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N > 1: weak = n x n cells per GPU (default, what the driver's scaling run uses); "
                         "strong = the one n x n plate split over the GPUs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
