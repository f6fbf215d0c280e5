"""Seeded problem set shared by the oracle tests and the GPU parity tests."""
import numpy as np

from oracle.rm_shell_ref import plate_mesh

CLAMP_X0 = lambda x: np.less(x[0], 3e-16)      # ex_simple_shell_opt.py:52-53


def make_problem(name):
    rng = np.random.default_rng(20261104)
    p = dict(name=name, ewm=False, ewp=False, uhat=None, bc=CLAMP_X0)
    if name == "plate_4x20":                       # BASELINE config 1
        nodes, quads = plate_mesh(20, 4, 10.0, 2.0)
        nn = nodes.shape[0]
        p.update(nodes=nodes, quads=quads, thickness=np.full(nn, 0.1), E=np.full(nn, 1e8), nu=np.full(nn, 0.3),
                 density=np.full(nn, 10.0), F_solid=np.tile([0.0, 0.0, 5.0], nn))
    elif name == "plate_varh":                     # CG1 thickness field, random E / nu
        nodes, quads = plate_mesh(7, 5, 3.0, 2.0)
        nn = nodes.shape[0]
        p.update(nodes=nodes, quads=quads, thickness=0.05 * (1 + 0.5 * rng.random(nn)),
                 E=1e6 * (1 + 0.2 * rng.random(nn)), nu=0.25 + 0.1 * rng.random(nn),
                 density=2.0 + rng.random(nn), F_solid=rng.standard_normal(3 * nn))
    elif name == "plate_dg0":                      # element-wise material and pressure
        nodes, quads = plate_mesh(6, 4, 3.0, 2.0)
        nc = quads.shape[0]
        p.update(nodes=nodes, quads=quads, ewm=True, ewp=True, thickness=0.05 * (1 + 0.5 * rng.random(nc)),
                 E=1e6 * (1 + 0.2 * rng.random(nc)), nu=0.25 + 0.1 * rng.random(nc),
                 density=2.0 + rng.random(nc), F_solid=rng.standard_normal(3 * nc))
    elif name == "warped":                         # non-planar, non-affine cells + mesh motion uhat
        nodes, quads = plate_mesh(5, 4, 2.0, 1.5, z_func=lambda X, Y: 0.15 * np.sin(1.3 * X) * (1 + Y) + 0.1 * X * Y)
        nn = nodes.shape[0]
        nodes = nodes + np.concatenate([0.05 * rng.standard_normal((nn, 2)), np.zeros((nn, 1))], 1) * \
            (nodes[:, :1] > 1e-9)
        p.update(nodes=nodes, quads=quads, thickness=0.05 * (1 + 0.5 * rng.random(nn)),
                 E=1e6 * (1 + 0.2 * rng.random(nn)), nu=0.25 + 0.1 * rng.random(nn),
                 density=2.0 + rng.random(nn), F_solid=rng.standard_normal(3 * nn),
                 uhat=0.02 * rng.standard_normal(3 * nn))
    elif name == "skew_planar":                    # planar but non-affine (trapezoids), rotated out of xy
        nodes, quads = plate_mesh(5, 4, 2.0, 1.5)
        nn = nodes.shape[0]
        nodes[:, :2] += 0.06 * rng.standard_normal((nn, 2)) * (nodes[:, :1] > 1e-9)
        c, s = np.cos(0.7), np.sin(0.7)
        R = np.array([[1, 0, 0], [0, c, -s], [0, s, c]])
        nodes = nodes @ R.T
        p.update(nodes=nodes, quads=quads, thickness=0.05 * (1 + 0.5 * rng.random(nn)),
                 E=1e6 * (1 + 0.2 * rng.random(nn)), nu=0.25 + 0.1 * rng.random(nn),
                 density=2.0 + rng.random(nn), F_solid=rng.standard_normal(3 * nn),
                 uhat=0.02 * rng.standard_normal(3 * nn))
    else:
        raise KeyError(name)
    return p


PROBLEMS = ["plate_4x20", "plate_varh", "plate_dg0", "warped", "skew_planar"]


def oracle_for(p, penalty=False, quadrature_degree=6):
    from oracle.rm_shell_ref import RMShellOracle
    o = RMShellOracle(p["nodes"], p["quads"], p["ewm"], p["ewp"], quadrature_degree)
    o.set_inputs(thickness=p["thickness"], E=p["E"], nu=p["nu"], density=p["density"], F_solid=p["F_solid"],
                 uhat=p["uhat"])
    (o.set_penalty_bc if penalty else o.set_strong_bc)(p["bc"])
    return o


def device_for(p, penalty=False, quadrature_degree=6):
    from femo_alpha_b200.shell_device import ShellDevice
    d = ShellDevice(p["nodes"], p["quads"], p["ewm"], p["ewp"], quadrature_degree)
    for k in ("thickness", "E", "nu", "density", "F_solid"):
        d.set_input(k, p[k])
    if p["uhat"] is not None:
        d.set_input("uhat", p["uhat"])
    if penalty:
        d.set_penalty_bc(np.asarray(p["bc"](p["nodes"].T)))
    else:
        on = np.asarray(p["bc"](d.block_coords().T))
        blocks = np.nonzero(on)[0]
        d.set_strong_bc(np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()))
    return d
