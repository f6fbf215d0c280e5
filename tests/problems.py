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


def scordelis_lo_roof(n, skew=0.12):
    """Scordelis-Lo roof (MacNeal & Harder 1985): cylindrical shell R = 25, L = 50, 80 degrees, t = 0.25, E = 4.32e8, nu = 0,
    gravity 90 per unit area, rigid diaphragms at the curved ends (u_x = u_z = 0); reference vertical displacement at the
    middle of the free edge 0.3024.  The FULL roof with n x n quads; `skew` (radians) bends the axial mesh lines sideways by
    skew * (1 - (phi/phi0)^2) * sin(2 pi y / L): the four vertices of a quad then lie on four different generators of the
    cylinder and the quad is out of plane (warped cells: per-point frames, dE2/dX terms), while the boundary, the crown
    mid-point and the free-edge mid-point stay where they are.  (Shifting nodes ALONG the axis would not do: two pairs of
    vertices on two parallel generators are always coplanar.)  Returns a dict like make_problem plus
    `bc_dofs(block_coords, nq2)` and the sampling point."""
    from oracle.rm_shell_ref import plate_mesh
    R, L, phi0 = 25.0, 50.0, np.deg2rad(40.0)
    P, quads = plate_mesh(n, n, 2.0 * phi0, L)              # parameter space (phi + phi0, y)
    phi, y = P[:, 0] - phi0, P[:, 1]
    phi = phi + skew * (1.0 - (phi / phi0) ** 2) * np.sin(2.0 * np.pi * y / L)
    nodes = np.stack([R * np.sin(phi), y, R * np.cos(phi)], 1)
    nn = nodes.shape[0]

    def bc_dofs(xb, nq2):
        ends = np.nonzero((np.abs(xb[:nq2, 1]) < 1e-9) | (np.abs(xb[:nq2, 1] - L) < 1e-9))[0]
        crown = int(np.argmin(np.abs(xb[:nq2, 0]) + np.abs(xb[:nq2, 1] - 0.5 * L)))
        # diaphragms: u_x = u_z = 0; one u_y at the crown mid-point takes out the rigid translation along the axis
        return np.unique(np.concatenate([3 * ends, 3 * ends + 2, [3 * crown + 1]]).astype(np.int64))

    return dict(nodes=nodes, quads=quads, thickness=np.full(nn, 0.25), E=np.full(nn, 4.32e8), nu=np.zeros(nn),
                density=np.ones(nn), F_solid=np.tile([0.0, 0.0, -90.0], nn), bc_dofs=bc_dofs,
                free_edge_mid=np.array([R * np.sin(phi0), 0.5 * L, R * np.cos(phi0)]), reference=0.3024)
