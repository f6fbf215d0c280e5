"""Multi-rank host logic on CPU (gloo, world_size 2 and 3): partition, owner/ghost numbering, halo lists and
the distributed PCG recurrence, checked against the oracle's global direct solve."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from tests.problems import make_problem, oracle_for

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _global_keys(topo, nv, ne, nc):
    """same key convention as femo_alpha_b200.partition for the *global* mesh numbering of the oracle"""
    nb = topo.nb
    key = np.empty(nb, dtype=np.int64)
    key[:topo.nq2] = np.arange(topo.nq2)               # vertices | edges (np.unique order) | centres
    key[topo.nq2:] = nv + ne + nc + np.arange(nv)
    return key


@pytest.mark.parametrize("world,name", [(2, "plate_4x20"), (3, "skew_planar"), (2, "plate_dg0")])
def test_distributed_pcg_matches_global_solve(tmp_path, world, name):
    port = _free_port()
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "dist_worker.py"), str(r), str(world),
                               str(port), name, str(tmp_path)], cwd=ROOT) for r in range(world)]
    for pr in procs:
        assert pr.wait(timeout=600) == 0
    p = make_problem(name)
    o = oracle_for(p)
    w = o.solve_state()
    t = o.topo
    gkey = _global_keys(t, t.nn, t.ne, t.nc)
    lookup = np.argsort(gkey)
    seen = np.zeros(t.nb, dtype=int)
    owned_cells = []
    tot_send = tot_recv = 0
    for r in range(world):
        d = np.load(os.path.join(tmp_path, "rank%d.npz" % r))
        pos = lookup[np.searchsorted(gkey[lookup], d["gkey"])]
        assert np.array_equal(gkey[pos], d["gkey"])
        x = d["x"].reshape(-1, 3)
        ref = w.reshape(-1, 3)[pos]
        # owned AND ghost entries (after the final halo exchange) match the global solution
        assert np.abs(x - ref).max() <= 1e-8 * np.abs(w).max()
        seen[pos[d["owner"] == r]] += 1
        owned_cells.append(d["cell_global"][:int(d["n_owned_cells"])])
        tot_send += int(d["send"]); tot_recv += int(d["recv"])
    assert (seen == 1).all()                                  # every block node has exactly one owner
    assert np.array_equal(np.sort(np.concatenate(owned_cells)), np.arange(t.nc))
    assert tot_send == tot_recv and tot_send > 0


def test_rcb_partition_is_balanced_and_covers():
    from femo_alpha_b200.partition import partition_cells_rcb
    rng = np.random.default_rng(5)
    c = rng.random((1000, 3))
    for world in (1, 2, 3, 4, 8):
        part = partition_cells_rcb(c, world)
        cnt = np.bincount(part, minlength=world)
        assert cnt.sum() == 1000 and cnt.min() >= 1000 // world - 1 and cnt.max() <= 1000 // world + 2


@pytest.mark.parametrize("world", [2, 3])
def test_distributed_hexahedral_solid_matches_global_solve(tmp_path, world):
    """build_local_part_hex (SURVEY 8 f4 on N GPUs): state, compliance, volume and the density gradient of the
    compliance of a distorted 6x4x3 box against the oracle's global solve."""
    from oracle.hex_ref import HexOracle
    from tests.dist_hex_worker import problem
    port = _free_port()
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "dist_hex_worker.py"), str(r), str(world),
                               str(port), str(tmp_path)], cwd=ROOT) for r in range(world)]
    for pr in procs:
        assert pr.wait(timeout=600) == 0
    nodes, cells, rho, f = problem()
    o = HexOracle(nodes, cells, E0=4.0, nu=0.3)
    o.set_inputs(density=rho, body_force=f)
    base = np.nonzero(nodes[:, 0] < 1e-12)[0]
    o.set_strong_bc((3 * base[:, None] + np.arange(3)).ravel())
    u = o.solve()
    g = o.total_dcompliance_drho()
    seen = np.zeros(o.nn, dtype=int); cells_seen = np.zeros(o.nc, dtype=int)
    tot_send = tot_recv = 0
    for r in range(world):
        d = np.load(os.path.join(tmp_path, "hex_rank%d.npz" % r))
        x = d["x"].reshape(-1, 3)
        assert np.abs(x - u.reshape(-1, 3)[d["gkey"]]).max() <= 1e-8 * np.abs(u).max()      # owned and ghost entries
        assert np.isclose(float(d["compliance"]), o.compliance(u), rtol=1e-9)
        assert np.isclose(float(d["volume"]), o.volume(), rtol=1e-12)
        n_own = int(d["n_owned_cells"])
        assert np.abs(d["g_own"] - g[d["cell_global"][:n_own]]).max() <= 1e-8 * np.abs(g).max()
        seen[d["gkey"][d["owner"] == r]] += 1
        cells_seen[d["cell_global"][:n_own]] += 1
        tot_send += int(d["send"]); tot_recv += int(d["recv"])
    assert (seen == 1).all() and (cells_seen == 1).all()
    assert tot_send == tot_recv and tot_send > 0


def test_distributed_two_grid_preconditioner(tmp_path):
    """NumPy prototype of the N-rank V-cycle (tests/dist_mg_worker.py): nested partition, local prolongation, restriction
    through a reverse halo exchange.  Same outer iteration count at world sizes 1, 2 and 3, far below Jacobi PCG, and the
    solution of the oracle's global solve."""
    from oracle.rm_shell_ref import RMShellOracle
    from tests.dist_mg_worker import global_keys, problem
    nodes_c, quads_c, nodes_f, quads_f = problem()
    o = RMShellOracle(nodes_f, quads_f, False, False)
    nn = nodes_f.shape[0]
    o.set_inputs(thickness=np.full(nn, 0.1), E=np.full(nn, 1e8), nu=np.full(nn, 0.3), density=np.full(nn, 10.0),
                 F_solid=np.tile([0.0, 0.0, 5.0], nn))
    o.set_strong_bc(lambda x: np.less(x[0], 3e-16))
    w = o.solve_state().reshape(-1, 3)
    gk = global_keys(o.topo)
    order = np.argsort(gk)
    counts = {}
    for world in (1, 2, 3):
        out = tmp_path / ("w%d" % world); out.mkdir()
        port = _free_port()
        procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "dist_mg_worker.py"), str(r), str(world),
                                   str(port), str(out)], cwd=ROOT) for r in range(world)]
        for pr in procs:
            assert pr.wait(timeout=900) == 0
        for r in range(world):
            d = np.load(os.path.join(out, "mg_rank%d.npz" % r))
            pos = order[np.searchsorted(gk[order], d["gkey"])]
            assert np.abs(d["x"].reshape(-1, 3) - w[pos]).max() <= 1e-7 * np.abs(w).max()
            counts.setdefault(world, (int(d["its"]), int(d["its_jacobi"])))
            assert counts[world] == (int(d["its"]), int(d["its_jacobi"]))
    its = [counts[k][0] for k in (1, 2, 3)]
    assert max(its) - min(its) <= 1, counts                    # the preconditioner is the same operator on any partition
    assert 0 < its[0] * 10 < counts[1][1], counts              # and an order of magnitude fewer iterations than Jacobi PCG
