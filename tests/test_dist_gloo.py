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
