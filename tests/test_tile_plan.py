"""Host plan of the tiled assembly kernel (femo_alpha_b200/csrc/tile_plan.hpp), checked on the CPU.

Every element contribution (cell, a, b) to the upper block triangle of dR/du must be referenced exactly once, by the
chunk that owns the destination, with the right moment offset / transposition / frame slot; the lower triangle is
covered through the transposed positions."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
import scipy.sparse as sp

from oracle.rm_shell_ref import ShellTopology, plate_mesh

HERE = os.path.dirname(os.path.abspath(__file__))
C, MOM_CELL, MOM_TT, MOM_UT = 16, 419, 225, 275
P32 = ctypes.POINTER(ctypes.c_int32)
PU32 = ctypes.POINTER(ctypes.c_uint32)


@pytest.fixture(scope="module")
def host_lib():
    out = os.path.join(HERE, "_build", "libtile_plan_host.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    src = os.path.join(HERE, "tile_plan_host.cpp")
    hdr = os.path.join(HERE, "..", "femo_alpha_b200", "csrc", "tile_plan.hpp")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.run(["g++", "-O2", "-fopenmp", "-shared", "-fPIC", "-o", out, src], check=True)
    lib = ctypes.CDLL(out)
    lib.tp_build.restype = ctypes.c_void_p
    return lib


def pattern(topo):
    cb = topo.cell_blocks
    rows = np.repeat(cb, 13, axis=1).ravel()
    cols = np.tile(cb, (1, 13)).ravel()
    A = sp.csr_matrix((np.ones(rows.size), (rows, cols)), shape=(topo.nb, topo.nb))
    A.sum_duplicates(); A.sort_indices()
    flat = cb.ravel()
    order = np.argsort(flat, kind="stable")
    nc_adj = order.astype(np.int32)                          # = cell*13 + a, ascending cell inside a node
    nc_ptr = np.zeros(topo.nb + 1, np.int32)
    np.cumsum(np.bincount(flat, minlength=topo.nb), out=nc_ptr[1:])
    return A.indptr.astype(np.int32), A.indices.astype(np.int32), nc_ptr, nc_adj


def build(lib, topo, order=None):
    rowptr, colidx, nc_ptr, nc_adj = pattern(topo)
    err = ctypes.create_string_buffer(256)
    o = None if order is None else np.ascontiguousarray(order, np.int32)
    h = lib.tp_build(ctypes.c_int64(topo.nc), ctypes.c_int64(topo.nb), rowptr.ctypes.data_as(P32),
                     colidx.ctypes.data_as(P32), nc_ptr.ctypes.data_as(P32), nc_adj.ctypes.data_as(P32),
                     None if o is None else o.ctypes.data_as(P32), err, 256)
    assert h, err.value
    sz = np.zeros(8, np.int64)
    lib.tp_sizes(ctypes.c_void_p(h), sz.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)))
    nch, nd, nr, nhi, nhc = (int(v) for v in sz[:5])
    P = dict(cell_order=np.zeros(nch * C, np.int32), dest_ptr=np.zeros(nch + 1, np.int32), dest_pos=np.zeros(nd, np.int32),
             dest_tpos=np.zeros(nd, np.int32), dest_info=np.zeros(nd, np.uint32), ref_ptr=np.zeros(nch + 1, np.int32),
             refs=np.zeros(nr, np.uint32), hitem_ptr=np.zeros(nch + 1, np.int32), hitems=np.zeros(nhi, np.uint32),
             hcell_ptr=np.zeros(nch + 1, np.int32), hcells=np.zeros(nhc, np.int32))
    lib.tp_copy(ctypes.c_void_p(h), *[P[k].ctypes.data_as(PU32 if P[k].dtype == np.uint32 else P32) for k in
                                      ("cell_order", "dest_ptr", "dest_pos", "dest_tpos", "dest_info", "ref_ptr", "refs",
                                       "hitem_ptr", "hitems", "hcell_ptr", "hcells")])
    lib.tp_free(ctypes.c_void_p(h))
    P.update(nchunks=nch, max_dests=int(sz[5]), max_hitems=int(sz[6]), max_hcells=int(sz[7]), rowptr=rowptr, colidx=colidx)
    return P


def sym_index(lo, hi, n):
    return lo * n - lo * (lo - 1) // 2 + (hi - lo)


def decode_and_check(P, topo):
    """Replay the plan: returns the set of (cell, a, b) contributions per BSR position."""
    cb = topo.cell_blocks
    rowptr, colidx = P["rowptr"], P["colidx"]
    row_of = np.repeat(np.arange(topo.nb), np.diff(rowptr))
    seen = {}
    for c in range(P["nchunks"]):
        hit = P["hitems"][P["hitem_ptr"][c]:P["hitem_ptr"][c + 1]]
        hcs = P["hcells"][P["hcell_ptr"][c]:P["hcell_ptr"][c + 1]]
        rb = P["ref_ptr"][c]
        used_items = set()
        for d in range(P["dest_ptr"][c], P["dest_ptr"][c + 1]):
            k = int(P["dest_pos"][d]); I, J = int(row_of[k]), int(colidx[k])
            assert I <= J
            tp = int(P["dest_tpos"][d])
            if I == J:
                assert tp == -1
            else:
                assert row_of[tp] == J and colidx[tp] == I
            start, n = int(P["dest_info"][d] & 0xFFFFF), int(P["dest_info"][d] >> 20)
            contribs, last_p = [], -1
            for r in range(n):
                w = int(P["refs"][rb + start + r])
                off, typ, tr, fs = w & 0xFFFF, (w >> 16) & 3, (w >> 18) & 1, w >> 19
                if off < C * MOM_CELL:
                    slot, rel = divmod(off, MOM_CELL)
                    assert fs == slot
                    p = c * C + slot
                    if rel < MOM_TT:
                        assert typ == 0 and rel % 5 == 0
                        lohi = [(lo, hi) for lo in range(9) for hi in range(lo, 9) if 5 * sym_index(lo, hi, 9) == rel][0]
                    elif rel < MOM_UT:
                        assert typ == 1 and (rel - MOM_TT) % 5 == 0
                        lohi = [(lo + 9, hi + 9) for lo in range(4) for hi in range(lo, 4)
                                if 5 * sym_index(lo, hi, 4) == rel - MOM_TT][0]
                    else:
                        assert typ == 2 and (rel - MOM_UT) % 4 == 0
                        q = (rel - MOM_UT) // 4
                        lohi = (q // 4, 9 + q % 4)
                else:
                    i, rem = divmod(off - C * MOM_CELL, 5)
                    assert rem == 0 and i < hit.size and i not in used_items
                    used_items.add(i)
                    p, code = int(hit[i] >> 8), int(hit[i] & 0xFF)
                    lohi = divmod(code, 13)
                    assert p // C != c and hcs[fs - C] == p
                    assert typ == (0 if lohi[1] < 9 else (1 if lohi[0] >= 9 else 2))
                a, b = (lohi[1], lohi[0]) if tr else lohi
                cell = int(P["cell_order"][p])
                assert cell >= 0 and cb[cell, a] == I and cb[cell, b] == J
                assert p > last_p                      # fixed ascending processing order
                if r == 0:
                    assert p // C == c                 # the owner holds the first contributor
                last_p = p
                contribs.append((cell, a, b))
            assert k not in seen
            seen[k] = contribs
        assert len(used_items) == hit.size
    return seen


@pytest.mark.parametrize("shape", [(3, 2), (9, 7), (20, 4), (37, 5)])
@pytest.mark.parametrize("order", ["identity", "morton", "random"])
def test_plan_covers_every_contribution_once(host_lib, shape, order):
    nodes, quads = plate_mesh(shape[0], shape[1], 3.0, 2.0)
    topo = ShellTopology(nodes, quads)
    perm = None
    if order == "random":
        perm = np.random.default_rng(5).permutation(topo.nc)
    elif order == "morton":
        perm = np.zeros(topo.nc, np.int32)
        cv = np.ascontiguousarray(topo.cv, np.int32)
        host_lib.tp_morton(ctypes.c_int64(topo.nc), topo.x.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
                           cv.ctypes.data_as(P32), ctypes.c_double(topo.cell_diameter().min() / np.sqrt(2.0)),
                           perm.ctypes.data_as(P32))
        assert np.array_equal(np.sort(perm), np.arange(topo.nc))
    P = build(host_lib, topo, perm)
    seen = decode_and_check(P, topo)
    cb = topo.cell_blocks
    # expected: all (cell, a, b) with block(a) <= block(b)
    exp = {}
    rowptr, colidx = P["rowptr"], P["colidx"]
    for cell in range(topo.nc):
        for a in range(13):
            for b in range(13):
                I, J = cb[cell, a], cb[cell, b]
                if I <= J:
                    k = rowptr[I] + int(np.searchsorted(colidx[rowptr[I]:rowptr[I + 1]], J))
                    exp.setdefault(k, set()).add((cell, a, b))
    assert set(seen) == set(exp)
    for k, v in seen.items():
        assert len(v) == len(set(v)) and set(v) == exp[k]


def test_plan_t_junction(host_lib):
    """Non-manifold mesh: three plates meeting along one edge line (wing-box rib/skin junction)."""
    n1, q1 = plate_mesh(4, 3, 4.0, 3.0)
    # second sheet: vertical wall attached along y = 0 (shares the x-line of nodes)
    n2, q2 = plate_mesh(4, 2, 4.0, 2.0)
    wall = np.stack([n2[:, 0], np.zeros(len(n2)), n2[:, 1]], 1)
    on_line = np.isclose(wall[:, 2], 0.0)
    remap = np.full(len(wall), -1)
    for i in np.nonzero(on_line)[0]:
        remap[i] = int(np.nonzero(np.isclose(n1[:, 0], wall[i, 0]) & np.isclose(n1[:, 1], 0.0))[0][0])
    new = np.nonzero(~on_line)[0]
    remap[new] = len(n1) + np.arange(new.size)
    nodes = np.concatenate([n1, wall[new]])
    quads = np.concatenate([q1, remap[q2]])
    topo = ShellTopology(nodes, quads)
    P = build(host_lib, topo, None)
    decode_and_check(P, topo)
