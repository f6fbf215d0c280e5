"""The shape-partial element maths of the CUDA kernels (femo_alpha_b200/csrc/shape_partials.cuh) run on
the CPU -- the same __host__ __device__ functions, compiled by nvcc into a host harness (tests/shape_host.cu) -- and
checked against the complex-step oracle (oracle/shape_ref.py).  No GPU needed; the GPU parity test of the kernels
themselves is tests/test_gpu_shape.py."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import shape_ref as S
from tests.problems import make_problem, oracle_for

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "..", "femo_alpha_b200", "csrc")


@pytest.fixture(scope="module")
def host_lib():
    out = os.path.join(HERE, "_build", "libshape_host.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    src = os.path.join(HERE, "shape_host.cu")
    deps = [src] + [os.path.join(CSRC, f) for f in ("shape_partials.cuh", "femo_internal.cuh", "tables.hpp", "penalty_shape.cuh")]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
        nvcc = "/usr/local/cuda/bin/nvcc" if os.path.exists("/usr/local/cuda/bin/nvcc") else "nvcc"
        subprocess.run([nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC",
                        "-o", out, src], check=True)
    return ctypes.CDLL(out)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def run(lib, o, op, w, lam=None, duhat=None, flag=0, m1d=None, pm=1e-6, prho=100.0):
    t = o.topo
    nc = t.nc
    per = lambda a: np.ascontiguousarray(np.repeat(a[:, None], 4, 1) if o.ewm else a[t.cv])
    xv = np.ascontiguousarray(t.x[t.cv]); uh = np.ascontiguousarray(o.uhat.reshape(-1, 3)[t.cv])
    f3 = o.f.reshape(-1, 3)
    fv = np.ascontiguousarray(np.repeat(f3[:, None, :], 4, 1) if o.ewp else f3[t.cv])
    we = np.ascontiguousarray(w[t.cell_dofs]); le = np.ascontiguousarray((w if lam is None else lam)[t.cell_dofs])
    du = np.ascontiguousarray((np.zeros(3 * t.nn) if duhat is None else duhat).reshape(-1, 3)[t.cv])
    nout = 39 if op == 1 else 12
    out = np.zeros((nc, nout))
    rc = lib.shape_host(op, o.m if m1d is None else m1d, ctypes.c_int64(nc), _p(xv), _p(uh), _p(we), _p(le), _p(per(o.h)),
                        _p(per(o.E)), _p(per(o.nu)), _p(per(o.rho)), _p(np.ascontiguousarray(o.hE)), _p(fv), _p(du), flag,
                        ctypes.c_double(pm), ctypes.c_double(prho), _p(out))
    assert rc == 0
    if op == 1:
        r = np.zeros(t.ndof); np.add.at(r, t.cell_dofs.ravel(), out.ravel())
        return r
    g = np.zeros((t.nn, 3)); np.add.at(g, t.cv.ravel(), out.reshape(-1, 3))
    return g.ravel()


def rel(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


@pytest.mark.parametrize("name", ["warped", "skew_planar", "plate_varh", "plate_dg0"])
def test_shape_partials_against_complex_step(host_lib, name):
    p = make_problem(name)
    o = oracle_for(p)
    rng = np.random.default_rng(0)
    n = o.topo.ndof
    w, lam, du = rng.standard_normal(n), rng.standard_normal(n), rng.standard_normal(3 * o.topo.nn)
    assert rel(run(host_lib, o, 0, w, lam, flag=1), S.dRduhat_T_apply(o, w, lam)) < 1e-11
    assert rel(run(host_lib, o, 1, w, duhat=du), S.dRduhat_apply(o, w, du)) < 1e-11
    assert rel(run(host_lib, o, 2, w, flag=0), S.doutput_duhat(o, "compliance", w)) < 1e-11
    assert rel(run(host_lib, o, 2, w, flag=1), S.doutput_duhat(o, "mass", w)) < 1e-11
    assert rel(0.5 * run(host_lib, o, 0, w, w, flag=0), S.doutput_duhat(o, "elastic_energy", w)) < 1e-11
    # p-norm of the von Mises stress (degree-4 rule = 3 points per direction); moderate exponent and scale keep the
    # integrand well inside the fp64 range for a random state
    pm, prho = 1e-6, 6.0
    t = o.topo
    area = sum(wq * S.qp_geometry(t.x[t.cv], np.zeros((t.nc, 4, 3)), X0, X1).detJ.sum() for X0, X1, wq in o.qp)
    got = run(host_lib, o, 3, w, flag=int(o.ewm), m1d=3, pm=pm, prho=prho) / area
    assert rel(got, S.dpnorm_duhat(o, w, pm, prho)) < 1e-10


def test_complex_step_oracle_against_finite_differences():
    p = make_problem("warped")
    o = oracle_for(p)
    rng = np.random.default_rng(1)
    w, du = rng.standard_normal(o.topo.ndof), rng.standard_normal(3 * o.topo.nn)
    assert rel(S.residual_at(o, w, o.uhat), o.assemble_residual(w)) < 1e-13
    eps = 1e-6
    fd = (S.residual_at(o, w, o.uhat + eps * du) - S.residual_at(o, w, o.uhat - eps * du)) / (2 * eps)
    assert rel(S.dRduhat_apply(o, w, du), fd) < 1e-8
    lam = rng.standard_normal(o.topo.ndof)
    assert np.isclose(lam @ S.dRduhat_apply(o, w, du), S.dRduhat_T_apply(o, w, lam) @ du, rtol=1e-12)


# ---- penalty-BC facet terms: shape derivative of Nanson's factor (csrc/penalty_shape.cuh) ---------------------------------
LOCAL_EDGES = np.array([(0, 1), (0, 2), (1, 3), (2, 3)])
EDGE_Q2 = np.array([(0, 4, 1), (0, 5, 2), (1, 6, 3), (2, 7, 3)])


def penalty_pairs(o):
    """(cell, local edge) pairs of the penalised edges and the dof triples of their 5 trace nodes -- penalty_build's lists"""
    t = o.topo
    sel = np.zeros(t.ne, dtype=bool); sel[o.pen_edges] = True
    cells, les = np.nonzero(sel[t.ce])
    blocks = np.stack([t.cell_blocks[cells, EDGE_Q2[les, 0]], t.cell_blocks[cells, EDGE_Q2[les, 1]],
                       t.cell_blocks[cells, EDGE_Q2[les, 2]], t.cell_blocks[cells, 9 + LOCAL_EDGES[les, 0]],
                       t.cell_blocks[cells, 9 + LOCAL_EDGES[les, 1]]], 1)
    dofs = (3 * blocks[:, :, None] + np.arange(3)).reshape(len(cells), 15)
    return cells, les.astype(np.int32), blocks, dofs


def run_penalty(lib, o, op, w, lam=None, duhat=None):
    t = o.topo
    cells, les, blocks, dofs = penalty_pairs(o)
    n = len(cells)
    xv = np.ascontiguousarray(t.x[t.cv[cells]]); uh = np.ascontiguousarray(o.uhat.reshape(-1, 3)[t.cv[cells]])
    du = np.ascontiguousarray((np.zeros(3 * t.nn) if duhat is None else duhat).reshape(-1, 3)[t.cv[cells]])
    we = np.ascontiguousarray(w[dofs]); le_ = np.ascontiguousarray((w if lam is None else lam)[dofs])
    out = np.zeros((n, {0: 12, 1: 15, 2: 25}[op]))
    assert lib.penalty_shape_host(op, ctypes.c_int64(n), _p(xv), _p(uh), _p(les), _p(np.ascontiguousarray(o.hE[cells])), _p(we),
                                  _p(le_), _p(du), _p(out)) == 0
    if op == 0:
        g = np.zeros((t.nn, 3)); np.add.at(g, t.cv[cells].ravel(), out.reshape(-1, 3))
        return g.ravel()
    if op == 1:
        r = np.zeros(t.ndof); np.add.at(r, dofs.ravel(), out.ravel())
        return r
    return out, blocks


@pytest.mark.parametrize("name", ["plate_4x20", "warped", "skew_planar"])
def test_penalty_shape_derivative_against_complex_step(host_lib, name):
    p = make_problem(name)
    o = oracle_for(p, penalty=True)
    rng = np.random.default_rng(2)
    if p["uhat"] is None:
        o.set_inputs(uhat=0.05 * rng.standard_normal(3 * o.topo.nn))
    n = o.topo.ndof
    w, lam, du = rng.standard_normal(n), rng.standard_normal(n), rng.standard_normal(3 * o.topo.nn)
    # the forward trace matrices reproduce the oracle's penalty matrix
    Me, blocks = run_penalty(host_lib, o, 2, w)
    import scipy.sparse as sp
    rows = np.repeat(blocks, 5, axis=1).ravel(); cols = np.tile(blocks, (1, 5)).ravel()
    Pb = sp.coo_matrix((Me.ravel(), (rows, cols)), shape=(o.topo.nb, o.topo.nb)).tocsr()
    P = o.penalty_matrix()
    assert abs(sp.kron(Pb, sp.identity(3)).tocsr() - P).max() < 1e-12 * abs(P).max()
    assert rel(run_penalty(host_lib, o, 1, w, duhat=du), S.dRpen_duhat_apply(o, w, du)) < 1e-11
    assert rel(run_penalty(host_lib, o, 0, w, lam), S.dRpen_duhat_T_apply(o, w, lam)) < 1e-11
