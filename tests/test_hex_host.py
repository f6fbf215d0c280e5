"""The element maths and index arithmetic of the hexahedral-solid kernels (femo_alpha_b200/csrc/hex_element.cuh,
hex_pattern.hpp) run on the CPU -- the same __host__ __device__ bodies, compiled by nvcc into a
host harness (tests/hex_host.cu) -- and checked against oracle/hex_ref.py.  No GPU needed; the GPU parity test of the
kernels through the C ABI is tests/test_gpu_hex.py."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
import scipy.sparse as sp

from oracle.hex_ref import HexOracle, box_mesh

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "..", "femo_alpha_b200", "csrc")
I64, F64 = ctypes.c_int64, ctypes.c_double


@pytest.fixture(scope="module")
def host_lib():
    out = os.path.join(HERE, "_build", "libhex_host.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    src = os.path.join(HERE, "hex_host.cu")
    deps = [src] + [os.path.join(CSRC, f) for f in ("hex_element.cuh", "hex_pattern.hpp")]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
        nvcc = "/usr/local/cuda/bin/nvcc" if os.path.exists("/usr/local/cuda/bin/nvcc") else "nvcc"
        subprocess.run([nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC",
                        "-o", out, src], check=True)
    lib = ctypes.CDLL(out)
    lib.hex_host_pattern.restype = ctypes.c_int64
    return lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


class Host:
    """the device arrays of a kind-1 context, built by the harness"""

    def __init__(self, lib, o):
        self.lib, self.o = lib, o
        self.cells = np.ascontiguousarray(o.cells, dtype=np.int32)
        self.xyz = np.ascontiguousarray(o.x)
        nb, nc = o.nn, o.nc
        nnzb = lib.hex_host_pattern(I64(nb), I64(nc), _p(self.cells), None, None, None, None, None)
        assert nnzb > 0
        self.nc_ptr = np.zeros(nb + 1, np.int32); self.nc_adj = np.zeros(nc * 8, np.int32)
        self.rowptr = np.zeros(nb + 1, np.int32); self.colidx = np.zeros(nnzb, np.int32); self.diag = np.zeros(nb, np.int32)
        assert lib.hex_host_pattern(I64(nb), I64(nc), _p(self.cells), _p(self.nc_ptr), _p(self.nc_adj), _p(self.rowptr),
                                    _p(self.colidx), _p(self.diag)) == nnzb
        self.geo = np.zeros(nc * 80); self.vol = np.zeros(nc)
        assert lib.hex_host_geom(I64(nc), _p(self.xyz), _p(self.cells), _p(self.geo), _p(self.vol)) == 0
        self.form = 1 if o.form == "elasticity" else 0

    def matrix(self, gradient_records=False):
        o = self.o
        vals = np.zeros(self.colidx.size * 9)
        rho = np.ascontiguousarray(o.rho)
        fn = self.lib.hex_host_assemble_g if gradient_records else self.lib.hex_host_assemble
        assert fn(self.form, F64(o.c1), F64(o.c2), F64(o.p), I64(o.nn), _p(self.geo), _p(self.cells),
                                          _p(rho), I64(o.nc), _p(self.nc_ptr), _p(self.nc_adj), _p(self.rowptr),
                                          _p(self.colidx), _p(vals)) == 0
        return sp.bsr_matrix((vals.reshape(-1, 3, 3), self.colidx, self.rowptr), shape=(o.ndof, o.ndof)).tocsr()

    def cellvec(self, mode, x, cellcoef=None, coef=1.0):
        o = self.o
        out = np.zeros(o.ndof)
        x = np.ascontiguousarray(x, dtype=float)
        cc = None if cellcoef is None else np.ascontiguousarray(cellcoef, dtype=float)
        assert self.lib.hex_host_cellvec_gather(mode, self.form, F64(o.c1), F64(o.c2), I64(o.nn), I64(o.nc), _p(self.geo),
                                                _p(self.cells), _p(x), _p(cc), F64(coef), _p(self.nc_ptr), _p(self.nc_adj),
                                                _p(out)) == 0
        return out

    def energy(self, u, lam):
        o = self.o
        out = np.zeros(o.nc)
        u = np.ascontiguousarray(u); lam = np.ascontiguousarray(lam); rho = np.ascontiguousarray(o.rho)
        assert self.lib.hex_host_energy(self.form, F64(o.c1), F64(o.c2), F64(o.p), I64(o.nc), _p(self.geo), _p(self.cells), _p(u),
                                        _p(lam), _p(rho), _p(out)) == 0
        return out


def rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / np.abs(np.asarray(b)).max()


def problem(form, shape=(3, 4, 2), distort=0.25):
    nodes, cells = box_mesh(*shape, distort=distort, seed=4)
    o = HexOracle(nodes, cells, form=form, E0=5.0, nu=0.3, simp_p=3.0)
    rng = np.random.default_rng(11)
    o.set_inputs(density=rng.uniform(0.05, 1.0, o.nc), body_force=rng.standard_normal((o.nn, 3)))
    return o, rng


@pytest.mark.parametrize("form", ["elasticity", "vector_poisson"])
def test_pattern_geometry_and_matrix(host_lib, form):
    o, rng = problem(form)
    h = Host(host_lib, o)
    K = o.stiffness()
    # pattern = node graph of the mesh, sorted rows, diagonal positions
    G = sp.csr_matrix((np.ones(o.nc * 64), (np.repeat(o.cells, 8, axis=1).ravel(), np.tile(o.cells, (1, 8)).ravel())),
                      shape=(o.nn, o.nn))
    G.sum_duplicates(); G.sort_indices()
    assert np.array_equal(h.rowptr, G.indptr) and np.array_equal(h.colidx, G.indices)
    assert np.array_equal(h.colidx[h.diag], np.arange(o.nn))
    assert rel(h.vol, o.vol) < 1e-13
    A = h.matrix()
    assert abs(A - K).max() < 1e-12 * abs(K).max()
    x = rng.standard_normal(o.ndof)
    assert rel(A @ x, K @ x) < 1e-12


@pytest.mark.parametrize("form", ["elasticity", "vector_poisson"])
def test_element_vectors_and_density_partials(host_lib, form):
    o, rng = problem(form, shape=(2, 3, 3))
    h = Host(host_lib, o)
    u = rng.standard_normal(o.ndof); lam = rng.standard_normal(o.ndof); d = rng.standard_normal(o.nc)
    assert rel(h.cellvec(0, o.f.ravel()), o.load()) < 1e-12                       # F = M f
    assert rel(h.cellvec(0, u, coef=-1.0), -(o.mass_matrix() @ u)) < 1e-12        # dR/df action
    assert rel(h.cellvec(1, u, cellcoef=o.rho ** o.p), o.stiffness() @ u) < 1e-12  # matrix-free K u
    assert rel(h.cellvec(1, u, cellcoef=o.p * o.rho ** (o.p - 1) * d), o.dRdrho(u, d)) < 1e-12
    assert rel(h.energy(u, lam), o.dRdrho_T(u, lam)) < 1e-12


def test_unstructured_valence(host_lib):
    # three hexahedra around an edge (valence-3 column): rows of different lengths, more than one block per lane stride
    th = np.deg2rad([0, 120, 240])
    ring = np.stack([np.cos(th), np.sin(th)], 1)
    mid = 0.5 * (ring + np.roll(ring, -1, axis=0))
    pts2 = np.vstack([[0, 0], ring, mid])            # 0 centre, 1-3 corners, 4-6 edge mids
    nodes = np.vstack([np.c_[pts2, np.zeros(7)], np.c_[pts2, np.ones(7)]])
    quads = [[0, 4 + 2, 1, 4 + 0], [0, 4 + 0, 2, 4 + 1], [0, 4 + 1, 3, 4 + 2]]
    cells = np.array([q + [v + 7 for v in q] for q in quads], dtype=np.int32)
    o = HexOracle(nodes, cells, E0=2.0, nu=0.2)
    o.set_inputs(density=np.array([0.3, 0.6, 1.0]))
    h = Host(host_lib, o)
    K = o.stiffness()
    assert abs(h.matrix() - K).max() < 1e-12 * abs(K).max()


def fan_mesh(n_sectors=10):
    """n hexahedra around one edge (extruded kites): the axis nodes see 8 n > 64 work items -> several staging passes."""
    th = 2 * np.pi * np.arange(n_sectors) / n_sectors
    ring = np.stack([np.cos(th), np.sin(th)], 1)
    outer = 1.8 * np.stack([np.cos(th + np.pi / n_sectors), np.sin(th + np.pi / n_sectors)], 1)
    pts2 = np.vstack([[0, 0], ring, outer])
    m = len(pts2)
    nodes = np.vstack([np.c_[pts2, np.zeros(m)], np.c_[pts2, np.full(m, 0.7)]])
    quads = [[0, 1 + k, 1 + n_sectors + k, 1 + (k + 1) % n_sectors] for k in range(n_sectors)]
    cells = np.array([q + [v + m for v in q] for q in quads], dtype=np.int32)
    return nodes, cells


def test_high_valence_rows_take_several_passes(host_lib):
    nodes, cells = fan_mesh(10)
    o = HexOracle(nodes, cells, E0=2.0, nu=0.2)
    o.set_inputs(density=np.linspace(0.2, 1.0, o.nc))
    h = Host(host_lib, o)
    assert (h.nc_ptr[1] - h.nc_ptr[0]) * 8 == 80
    K = o.stiffness()
    assert abs(h.matrix() - K).max() < 1e-12 * abs(K).max()


@pytest.mark.parametrize("form", ["elasticity", "vector_poisson"])
def test_gradient_record_variant(host_lib, form):
    """FEMO_HEX_GRAD=1 (k_hex_geom2 + k_hex_assemble_g): same matrix from the gradient records, on a box, on the
    valence-3 column and on the ten-cell fan (two chunks of sub-passes)."""
    o, rng = problem(form)
    K = o.stiffness()
    assert abs(Host(host_lib, o).matrix(gradient_records=True) - K).max() < 1e-12 * abs(K).max()
    nodes, cells = fan_mesh(10)
    o = HexOracle(nodes, cells, form=form, E0=2.0, nu=0.2)
    o.set_inputs(density=np.linspace(0.2, 1.0, o.nc))
    K = o.stiffness()
    assert abs(Host(host_lib, o).matrix(gradient_records=True) - K).max() < 1e-12 * abs(K).max()
