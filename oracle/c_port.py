"""ctypes wrapper of oracle/rm_shell_c.c (C + OpenMP CPU baseline). TEST/BASELINE INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "librm_shell_c.so")
_lib = None


def build():
    src = os.path.join(_HERE, "rm_shell_c.c")
    if not os.path.exists(_SO) or os.path.getmtime(src) > os.path.getmtime(_SO):
        # -march=x86-64-v3 (not native): the .so travels from the build container to the GPU box
        subprocess.check_call(["gcc", "-O3", "-march=x86-64-v3", "-fopenmp", "-fPIC", "-std=gnu11", "-shared",
                               "-o", _SO, src, "-lm"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.rmc_assemble_csr.restype = C.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def use_all_cores():
    """Let the OpenMP loops of both ports use every core this process may run on, whatever OMP_NUM_THREADS says
    (torch.distributed.run exports OMP_NUM_THREADS=1 to its workers).  Returns the thread count."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lib().rmc_set_threads(C.c_int(n))
    hexlib().hexc_set_threads(C.c_int(n))
    return n


def element_matrices(topo, h, E, nu, uhat, ewm, m=4):
    cv = np.ascontiguousarray(topo.cv, dtype=np.int32)
    out = np.empty((topo.nc, 39, 39))
    uh = None if uhat is None else np.ascontiguousarray(uhat, dtype=np.float64)
    lib().rmc_element_matrices(C.c_int64(topo.nc), _p(np.ascontiguousarray(topo.x)), _p(cv), _p(h), _p(E), _p(nu),
                               _p(uh), C.c_int(int(ewm)), C.c_int(m), _p(out))
    return out


class CsrAssembler:
    """Pattern + element->CSR map built once (as dolfinx does in create_matrix); `assemble` is the timed part."""

    def __init__(self, topo):
        import scipy.sparse as sp
        self.topo = topo
        rows = np.repeat(topo.cell_dofs, 39, axis=1).ravel()
        cols = np.tile(topo.cell_dofs, (1, 39)).ravel()
        n = topo.ndof
        key = rows.astype(np.int64) * n + cols
        ukey, inv = np.unique(key, return_inverse=True)
        self.emap = np.ascontiguousarray(inv.astype(np.int64))
        self.nnz = ukey.size
        r = (ukey // n).astype(np.int64)
        self.indices = np.ascontiguousarray((ukey % n).astype(np.int32))
        self.indptr = np.zeros(n + 1, dtype=np.int64)
        np.add.at(self.indptr, r + 1, 1)
        self.indptr = np.cumsum(self.indptr)
        self.vals = np.zeros(self.nnz)
        self.cv = np.ascontiguousarray(topo.cv, dtype=np.int32)
        self.x = np.ascontiguousarray(topo.x)

    def assemble(self, h, E, nu, uhat=None, ewm=False, m=4):
        uh = None if uhat is None else np.ascontiguousarray(uhat, dtype=np.float64)
        return lib().rmc_assemble_csr(C.c_int64(self.topo.nc), _p(self.x), _p(self.cv), _p(h), _p(E), _p(nu), _p(uh),
                                      C.c_int(int(ewm)), C.c_int(m), _p(self.emap), C.c_int64(self.nnz), _p(self.vals))

    def matrix(self):
        import scipy.sparse as sp
        return sp.csr_matrix((self.vals, self.indices, self.indptr), shape=(self.topo.ndof,) * 2)


# ---- oracle/hex_c.c: generic 3-D forms on hexahedra (config 5) -----------------------------------------------------------
_HEX_SO = os.path.join(_HERE, "libhex_c.so")
_hexlib = None


def build_hex():
    src = os.path.join(_HERE, "hex_c.c")
    if not os.path.exists(_HEX_SO) or os.path.getmtime(src) > os.path.getmtime(_HEX_SO):
        subprocess.check_call(["gcc", "-O3", "-march=x86-64-v3", "-fopenmp", "-fPIC", "-std=gnu11", "-shared",
                               "-o", _HEX_SO, src, "-lm"])
    return _HEX_SO


def hexlib():
    global _hexlib
    if _hexlib is None:
        _hexlib = C.CDLL(build_hex())
    return _hexlib


class HexCsrAssembler:
    """Pattern + element -> CSR map of a HexOracle built once (as dolfinx does in create_matrix); `assemble` is the
    timed part (hexc_assemble_csr, OpenMP over the cells)."""

    def __init__(self, o):
        self.o = o
        ed = (3 * o.cells[:, :, None] + np.arange(3)[None, None, :]).reshape(o.nc, 24)
        rows = np.repeat(ed, 24, axis=1).ravel()
        cols = np.tile(ed, (1, 24)).ravel()
        n = o.ndof
        ukey, inv = np.unique(rows.astype(np.int64) * n + cols, return_inverse=True)
        self.emap = np.ascontiguousarray(inv.astype(np.int64))
        self.nnz = ukey.size
        self.indices = np.ascontiguousarray((ukey % n).astype(np.int32))
        self.indptr = np.zeros(n + 1, dtype=np.int64)
        np.add.at(self.indptr, (ukey // n) + 1, 1)
        self.indptr = np.cumsum(self.indptr)
        self.vals = np.zeros(self.nnz)
        self.xyz = np.ascontiguousarray(o.x)
        self.cells = np.ascontiguousarray(o.cells, dtype=np.int32)

    def assemble(self):
        o = self.o
        rho = np.ascontiguousarray(o.rho)
        hexlib().hexc_assemble_csr(C.c_int64(o.nc), _p(self.xyz), _p(self.cells), _p(rho), C.c_double(o.p),
                                   C.c_int(1 if o.form == "elasticity" else 0), C.c_double(o.c1), C.c_double(o.c2),
                                   _p(self.emap), C.c_int64(self.nnz), _p(self.vals))
        return self.vals

    def matrix(self):
        import scipy.sparse as sp
        return sp.csr_matrix((self.assemble().copy(), self.indices, self.indptr), shape=(self.o.ndof, self.o.ndof))
