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
