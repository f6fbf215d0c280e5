"""Free functions of femo_alpha/fea/utils_dolfinx.py re-stated on top of libfemo_b200.

The reference hands UFL forms to dolfinx; here a "form" is a small symbolic handle
(`ResidualForm`, `OutputForm`, `DerivativeForm`) that names a kernel family of the CUDA
library, and a "function" is a host mirror bound to a device buffer.  Names, argument
meaning and error behaviour follow the reference (file:line in each docstring) so scripts
written against `femo_alpha.fea.utils_dolfinx` keep working for the RM-shell path.
"""
from __future__ import annotations

import numpy as np

from ..mesh import (ShellMesh, createRectangleMesh, createUnitSquareMesh, reconstructFEAMesh,  # noqa: F401
                    readPickleMesh, refine_uniform)

DOLFIN_EPS = 3E-16


# ---- spaces / functions -----------------------------------------------------------------
class FunctionSpace:
    """Descriptor of one of the spaces of RMShellPDE (rm_shell_pde.py:25-47)."""

    def __init__(self, mesh, kind, size, device=None, label=None):
        self.mesh, self.kind, self.size, self.device, self.label = mesh, kind, int(size), device, label
        self.num_sub_spaces = 2 if kind == "W" else (3 if kind in ("VF", "VU") else 0)

    def sub(self, i):
        if self.kind != "W":
            raise ValueError("sub() is only provided for the mixed state space W")
        return _SubSpace(self, i)

    def collapse(self):
        return self, np.arange(self.size)

    def tabulate_dof_coordinates(self):
        if self.kind in ("W", "W3"):
            return np.repeat(self.device.block_coords(), 3, axis=0)
        raise NotImplementedError


class _SubSpace:
    def __init__(self, parent, i):
        self.parent, self.index = parent, i

    def dof_range(self):
        d = self.parent.device
        return (0, 3 * d.nq2) if self.index == 0 else (3 * d.nq2, d.ndof)

    def collapse(self):
        lo, hi = self.dof_range()
        return FunctionSpace(self.parent.mesh, "Wsub%d" % self.index, hi - lo, self.parent.device), np.arange(lo, hi)


class _XArray:
    def __init__(self, n):
        self.array = np.zeros(n)

    def scatter_forward(self):
        pass


class _VectorView:
    """Just enough of petsc4py.Vec for the reference's call sites (getArray / set / [:] / array)."""

    def __init__(self, fn):
        self._fn = fn

    def getArray(self):
        return self._fn.x.array

    @property
    def array(self):
        return self._fn.x.array

    def set(self, v):
        self._fn.x.array[:] = v
        self._fn._push()

    def setArray(self, a):
        self.__setitem__(slice(None), a)

    def __setitem__(self, k, v):
        self._fn.x.array[k] = v
        self._fn._push()

    def __getitem__(self, k):
        return self._fn.x.array[k]

    def assemble(self):
        pass

    def ghostUpdate(self, *a, **k):
        pass


class Function:
    """Host mirror of a device vector (dolfinx.fem.Function stand-in)."""

    def __init__(self, V, name=None):
        self.function_space = V
        self.x = _XArray(V.size)
        self.name = name
        self._role = None            # ('input', name) | ('state',) | None

    @property
    def vector(self):
        return _VectorView(self)

    def _push(self):
        dev = self.function_space.device
        if dev is None or self._role is None:
            return
        if self._role[0] == "input":
            dev.set_input(self._role[1], self.x.array)
        elif self._role[0] == "state":
            dev.set_state(self.x.array)

    def sub(self, i):
        return _SubFunction(self, i)

    def split(self):
        return self.sub(0).collapse(), self.sub(1).collapse()


class _SubFunction:
    def __init__(self, fn, i):
        self.fn, self.i = fn, i

    def collapse(self):
        V, idx = self.fn.function_space.sub(self.i).collapse()
        out = Function(V)
        out.x.array[:] = self.fn.x.array[idx]
        return out


def createFunction(function):
    """utils_dolfinx.py:335-336"""
    return Function(function.function_space)


def getFuncArray(v):
    """utils_dolfinx.py:174-178"""
    return v.x.array


def setFuncArray(v, v_array):
    """utils_dolfinx.py:180-186"""
    v.x.array[:] = np.asarray(v_array, dtype=np.float64).ravel()
    v._push()


def update(v, v_values):
    """utils_dolfinx.py:319-330: a length-1 array broadcasts."""
    v_values = np.asarray(v_values, dtype=np.float64).ravel()
    if len(v_values) == 1:
        v.x.array[:] = v_values[0]
        v._push()
    else:
        if v_values.size != v.x.array.size:
            raise ValueError("update: got %d values for a function with %d dofs" % (v_values.size, v.x.array.size))
        setFuncArray(v, v_values)


# ---- boundary conditions ------------------------------------------------------------------
class DirichletBC:
    def __init__(self, dofs, value=0.0):
        self._dofs = np.asarray(dofs, dtype=np.int32)
        self.value = value

    def dof_indices(self):
        return self._dofs, len(self._dofs)


def locate_dofs_geometrical(V, marker):
    """dolfinx.fem.locate_dofs_geometrical on W.sub(0) / W.sub(1) (rm_shell_model.py:168-173):
    dofs whose coordinates satisfy `marker(x)` (x has shape (3, n))."""
    sub = V[0] if isinstance(V, (tuple, list)) else V
    if isinstance(sub, FunctionSpace) and sub.kind == "W3":       # solid forms (fea/named_forms.py): 3 dofs per node
        blocks = np.nonzero(np.asarray(marker(sub.device.block_coords().T)).astype(bool))[0]
        return np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()).astype(np.int32)
    if not isinstance(sub, _SubSpace):
        raise ValueError("locate_dofs_geometrical expects W.sub(i) (optionally with its collapsed space)")
    dev = sub.parent.device
    xc = dev.block_coords()
    lo, hi = (0, dev.nq2) if sub.index == 0 else (dev.nq2, dev.nb)
    on = np.asarray(marker(xc[lo:hi].T)).astype(bool)
    blocks = lo + np.nonzero(on)[0]
    return np.sort((3 * blocks[:, None] + np.arange(3)[None, :]).ravel()).astype(np.int32)


def dirichletbc(ubc, dofs, V=None):
    return DirichletBC(dofs[0] if isinstance(dofs, (tuple, list)) else dofs)


class Measure:
    """Tagged facet measure created by createCustomMeasure (utils_dolfinx.py:555-565)."""

    def __init__(self, mesh, measure, marker, tag):
        self.mesh, self.measure, self.marker, self.tag = mesh, measure, marker, tag
        self.metadata = {"quadrature_degree": 4}

    def __call__(self, tag):
        return self


def createCustomMeasure(mesh, dim, SubdomainFunc, measure: str, tag: int):
    return Measure(mesh, measure, SubdomainFunc, tag)


# ---- forms -------------------------------------------------------------------------------
class ResidualForm:
    kind = "residual"

    def __init__(self, device, state, inputs, penalty=False, bc_marker=None):
        self.device, self.state, self.inputs, self.penalty, self.bc_marker = device, state, inputs, penalty, bc_marker


class OutputForm:
    kind = "output"

    def __init__(self, device, name, state=None):
        self.device, self.name, self.state = device, name, state


class DerivativeForm:
    kind = "derivative"

    def __init__(self, form, function):
        self.form, self.function = form, function


def derivative(form, function, du=None):
    return DerivativeForm(form, function)


def computePartials(form, function):
    """utils_dolfinx.py:332-333"""
    return derivative(form, function)


def _arg_of(function):
    role = function._role
    if role is None:
        raise ValueError("function is not registered with the FEA (add_input/add_state)")
    return "disp_solid" if role[0] == "state" else role[1]


# ---- operators returned by assembleMatrix -------------------------------------------------
class JacobianMat:
    """dR/du on the device (BSR 3x3); `bcs` -> dolfinx BC convention applied on the fly."""

    def __init__(self, device, with_bc):
        self.device, self.with_bc = device, bool(with_bc)

    def mult(self, x):
        return self.device.jacobian_apply(x, self.with_bc)

    multTranspose = mult      # symmetric

    def getSizes(self):
        return self.device.ndof, self.device.ndof

    def getValuesCSR(self):
        A = self.device.get_csr(self.with_bc)
        return A.indptr, A.indices, A.data

    def toscipy(self):
        return self.device.get_csr(self.with_bc)


class PartialMat:
    """dR/d(arg) as a matrix-free operator (the reference assembles a rectangular AIJ matrix,
    state_operation.py:282-286)."""

    def __init__(self, device, arg):
        self.device, self.arg = device, arg

    def mult(self, df):
        return self.device.dRdf_apply(self.arg, df)

    def multTranspose(self, lam):
        return self.device.dRdf_T_apply(self.arg, lam)


class KSP:
    """Stand-in for the PETSc KSP(preonly + LU/MUMPS) objects of setUpKSP_MUMPS
    (utils_dolfinx.py:514-531): block-Jacobi PCG on the device."""

    def __init__(self, A, rtol=1e-12, max_it=200000):
        self.A, self.rtol, self.max_it = A, rtol, max_it
        self.its, self.relres = 0, 0.0

    def solve(self, b, x):
        """Solves A x = b. `b`, `x` are Functions or their `.vector` views (PETSc argument order)."""
        bf = b._fn if isinstance(b, _VectorView) else b
        xf = x._fn if isinstance(x, _VectorView) else x
        sol = self.A.device.solve_forward(bf.x.array, self.rtol, self.max_it)
        self.its, self.relres = self.A.device.last_solve["iters"], self.A.device.last_solve["relres"]
        xf.x.array[:] = sol


def setUpKSP_MUMPS(A):
    return KSP(A)


def assembleMatrix(M, bcs=[]):
    """utils_dolfinx.py:200-206"""
    if not isinstance(M, DerivativeForm) or M.form.kind != "residual":
        raise TypeError("assembleMatrix: only derivative(residual_form, function) is supported on this path")
    dev = M.form.device
    arg = _arg_of(M.function)
    if arg == "disp_solid":
        dev.assemble_jacobian()
        return JacobianMat(dev, with_bc=len(bcs) > 0)
    return PartialMat(dev, arg)


def assembleVector(v):
    """utils_dolfinx.py:194-198"""
    if isinstance(v, ResidualForm):
        return v.device.assemble_residual(v.state.x.array)
    if isinstance(v, DerivativeForm) and v.form.kind == "output":
        if v.form.name == "pnorm_stress":
            return v.form.device.pnorm_stress_grad(_arg_of(v.function), getattr(v.form, "tag", -1))
        return v.form.device.output_grad(v.form.name, _arg_of(v.function))
    raise TypeError("assembleVector: unsupported form")


def assembleScalar(c):
    """utils_dolfinx.py:188-192"""
    if not isinstance(c, OutputForm):
        raise TypeError("assembleScalar: unsupported form")
    if c.name == "pnorm_stress":
        return c.device.pnorm_stress(getattr(c, "tag", -1))
    return c.device.output(c.name)


def assembleSystem(J, F, bcs=[]):
    """utils_dolfinx.py:208-221: (A with BCs, b with lifting + set_bc)."""
    A = assembleMatrix(J, bcs=bcs if len(bcs) else [None])
    A.with_bc = len(bcs) > 0
    b = assembleVector(F)
    for bc in bcs:
        idx = bc.dof_indices()[0]
        b[idx] = F.state.x.array[idx]          # set_bc with homogeneous g: x - g
    return A, b


def assemble(f, dim=0, bcs=[]):
    """utils_dolfinx.py:223-232"""
    if dim == 0:
        return assembleScalar(f)
    elif dim == 1:
        return assembleVector(f)
    elif dim == 2:
        return convertToDense(assembleMatrix(f, bcs=bcs))
    else:
        return TypeError('Invalid type for assembly.')


def convertToDense(A):
    """utils_dolfinx.py:309-316 (debugging only)"""
    return A.toscipy().toarray()


def computeMatVecProductFwd(A, x):
    """utils_dolfinx.py:275-283: y = A x (x is a Function)"""
    return A.mult(x.x.array)


def computeMatVecProductBwd(A, R):
    """utils_dolfinx.py:294-306: y = A^T R (R is a Function)"""
    return A.multTranspose(R.x.array)


# ---- mesh readers (utils_dolfinx.py:34-50,88-142,653-668): host side, femo_alpha_b200/mesh_io.py --------------------
from ..mesh_io import readFEAMesh, import_mesh, writeFEAMesh, read_msh, readMshShellMesh, readMshSolidMesh   # noqa: E402,F401
from ..mesh import reconstructFEAMesh, createRectangleMesh, createUnitSquareMesh          # noqa: E402,F401
