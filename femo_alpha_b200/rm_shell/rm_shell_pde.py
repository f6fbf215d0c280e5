"""RMShellPDE: function spaces and "forms" of the Reissner-Mindlin shell -- drop-in for
femo_alpha/rm_shell/rm_shell_pde.py:21-293.  The forms are handles naming CUDA kernel families
(the physics, restated from rm_shell/linear_shell_fenicsx/, lives in femo_alpha_b200/csrc).
"""
from __future__ import annotations

import os

import numpy as np
import scipy.sparse as sp

from ..fea.utils_dolfinx import FunctionSpace, Function, ResidualForm, OutputForm
from ..shell_device import ShellDevice


class FieldForm(OutputForm):
    """Field output (von Mises stress) projected cell-wise onto DG1 (rm_shell_model.py:234-239)."""
    kind = "field"

    def __init__(self, device, name, mesh, surface="Top"):
        super().__init__(device, name)
        self.mesh, self.surface = mesh, surface

    def field_space(self, function_space):
        if tuple(function_space) != ("DG", 1):
            raise NotImplementedError("field outputs are projected onto DG1 only")
        return FunctionSpace(self.mesh, "DG1", 4 * self.device.nc, self.device)

    def project(self):
        return self.device.stress_field(self.surface)


class RMShellPDE:
    def __init__(self, mesh, element_wise_material=False, elementwise_pressure=False, quadrature_degree=6, device=0):
        self.mesh = mesh
        self.element_wise_material = element_wise_material
        self.elementwise_pressure = elementwise_pressure
        self.device = dev = ShellDevice(mesh.geometry.x, mesh.cells, element_wise_material, elementwise_pressure,
                                        quadrature_degree, device)
        # a mesh produced by refine_uniform remembers its parents: nested hierarchy -> multigrid-preconditioned solves
        self.hierarchy = None
        chain = getattr(mesh, "mg_meshes", None)
        if chain and len(chain) >= 2 and os.environ.get("FEMO_MG", "1") != "0":
            n_coarse = 3 * (2 * chain[0][0].shape[0] + 3 * chain[0][1].shape[0])      # upper bound of the coarsest dof count
            if n_coarse <= 400000:       # the coarsest level is factorised by the sparse direct solver (csrc/coarse.cu)
                from ..multigrid import ShellHierarchy
                self.hierarchy = ShellHierarchy.for_device(dev, chain, quadrature_degree, device,
                                                           element_wise_material=element_wise_material,
                                                           elementwise_pressure=elementwise_pressure)
        # no hierarchy (a mesh read from a file, e.g. the reference's wing box): sparse direct solves like the reference's
        # MUMPS, as long as the factor fits; decided at the first solve, Krylov fallback otherwise (FEMO_DIRECT=0: off)
        if self.hierarchy is None and dev.ndof <= 600000 and os.environ.get("FEMO_DIRECT", "1") != "0":
            dev.enable_direct(refine=2, lazy=True)
        nn, nc = dev.nn, dev.nc
        self.W = FunctionSpace(mesh, "W", dev.ndof, dev, "CG2CG1")
        self.VT = FunctionSpace(mesh, "VT", nc if element_wise_material else nn, dev,
                                "DG0" if element_wise_material else "CG1")
        self.VF = FunctionSpace(mesh, "VF", 3 * (nc if elementwise_pressure else nn), dev,
                                "DG0^3" if elementwise_pressure else "CG1^3")
        self.VU = FunctionSpace(mesh, "VU", 3 * nn, dev, "CG1^3")
        self._bf_sup_sizes = None

    @property
    def bf_sup_sizes(self):
        """assemble_vector(v*dx) for the scalar load space (rm_shell_pde.py:46-47): basis-function support sizes."""
        if self._bf_sup_sizes is None:
            self._bf_sup_sizes = self.device.support_sizes()
        return self._bf_sup_sizes

    # ---- forms (rm_shell_pde.py:50-166) ------------------------------------------------------
    def pdeRes(self, h, w, uhat, f, E, nu, penalty=False, dss=None, dSS=None, g=None):
        marker = None
        if penalty:
            if dss is None or getattr(dss, "marker", None) is None:
                raise ValueError("penalty=True needs the custom ds/dS measures of createCustomMeasure")
            marker = dss.marker
            self.device.set_penalty_bc(np.asarray(marker(self.mesh.geometry.x.T)).astype(np.uint8))
        self.elastic_model = True
        return ResidualForm(self.device, w, dict(thickness=h, F_solid=f, E=E, nu=nu, uhat=uhat), penalty, marker)

    def compliance(self, u_mid, uhat, h, f):
        return OutputForm(self.device, "compliance")

    def mass(self, uhat, h, rho):
        return OutputForm(self.device, "mass")

    def volume(self, uhat, h):
        return OutputForm(self.device, "volume")

    def elastic_energy(self, w, uhat, h, E):
        return OutputForm(self.device, "elastic_energy")

    def pnorm_stress(self, w, uhat, h, E, nu, dx, m=1e-6, rho=100, alpha=None, regularization=False):
        if regularization:
            raise NotImplementedError("pnorm_stress(regularization=True) is not provided")
        return OutputForm(self.device, "pnorm_stress")

    def von_Mises_stress(self, w, uhat, h, E, nu, surface='Top'):
        if surface not in ('Top', 'Mid', 'Bot'):
            raise TypeError("Unsupported surface type for stress computation.")
        return FieldForm(self.device, "stress", self.mesh, surface)

    # ---- maps (rm_shell_pde.py:194-293) --------------------------------------------------------
    def construct_force_to_pressure_map(self):
        """Vector load-space mass matrix int Pv.w dx as SciPy CSR (rm_shell_pde.py:194-209)."""
        return self.device.load_mass_matrix()

    def construct_nodal_disp_map(self):
        """(3 nn x ndof) CSR picking u_mid at the mesh vertices, stacked by component
        (rm_shell_pde.py:212-221): for nodal Q2 these are the vertex dofs."""
        nn, ndof = self.device.nn, self.device.ndof
        rows = np.arange(3 * nn)
        cols = (3 * np.arange(nn)[None, :] + np.arange(3)[:, None]).ravel()
        return sp.csr_matrix((np.ones(3 * nn), (rows, cols)), shape=(3 * nn, ndof))

    def compute_nodal_disp(self, func):
        return func.x.array.reshape(-1, 3)[:self.device.nn].copy()
