"""CPU oracle for the shape partials dR/duhat and dOutput/duhat -- SURVEY 8 row f3.

TEST INFRASTRUCTURE ONLY (same rules as oracle/rm_shell_ref.py).  PARITY UNPINNED (same reasons).

The reference obtains these partials by UFL differentiation of the residual / output forms w.r.t. the mesh
displacement function `uhat` (csdl_alpha_opt/state_operation.py:242-258,282-286; output_operation.py:64-69);
`uhat` enters through F = I + grad(uhat), gradx(.) = grad(.) F^-1 and J = det F
(rm_shell/linear_shell_fenicsx/kinematics.py:12-44; linear_shell_model.py:220,242,276,296,320).

Independent of the hand-derived formulas of the CUDA kernels, the derivatives are taken here by the
COMPLEX-STEP method on the oracle's own residual: every operation between `uhat` and the residual is analytic
(det, inverse, products), so Im R(uhat + i eps d) / eps is the exact directional derivative to round-off
(eps = 1e-30, no subtractive cancellation).
"""
from __future__ import annotations

import numpy as np

from .rm_shell_ref import clt_matrix, qp_geometry, strain_operator

EPS = 1e-30


def _cell_terms(o, w, lam, uh_cells):
    """Per-cell residual vectors R_e (nc,39) and, if lam is given, scalars lam_e . R_e, for cell-local copies of
    uhat `uh_cells` (nc,4,3), real or complex.  No penalty terms (strong BCs)."""
    t = o.topo
    xv = t.x[t.cv]
    we = w[t.cell_dofs]
    Re = np.zeros((t.nc, 39), dtype=uh_cells.dtype)
    for X0, X1, wq in o.qp:
        g = qp_geometry(xv, uh_cells, X0, X1)
        B = strain_operator(g)
        D = clt_matrix(o._mat_at(o.h, g.N1), o._mat_at(o.E, g.N1), o._mat_at(o.nu, g.N1), o.hE, g.Ju, "value")
        Re += np.einsum("c,csp,cst,ctq,cq->cp", wq * g.detJ, B, D, B, we)
        fq = o._f_at(g.N1)
        Re[:, :27] -= ((wq * g.detJ * g.Ju)[:, None, None] * g.phi2[None, :, None] * fq[:, None, :]).reshape(t.nc, 27)
    if lam is None:
        return Re
    return Re, np.einsum("cp,cp->c", Re, lam[t.cell_dofs])


def residual_at(o, w, uhat):
    t = o.topo
    Re = _cell_terms(o, w, None, np.asarray(uhat).reshape(-1, 3)[t.cv])
    R = np.zeros(t.ndof, dtype=Re.dtype)
    np.add.at(R, t.cell_dofs.ravel(), Re.ravel())
    return R


def dRduhat_apply(o, w, duhat):
    """(dR/duhat) duhat  (n_dof)."""
    uh = o.uhat.astype(complex) + 1j * EPS * np.asarray(duhat, dtype=np.float64)
    return residual_at(o, w, uh).imag / EPS


def _vertex_gradient(o, cell_scalar):
    """d/duhat of sum_c s_c(uhat_c): perturb one of the 12 cell-local uhat entries in all cells at once."""
    t = o.topo
    uh0 = o.uhat.reshape(-1, 3)[t.cv].astype(complex)
    g = np.zeros((t.nn, 3))
    for v in range(4):
        for i in range(3):
            uh = uh0.copy()
            uh[:, v, i] += 1j * EPS
            np.add.at(g[:, i], t.cv[:, v], cell_scalar(uh).imag / EPS)
    return g.ravel()


def dRduhat_T_apply(o, w, lam):
    """(dR/duhat)^T lam  (3 nn)."""
    return _vertex_gradient(o, lambda uh: _cell_terms(o, w, lam, uh)[1])


def _cell_output(o, name, w, uh_cells):
    t = o.topo
    xv = t.x[t.cv]
    out = np.zeros(t.nc, dtype=uh_cells.dtype)
    if name == "elastic_energy":
        f = o.f.copy(); o.f = np.zeros_like(f)
        out = 0.5 * _cell_terms(o, w, w, uh_cells)[1]
        o.f = f
        return out
    for X0, X1, wq in o.qp:
        g = qp_geometry(xv, uh_cells, X0, X1)
        if name == "compliance":                       # rm_shell_pde.py:85-89 (the regulariser has no uhat)
            u = np.einsum("a,cai->ci", g.phi2, w[t.cell_dofs[:, :27]].reshape(t.nc, 9, 3))
            out += wq * g.detJ * g.Ju * np.einsum("ci,ci->c", u, u)
        elif name == "mass":                           # rm_shell_pde.py:101-102
            out += wq * g.detJ * g.Ju * o._mat_at(o.rho, g.N1) * o._mat_at(o.h, g.N1)
        else:
            raise KeyError(name)
    return out


def doutput_duhat(o, name, w):
    """d output / d uhat (3 nn) for compliance, mass, elastic_energy."""
    return _vertex_gradient(o, lambda uh: _cell_output(o, name, w, uh))


def _cell_pnorm(o, w, uh_cells, m, rho, cells=None):
    """Per-cell  int (m vm)^rho J dx  (degree-4 rule), uhat cell-local and possibly complex (rm_shell_pde.py:112-121)."""
    from .rm_shell_ref import _von_mises_at, gauss01
    t = o.topo
    xv = t.x[t.cv]
    we = w[t.cell_dofs]
    pick = (lambda a: a) if o.ewm else (lambda a: a[t.cv])
    gx, gw = gauss01(3)
    mask = np.ones(t.nc) if cells is None else np.isin(np.arange(t.nc), cells).astype(float)
    out = np.zeros(t.nc, dtype=uh_cells.dtype)
    for i in range(3):
        for j in range(3):
            g = qp_geometry(xv, uh_cells, gx[i], gx[j])
            vm = _von_mises_at(g, we.astype(uh_cells.dtype), pick(o.h), pick(o.E), pick(o.nu), o.ewm)
            out += mask * gw[i] * gw[j] * g.detJ * g.Ju * (m * vm) ** rho
    return out


def dpnorm_duhat(o, w, m=1e-6, rho=100, cells=None):
    """d pnorm_stress / d uhat (3 nn); the area normalisation alpha is a constant of the form (pde:122-127)."""
    t = o.topo
    xv = t.x[t.cv]
    mask = np.ones(t.nc) if cells is None else np.isin(np.arange(t.nc), cells).astype(float)
    area = 0.0
    for X0, X1, wq in o.qp:
        area += np.sum(mask * wq * qp_geometry(xv, np.zeros_like(xv), X0, X1).detJ)
    return _vertex_gradient(o, lambda uh: _cell_pnorm(o, w, uh, m, rho, cells)) / area


# ---- penalty-BC facet terms (linear_shell_model.py:323-333): Nanson's factor |J F^-T N| depends on uhat -----------------
def penalty_residual_at(o, w, uhat):
    """R_pen(w; uhat) = P(uhat) w with the oracle's own penalty matrix, uhat real or complex."""
    saved = o.uhat
    try:
        o.uhat = np.asarray(uhat).ravel()
        return o.penalty_matrix() @ w
    finally:
        o.uhat = saved


def dRpen_duhat_apply(o, w, duhat):
    """(dR_pen/duhat) duhat  (n_dof), complex step."""
    uh = o.uhat.astype(complex) + 1j * EPS * np.asarray(duhat, dtype=np.float64)
    return penalty_residual_at(o, w, uh).imag / EPS


def dRpen_duhat_T_apply(o, w, lam):
    """(dR_pen/duhat)^T lam  (3 nn): one complex step per uhat entry of the cells that carry a penalised edge."""
    t = o.topo
    sel = np.zeros(t.ne, dtype=bool); sel[o.pen_edges] = True
    verts = np.unique(t.cv[sel[t.ce].any(axis=1)])
    g = np.zeros(3 * t.nn)
    for v in verts:
        for i in range(3):
            uh = o.uhat.astype(complex)
            uh[3 * v + i] += 1j * EPS
            g[3 * v + i] = lam @ (penalty_residual_at(o, w, uh).imag / EPS)
    return g
