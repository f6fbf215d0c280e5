"""SymPy cross-check of the hand-derived element maths on ONE cell (TEST INFRASTRUCTURE ONLY, parity unpinned).

The reference never writes a B-matrix: it states the ENERGY in UFL and lets UFL differentiate it
(femo_alpha/rm_shell/linear_shell_fenicsx/linear_shell_model.py:220-321, kinematics.py:12-106; fea_dolfinx.py:81-82
`derivative(residual_form, function)`).  This module does the same with SymPy instead of UFL:

  * geometry, frames, pseudo-inverse, F(uhat), its inverse and determinant are SymPy expressions of the reference
    point (X0, X1); every spatial derivative -- including the derivative of the cell normal inside
    grad(cross(E2, theta)) on a warped quad -- is taken by `sympy.diff`, not by a formula;
  * the basis functions are built here from their nodes (Lagrange interpolation), their derivatives by `sympy.diff`;
  * the energy density is the literal formula in the 11 strain-like quantities the UFL names (t_gu, kappa before
    symmetrisation, gamma, theta.E2); its gradient and Hessian with respect to them, and their thickness derivative,
    are `sympy.diff` results; the strain quantities are linear in the 39 element dofs and their Jacobian is again a
    `sympy` derivative.  K_e, R_e and dR_e/dh_b follow by the chain rule through that linear map.

Nothing is shared with oracle/rm_shell_ref.py (strain_operator, clt_matrix, basis tables) except the element dof order
(9 Q2 nodes x 3, then 4 Q1 nodes x 3; Q2 nodes: 4 vertices in tensor order, 4 edge mid-points, centre) and the Gauss rule.
tests/test_oracle.py::test_sympy_* compares the two to 1e-13 on a distorted, warped cell with uhat != 0."""
import numpy as np
import sympy as sp

SHEAR_K = sp.Rational(833, 1000)          # linear_shell_model.py:149 (0.833, not 5/6)
X0, X1 = sp.symbols("X0 X1", real=True)
# reference positions of the local nodes: Q1 vertices in tensor order; Q2 = vertices, edge mid-points, centre
_Q1_REF = [(0, 0), (1, 0), (0, 1), (1, 1)]
_Q2_REF = _Q1_REF + [(sp.Rational(1, 2), 0), (0, sp.Rational(1, 2)), (1, sp.Rational(1, 2)), (sp.Rational(1, 2), 1),
                     (sp.Rational(1, 2), sp.Rational(1, 2))]


def _lagrange_1d(nodes, k, x):
    e = sp.Integer(1)
    for j, xj in enumerate(nodes):
        if j != k:
            e *= (x - xj) / (nodes[k] - xj)
    return e


def _tensor_basis(ref, nodes_1d):
    return [_lagrange_1d(nodes_1d, nodes_1d.index(a), X0) * _lagrange_1d(nodes_1d, nodes_1d.index(b), X1) for a, b in ref]


N1 = _tensor_basis(_Q1_REF, [0, 1])
PHI2 = _tensor_basis(_Q2_REF, [0, sp.Rational(1, 2), 1])


def _num(a):
    """exact rationals of the double-precision inputs: the symbolic side sees the same numbers as the oracle"""
    a = np.asarray(a, dtype=np.float64)
    return sp.Matrix(a.shape[0], a.shape[1], lambda i, j: sp.Rational(float(a[i][j])))


class SymbolicCell:
    """One quad (vertices `xv` (4,3) in tensor order) with mesh motion `uh` (4,3).

    Kept symbolic in (X0, X1): the position, the Jacobian, the normal E2(X), the fields u(X), theta(X), uhat(X) and
    beta(X) = cross(E2(X), theta(X)); their reference derivatives are `jacobian` (= sympy.diff) results.  `strains_at`
    substitutes a quadrature point (30 digits) and only then carries out the matrix algebra of kinematics.py (products
    with the pseudo-inverse and with F^-1): pushing symbolic inverses through every expression costs minutes."""

    def __init__(self, xv, uh):
        self.w = sp.symbols("w0:39", real=True)
        xvm, uhm = _num(xv), _num(uh)
        self.Xs = Xs = sp.Matrix([X0, X1])
        x = sum((N1[v] * xvm[v, :].T for v in range(4)), sp.zeros(3, 1))
        uhat = sum((N1[v] * uhm[v, :].T for v in range(4)), sp.zeros(3, 1))
        self.Jm = x.jacobian(Xs)                                   # 3 x 2
        n = self.Jm[:, 0].cross(self.Jm[:, 1])
        self.E2 = n / sp.sqrt(n.dot(n))                            # CellNormal, a function of X on a warped quad
        w = self.w
        self.u = sum((PHI2[a] * sp.Matrix(w[3 * a:3 * a + 3]) for a in range(9)), sp.zeros(3, 1))
        self.theta = sum((N1[v] * sp.Matrix(w[27 + 3 * v:30 + 3 * v]) for v in range(4)), sp.zeros(3, 1))
        self.du_dX = self.u.jacobian(Xs)
        self.duhat_dX = uhat.jacobian(Xs)
        self.beta = self.E2.cross(self.theta)
        self.dbeta_dX = self.beta.jacobian(Xs)                     # carries dE2/dX: taken by sympy, not by a formula

    def strains_at(self, q0, q1, digits=30):
        """(B (11 x 39), detJ, Ju) at one reference point"""
        pt = {X0: sp.Rational(float(q0)), X1: sp.Rational(float(q1))}
        ev = lambda M: M.subs(pt).evalf(digits)
        Jm, E2 = ev(self.Jm), ev(self.E2)
        Gm = Jm.T * Jm
        detJ = sp.sqrt(Gm.det())
        Kp = Gm.inv() * Jm.T                                       # pseudo-inverse: grad f = df/dX Kp
        E0 = Jm[:, 0] / sp.sqrt(Jm[:, 0].dot(Jm[:, 0]))            # kinematics.py:66-67
        E1 = E2.cross(E0)
        T = sp.Matrix([E0.T, E1.T])                                # 2 x 3, kinematics.py:72-82
        F = sp.eye(3) + ev(self.duhat_dX) * Kp                     # kinematics.py:34-44
        Finv, Ju = F.inv(), F.det()                                # kinematics.py:24-32
        gu = ev(self.du_dX) * Kp * Finv                            # gradx(u_mid), kinematics.py:12-21
        t_gu = T * gu * T.T                                        # model:220-222
        kap_m = T * (ev(self.dbeta_dX) * Kp * Finv) * T.T          # model:241-243
        gamma = T * (-ev(self.beta)) + T * (gu.T * E2)             # model:245-258
        th_n = ev(self.theta).dot(E2)                              # model:288-289
        strains = sp.Matrix([t_gu[0, 0], t_gu[0, 1], t_gu[1, 0], t_gu[1, 1], kap_m[0, 0], kap_m[0, 1], kap_m[1, 0],
                             kap_m[1, 1], gamma[0], gamma[1], th_n]).expand()
        B = strains.jacobian(sp.Matrix(self.w))                    # the strain-like quantities are linear in w
        return np.array(B.tolist(), dtype=np.float64), float(detJ), float(Ju)


def literal_energy():
    """(e, s, h, E, nu, hE, Ju): the energy density of linear_shell_model.py:260-296 in the strain-like symbols s"""
    s = sp.symbols("s0:11", real=True)
    h, E, nu, hE, Ju = sp.symbols("h E nu hE Ju", positive=True)
    t_gu = sp.Matrix([[s[0], s[1]], [s[2], s[3]]])
    kap_m = sp.Matrix([[s[4], s[5]], [s[6], s[7]]])
    gamma = sp.Matrix([s[8], s[9]])
    eps = (t_gu + t_gu.T) / 2                                       # model:232-239 (offset = 0)
    kappa = (kap_m + kap_m.T) / 2                                   # model:241-243
    voigt = lambda S: sp.Matrix([S[0, 0], S[1, 1], 2 * S[0, 1]])    # kinematics.py:94-106
    C = E / (1 - nu ** 2) * sp.Matrix([[1, nu, 0], [nu, 1, 0], [0, 0, (1 - nu) / 2]])      # model:136-157
    G = E / 2 / (1 + nu)
    A, D, A_s = h * C, h ** 3 / 12 * C, SHEAR_K * G * h * sp.eye(2)
    N, M, Q = A * voigt(eps), D * voigt(kappa), A_s * gamma         # model:260-273
    drill = (t_gu[0, 1] - t_gu[1, 0]) / 2 + s[10]                   # model:288-289
    e = sp.Rational(1, 2) * Q.dot(gamma) * Ju                       # shear     model:275-276
    e += sp.Rational(1, 2) * N.dot(voigt(eps))                      # membrane  model:278-279 (no Ju: SURVEY B3)
    e += sp.Rational(1, 2) * M.dot(voigt(kappa))                    # bending   model:281-282 (no Ju)
    e += sp.Rational(1, 2) * (E * h ** 3 * drill / hE ** 2) * drill * Ju      # drilling  model:284-296
    return e, s, h, E, nu, hE, Ju


def element_quantities(xv, uh, h_v, E_v, nu_v, f_v, w0, m=4):
    """K_e (39,39), R_e (39,), dR_e/dh_b (39,4) of one cell from SymPy derivatives.  h_v, E_v, nu_v: nodal values (CG1);
    f_v (4,3): nodal load; w0 (39,): element state; m x m Gauss points."""
    cell = SymbolicCell(xv, uh)
    e, s, h, E, nu, hE, Ju = literal_energy()
    svec = sp.Matrix(s)
    grad_e = sp.Matrix([sp.diff(e, si) for si in s])                # de/ds
    H = grad_e.jacobian(svec)                                       # d2e/ds2 (11 x 11), free of s: e is quadratic
    dH_dh = H.diff(h)
    args = (h, E, nu, hE, Ju)
    fH, fdH = sp.lambdify(args, H, "numpy"), sp.lambdify(args, dH_dh, "numpy")
    fN1 = sp.lambdify((X0, X1), N1, "numpy")
    # load form  - f . u_mid J(uhat) dx  (model:308-321): its w-derivative by sympy
    fsym = sp.symbols("f0:3", real=True)
    load_vec = sp.Matrix([sp.diff(sum(fsym[i] * cell.u[i] for i in range(3)), wi) for wi in cell.w])
    fload = sp.lambdify((X0, X1) + fsym, load_vec, "numpy")
    xv = np.asarray(xv, dtype=np.float64)
    hE_val = max(np.linalg.norm(xv[i] - xv[j]) for i in range(4) for j in range(4))         # CellDiameter
    gx, gw = np.polynomial.legendre.leggauss(m)
    gx, gw = 0.5 * (gx + 1.0), 0.5 * gw
    K = np.zeros((39, 39)); R = np.zeros(39); dRdh = np.zeros((39, 4))
    w0 = np.asarray(w0, dtype=np.float64)
    for i in range(m):
        for j in range(m):
            q0, q1 = float(gx[i]), float(gx[j])
            B, detJ, Ju_q = cell.strains_at(q0, q1)
            Nq = np.asarray(fN1(q0, q1), dtype=np.float64)
            hq, Eq, nuq = Nq @ h_v, Nq @ E_v, Nq @ nu_v
            wt = gw[i] * gw[j] * detJ
            Hq = np.asarray(fH(hq, Eq, nuq, hE_val, Ju_q), dtype=np.float64)
            dHq = np.asarray(fdH(hq, Eq, nuq, hE_val, Ju_q), dtype=np.float64)
            K += wt * B.T @ Hq @ B
            R += wt * B.T @ (Hq @ (B @ w0))
            R -= wt * Ju_q * np.asarray(fload(q0, q1, *(Nq @ np.asarray(f_v))), dtype=np.float64).ravel()
            dRdh += wt * np.outer(B.T @ (dHq @ (B @ w0)), Nq)
    return K, R, dRdh
