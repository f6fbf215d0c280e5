"""Independent, literal re-evaluation of the RM-shell energy on ONE cell.

TEST INFRASTRUCTURE ONLY (see oracle/rm_shell_ref.py header; parity unpinned).

`energy_density` evaluates the reference's UFL expressions line by line with
3x3 NumPy matrices and np.cross for a *given* element DOF vector -- no B
matrices, no vectorisation -- so it shares no code with
`rm_shell_ref.strain_operator`.  The energy is quadratic in w, hence the exact
element stiffness follows by polarisation:
    K_ij = Pi(e_i + e_j) - Pi(e_i) - Pi(e_j)
which the tests compare with `RMShellOracle.element_matrices`.

Follows /root/reference/femo_alpha/rm_shell/linear_shell_fenicsx/
linear_shell_model.py:136-157,220-296 and kinematics.py:12-106.
"""
import numpy as np

from .rm_shell_ref import basis_q1, basis_q2, gauss01, SHEAR_K


def _fields(xv, uh, we, X0, X1, eps_fd=None):
    N1, dN1 = basis_q1(X0, X1)
    phi2, dphi2 = basis_q2(X0, X1)
    Jm = xv.T @ dN1                                            # Jacobian(mesh) 3x2
    Kp = np.linalg.inv(Jm.T @ Jm) @ Jm.T                       # pseudo-inverse
    detJ = np.sqrt(np.linalg.det(Jm.T @ Jm))
    n = np.cross(Jm[:, 0], Jm[:, 1])
    E2 = n / np.linalg.norm(n)                                 # CellNormal
    E0 = Jm[:, 0] / np.linalg.norm(Jm[:, 0])                   # unit(A0)            kin:66-67
    E1 = np.cross(E2, E0)                                      #                     kin:68
    T = np.array([E0, E1])                                     # E01                 kin:72-82
    ue = we[:27].reshape(9, 3); te = we[27:].reshape(4, 3)
    theta = te.T @ N1
    grad_u = (ue.T @ dphi2) @ Kp                               # grad(u_mid)
    F = np.eye(3) + (uh.T @ dN1) @ Kp                          # F(uhat)             kin:34-44
    Finv = np.linalg.inv(F)
    Ju = np.linalg.det(F)                                      # J(uhat)             kin:24-32
    return N1, dN1, phi2, dphi2, Jm, Kp, detJ, E0, E1, E2, T, theta, grad_u, F, Finv, Ju, ue, te


def _beta(xv, te, X0, X1):
    """cross(E2, theta) as a function of the reference point (for its gradient)."""
    N1, dN1 = basis_q1(X0, X1)
    Jm = xv.T @ dN1
    n = np.cross(Jm[:, 0], Jm[:, 1])
    return np.cross(n / np.linalg.norm(n), te.T @ N1)


def energy_density(xv, uh, we, X0, X1, h, E, nu, hE):
    """Energy density times surface measure at one reference point (literal UFL)."""
    (N1, dN1, phi2, dphi2, Jm, Kp, detJ, E0, E1, E2, T, theta, grad_u, F, Finv, Ju, ue, te) = \
        _fields(xv, uh, we, X0, X1)
    gradu = grad_u @ Finv                                      # gradx(u_mid, uhat)  model:220
    t_gu = T @ gradu @ T.T                                     # gradv_local         model:222
    # grad(cross(E2, theta)): reference derivative by central differences of the
    # *exact* field (E2 varies with X on warped quads), then pull-back.  The
    # field is smooth; step 1e-6 gives ~1e-10 relative accuracy, enough to
    # detect a missing/incorrect dE2/dX term (effect O(1) on warped cells).
    hh = 1e-6
    dbeta = np.stack([(_beta(xv, te, X0 + hh, X1) - _beta(xv, te, X0 - hh, X1)) / (2 * hh),
                      (_beta(xv, te, X0, X1 + hh) - _beta(xv, te, X0, X1 - hh)) / (2 * hh)], 1)
    gradx_beta = (dbeta @ Kp) @ Finv
    kap_m = T @ gradx_beta @ T.T
    kappa = 0.5 * (kap_m + kap_m.T)                            # model:241-243
    eps = 0.5 * (t_gu + t_gu.T)                                # model:232-239
    dudxi2 = -np.cross(E2, theta)
    gamma = T @ dudxi2 + T @ (E2 @ gradu)                      # model:245-258
    G = E / 2 / (1 + nu)
    C = (E / (1.0 - nu * nu)) * np.array([[1.0, nu, 0.0], [nu, 1.0, 0.0], [0.0, 0.0, 0.5 * (1.0 - nu)]])
    A, D, A_s = h * C, h ** 3 / 12 * C, SHEAR_K * G * h * np.eye(2)     # model:136-157
    v = lambda S: np.array([S[0, 0], S[1, 1], 2.0 * S[0, 1]])           # voigt2D kin:94-106
    N, M, Q = A @ v(eps), D @ v(kappa), A_s @ gamma                     # model:260-273
    drill = (t_gu[0, 1] - t_gu[1, 0]) / 2 + theta @ E2                  # model:288-289
    alpha = E * h ** 3
    e = 0.5 * (Q @ gamma) * Ju                                          # shearEnergy     model:275-276
    e += 0.5 * (N @ v(eps))                                             # membraneEnergy  model:278-279
    e += 0.5 * (M @ v(kappa))                                           # bendingEnergy   model:281-282
    e += 0.5 * (alpha * drill / hE ** 2) * drill * Ju                   # drillingEnergy  model:284-296
    return e * detJ


def cell_energy(xv, uh, we, h_v, E_v, nu_v, m=4):
    """Pi on one cell. xv, uh: (4,3) tensor vertex order; h_v/E_v/nu_v: 4 nodal values (CG1)."""
    hE = max(np.linalg.norm(xv[i] - xv[j]) for i in range(4) for j in range(4))
    gx, gw = gauss01(m)
    tot = 0.0
    for i in range(m):
        for j in range(m):
            N1, _ = basis_q1(gx[i], gx[j])
            tot += gw[i] * gw[j] * energy_density(xv, uh, we, gx[i], gx[j], h_v @ N1, E_v @ N1, nu_v @ N1, hE)
    return tot


def cell_stiffness_by_polarisation(xv, uh, h_v, E_v, nu_v, m=4):
    I = np.eye(39)
    Pi = np.array([cell_energy(xv, uh, I[i], h_v, E_v, nu_v, m) for i in range(39)])
    K = np.zeros((39, 39))
    for i in range(39):
        K[i, i] = 2.0 * Pi[i]
        for j in range(i + 1, 39):
            K[i, j] = K[j, i] = cell_energy(xv, uh, I[i] + I[j], h_v, E_v, nu_v, m) - Pi[i] - Pi[j]
    return K
