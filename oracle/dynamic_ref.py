"""CPU oracle for the time-stepped (dynamic) RM-shell path -- SURVEY 8 row f1, BASELINE config 4.

TEST INFRASTRUCTURE ONLY (same rules as oracle/rm_shell_ref.py: only tests/, smoke() and the
cpu_baseline / --impl reference legs of bench.py may use it, and only as the checker).

PARITY UNPINNED.  femo_alpha/dynamic_rm_shell takes its element classes (DynamicElasticModel,
ShellElement, MaterialModel ...) from `from shell_analysis_fenicsx import *` (plate_sim.py:17), a package
that is neither vendored under /root/reference nor pinned anywhere; dolfinx / PETSc / MUMPS are absent as
for the static path.  The physics restated here is the reference's in-tree analogue (the inertia form of
rm_shell/linear_shell_fenicsx/linear_shell_model.py:335-348 and the energy of :260-306, i.e. the static
oracle's K); the time integration, the residual bookkeeping and the adjoint bookkeeping follow the
reference's own files line by line (file:line relative to /root/reference/femo_alpha/dynamic_rm_shell):

  * implicit midpoint rule                          plate_sim.py:135-140
      w_mid = (w + w_old)/2,  wdot = 2/dt (w - w_old) - wdot_old,  wddot = (wdot - wdot_old)/dt
  * residual  R = M wddot + K w_mid - F(f_i) - W     plate_sim.py:190-215 (self weight W: :203-213)
  * time loop, zero initial conditions, load history (constant after its last row)   plate_sim.py:281-361
  * state history layout: level-major [w_0; w_1; ...] (ravel order 'F' of (ndof, levels))   utils.py:9-16
  * dR/dw_cur, dR/dw_old, dR/dwdot_old with BC rows zeroed    state_operation_dynamic.py:478-502
  * reverse sweep with the O(N^2) right-hand-side store        state_operation_dynamic.py:590-640 (numpy mode)
  * dR_i/dthickness at (w_i, w_{i-1}, wdot_{i-1})             state_operation_dynamic.py:385-416, plate_sim.py:400-416
  * total strain energy = sum over time levels, its partials  total_strain_energy_operation.py:56-125
  * volume = int t dx                                         volume_operation.py:57-70

Mass form (linear_shell_model.py:335-348):  rho h (u, du) J dx + rho h h_E^2 (theta, dtheta) J dx.
It is integrated with the 4x4 Gauss rule (the reference uses the default `dx`, i.e. an estimated degree
that is exact for these integrands), independent of the `quad_deg` of the energy.

One reference inconsistency is NOT reproduced: the forward-mode chain-rule sum of
state_operation_dynamic.py:253-257,531-536 uses (-1)^j where the reverse mode (and the recurrence) have
(-1)^(j-1); the reverse mode -- the one every gradient goes through -- is the one restated.
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from .rm_shell_ref import RMShellOracle, gauss01, qp_geometry


class DynamicShellOracle:
    def __init__(self, base: RMShellOracle, dt: float, n_steps: int, add_self_weight=False, g_factor=None):
        self.o = base
        self.dt = float(dt)
        self.n_steps = int(n_steps)
        self.levels = self.n_steps + 1
        self.add_self_weight = bool(add_self_weight)
        self.g = (-1.0 if g_factor is None else float(g_factor)) * 9.81       # plate_sim.py:205-209
        gx, gw = gauss01(4)
        self.mqp = [(gx[i], gx[j], gw[i] * gw[j]) for i in range(4) for j in range(4)]

    # ---- mass matrix and its thickness derivative --------------------------------
    def _mass_elements(self, coef_at):
        """(nc,9,9) and (nc,4,4) scalar element matrices  int coef phi_a phi_b,  int coef hE^2 N_a N_b."""
        o, t = self.o, self.o.topo
        xv = t.x[t.cv]; uh = o.uhat.reshape(-1, 3)[t.cv]
        Mu = np.zeros((t.nc, 9, 9)); Mt = np.zeros((t.nc, 4, 4))
        for X0, X1, w in self.mqp:
            g = qp_geometry(xv, uh, X0, X1)
            c = coef_at(g.N1) * w * g.detJ * g.Ju
            Mu += c[:, None, None] * np.outer(g.phi2, g.phi2)[None]
            Mt += (c * o.hE ** 2)[:, None, None] * np.outer(g.N1, g.N1)[None]
        return Mu, Mt

    def _scatter_mass(self, Mu, Mt):
        t = self.o.topo
        rows, cols, vals = [], [], []
        for i in range(3):
            du = 3 * t.cell_blocks[:, :9] + i
            dth = 3 * t.cell_blocks[:, 9:] + i
            rows += [np.repeat(du, 9, axis=1).ravel(), np.repeat(dth, 4, axis=1).ravel()]
            cols += [np.tile(du, (1, 9)).ravel(), np.tile(dth, (1, 4)).ravel()]
            vals += [Mu.ravel(), Mt.ravel()]
        M = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))),
                          shape=(t.ndof, t.ndof)).tocsr()
        M.sum_duplicates()
        return M

    def mass_matrix(self):
        o = self.o
        return self._scatter_mass(*self._mass_elements(lambda N1: o._mat_at(o.rho, N1) * o._mat_at(o.h, N1)))

    def dmass_dh_action(self, a):
        """d(M a)/dh as a sparse (ndof, n_mat) matrix."""
        o, t = self.o, self.o.topo
        xv = t.x[t.cv]; uh = o.uhat.reshape(-1, 3)[t.cv]
        ae = a[t.cell_dofs]
        nbp = 1 if o.ewm else 4
        act = np.zeros((t.nc, 39, nbp))
        for X0, X1, w in self.mqp:
            g = qp_geometry(xv, uh, X0, X1)
            c = o._mat_at(o.rho, g.N1) * w * g.detJ * g.Ju
            uq = np.einsum("a,cai->ci", g.phi2, ae[:, :27].reshape(t.nc, 9, 3))
            tq = np.einsum("a,cai->ci", g.N1, ae[:, 27:].reshape(t.nc, 4, 3))
            r = np.zeros((t.nc, 39))
            r[:, :27] = (c[:, None, None] * g.phi2[None, :, None] * uq[:, None, :]).reshape(t.nc, 27)
            r[:, 27:] = ((c * o.hE ** 2)[:, None, None] * g.N1[None, :, None] * tq[:, None, :]).reshape(t.nc, 12)
            wt = np.ones(1) if o.ewm else g.N1
            act += r[:, :, None] * wt[None, None, :]
        pcols = (np.arange(t.nc)[:, None] if o.ewm else t.cv)
        rows = np.repeat(t.cell_dofs, nbp, axis=1).ravel()
        cols = np.tile(pcols, (1, 39)).ravel()
        return sp.coo_matrix((act.ravel(), (rows, cols)), shape=(t.ndof, o.n_mat)).tocsr()

    # ---- self weight  int rho t g du_z dx  (plate_sim.py:203-213) ---------------------
    def _weight_elements(self):
        o, t = self.o, self.o.topo
        xv = t.x[t.cv]; uh = o.uhat.reshape(-1, 3)[t.cv]
        nbp = 1 if o.ewm else 4
        act = np.zeros((t.nc, 9, nbp))
        for X0, X1, w in self.mqp:
            g = qp_geometry(xv, uh, X0, X1)
            c = o._mat_at(o.rho, g.N1) * self.g * w * g.detJ
            wt = np.ones(1) if o.ewm else g.N1
            act += c[:, None, None] * g.phi2[None, :, None] * wt[None, None, :]
        return act

    def self_weight(self):
        t = self.o.topo
        W = np.zeros(t.ndof)
        if self.add_self_weight:
            np.add.at(W, 3 * t.cell_blocks[:, :9].ravel() + 2, self._weight_plain().ravel())
        return W

    def _weight_plain(self):
        o, t = self.o, self.o.topo
        xv = t.x[t.cv]; uh = o.uhat.reshape(-1, 3)[t.cv]
        val = np.zeros((t.nc, 9))
        for X0, X1, w in self.mqp:
            g = qp_geometry(xv, uh, X0, X1)
            c = o._mat_at(o.rho, g.N1) * o._mat_at(o.h, g.N1) * self.g * w * g.detJ
            val += c[:, None] * g.phi2[None, :]
        return val

    def dself_weight_dh(self):
        o, t = self.o, self.o.topo
        if not self.add_self_weight:
            return sp.csr_matrix((t.ndof, o.n_mat))
        act = self._weight_elements()                     # (nc,9,nbp): int rho g phi_a N_b
        nbp = act.shape[2]
        pcols = (np.arange(t.nc)[:, None] if o.ewm else t.cv)
        rows = np.repeat(3 * t.cell_blocks[:, :9] + 2, nbp, axis=1).ravel()
        cols = np.tile(pcols, (1, 9)).ravel()
        return sp.coo_matrix((act.ravel(), (rows, cols)), shape=(t.ndof, o.n_mat)).tocsr()

    # ---- operators of one time step -------------------------------------------------
    def operators(self):
        """K, M and the BC-modified effective matrix  A = (2/dt^2) M + K/2  (rows+cols zeroed, unit diagonal)."""
        o = self.o
        self.K = o.assemble_dRdu(with_penalty=False)
        self.M = self.mass_matrix()
        c = 2.0 / self.dt ** 2
        self.Acur = (c * self.M + 0.5 * self.K).tocsr()        # dR/dw_cur
        self.Bold = (-c * self.M + 0.5 * self.K).tocsr()       # dR/dw_old
        self.Cdot = (-(2.0 / self.dt) * self.M).tocsr()        # dR/dwdot_old
        self.A = o.apply_bc_matrix(self.Acur)
        self.lu = spla.splu(self.A.tocsc())
        self.keep = np.ones(o.topo.ndof); self.keep[o.bc_dofs] = 0.0
        return self

    def load(self, f_row):
        o = self.o
        o.set_inputs(F_solid=f_row)
        return o.assemble_load()

    # ---- forward time loop (plate_sim.py:281-361) --------------------------------------
    def solve(self, f_history):
        o = self.o
        f_history = np.atleast_2d(np.asarray(f_history, dtype=np.float64))
        self.operators()
        n = o.topo.ndof
        W = self.self_weight()
        w_hist = np.zeros((self.levels, n))
        energy = np.zeros(self.levels)
        v_old = np.zeros(n)
        for i in range(1, self.levels):
            F = self.load(f_history[min(i, f_history.shape[0] - 1)])
            rhs = F + W - self.Bold @ w_hist[i - 1] - self.Cdot @ v_old
            rhs[o.bc_dofs] = 0.0
            w_hist[i] = self.lu.solve(rhs)
            v_old = 2.0 / self.dt * (w_hist[i] - w_hist[i - 1]) - v_old
            energy[i] = 0.5 * w_hist[i] @ (self.K @ w_hist[i])
        self.w_hist, self.energy = w_hist, energy
        return w_hist

    def residual_history(self, w_hist, f_history):
        """R_i(w_i, w_{i-1}, wdot_{i-1}) for all levels (R_0 = w_0), BC rows = w_i[bc]."""
        o = self.o
        f_history = np.atleast_2d(np.asarray(f_history, dtype=np.float64))
        R = np.zeros_like(w_hist)
        R[0] = w_hist[0]
        v_old = np.zeros(w_hist.shape[1])
        W = self.self_weight()
        for i in range(1, self.levels):
            F = self.load(f_history[min(i, f_history.shape[0] - 1)])
            r = self.Acur @ w_hist[i] + self.Bold @ w_hist[i - 1] + self.Cdot @ v_old - F - W
            r[o.bc_dofs] = w_hist[i][o.bc_dofs]
            R[i] = r
            v_old = 2.0 / self.dt * (w_hist[i] - w_hist[i - 1]) - v_old
        return R

    # ---- outputs -------------------------------------------------------------------------
    def total_strain_energy(self, w_hist):
        return float(sum(0.5 * w @ (self.K @ w) for w in w_hist))

    def dtotal_strain_energy_dw(self, w_hist):
        return np.stack([self.K @ w for w in w_hist])

    def dtotal_strain_energy_dh(self, w_hist):
        o = self.o
        g = np.zeros(o.n_mat)
        for w in w_hist:
            g += 0.5 * (o.assemble_dRdparam(w, "dh").T @ w)
        return g

    def volume(self):
        o = self.o
        rho = o.rho.copy()
        o.rho = np.ones_like(rho)
        uh = o.uhat.copy(); o.uhat = np.zeros_like(uh)      # volume_operation.py:68 uses the plain dx
        v = o.mass()
        o.rho, o.uhat = rho, uh
        return v

    def dvolume_dh(self):
        o = self.o
        rho = o.rho.copy()
        o.rho = np.ones_like(rho)
        uh = o.uhat.copy(); o.uhat = np.zeros_like(uh)
        g = o.dmass_dh()
        o.rho, o.uhat = rho, uh
        return g

    # ---- reverse sweep, literal restatement (state_operation_dynamic.py:590-640) ------------
    def adjoint_reference(self, d_out):
        """d_out (levels, ndof) -> adjoint history, with the reference's O(N^2) right-hand-side store.

        The reference solves with the rows-only BC matrix; the BC entries of the adjoints do not reach any
        gradient (all derivative matrices have their BC rows zeroed); here they are zeroed, as the static path
        does explicitly (csdl_alpha_opt/state_operation.py:217-218)."""
        o, dt, L = self.o, self.dt, self.levels
        keep = self.keep
        Bz = sp.diags(keep) @ self.Bold          # zeroRows(bc_dofs, diag=0)
        Cz = sp.diags(keep) @ self.Cdot
        n = o.topo.ndof
        rhs_store = [[np.zeros(n) for _ in range(L)] for _ in range(L)]
        lam = np.zeros((L, n))
        for i in range(L - 1, -1, -1):
            add = np.zeros(n)
            for j in range(i + 1, L):
                add -= rhs_store[i][j]
            if i == 0:
                lam[i] = 0.0
            else:
                lam[i] = keep * self.lu.solve(keep * (d_out[i] + add))
            Ct_l = Cz.T @ lam[i]
            B1 = (Bz + (2.0 / dt) * Cz)
            for j in range(i - 1, -1, -1):
                if j == i - 1 and j > 0:
                    rhs_store[j][i] += B1.T @ lam[i]
                elif j == i - 1 and j == 0:
                    rhs_store[j][i] += Bz.T @ lam[i]
                elif j < i - 1 and j == 0:
                    rhs_store[j][i] += ((-1.0) ** j) * (2.0 / dt) * Ct_l
                elif i % 2 != 0:
                    rhs_store[j][i] += ((-1.0) ** j) * (4.0 / dt) * Ct_l
                else:
                    rhs_store[j][i] += ((-1.0) ** (j + 1)) * (4.0 / dt) * Ct_l
        return lam

    def adjoint_recurrence(self, d_out):
        """The same adjoints from the two-field (w, wdot) recurrence: O(N) vectors, one solve per level."""
        o, dt, L = self.o, self.dt, self.levels
        keep = self.keep
        n = o.topo.ndof
        lam = np.zeros((L, n))
        mu_next = np.zeros(n); lam_next = np.zeros(n)
        for i in range(L - 1, 0, -1):
            if i == L - 1:
                mu = np.zeros(n)
                rhs = d_out[i].copy()
            else:
                ln = keep * lam_next
                mu = -(self.Cdot.T @ ln) - mu_next
                rhs = d_out[i] - self.Bold.T @ ln + (2.0 / dt) * (mu - mu_next)
            lam[i] = keep * self.lu.solve(keep * rhs)
            lam_next, mu_next = lam[i], mu
        return lam

    # ---- thickness partial of the residual history ---------------------------------------------
    def dRdh_T(self, w_hist, lam_hist):
        """sum_i (dR_i/dh)^T lam_i  (BC rows of dR/dh zeroed, level 0 contributes nothing)."""
        o, dt = self.o, self.dt
        g = np.zeros(o.n_mat)
        v_old = np.zeros(w_hist.shape[1])
        dW = self.dself_weight_dh()
        for i in range(1, self.levels):
            wm = 0.5 * (w_hist[i] + w_hist[i - 1])
            acc = 2.0 / dt ** 2 * (w_hist[i] - w_hist[i - 1]) - 2.0 / dt * v_old
            D = o.assemble_dRdparam(wm, "dh") + self.dmass_dh_action(acc) - dW
            g += D.T @ (self.keep * lam_hist[i])
            v_old = 2.0 / dt * (w_hist[i] - w_hist[i - 1]) - v_old
        return g

    def dRdh(self, w_hist, dh):
        o, dt = self.o, self.dt
        out = np.zeros_like(w_hist)
        v_old = np.zeros(w_hist.shape[1])
        dW = self.dself_weight_dh()
        for i in range(1, self.levels):
            wm = 0.5 * (w_hist[i] + w_hist[i - 1])
            acc = 2.0 / dt ** 2 * (w_hist[i] - w_hist[i - 1]) - 2.0 / dt * v_old
            D = o.assemble_dRdparam(wm, "dh") + self.dmass_dh_action(acc) - dW
            out[i] = self.keep * (D @ dh)
            v_old = 2.0 / dt * (w_hist[i] - w_hist[i - 1]) - v_old
        return out

    # ---- products with the Jacobian of the residual history w.r.t. the state history ------------
    def dRdw_apply(self, x_hist, transpose=False):
        """(dR/dw) x or (dR/dw)^T x, wdot eliminated (block lower triangular; level 0 = identity)."""
        dt, L = self.dt, self.levels
        keep = self.keep
        Az = sp.diags(keep) @ self.Acur @ sp.diags(keep) + sp.diags(1.0 - keep)
        Bz = sp.diags(keep) @ self.Bold
        Cz = sp.diags(keep) @ self.Cdot
        y = np.zeros_like(x_hist)

        def coef(i, j):            # d wdot_{i-1} / d w_j  (j <= i-1)
            if j == i - 1:
                return 2.0 / dt if j > 0 else 2.0 / dt
            k = i - j
            c = (-1.0) ** (k - 1) * (4.0 / dt)
            return c / 2.0 if j == 0 else c
        for i in range(L):
            for j in range(i + 1):
                if i == 0:
                    blk = sp.identity(x_hist.shape[1], format="csr")
                elif j == i:
                    blk = Az
                elif j == i - 1:
                    blk = Bz + (coef(i, j) * Cz if i - 1 > 0 else 0 * Cz)
                else:
                    blk = coef(i, j) * Cz
                if transpose:
                    y[j] += blk.T @ x_hist[i]
                else:
                    y[i] += blk @ x_hist[j]
        return y

    def forward_recurrence(self, d_res):
        """Solve (dR/dw) x = d_res level by level (apply_inverse_jacobian 'fwd', state_operation_dynamic.py:516-560,
        with the consistent chain-rule signs): O(N) vectors."""
        o, dt, L = self.o, self.dt, self.levels
        keep = self.keep
        x = np.zeros_like(d_res)
        x[0] = d_res[0]
        dv = np.zeros(d_res.shape[1])
        for i in range(1, L):
            rhs = d_res[i] - keep * (self.Bold @ x[i - 1]) - keep * (self.Cdot @ dv)
            xi = self.lu.solve(keep * rhs)
            xi[o.bc_dofs] = d_res[i][o.bc_dofs]
            x[i] = xi
            dv = 2.0 / dt * (x[i] - x[i - 1]) - dv
        return x
