// Strains and the von Mises stress of the RM shell at one quadrature point (host + device).
// Restates ShellStressRM.vonMisesStress (femo_alpha/rm_shell/linear_shell_fenicsx/linear_shell_model.py:393-467) and the
// strain definitions (:220-258,288-289); used by the stress functionals in outputs.cu and their shape partials.
#pragma once
#include "femo_internal.cuh"

#ifdef __CUDACC__

// strains [e0,e1,e2,k0,k1,k2,g0,g1,dr] of element vector `we` (39, shared) at one point
FEMO_HD void strains_at(const QpGeom& g, const Tables* tab, int q, const double* we, double* s) {
  const double k00 = dot(g.Kx0, g.E0), k01 = dot(g.Kx0, g.E1), k10 = dot(g.Kx1, g.E0), k11 = dot(g.Kx1, g.E1);
  const V3 gq0 = k00 * g.dE2_0 + k10 * g.dE2_1, gq1 = k01 * g.dE2_0 + k11 * g.dE2_1;
#pragma unroll
  for (int i = 0; i < 9; ++i) s[i] = 0.0;
  for (int a = 0; a < 9; ++a) {
    double d0 = tab->dphi2[q][a][0] * k00 + tab->dphi2[q][a][1] * k10, d1 = tab->dphi2[q][a][0] * k01 + tab->dphi2[q][a][1] * k11;
    URows R = u_rows(d0, d1, g.E0, g.E1, g.E2);
    V3 u = ldv3(we + 3 * a);
    s[0] += dot(R.e0, u); s[1] += dot(R.e1, u); s[2] += dot(R.e2, u);
    s[6] += dot(R.g0, u); s[7] += dot(R.g1, u); s[8] += dot(R.dr, u);
  }
  for (int b = 0; b < 4; ++b) {
    double d0 = tab->dn1[q][b][0] * k00 + tab->dn1[q][b][1] * k10, d1 = tab->dn1[q][b][0] * k01 + tab->dn1[q][b][1] * k11;
    TRows R = t_rows(tab->n1[q][b], d0, d1, g.E0, g.E1, g.E2, gq0, gq1);
    V3 t = ldv3(we + 27 + 3 * b);
    s[3] += dot(R.k0, t); s[4] += dot(R.k1, t); s[5] += dot(R.k2, t);
    s[6] += dot(R.g0, t); s[7] += dot(R.g1, t); s[8] += dot(R.dr, t);
  }
}


// von Mises stress at one point (linear_shell_model.py:393-467): in-plane stress of
// u(xi2) = u_mid - xi2 cross(E2, theta) at xi2 = xi_sign*h/2 (grad of xi2 included for CG1 thickness)
struct VmPoint {
  double vm, t[3];        // t_j = d vm / d eps_j
  double xi, gh0, gh1, b0, b1, c0, nu;
  double e[3], kap[3];
};
FEMO_HD void vm_at(const QpGeom& g, const Tables* tab, int q, const double* we, double hq, double Eq,
                                      double nuq, const double* hv, int ewm, double xi_sign, VmPoint& P) {
  double s[9];
  strains_at(g, tab, q, we, s);
  const double k00 = dot(g.Kx0, g.E0), k01 = dot(g.Kx0, g.E1), k10 = dot(g.Kx1, g.E0), k11 = dot(g.Kx1, g.E1);
  V3 th = v3(0, 0, 0);
  for (int b = 0; b < 4; ++b) th = th + tab->n1[q][b] * ldv3(we + 27 + 3 * b);
  P.b0 = -dot(g.E1, th); P.b1 = dot(g.E0, th);
  P.gh0 = P.gh1 = 0.0;
  if (!ewm)
    for (int v = 0; v < 4; ++v) {
      P.gh0 += 0.5 * xi_sign * hv[v] * (tab->dn1[q][v][0] * k00 + tab->dn1[q][v][1] * k10);
      P.gh1 += 0.5 * xi_sign * hv[v] * (tab->dn1[q][v][0] * k01 + tab->dn1[q][v][1] * k11);
    }
  P.xi = 0.5 * xi_sign * hq;
  P.kap[0] = s[3]; P.kap[1] = s[4]; P.kap[2] = s[5];
  P.e[0] = s[0] - P.xi * s[3] - P.b0 * P.gh0;
  P.e[1] = s[1] - P.xi * s[4] - P.b1 * P.gh1;
  P.e[2] = s[2] - P.xi * s[5] - (P.b0 * P.gh1 + P.b1 * P.gh0);
  P.c0 = Eq / (1.0 - nuq * nuq); P.nu = nuq;
  double s0 = P.c0 * (P.e[0] + nuq * P.e[1]), s1 = P.c0 * (nuq * P.e[0] + P.e[1]), s2 = P.c0 * 0.5 * (1 - nuq) * P.e[2];
  P.vm = sqrt(s0 * s0 - s0 * s1 + s1 * s1 + 3.0 * s2 * s2);
  double iv = P.vm > 0 ? 0.5 / P.vm : 0.0;
  double a0 = (2 * s0 - s1) * iv, a1 = (2 * s1 - s0) * iv, a2 = 6 * s2 * iv;
  P.t[0] = P.c0 * (a0 + nuq * a1); P.t[1] = P.c0 * (nuq * a0 + a1); P.t[2] = P.c0 * 0.5 * (1 - nuq) * a2;
}

#endif  // __CUDACC__
