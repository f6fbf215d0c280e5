// Shape partials of the RM-shell residual: (dR/duhat)^T lam and (dR/duhat) duhat, per quadrature point -- SURVEY 8 f3.
//
// Replaces the UFL differentiation of the residual w.r.t. the mesh displacement `uhat`
// (femo_alpha/csdl_alpha_opt/state_operation.py:242-258,282-286).  `uhat` enters the forms only through
//     F = I + grad(uhat),   gradx(.) = grad(.) F^-1,   J = det F
// (rm_shell/linear_shell_fenicsx/kinematics.py:12-44; linear_shell_model.py:220,242,276,296,320); the local frame,
// the surface measure and h_E belong to the undeformed mesh.  With  k_ab = (Kinv_a F^-1) . E_b  (2x2: the in-plane
// gradient map) every strain is linear in k, and the shear / drilling stiffnesses and the load carry one factor J:
//     s(uhat) = lam^T R = sum_q wd [ eps(lam; k)^T D(J) eps(w; k) - J f . lam_u ]
//     d k_ab = -(Kx_a . duhat_v) (gN_v F^-1 E_b),        dJ = J duhat_v . (F^-T gN_v),     gN_v = dN_v/dX Kinv
// Host + device (FEMO_HD): tests/shape_host.cu runs the same functions on the CPU against a complex-step checker.
#pragma once
#include "femo_internal.cuh"
#include "stress_point.cuh"

#ifdef __CUDACC__

struct KMap { double k00, k01, k10, k11; };

FEMO_HD KMap kmap_of(const QpGeom& g) {
  KMap k;
  k.k00 = dot(g.Kx0, g.E0); k.k01 = dot(g.Kx0, g.E1); k.k10 = dot(g.Kx1, g.E0); k.k11 = dot(g.Kx1, g.E1);
  return k;
}

// strain rows of one displacement / rotation node for an arbitrary gradient map k; `nscale` multiplies the rows
// that do not depend on k (the pure shape-function terms of shear and drilling)
FEMO_HD URows u_rows_k(const QpGeom& g, double dp0, double dp1, const KMap& k) {
  return u_rows(dp0 * k.k00 + dp1 * k.k10, dp0 * k.k01 + dp1 * k.k11, g.E0, g.E1, g.E2);
}
FEMO_HD TRows t_rows_k(const QpGeom& g, double N, double dn0, double dn1, const KMap& k, double nscale) {
  const V3 gq0 = k.k00 * g.dE2_0 + k.k10 * g.dE2_1, gq1 = k.k01 * g.dE2_0 + k.k11 * g.dE2_1;
  TRows R = t_rows(N, dn0 * k.k00 + dn1 * k.k10, dn0 * k.k01 + dn1 * k.k11, g.E0, g.E1, g.E2, gq0, gq1);
  R.g0 = nscale * R.g0; R.g1 = nscale * R.g1; R.dr = nscale * R.dr;
  return R;
}

// strains [e0,e1,e2,k0,k1,k2,g0,g1,dr] of the element vector `we` (39) for the gradient map k
FEMO_HD void strains_k(const QpGeom& g, const Tables* tab, int q, const double* we, const KMap& k, double nscale, double* s) {
#pragma unroll
  for (int i = 0; i < 9; ++i) s[i] = 0.0;
  for (int a = 0; a < 9; ++a) {
    URows R = u_rows_k(g, tab->dphi2[q][a][0], tab->dphi2[q][a][1], k);
    V3 u = ldv3(we + 3 * a);
    s[0] += dot(R.e0, u); s[1] += dot(R.e1, u); s[2] += dot(R.e2, u);
    s[6] += dot(R.g0, u); s[7] += dot(R.g1, u); s[8] += dot(R.dr, u);
  }
  for (int b = 0; b < 4; ++b) {
    TRows R = t_rows_k(g, tab->n1[q][b], tab->dn1[q][b][0], tab->dn1[q][b][1], k, nscale);
    V3 t = ldv3(we + 27 + 3 * b);
    s[3] += dot(R.k0, t); s[4] += dot(R.k1, t); s[5] += dot(R.k2, t);
    s[6] += dot(R.g0, t); s[7] += dot(R.g1, t); s[8] += dot(R.dr, t);
  }
}

FEMO_HD void apply_D9(const Coefs& c, const double* s, double* sig) {
  sig[0] = c.cm11 * s[0] + c.cm12 * s[1]; sig[1] = c.cm12 * s[0] + c.cm11 * s[1]; sig[2] = c.cm33 * s[2];
  sig[3] = c.cb11 * s[3] + c.cb12 * s[4]; sig[4] = c.cb12 * s[3] + c.cb11 * s[4]; sig[5] = c.cb33 * s[5];
  sig[6] = c.cs * s[6]; sig[7] = c.cs * s[7]; sig[8] = c.cd * s[8];
}

// per vertex v: gN_v, a_b = gN_v . (F^-1 E_b), c = F^-T gN_v
struct ShapeVert { double a0, a1; V3 c; };
FEMO_HD ShapeVert shape_vert(const QpGeom& g, double dN0, double dN1) {
  const V3 gN = dN0 * g.Ki0 + dN1 * g.Ki1;
  const double* Fi = g.Finv;
  const V3 fe0 = v3(Fi[0] * g.E0.x + Fi[1] * g.E0.y + Fi[2] * g.E0.z, Fi[3] * g.E0.x + Fi[4] * g.E0.y + Fi[5] * g.E0.z,
                    Fi[6] * g.E0.x + Fi[7] * g.E0.y + Fi[8] * g.E0.z);
  const V3 fe1 = v3(Fi[0] * g.E1.x + Fi[1] * g.E1.y + Fi[2] * g.E1.z, Fi[3] * g.E1.x + Fi[4] * g.E1.y + Fi[5] * g.E1.z,
                    Fi[6] * g.E1.x + Fi[7] * g.E1.y + Fi[8] * g.E1.z);
  ShapeVert r;
  r.a0 = dot(gN, fe0); r.a1 = dot(gN, fe1);
  r.c = v3(Fi[0] * gN.x + Fi[3] * gN.y + Fi[6] * gN.z, Fi[1] * gN.x + Fi[4] * gN.y + Fi[7] * gN.z,
           Fi[2] * gN.x + Fi[5] * gN.y + Fi[8] * gN.z);
  return r;
}

// d(lam^T R)/d uhat_v at one quadrature point, v = 0..3 -> out12 (weight wd included).
//   we / le: element vectors of w and lam;  fq: load density at the point (only if with_load);  cf: D at (h,E,nu,hE,Ju)
FEMO_HD void shape_T_qp(const QpGeom& g, const Tables* tab, int q, const double* we, const double* le, const Coefs& cf,
                        V3 fq, double wd, int with_load, double* out12) {
  const KMap k = kmap_of(g);
  double sw[9], sl[9], sigw[9], sigl[9];
  strains_k(g, tab, q, we, k, 1.0, sw);
  strains_k(g, tab, q, le, k, 1.0, sl);
  apply_D9(cf, sw, sigw);
  apply_D9(cf, sl, sigl);
  double Gk[4];
#pragma unroll
  for (int ab = 0; ab < 4; ++ab) {
    KMap e; e.k00 = (ab == 0); e.k01 = (ab == 1); e.k10 = (ab == 2); e.k11 = (ab == 3);
    double ew[9], el[9];
    strains_k(g, tab, q, we, e, 0.0, ew);
    strains_k(g, tab, q, le, e, 0.0, el);
    double t = 0;
#pragma unroll
    for (int i = 0; i < 9; ++i) t += el[i] * sigw[i] + ew[i] * sigl[i];
    Gk[ab] = wd * t;
  }
  // d/dJ: shear and drilling stiffness are proportional to J, the load term is -J f . lam_u
  double GJ = (sl[6] * sigw[6] + sl[7] * sigw[7] + sl[8] * sigw[8]) / g.Ju;
  if (with_load) {
    V3 lu = v3(0, 0, 0);
    for (int a = 0; a < 9; ++a) lu = lu + tab->phi2[q][a] * ldv3(le + 3 * a);
    GJ -= dot(fq, lu);
  }
  GJ *= wd * g.Ju;
  const double dN0[4] = {-(1 - tab->X1[q]), (1 - tab->X1[q]), -tab->X1[q], tab->X1[q]};
  const double dN1[4] = {-(1 - tab->X0[q]), -tab->X0[q], (1 - tab->X0[q]), tab->X0[q]};
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    const ShapeVert sv = shape_vert(g, dN0[v], dN1[v]);
    const double c0 = -(Gk[0] * sv.a0 + Gk[1] * sv.a1), c1 = -(Gk[2] * sv.a0 + Gk[3] * sv.a1);
    const V3 r = c0 * g.Kx0 + c1 * g.Kx1 + GJ * sv.c;
    out12[3 * v] = r.x; out12[3 * v + 1] = r.y; out12[3 * v + 2] = r.z;
  }
}

// (dR_e/duhat) duhat at one quadrature point -> out39 (weight wd included).  duh: the perturbation at the 4 vertices
FEMO_HD void shape_fwd_qp(const QpGeom& g, const Tables* tab, int q, const double* we, const Coefs& cf, V3 fq, double wd,
                          const V3* duh, double* out39) {
  const KMap k = kmap_of(g);
  const double dN0[4] = {-(1 - tab->X1[q]), (1 - tab->X1[q]), -tab->X1[q], tab->X1[q]};
  const double dN1[4] = {-(1 - tab->X0[q]), -tab->X0[q], (1 - tab->X0[q]), tab->X0[q]};
  KMap dk; dk.k00 = dk.k01 = dk.k10 = dk.k11 = 0.0;
  double dJ = 0.0;
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    const ShapeVert sv = shape_vert(g, dN0[v], dN1[v]);
    const double x0 = dot(g.Kx0, duh[v]), x1 = dot(g.Kx1, duh[v]);
    dk.k00 -= x0 * sv.a0; dk.k01 -= x0 * sv.a1; dk.k10 -= x1 * sv.a0; dk.k11 -= x1 * sv.a1;
    dJ += g.Ju * dot(sv.c, duh[v]);
  }
  double sw[9], sigw[9], ew[9], sig1[9];
  strains_k(g, tab, q, we, k, 1.0, sw);
  apply_D9(cf, sw, sigw);
  strains_k(g, tab, q, we, dk, 0.0, ew);
  apply_D9(cf, ew, sig1);
  const double rJ = dJ / g.Ju;
  sig1[6] += rJ * sigw[6]; sig1[7] += rJ * sigw[7]; sig1[8] += rJ * sigw[8];
#pragma unroll
  for (int i = 0; i < 9; ++i) { sigw[i] *= wd; sig1[i] *= wd; }
  for (int a = 0; a < 9; ++a) {
    const double dp0 = tab->dphi2[q][a][0], dp1 = tab->dphi2[q][a][1];
    URows R = u_rows_k(g, dp0, dp1, k), Q = u_rows_k(g, dp0, dp1, dk);
    V3 r = sig1[0] * R.e0 + sig1[1] * R.e1 + sig1[2] * R.e2 + sig1[6] * R.g0 + sig1[7] * R.g1 + sig1[8] * R.dr +
           sigw[0] * Q.e0 + sigw[1] * Q.e1 + sigw[2] * Q.e2 + sigw[6] * Q.g0 + sigw[7] * Q.g1 + sigw[8] * Q.dr;
    r = r - (wd * dJ * tab->phi2[q][a]) * fq;
    out39[3 * a] = r.x; out39[3 * a + 1] = r.y; out39[3 * a + 2] = r.z;
  }
  for (int b = 0; b < 4; ++b) {
    const double N = tab->n1[q][b], dn0 = tab->dn1[q][b][0], dn1 = tab->dn1[q][b][1];
    TRows R = t_rows_k(g, N, dn0, dn1, k, 1.0), Q = t_rows_k(g, N, dn0, dn1, dk, 0.0);
    V3 r = sig1[3] * R.k0 + sig1[4] * R.k1 + sig1[5] * R.k2 + sig1[6] * R.g0 + sig1[7] * R.g1 + sig1[8] * R.dr +
           sigw[3] * Q.k0 + sigw[4] * Q.k1 + sigw[5] * Q.k2 + sigw[6] * Q.g0 + sigw[7] * Q.g1 + sigw[8] * Q.dr;
    out39[27 + 3 * b] = r.x; out39[27 + 3 * b + 1] = r.y; out39[27 + 3 * b + 2] = r.z;
  }
}

// d/duhat_v of  sum_q wd J c(q)  (compliance c = u.u, mass c = rho h): wd c J F^-T gN_v  -> out12
FEMO_HD void shape_measure_qp(const QpGeom& g, const Tables* tab, int q, double wdc, double* out12) {
  const double dN0[4] = {-(1 - tab->X1[q]), (1 - tab->X1[q]), -tab->X1[q], tab->X1[q]};
  const double dN1[4] = {-(1 - tab->X0[q]), -tab->X0[q], (1 - tab->X0[q]), tab->X0[q]};
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    const ShapeVert sv = shape_vert(g, dN0[v], dN1[v]);
    const V3 r = (wdc * g.Ju) * sv.c;
    out12[3 * v] = r.x; out12[3 * v + 1] = r.y; out12[3 * v + 2] = r.z;
  }
}

// d/duhat_v of  wd J (m sigma_vm)^rho  at one quadrature point (the integrand of pnorm_stress, rm_shell_pde.py:112-128;
// the area normalisation alpha is a constant of the form).  sigma_vm (top / mid / bottom surface) depends on uhat
// through the membrane and bending strains and through grad(h) (CG1 thickness), all linear in the gradient map k.
//   hv: nodal thickness (ignored when ewm), P: the point evaluated by vm_at at the actual k
FEMO_HD void shape_pnorm_qp(const QpGeom& g, const Tables* tab, int q, const double* we, const double* hv, int ewm,
                            double xi_sign, const VmPoint& P, double wd, double sm, double srho, double* out12) {
  const double pw = pow(sm * P.vm, srho);
  const double G = (P.vm > 0) ? wd * g.Ju * srho * sm * pow(sm * P.vm, srho - 1.0) : 0.0;
  double Gk[4];
#pragma unroll
  for (int ab = 0; ab < 4; ++ab) {
    KMap e; e.k00 = (ab == 0); e.k01 = (ab == 1); e.k10 = (ab == 2); e.k11 = (ab == 3);
    double s[9];
    strains_k(g, tab, q, we, e, 0.0, s);
    double gh0 = 0.0, gh1 = 0.0;
    if (!ewm)
      for (int v = 0; v < 4; ++v) {
        gh0 += 0.5 * xi_sign * hv[v] * (tab->dn1[q][v][0] * e.k00 + tab->dn1[q][v][1] * e.k10);
        gh1 += 0.5 * xi_sign * hv[v] * (tab->dn1[q][v][0] * e.k01 + tab->dn1[q][v][1] * e.k11);
      }
    const double de0 = s[0] - P.xi * s[3] - P.b0 * gh0;
    const double de1 = s[1] - P.xi * s[4] - P.b1 * gh1;
    const double de2 = s[2] - P.xi * s[5] - (P.b0 * gh1 + P.b1 * gh0);
    Gk[ab] = G * (P.t[0] * de0 + P.t[1] * de1 + P.t[2] * de2);
  }
  const double GJ = wd * pw * g.Ju;
  const double dN0[4] = {-(1 - tab->X1[q]), (1 - tab->X1[q]), -tab->X1[q], tab->X1[q]};
  const double dN1[4] = {-(1 - tab->X0[q]), -tab->X0[q], (1 - tab->X0[q]), tab->X0[q]};
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    const ShapeVert sv = shape_vert(g, dN0[v], dN1[v]);
    const double c0 = -(Gk[0] * sv.a0 + Gk[1] * sv.a1), c1 = -(Gk[2] * sv.a0 + Gk[3] * sv.a1);
    const V3 r = c0 * g.Kx0 + c1 * g.Kx1 + GJ * sv.c;
    out12[3 * v] = r.x; out12[3 * v + 1] = r.y; out12[3 * v + 2] = r.z;
  }
}

#endif  // __CUDACC__
