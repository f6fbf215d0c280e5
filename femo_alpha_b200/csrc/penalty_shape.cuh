// Shape derivative of the penalty-BC facet terms (linear_shell_model.py:323-333): the facet measure carries
// Nanson's factor sc = |J F^-T N| of the mesh displacement uhat, so
//     R_pen(w; uhat) = sum_(edge,cell) sum_k  coef_k sc_k(uhat) (tr (x) tr) w ,      coef_k = gw_k L beta / h_E ,
// and with F = I + sum_v uhat_v (x) G_v (G_v the surface gradient of the vertex shape function),  b = F^-T N,
// c_v = F^-T G_v:
//     d sc / d uhat_v = J ( |b| c_v - (b . c_v) b / |b| ) .
// __host__ __device__ bodies shared by penalty_shape.cu and the CPU harness tests/shape_host.cu.
#pragma once
#include "femo_internal.cuh"

struct PenQp {
  double coef;      // gw L beta / hE   (without Nanson's factor)
  double sc;        // Nanson's factor
  V3 dsc[4];        // d sc / d uhat_v, tensor vertex order
  double tr[5];     // traces of (q2 va, q2 mid, q2 vb, q1 va, q1 vb)
};

FEMO_HD void penalty_qp(const V3* xv, const V3* uh, int le, int k, double hE, PenQp& o) {
  const int LE[4][2] = {{0, 1}, {0, 2}, {1, 3}, {2, 3}};
  const double gx[3] = {0.5 - 0.5 * 0.7745966692414834, 0.5, 0.5 + 0.5 * 0.7745966692414834};
  const double gw[3] = {5.0 / 18.0, 8.0 / 18.0, 5.0 / 18.0};   // degree-4 facet rule (utils_dolfinx.py:556)
  V3 t = xv[LE[le][1]] - xv[LE[le][0]];
  const double L = sqrt(dot(t, t));
  t = (1.0 / L) * t;
  const double s = gx[k];
  const double X0 = (le == 0 || le == 3) ? s : (le == 2 ? 1.0 : 0.0);
  const double X1 = (le == 1 || le == 2) ? s : (le == 3 ? 1.0 : 0.0);
  QpGeom g;
  qp_geometry(xv, uh, X0, X1, g);
  const V3 Nf = cross(t, g.E2);
  const double* Fi = g.Finv;
  const V3 b = v3(Fi[0] * Nf.x + Fi[3] * Nf.y + Fi[6] * Nf.z, Fi[1] * Nf.x + Fi[4] * Nf.y + Fi[7] * Nf.z,
                  Fi[2] * Nf.x + Fi[5] * Nf.y + Fi[8] * Nf.z);
  const double nb = sqrt(dot(b, b));
  o.sc = g.Ju * nb;
  o.coef = gw[k] * L * FEMO_BETA / hE;
  const double dN0[4] = {-(1 - X1), (1 - X1), -X1, X1};
  const double dN1[4] = {-(1 - X0), -X0, (1 - X0), X0};
  for (int v = 0; v < 4; ++v) {
    const V3 cv = dN0[v] * g.Kx0 + dN1[v] * g.Kx1;          // F^-T G_v
    o.dsc[v] = g.Ju * (nb * cv - (dot(b, cv) / nb) * b);
  }
  o.tr[0] = 2.0 * (s - 0.5) * (s - 1.0); o.tr[1] = 4.0 * s * (1.0 - s); o.tr[2] = 2.0 * s * (s - 0.5);
  o.tr[3] = 1.0 - s; o.tr[4] = s;
}

// one (edge, cell) pair: we, le_ = the 5 trace-node dof triples of the state and of lambda; out12 = (dR_pen/duhat)^T lam
FEMO_HD void penalty_shape_T_pair(const V3* xv, const V3* uh, int le, double hE, const double* we, const double* le_,
                                  double* out12) {
  for (int i = 0; i < 12; ++i) out12[i] = 0.0;
  for (int k = 0; k < 3; ++k) {
    PenQp q;
    penalty_qp(xv, uh, le, k, hE, q);
    double S = 0.0;
    for (int i = 0; i < 5; ++i)
      for (int j = 0; j < 5; ++j)
        if ((i < 3) == (j < 3))
          S += q.tr[i] * q.tr[j] * (le_[3 * i] * we[3 * j] + le_[3 * i + 1] * we[3 * j + 1] + le_[3 * i + 2] * we[3 * j + 2]);
    for (int v = 0; v < 4; ++v) {
      out12[3 * v] += q.coef * S * q.dsc[v].x; out12[3 * v + 1] += q.coef * S * q.dsc[v].y; out12[3 * v + 2] += q.coef * S * q.dsc[v].z;
    }
  }
}

// one pair: duh = the 4 vertex perturbations of uhat; out15 = (dR_pen/duhat) duhat on the 5 trace nodes
FEMO_HD void penalty_shape_fwd_pair(const V3* xv, const V3* uh, int le, double hE, const double* we, const V3* duh,
                                    double* out15) {
  for (int i = 0; i < 15; ++i) out15[i] = 0.0;
  for (int k = 0; k < 3; ++k) {
    PenQp q;
    penalty_qp(xv, uh, le, k, hE, q);
    double d = 0.0;
    for (int v = 0; v < 4; ++v) d += dot(q.dsc[v], duh[v]);
    for (int i = 0; i < 5; ++i)
      for (int j = 0; j < 5; ++j)
        if ((i < 3) == (j < 3)) {
          const double m = q.coef * d * q.tr[i] * q.tr[j];
          out15[3 * i] += m * we[3 * j]; out15[3 * i + 1] += m * we[3 * j + 1]; out15[3 * i + 2] += m * we[3 * j + 2];
        }
  }
}
