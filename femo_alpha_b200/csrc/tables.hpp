// Reference-element tables of the Q2 x Q1 quadrilateral (dolfinx tensor vertex order): Gauss-Legendre tensor rule
// on [0,1]^2, Q2 / Q1 basis values and reference gradients.  Host code shared by api.cu and the CPU test harnesses.
#pragma once
#include <cmath>
#include <cstring>
#include "femo_internal.cuh"

inline void gauss01(int m, double* x, double* w) {
  // Gauss-Legendre on [0,1] by Newton iteration on P_m
  for (int i = 0; i < m; ++i) {
    double z = cos(M_PI * (i + 0.75) / (m + 0.5)), pp = 1;
    for (int it = 0; it < 100; ++it) {
      double p1 = 1, p2 = 0;
      for (int j = 0; j < m; ++j) { double p3 = p2; p2 = p1; p1 = ((2.0 * j + 1.0) * z * p2 - j * p3) / (j + 1.0); }
      pp = m * (z * p1 - p2) / (z * z - 1.0);
      double dz = p1 / pp;
      z -= dz;
      if (fabs(dz) < 1e-16) break;
    }
    x[m - 1 - i] = 0.5 * (z + 1.0);
    w[m - 1 - i] = 1.0 / ((1.0 - z * z) * pp * pp);
  }
}

inline void lag2(double x, double* v, double* d) {
  v[0] = 2.0 * (x - 0.5) * (x - 1.0); v[1] = 4.0 * x * (1.0 - x); v[2] = 2.0 * x * (x - 0.5);
  d[0] = 4.0 * x - 3.0; d[1] = 4.0 - 8.0 * x; d[2] = 4.0 * x - 1.0;
}

inline void build_tables(Tables& T, int m) {
  static const int IJ[9][2] = {{0, 0}, {2, 0}, {0, 2}, {2, 2}, {1, 0}, {0, 1}, {2, 1}, {1, 2}, {1, 1}};
  memset(&T, 0, sizeof(T));
  T.m = m; T.nq = m * m;
  double gx[FEMO_MAX_Q1D], gw[FEMO_MAX_Q1D];
  gauss01(m, gx, gw);
  for (int i = 0; i < m; ++i)
    for (int j = 0; j < m; ++j) {
      int q = i * m + j;
      double X0 = gx[i], X1 = gx[j];
      T.X0[q] = X0; T.X1[q] = X1; T.W[q] = gw[i] * gw[j];
      double v0[3], d0[3], v1[3], d1[3];
      lag2(X0, v0, d0); lag2(X1, v1, d1);
      for (int a = 0; a < 9; ++a) {
        T.phi2[q][a] = v0[IJ[a][0]] * v1[IJ[a][1]];
        T.dphi2[q][a][0] = d0[IJ[a][0]] * v1[IJ[a][1]];
        T.dphi2[q][a][1] = v0[IJ[a][0]] * d1[IJ[a][1]];
      }
      double N[4] = {(1 - X0) * (1 - X1), X0 * (1 - X1), (1 - X0) * X1, X0 * X1};
      double dN[4][2] = {{-(1 - X1), -(1 - X0)}, {(1 - X1), -X0}, {-X1, (1 - X0)}, {X1, X0}};
      for (int v = 0; v < 4; ++v) { T.n1[q][v] = N[v]; T.dn1[q][v][0] = dN[v][0]; T.dn1[q][v][1] = dN[v][1]; }
    }
}
