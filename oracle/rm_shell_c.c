/* CPU restatement (C + OpenMP) of the RM-shell tangent-Jacobian assembly -- the timed CPU
 * baseline ("port") of bench.py.  TEST / BASELINE INFRASTRUCTURE ONLY: nothing under
 * femo_alpha_b200/ links or calls this.  PARITY UNPINNED (see oracle/rm_shell_ref.py header);
 * this file is validated against that NumPy oracle in tests/test_oracle_c.py.
 *
 * Follows /root/reference/femo_alpha/rm_shell/linear_shell_fenicsx/linear_shell_model.py:136-157
 * (CLT), :220-258 (strains), :275-296 (energies) and kinematics.py:12-106, in the way the
 * reference's tool-chain evaluates them: a dense strain operator B (9x39) per quadrature
 * point and K_e += w B^T D B, then a scatter-add into CSR (dolfinx MatSetValuesLocal
 * analogue, fea/utils_dolfinx.py:200-206).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define SHEAR_K 0.833

static const int Q2_IJ[9][2] = {{0,0},{2,0},{0,2},{2,2},{1,0},{0,1},{2,1},{1,2},{1,1}};

static void lag2(double x, double* v, double* d) {
  v[0] = 2.0*(x-0.5)*(x-1.0); v[1] = 4.0*x*(1.0-x); v[2] = 2.0*x*(x-0.5);
  d[0] = 4.0*x-3.0; d[1] = 4.0-8.0*x; d[2] = 4.0*x-1.0;
}
static void cross3(const double* a, const double* b, double* c) {
  c[0] = a[1]*b[2]-a[2]*b[1]; c[1] = a[2]*b[0]-a[0]*b[2]; c[2] = a[0]*b[1]-a[1]*b[0];
}
static double dot3(const double* a, const double* b) { return a[0]*b[0]+a[1]*b[1]+a[2]*b[2]; }

static void gauss01(int m, double* x, double* w) {
  for (int i = 0; i < m; ++i) {
    double z = cos(M_PI*(i+0.75)/(m+0.5)), pp = 1;
    for (int it = 0; it < 100; ++it) {
      double p1 = 1, p2 = 0;
      for (int j = 0; j < m; ++j) { double p3 = p2; p2 = p1; p1 = ((2.0*j+1.0)*z*p2 - j*p3)/(j+1.0); }
      pp = m*(z*p1-p2)/(z*z-1.0);
      double dz = p1/pp; z -= dz;
      if (fabs(dz) < 1e-16) break;
    }
    x[m-1-i] = 0.5*(z+1.0); w[m-1-i] = 1.0/((1.0-z*z)*pp*pp);
  }
}

/* K_e (39x39, row-major) of one cell.  xv, uh: 4x3 tensor-ordered; hv/Ev/nuv: 4 nodal values (CG1)
 * or the cell value replicated (DG0). */
static void element_matrix(const double xv[4][3], const double uh[4][3], const double* hv, const double* Ev,
                           const double* nuv, int m, const double* gx, const double* gw, double* Ke) {
  memset(Ke, 0, 39*39*sizeof(double));
  double hE = 0;
  for (int i = 0; i < 4; ++i) for (int j = i+1; j < 4; ++j) {
    double d[3] = {xv[i][0]-xv[j][0], xv[i][1]-xv[j][1], xv[i][2]-xv[j][2]};
    double l = sqrt(dot3(d,d)); if (l > hE) hE = l;
  }
  for (int q0 = 0; q0 < m; ++q0) for (int q1 = 0; q1 < m; ++q1) {
    double X0 = gx[q0], X1 = gx[q1], wq = gw[q0]*gw[q1];
    double N[4] = {(1-X0)*(1-X1), X0*(1-X1), (1-X0)*X1, X0*X1};
    double dN[4][2] = {{-(1-X1),-(1-X0)},{(1-X1),-X0},{-X1,(1-X0)},{X1,X0}};
    double v0[3], d0[3], v1[3], d1[3]; lag2(X0, v0, d0); lag2(X1, v1, d1);
    double dphi[9][2];
    for (int a = 0; a < 9; ++a) { dphi[a][0] = d0[Q2_IJ[a][0]]*v1[Q2_IJ[a][1]]; dphi[a][1] = v0[Q2_IJ[a][0]]*d1[Q2_IJ[a][1]]; }
    double a0[3] = {0,0,0}, a1[3] = {0,0,0}, u0[3] = {0,0,0}, u1[3] = {0,0,0};
    for (int v = 0; v < 4; ++v) for (int i = 0; i < 3; ++i) {
      a0[i] += dN[v][0]*xv[v][i]; a1[i] += dN[v][1]*xv[v][i];
      u0[i] += dN[v][0]*uh[v][i]; u1[i] += dN[v][1]*uh[v][i];
    }
    double g00 = dot3(a0,a0), g01 = dot3(a0,a1), g11 = dot3(a1,a1), dg = g00*g11-g01*g01;
    double detJ = sqrt(dg), Ki[2][3];
    for (int i = 0; i < 3; ++i) { Ki[0][i] = (g11*a0[i]-g01*a1[i])/dg; Ki[1][i] = (g00*a1[i]-g01*a0[i])/dg; }
    double n[3], E0[3], E1[3], E2[3]; cross3(a0, a1, n);
    double nl = sqrt(dot3(n,n));
    for (int i = 0; i < 3; ++i) { E2[i] = n[i]/nl; E0[i] = a0[i]/sqrt(g00); }
    cross3(E2, E0, E1);
    double c[3], dn0[3], dn1[3], dE2[2][3];
    for (int i = 0; i < 3; ++i) c[i] = xv[0][i]-xv[1][i]-xv[2][i]+xv[3][i];
    cross3(a0, c, dn0); cross3(c, a1, dn1);
    double p0 = dot3(dn0,E2), p1 = dot3(dn1,E2);
    for (int i = 0; i < 3; ++i) { dE2[0][i] = (dn0[i]-p0*E2[i])/nl; dE2[1][i] = (dn1[i]-p1*E2[i])/nl; }
    /* F = I + grad(uhat), Finv, J */
    double F[9], Fi[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) {
      const double* ui0 = u0; const double* ui1 = u1;
      F[3*i+j] = (i==j) + ui0[i]*Ki[0][j] + ui1[i]*Ki[1][j];
    }
    double c00 = F[4]*F[8]-F[5]*F[7], c01 = F[5]*F[6]-F[3]*F[8], c02 = F[3]*F[7]-F[4]*F[6];
    double Ju = F[0]*c00+F[1]*c01+F[2]*c02;
    Fi[0] = c00/Ju; Fi[1] = (F[2]*F[7]-F[1]*F[8])/Ju; Fi[2] = (F[1]*F[5]-F[2]*F[4])/Ju;
    Fi[3] = c01/Ju; Fi[4] = (F[0]*F[8]-F[2]*F[6])/Ju; Fi[5] = (F[2]*F[3]-F[0]*F[5])/Ju;
    Fi[6] = c02/Ju; Fi[7] = (F[1]*F[6]-F[0]*F[7])/Ju; Fi[8] = (F[0]*F[4]-F[1]*F[3])/Ju;
    double Kx[2][3];
    for (int a = 0; a < 2; ++a) for (int j = 0; j < 3; ++j) Kx[a][j] = Ki[a][0]*Fi[j]+Ki[a][1]*Fi[3+j]+Ki[a][2]*Fi[6+j];
    double k[2][2] = {{dot3(Kx[0],E0), dot3(Kx[0],E1)}, {dot3(Kx[1],E0), dot3(Kx[1],E1)}};
    double gq[2][3];
    for (int q = 0; q < 2; ++q) for (int i = 0; i < 3; ++i) gq[q][i] = k[0][q]*dE2[0][i]+k[1][q]*dE2[1][i];
    /* strain operator B (9 x 39) */
    double B[9][39]; memset(B, 0, sizeof(B));
    for (int a = 0; a < 9; ++a) {
      double da0 = dphi[a][0]*k[0][0]+dphi[a][1]*k[1][0], da1 = dphi[a][0]*k[0][1]+dphi[a][1]*k[1][1];
      for (int i = 0; i < 3; ++i) {
        int col = 3*a+i;
        B[0][col] = da0*E0[i]; B[1][col] = da1*E1[i]; B[2][col] = da1*E0[i]+da0*E1[i];
        B[6][col] = da0*E2[i]; B[7][col] = da1*E2[i]; B[8][col] = 0.5*(da1*E0[i]-da0*E1[i]);
      }
    }
    for (int b = 0; b < 4; ++b) {
      double db0 = dN[b][0]*k[0][0]+dN[b][1]*k[1][0], db1 = dN[b][0]*k[0][1]+dN[b][1]*k[1][1];
      double x00[3], x11[3], x01[3], x10[3];
      cross3(E0, gq[0], x00); cross3(E1, gq[1], x11); cross3(E0, gq[1], x01); cross3(E1, gq[0], x10);
      for (int i = 0; i < 3; ++i) {
        int col = 27+3*b+i;
        double r00 = N[b]*x00[i]-db0*E1[i], r11 = db1*E0[i]+N[b]*x11[i];
        double r01 = N[b]*x01[i]-db1*E1[i], r10 = db0*E0[i]+N[b]*x10[i];
        B[3][col] = r00; B[4][col] = r11; B[5][col] = r01+r10;
        B[6][col] = N[b]*E1[i]; B[7][col] = -N[b]*E0[i]; B[8][col] = N[b]*E2[i];
      }
    }
    double h = 0, E = 0, nu = 0;
    for (int v = 0; v < 4; ++v) { h += N[v]*hv[v]; E += N[v]*Ev[v]; nu += N[v]*nuv[v]; }
    double c0 = E/(1-nu*nu), G = E/(2*(1+nu)), D[9][9]; memset(D, 0, sizeof(D));
    double C[3][3] = {{c0, c0*nu, 0}, {c0*nu, c0, 0}, {0, 0, c0*0.5*(1-nu)}};
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { D[i][j] = h*C[i][j]; D[3+i][3+j] = h*h*h/12.0*C[i][j]; }
    D[6][6] = D[7][7] = SHEAR_K*G*h*Ju; D[8][8] = E*h*h*h/(hE*hE)*Ju;
    double DB[9][39], wd = wq*detJ;
    for (int r = 0; r < 9; ++r) for (int col = 0; col < 39; ++col) {
      double s = 0; for (int t = 0; t < 9; ++t) s += D[r][t]*B[t][col];
      DB[r][col] = wd*s;
    }
    for (int i = 0; i < 39; ++i) for (int r = 0; r < 9; ++r) {
      double bi = B[r][i]; if (bi == 0.0) continue;
      double* row = Ke+39*i; const double* db = DB[r];
      for (int j = 0; j < 39; ++j) row[j] += bi*db[j];
    }
  }
}

static void gather_cell(int64_t c, const double* xyz, const int32_t* cv, const double* h, const double* E,
                        const double* nu, const double* uhat, int ewm, double xv[4][3], double uh[4][3],
                        double* hv, double* Ev, double* nuv) {
  for (int v = 0; v < 4; ++v) {
    int32_t id = cv[4*c+v];
    for (int i = 0; i < 3; ++i) { xv[v][i] = xyz[3*(int64_t)id+i]; uh[v][i] = uhat ? uhat[3*(int64_t)id+i] : 0.0; }
    if (ewm) { hv[v] = h[c]; Ev[v] = E[c]; nuv[v] = nu[c]; }
    else { hv[v] = h[id]; Ev[v] = E[id]; nuv[v] = nu[id]; }
  }
}

/* element matrices of cells [0,nc) into Ke_out (nc*1521) -- validation against the NumPy oracle */
void rmc_element_matrices(int64_t nc, const double* xyz, const int32_t* cv, const double* h, const double* E,
                          const double* nu, const double* uhat, int ewm, int m, double* Ke_out) {
  double gx[8], gw[8]; gauss01(m, gx, gw);
#pragma omp parallel for schedule(static)
  for (int64_t c = 0; c < nc; ++c) {
    double xv[4][3], uh[4][3], hv[4], Ev[4], nuv[4];
    gather_cell(c, xyz, cv, h, E, nu, uhat, ewm, xv, uh, hv, Ev, nuv);
    element_matrix(xv, uh, hv, Ev, nuv, m, gx, gw, Ke_out+1521*c);
  }
}

/* assemble dR/du into CSR values (pattern precomputed): emap[c*1521 + 39*i + j] = CSR position.
 * Returns the number of OpenMP threads used. */
int rmc_assemble_csr(int64_t nc, const double* xyz, const int32_t* cv, const double* h, const double* E,
                     const double* nu, const double* uhat, int ewm, int m, const int64_t* emap, int64_t nnz,
                     double* vals) {
  double gx[8], gw[8]; gauss01(m, gx, gw);
  memset(vals, 0, sizeof(double)*nnz);
  int nthreads = 1;
#pragma omp parallel
  {
#ifdef _OPENMP
#pragma omp single
    nthreads = omp_get_num_threads();
#endif
    double Ke[1521];
#pragma omp for schedule(static)
    for (int64_t c = 0; c < nc; ++c) {
      double xv[4][3], uh[4][3], hv[4], Ev[4], nuv[4];
      gather_cell(c, xyz, cv, h, E, nu, uhat, ewm, xv, uh, hv, Ev, nuv);
      element_matrix(xv, uh, hv, Ev, nuv, m, gx, gw, Ke);
      const int64_t* em = emap+1521*c;
      for (int k = 0; k < 1521; ++k) {
#pragma omp atomic
        vals[em[k]] += Ke[k];
      }
    }
  }
  return nthreads;
}

/* y = A x for CSR (one PCG-iteration building block of the CPU baseline) */
void rmc_csr_spmv(int64_t n, const int64_t* indptr, const int32_t* indices, const double* vals, const double* x,
                  double* y) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    double s = 0;
    for (int64_t k = indptr[i]; k < indptr[i+1]; ++k) s += vals[k]*x[indices[k]];
    y[i] = s;
  }
}

/* number of OpenMP threads of the following calls (bench.py sets it to the host cores it may use: under
 * torch.distributed.run the environment carries OMP_NUM_THREADS=1, which would make the CPU arm single-threaded) */
void rmc_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}
