/* C + OpenMP restatement of oracle/hex_ref.py (generic 3-D forms on Q1 hexahedra): the timed CPU
 * baseline of scripts/hex_bench.py.  TEST / BASELINE INFRASTRUCTURE ONLY -- nothing under femo_alpha_b200/ links this.
 * PARITY UNPINNED like the Python oracle (no reference form exists for this configuration); checked against hex_ref.py in
 * tests/test_oracle_hex.py.  The assembly follows what dolfinx + PETSc do for the reference (fea/utils_dolfinx.py:200-206):
 * one dense element matrix per cell (2x2x2 Gauss), added into a pre-built CSR pattern through a precomputed
 * element -> CSR position map (MatSetValuesLocal without the per-row search), cells distributed over threads with
 * atomic adds. */
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include <stdint.h>
#include <string.h>

static const double G = 0.57735026918962576451;
static const double SX[8] = {-1, 1, 1, -1, -1, 1, 1, -1}, SY[8] = {-1, -1, 1, 1, -1, -1, 1, 1}, SZ[8] = {-1, -1, -1, -1, 1, 1, 1, 1};

/* element matrix (24 x 24, dof 3 a + i) of one cell, scaled by s; form 1 elasticity (c1 = lam, c2 = mu), 0 vector Poisson (c1 = kappa) */
static void hex_ke(const double* xyz, const int32_t* cn, int form, double c1, double c2, double s, double* Ke) {
  memset(Ke, 0, sizeof(double) * 576);
  for (int q = 0; q < 8; ++q) {
    const double xi = (q & 1) ? G : -G, eta = (q & 2) ? G : -G, zeta = (q & 4) ? G : -G;
    double dN[8][3], J[9] = {0};
    for (int a = 0; a < 8; ++a) {
      const double fx = 1 + SX[a] * xi, fy = 1 + SY[a] * eta, fz = 1 + SZ[a] * zeta;
      dN[a][0] = 0.125 * SX[a] * fy * fz; dN[a][1] = 0.125 * fx * SY[a] * fz; dN[a][2] = 0.125 * fx * fy * SZ[a];
      const double* x = xyz + 3 * (int64_t)cn[a];
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) J[3 * i + j] += x[i] * dN[a][j];
    }
    const double c00 = J[4] * J[8] - J[5] * J[7], c01 = J[5] * J[6] - J[3] * J[8], c02 = J[3] * J[7] - J[4] * J[6];
    const double det = J[0] * c00 + J[1] * c01 + J[2] * c02, id = 1.0 / det;
    const double Ji[9] = {c00 * id, (J[2] * J[7] - J[1] * J[8]) * id, (J[1] * J[5] - J[2] * J[4]) * id,
                          c01 * id, (J[0] * J[8] - J[2] * J[6]) * id, (J[2] * J[3] - J[0] * J[5]) * id,
                          c02 * id, (J[1] * J[6] - J[0] * J[7]) * id, (J[0] * J[4] - J[1] * J[3]) * id};   /* [j][i] = dxi_j/dx_i */
    double g[8][3];
    for (int a = 0; a < 8; ++a)
      for (int i = 0; i < 3; ++i) g[a][i] = Ji[i] * dN[a][0] + Ji[3 + i] * dN[a][1] + Ji[6 + i] * dN[a][2];
    const double w = s * det;
    for (int a = 0; a < 8; ++a)
      for (int b = 0; b < 8; ++b) {
        const double d = g[a][0] * g[b][0] + g[a][1] * g[b][1] + g[a][2] * g[b][2];
        double* K = Ke + (3 * a) * 24 + 3 * b;
        if (form == 1) {
          for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) K[r * 24 + c] += w * (c1 * g[a][r] * g[b][c] + c2 * g[b][r] * g[a][c]);
          K[0] += w * c2 * d; K[25] += w * c2 * d; K[50] += w * c2 * d;
        } else {
          K[0] += w * c1 * d; K[25] += w * c1 * d; K[50] += w * c1 * d;
        }
      }
  }
}

void hexc_element_matrices(int64_t nc, const double* xyz, const int32_t* cells, const double* rho, double p, int form, double c1,
                           double c2, double* Ke) {
#pragma omp parallel for schedule(static)
  for (int64_t e = 0; e < nc; ++e) hex_ke(xyz, cells + 8 * e, form, c1, c2, pow(rho[e], p), Ke + e * 576);
}

/* values of a pre-patterned CSR matrix: emap[e * 576 + k] = position of entry k of cell e */
int hexc_assemble_csr(int64_t nc, const double* xyz, const int32_t* cells, const double* rho, double p, int form, double c1,
                      double c2, const int64_t* emap, int64_t nnz, double* vals) {
  memset(vals, 0, sizeof(double) * nnz);
#pragma omp parallel
  {
    double Ke[576];
#pragma omp for schedule(static)
    for (int64_t e = 0; e < nc; ++e) {
      hex_ke(xyz, cells + 8 * e, form, c1, c2, pow(rho[e], p), Ke);
      const int64_t* m = emap + e * 576;
      for (int k = 0; k < 576; ++k) {
#pragma omp atomic
        vals[m[k]] += Ke[k];
      }
    }
  }
  return 0;
}

/* number of OpenMP threads of the following calls (bench.py sets it to the host cores it may use: under
 * torch.distributed.run the environment carries OMP_NUM_THREADS=1, which would make the CPU arm single-threaded) */
void hexc_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}
