// Trilinear hexahedron (Q1, 2x2x2 Gauss) element mathematics of the generic 3-D forms (SURVEY 8 f4 / BASELINE config 5):
// per-thread bodies shared by the kernels of hex.cu and by the host harness tests/hex_host.cu (nvcc compiles them for
// both sides).  The reference has no in-tree form for this configuration -- its generic FEA takes any UFL form
// (femo_alpha/fea/fea_dolfinx.py:71-116); the two forms here are the ones SURVEY 8d names for config 5:
//   form 1  linear elasticity   a(u, v) = int  s_e ( lam div u div v + 2 mu eps(u) : eps(v) ) dx
//   form 0  vector Poisson      a(u, v) = int  s_e  kappa grad u : grad v dx       (three independent components)
// with the SIMP cell factor s_e = rho_e^p and the load  l(v) = int f . v dx  of a nodal (Q1) field f.
// Node order of a cell: VTK_HEXAHEDRON (bottom face counter-clockwise, then the top face).
#pragma once
#include <cstdint>

#ifndef FEMO_HD
#define FEMO_HD __host__ __device__ __forceinline__
#endif

namespace femo_hex {

constexpr int GEO = 10;          // per (cell, point): inverse Jacobian (9, [j][i] = d xi_j / d x_i) and weight * det J

// reference values at Gauss point q (bit 0/1/2 of q = sign of xi/eta/zeta) of shape function a
FEMO_HD void ref_shape(int a, int q, double* N, double dN[3]) {
  const double g = 0.57735026918962576451;
  const double sx = ((a + 1) & 2) ? 1.0 : -1.0;      // a = 1, 2, 5, 6
  const double sy = (a & 2) ? 1.0 : -1.0;            // a = 2, 3, 6, 7
  const double sz = (a & 4) ? 1.0 : -1.0;            // a = 4 .. 7
  const double fx = 1.0 + sx * ((q & 1) ? g : -g);
  const double fy = 1.0 + sy * ((q & 2) ? g : -g);
  const double fz = 1.0 + sz * ((q & 4) ? g : -g);
  *N = 0.125 * fx * fy * fz;
  dN[0] = 0.125 * sx * fy * fz;
  dN[1] = 0.125 * fx * sy * fz;
  dN[2] = 0.125 * fx * fy * sz;
}

// geometry record of (cell, q); returns det J
FEMO_HD double geom_qp(const double* __restrict__ xyz, const int32_t* __restrict__ cn, int q, double* __restrict__ rec) {
  double J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};         // J[i][j] = d x_i / d xi_j
  for (int a = 0; a < 8; ++a) {
    double N, dN[3];
    ref_shape(a, q, &N, dN);
    const double* x = xyz + 3 * (int64_t)cn[a];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) J[3 * i + j] += x[i] * dN[j];
  }
  const double c00 = J[4] * J[8] - J[5] * J[7], c01 = J[5] * J[6] - J[3] * J[8], c02 = J[3] * J[7] - J[4] * J[6];
  const double det = J[0] * c00 + J[1] * c01 + J[2] * c02;
  const double id = 1.0 / det;
  // inverse = adj / det, stored as rec[3 j + i] = d xi_j / d x_i
  rec[0] = c00 * id;                         rec[1] = (J[2] * J[7] - J[1] * J[8]) * id; rec[2] = (J[1] * J[5] - J[2] * J[4]) * id;
  rec[3] = c01 * id;                         rec[4] = (J[0] * J[8] - J[2] * J[6]) * id; rec[5] = (J[2] * J[3] - J[0] * J[5]) * id;
  rec[6] = c02 * id;                         rec[7] = (J[1] * J[6] - J[0] * J[7]) * id; rec[8] = (J[0] * J[4] - J[1] * J[3]) * id;
  rec[9] = det;                              // Gauss weights are 1
  return det;
}

// physical gradient of shape function a at the point of `rec`
FEMO_HD void phys_grad(const double* __restrict__ rec, int a, int q, double* N, double g[3]) {
  double dN[3];
  ref_shape(a, q, N, dN);
  for (int i = 0; i < 3; ++i) g[i] = rec[i] * dN[0] + rec[3 + i] * dN[1] + rec[6 + i] * dN[2];
}

// block of the form from the moment matrix M = sum_q w g_a (x) g_b:  lam M + mu M^T + mu tr(M) I   or   kappa tr(M) I
FEMO_HD void block_from_moments(int form, double c1, double c2, double s, const double M[9], double B[9]) {
  const double tr = M[0] + M[4] + M[8];
  if (form == 1) {
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) B[3 * r + c] = s * (c1 * M[3 * r + c] + c2 * M[3 * c + r] + ((r == c) ? c2 * tr : 0.0));
  } else {
    for (int k = 0; k < 9; ++k) B[k] = 0.0;
    B[0] = B[4] = B[8] = s * c1 * tr;
  }
}

// contributions of one cell to the two 3x3 blocks (row = local node a, columns = local nodes b and b + 1), SIMP factor
// s included: the geometry record and the row gradient are shared by the pair.  `cg` = the 8 records of the cell
// (global memory, or the copy k_hex_assemble stages in shared memory)
FEMO_HD void item_pair_body(int form, double c1, double c2, const double* __restrict__ cg, double s, int a, int b,
                            double B0[9], double B1[9]) {
  double M0[9], M1[9];
  for (int k = 0; k < 9; ++k) { M0[k] = 0.0; M1[k] = 0.0; }
  for (int q = 0; q < 8; ++q) {
    const double* rec = cg + q * GEO;
    double N, ga[3], g0[3], g1[3];
    phys_grad(rec, a, q, &N, ga);
    phys_grad(rec, b, q, &N, g0);
    phys_grad(rec, b + 1, q, &N, g1);
    const double w = rec[9];
    for (int r = 0; r < 3; ++r) {
      const double wg = w * ga[r];
      for (int c = 0; c < 3; ++c) { M0[3 * r + c] += wg * g0[c]; M1[3 * r + c] += wg * g1[c]; }
    }
  }
  block_from_moments(form, c1, c2, s, M0, B0);
  block_from_moments(form, c1, c2, s, M1, B1);
}

// ---- variant with gradient records (FEMO_HEX_GRAD=1, k_hex_assemble_g): per (cell, point) the physical gradients of the
// 8 shape functions and weight * det J, so that an item costs 7 loads and 12 fp64 instructions per point instead of three
// inverse-Jacobian products.  2.6x the record memory of the default path.
constexpr int GEO2 = 26;         // 8 x 3 gradients, weight * det J, one pad (16-byte multiples)

FEMO_HD void geom2_qp(const double* __restrict__ rec, int q, double* __restrict__ rec2) {
  for (int a = 0; a < 8; ++a) {
    double N, g[3];
    phys_grad(rec, a, q, &N, g);
    rec2[3 * a] = g[0]; rec2[3 * a + 1] = g[1]; rec2[3 * a + 2] = g[2];
  }
  rec2[24] = rec[9];
  rec2[25] = 0.0;
}

// contribution of one cell to the block (row = local node a, column = local node b); `cg` = the 8 gradient records of the cell
FEMO_HD void item_body_g(int form, double c1, double c2, const double* __restrict__ cg, double s, int a, int b, double B[9]) {
  double M[9];
  for (int k = 0; k < 9; ++k) M[k] = 0.0;
  for (int q = 0; q < 8; ++q) {
    const double* rec = cg + q * GEO2;
    const double w = rec[24];
    for (int r = 0; r < 3; ++r) {
      const double wg = w * rec[3 * a + r];
      for (int c = 0; c < 3; ++c) M[3 * r + c] += wg * rec[3 * b + c];
    }
  }
  block_from_moments(form, c1, c2, s, M, B);
}

// Row assembly, the algorithm of k_hex_assemble (one warp per row node i): the 8 n_cells(i) work items (cell around i,
// local column b), item t = 8 (cell position in the adjacency of i) + b, are evaluated HEX_PASS at a time -- lane l takes
// the pair (2 l, 2 l + 1) of the pass, two columns of one cell -- into a staging area; then the lane that owns a
// destination block sums the staged items of its column in item (= cell-ascending) order.  Deterministic, no atomics.
constexpr int HEX_PASS = 64;

// items t, t + 1 (t even) of row i -> staged blocks and columns; returns false past the end
FEMO_HD bool row_item_pair(int form, double c1, double c2, const double* __restrict__ geo, const int32_t* __restrict__ cells,
                           const double* __restrict__ scale, const int32_t* __restrict__ nc_ptr,
                           const int32_t* __restrict__ nc_adj, int64_t i, int t, double B0[9], double B1[9], int32_t col[2]) {
  const int32_t e = nc_ptr[i] + (t >> 3);
  if (e >= nc_ptr[i + 1]) return false;
  const int32_t ca = nc_adj[e];
  const int64_t cell = ca >> 3;
  const int b = t & 7;
  col[0] = cells[8 * cell + b];
  col[1] = cells[8 * cell + b + 1];
  item_pair_body(form, c1, c2, geo + cell * 8 * GEO, scale[cell], ca & 7, b, B0, B1);
  return true;
}

// sum of the staged items [0, n_staged) whose column is `col`, in ascending item order; staging is component-major:
// s_val[k * HEX_PASS + t].  Two phases so that a warp does not run the 9-term body once per item with a single lane
// active: (1) a 64-bit match mask, every lane busy; (2) a loop over the set bits (1, 2, 4 or 8 on a structured mesh).
// s_col must hold HEX_PASS readable entries (16-byte aligned on the device); entries past n_staged are ignored.
FEMO_HD void row_collect(const double* __restrict__ s_val, const int32_t* __restrict__ s_col, int n_staged, int32_t col,
                         double acc[9]) {
  unsigned long long m = 0ull;
#ifdef __CUDA_ARCH__
  const int4* c4 = reinterpret_cast<const int4*>(s_col);
#pragma unroll
  for (int j = 0; j < HEX_PASS / 4; ++j) {
    const int4 v = c4[j];
    const unsigned long long bits = (unsigned long long)((v.x == col) | ((v.y == col) << 1) | ((v.z == col) << 2) | ((v.w == col) << 3));
    m |= bits << (4 * j);
  }
#else
  for (int t = 0; t < HEX_PASS; ++t) m |= (unsigned long long)(s_col[t] == col) << t;
#endif
  if (n_staged < HEX_PASS) m &= (1ull << n_staged) - 1ull;
  for (int k = 0; k < 9; ++k) acc[k] = 0.0;
  while (m) {
#ifdef __CUDA_ARCH__
    const int t = __ffsll((long long)m) - 1;
#else
    int t = 0;
    while (!((m >> t) & 1ull)) ++t;
#endif
    m &= m - 1ull;
    for (int k = 0; k < 9; ++k) acc[k] += s_val[k * HEX_PASS + t];
  }
}

// displacement gradient G[i][k] = sum_a u_a[i] d N_a / d x_k of one cell at point q
FEMO_HD void cell_grad(const double* __restrict__ rec, const int32_t* __restrict__ cn, const double* __restrict__ u, int q,
                       double G[9]) {
  for (int k = 0; k < 9; ++k) G[k] = 0.0;
  for (int a = 0; a < 8; ++a) {
    double N, g[3];
    phys_grad(rec, a, q, &N, g);
    const double* ua = u + 3 * (int64_t)cn[a];
    for (int i = 0; i < 3; ++i)
      for (int k = 0; k < 3; ++k) G[3 * i + k] += ua[i] * g[k];
  }
}

// unit-density "stress" of the form: S = lam tr(G) I + mu (G + G^T)   or   kappa G
FEMO_HD void stress0(int form, double c1, double c2, const double G[9], double S[9]) {
  if (form == 1) {
    const double tr = c1 * (G[0] + G[4] + G[8]);
    for (int i = 0; i < 3; ++i)
      for (int k = 0; k < 3; ++k) S[3 * i + k] = c2 * (G[3 * i + k] + G[3 * k + i]) + ((i == k) ? tr : 0.0);
  } else {
    for (int k = 0; k < 9; ++k) S[k] = c1 * G[k];
  }
}

// a0_e(lam, u) = int_e S0(u) : grad(lam) dx   (unit density)
FEMO_HD double cell_energy(int form, double c1, double c2, const double* __restrict__ geo, const int32_t* __restrict__ cells,
                           int64_t cell, const double* __restrict__ u, const double* __restrict__ lm) {
  const int32_t* cn = cells + 8 * cell;
  double acc = 0.0;
  for (int q = 0; q < 8; ++q) {
    const double* rec = geo + (cell * 8 + q) * GEO;
    double G[9], L[9], S[9];
    cell_grad(rec, cn, u, q, G);
    cell_grad(rec, cn, lm, q, L);
    stress0(form, c1, c2, G, S);
    double s = 0.0;
    for (int k = 0; k < 9; ++k) s += S[k] * L[k];
    acc += rec[9] * s;
  }
  return acc;
}

// element vector  ye[3 a + i] = coef * int_e S0(u)[i][:] . grad N_a dx  (= coef * K0_e u_e)
FEMO_HD void cell_apply(int form, double c1, double c2, const double* __restrict__ geo, const int32_t* __restrict__ cells,
                        int64_t cell, const double* __restrict__ u, double coef, double* __restrict__ ye) {
  const int32_t* cn = cells + 8 * cell;
  for (int k = 0; k < 24; ++k) ye[k] = 0.0;
  for (int q = 0; q < 8; ++q) {
    const double* rec = geo + (cell * 8 + q) * GEO;
    double G[9], S[9];
    cell_grad(rec, cn, u, q, G);
    stress0(form, c1, c2, G, S);
    const double w = coef * rec[9];
    for (int a = 0; a < 8; ++a) {
      double N, g[3];
      phys_grad(rec, a, q, &N, g);
      for (int i = 0; i < 3; ++i) ye[3 * a + i] += w * (S[3 * i] * g[0] + S[3 * i + 1] * g[1] + S[3 * i + 2] * g[2]);
    }
  }
}

// element vector of the consistent mass product  ye[3 a + i] = coef * int_e N_a (sum_b N_b f_b[i]) dx
FEMO_HD void cell_mass_apply(const double* __restrict__ geo, const int32_t* __restrict__ cells, int64_t cell,
                             const double* __restrict__ f, double coef, double* __restrict__ ye) {
  const int32_t* cn = cells + 8 * cell;
  for (int k = 0; k < 24; ++k) ye[k] = 0.0;
  for (int q = 0; q < 8; ++q) {
    const double* rec = geo + (cell * 8 + q) * GEO;
    double N[8], dN[3], fq[3] = {0, 0, 0};
    for (int a = 0; a < 8; ++a) {
      ref_shape(a, q, &N[a], dN);
      const double* fa = f + 3 * (int64_t)cn[a];
      for (int i = 0; i < 3; ++i) fq[i] += N[a] * fa[i];
    }
    const double w = coef * rec[9];
    for (int a = 0; a < 8; ++a)
      for (int i = 0; i < 3; ++i) ye[3 * a + i] += w * N[a] * fq[i];
  }
}

// nodal gather of element vectors (adjacency order = cell-ascending): deterministic, no atomics
FEMO_HD void node_gather(const int32_t* __restrict__ nc_ptr, const int32_t* __restrict__ nc_adj,
                         const double* __restrict__ cellvec, int64_t i, double out[3]) {
  out[0] = out[1] = out[2] = 0.0;
  for (int32_t e = nc_ptr[i]; e < nc_ptr[i + 1]; ++e) {
    const double* v = cellvec + 3 * (int64_t)nc_adj[e];
    out[0] += v[0]; out[1] += v[1]; out[2] += v[2];
  }
}

}  // namespace femo_hex
