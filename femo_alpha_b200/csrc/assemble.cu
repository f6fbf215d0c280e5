// Deterministic assembly of the RM-shell tangent Jacobian dR/du into BSR(3x3).
//
// Replaces assembleMatrix(dR_du) / NonlinearProblem.J (femo_alpha/fea/utils_dolfinx.py:200-221,
// csdl_alpha_opt/state_operation.py:289-294), i.e. FFCx element kernels + dolfinx
// MatSetValuesLocal, for the forms of rm_shell/linear_shell_fenicsx/linear_shell_model.py:220-333.
//
// Two kernels:
//   k_cell_records   one thread per (cell, quadrature point): geometry + CLT coefficients
//   k_gather<PLANAR> one thread per element-block CONTRIBUTION, in destination order (the
//                    precomputed cell->nnz permutation); contributions to the same BSR block
//                    are summed in a fixed order through shared memory, each block of the
//                    matrix is written exactly once -> bitwise reproducible, no atomics.
//   k_warped_elem + k_warped_gather   warped meshes: element matrices per group of 32 cells, then one thread per block
#include <algorithm>
#include <vector>
#include "femo_internal.cuh"

namespace {

// ------------------------------------------------------------------------------------------
__global__ void k_cell_records(const double* __restrict__ xyz, const int32_t* __restrict__ cv,
                               const double* __restrict__ uhat, const double* __restrict__ h,
                               const double* __restrict__ E, const double* __restrict__ nu,
                               const double* __restrict__ hE, const Tables* __restrict__ tab, int64_t nc, int nq,
                               int ewm, int nqf, int stride, double* __restrict__ rec) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= nc * nq) return;
  int64_t cell = i / nq; int q = (int)(i - cell * nq);
  V3 xv[4], uh[4];
  int32_t vid[4];
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    vid[v] = cv[cell * 4 + v];
    xv[v] = ldv3(xyz + 3 * (int64_t)vid[v]);
    if (uhat) uh[v] = ldv3(uhat + 3 * (int64_t)vid[v]);
  }
  QpGeom g;
  qp_geometry(xv, uhat ? uh : nullptr, tab->X0[q], tab->X1[q], g);
  double hq, Eq, nuq;
  if (ewm) { hq = h[cell]; Eq = E[cell]; nuq = nu[cell]; }
  else {
    hq = Eq = nuq = 0;
#pragma unroll
    for (int v = 0; v < 4; ++v) { double N = tab->n1[q][v]; hq += N * h[vid[v]]; Eq += N * E[vid[v]]; nuq += N * nu[vid[v]]; }
  }
  Coefs cf = clt_coefs(hq, Eq, nuq, hE[cell], g.Ju, 0);
  double wd = tab->W[q] * g.detJ;
  double* r = rec + cell * stride + REC_HEADER + q * nqf;
  V3 E0 = g.E0, E1 = g.E1;
  if (nqf == REC_NQF_PLANAR) {
    // planar cell: any in-plane orthonormal frame gives the same energy (isotropic CLT), so one
    // constant frame per cell (taken at the cell centre) is used at every quadrature point.
    QpGeom gc;
    qp_geometry(xv, nullptr, 0.5, 0.5, gc);
    E0 = gc.E0; E1 = gc.E1;
    if (q == 0) {
      double* hd = rec + cell * stride;
      hd[0] = gc.E0.x; hd[1] = gc.E0.y; hd[2] = gc.E0.z;
      hd[3] = gc.E1.x; hd[4] = gc.E1.y; hd[5] = gc.E1.z;
      hd[6] = gc.E2.x; hd[7] = gc.E2.y; hd[8] = gc.E2.z;
    }
  }
  r[REC_K00] = dot(g.Kx0, E0); r[REC_K01] = dot(g.Kx0, E1);
  r[REC_K10] = dot(g.Kx1, E0); r[REC_K11] = dot(g.Kx1, E1);
  r[REC_CM11] = wd * cf.cm11; r[REC_CM12] = wd * cf.cm12; r[REC_CM33] = wd * cf.cm33;
  r[REC_HB] = hq * hq / 12.0;
  r[REC_CS] = wd * cf.cs; r[REC_CD] = wd * cf.cd;
  if (nqf == REC_NQF_WARPED) {
    double* e = r + REC_E0;
    e[0] = g.E0.x; e[1] = g.E0.y; e[2] = g.E0.z; e[3] = g.E1.x; e[4] = g.E1.y; e[5] = g.E1.z;
    e[6] = g.E2.x; e[7] = g.E2.y; e[8] = g.E2.z;
    double k00 = r[REC_K00], k01 = r[REC_K01], k10 = r[REC_K10], k11 = r[REC_K11];
    V3 g0 = k00 * g.dE2_0 + k10 * g.dE2_1, g1 = k01 * g.dE2_0 + k11 * g.dE2_1;
    double* gg = r + REC_G0;
    gg[0] = g0.x; gg[1] = g0.y; gg[2] = g0.z; gg[3] = g1.x; gg[4] = g1.y; gg[5] = g1.z;
  }
}

// ------------------------------------------------------------------------------------------
struct SmemTab {
  double dphi2[FEMO_MAX_NQ][9][2];
  double n1[FEMO_MAX_NQ][4];
  double dn1[FEMO_MAX_NQ][4][2];
};

// 3x3 block (a,b) of one element matrix, planar cell: local-frame accumulation, one rotation
__device__ __forceinline__ void block_planar(const double* __restrict__ rc, const SmemTab& T, int nq, int a, int b,
                                             double* K) {
  const V3 E0 = ldv3(rc), E1 = ldv3(rc + 3), E2 = ldv3(rc + 6);
  const double* r = rc + REC_HEADER;
#pragma unroll
  for (int i = 0; i < 9; ++i) K[i] = 0.0;
  if (a < 9 && b < 9) {
    double m00 = 0, m01 = 0, m10 = 0, m11 = 0, s = 0;
    for (int q = 0; q < nq; ++q, r += REC_NQF_PLANAR) {
      double k00 = r[REC_K00], k01 = r[REC_K01], k10 = r[REC_K10], k11 = r[REC_K11];
      double da0 = T.dphi2[q][a][0] * k00 + T.dphi2[q][a][1] * k10, da1 = T.dphi2[q][a][0] * k01 + T.dphi2[q][a][1] * k11;
      double db0 = T.dphi2[q][b][0] * k00 + T.dphi2[q][b][1] * k10, db1 = T.dphi2[q][b][0] * k01 + T.dphi2[q][b][1] * k11;
      double P00 = da0 * db0, P01 = da0 * db1, P10 = da1 * db0, P11 = da1 * db1;
      double cm11 = r[REC_CM11], cm12 = r[REC_CM12], cm33 = r[REC_CM33], cd4 = 0.25 * r[REC_CD], cs = r[REC_CS];
      m00 += cm11 * P00 + (cm33 + cd4) * P11;
      m01 += cm12 * P01 + (cm33 - cd4) * P10;
      m10 += cm12 * P10 + (cm33 - cd4) * P01;
      m11 += cm11 * P11 + (cm33 + cd4) * P00;
      s += cs * (P00 + P11);
    }
    outer_acc(K, m00 * E0 + m10 * E1, E0);
    outer_acc(K, m01 * E0 + m11 * E1, E1);
    outer_acc(K, s * E2, E2);
  } else if (a >= 9 && b >= 9) {
    int ta = a - 9, tb = b - 9;
    double t00 = 0, t01 = 0, t10 = 0, t11 = 0, t22 = 0;
    for (int q = 0; q < nq; ++q, r += REC_NQF_PLANAR) {
      double k00 = r[REC_K00], k01 = r[REC_K01], k10 = r[REC_K10], k11 = r[REC_K11];
      double da0 = T.dn1[q][ta][0] * k00 + T.dn1[q][ta][1] * k10, da1 = T.dn1[q][ta][0] * k01 + T.dn1[q][ta][1] * k11;
      double db0 = T.dn1[q][tb][0] * k00 + T.dn1[q][tb][1] * k10, db1 = T.dn1[q][tb][0] * k01 + T.dn1[q][tb][1] * k11;
      double NN = T.n1[q][ta] * T.n1[q][tb];
      double hb = r[REC_HB];
      double cb11 = hb * r[REC_CM11], cb12 = hb * r[REC_CM12], cb33 = hb * r[REC_CM33], cs = r[REC_CS];
      double P00 = da0 * db0, P01 = da0 * db1, P10 = da1 * db0, P11 = da1 * db1;
      t00 += cb11 * P11 + cb33 * P00 + cs * NN;
      t11 += cb11 * P00 + cb33 * P11 + cs * NN;
      t01 -= cb12 * P10 + cb33 * P01;
      t10 -= cb12 * P01 + cb33 * P10;
      t22 += r[REC_CD] * NN;
    }
    outer_acc(K, t00 * E0 + t10 * E1, E0);
    outer_acc(K, t01 * E0 + t11 * E1, E1);
    outer_acc(K, t22 * E2, E2);
  } else {
    // displacement node ua, rotation node tb; the (theta,u) block is the transpose
    bool tr = (a >= 9);
    int ua = tr ? b : a, tb = (tr ? a : b) - 9;
    double A0 = 0, A1 = 0, D0 = 0, D1 = 0;
    for (int q = 0; q < nq; ++q, r += REC_NQF_PLANAR) {
      double k00 = r[REC_K00], k01 = r[REC_K01], k10 = r[REC_K10], k11 = r[REC_K11];
      double da0 = T.dphi2[q][ua][0] * k00 + T.dphi2[q][ua][1] * k10, da1 = T.dphi2[q][ua][0] * k01 + T.dphi2[q][ua][1] * k11;
      double N = T.n1[q][tb];
      double cs = r[REC_CS] * N, cd = r[REC_CD] * N;
      A0 += cs * da0; A1 += cs * da1; D0 += cd * da0; D1 += cd * da1;
    }
    // K[u_i, th_j] = A0 E2 E1^T - A1 E2 E0^T + 0.5 D1 E0 E2^T - 0.5 D0 E1 E2^T
    double Kl[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    outer_acc(Kl, E2, A0 * E1 - A1 * E0);
    outer_acc(Kl, 0.5 * (D1 * E0 - D0 * E1), E2);
    if (!tr) {
#pragma unroll
      for (int i = 0; i < 9; ++i) K[i] = Kl[i];
    } else {
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) K[3 * i + j] = Kl[3 * j + i];
    }
  }
}

// general (warped / per-point frame) version (host + device: tests/warped_host.cu checks warped_uu / warped_ut / warped_tt against it)
__host__ __device__ __forceinline__ void block_general(const double* __restrict__ rc, const SmemTab& T, int nq, int a, int b,
                                              double* K) {
  const double* r = rc + REC_HEADER;
#pragma unroll
  for (int i = 0; i < 9; ++i) K[i] = 0.0;
  for (int q = 0; q < nq; ++q, r += REC_NQF_WARPED) {
    double k00 = r[REC_K00], k01 = r[REC_K01], k10 = r[REC_K10], k11 = r[REC_K11];
    V3 E0 = ldv3(r + REC_E0), E1 = ldv3(r + REC_E0 + 3), E2 = ldv3(r + REC_E0 + 6);
    double cm11 = r[REC_CM11], cm12 = r[REC_CM12], cm33 = r[REC_CM33], hb = r[REC_HB], cs = r[REC_CS], cd = r[REC_CD];
    if (a < 9 && b < 9) {
      double da0 = T.dphi2[q][a][0] * k00 + T.dphi2[q][a][1] * k10, da1 = T.dphi2[q][a][0] * k01 + T.dphi2[q][a][1] * k11;
      double db0 = T.dphi2[q][b][0] * k00 + T.dphi2[q][b][1] * k10, db1 = T.dphi2[q][b][0] * k01 + T.dphi2[q][b][1] * k11;
      URows A = u_rows(da0, da1, E0, E1, E2), B = u_rows(db0, db1, E0, E1, E2);
      outer_acc(K, A.e0, cm11 * B.e0 + cm12 * B.e1);
      outer_acc(K, A.e1, cm12 * B.e0 + cm11 * B.e1);
      outer_acc(K, A.e2, cm33 * B.e2);
      outer_acc(K, A.g0, cs * B.g0);
      outer_acc(K, A.g1, cs * B.g1);
      outer_acc(K, A.dr, cd * B.dr);
    } else if (a >= 9 && b >= 9) {
      int ta = a - 9, tb = b - 9;
      V3 g0 = ldv3(r + REC_G0), g1 = ldv3(r + REC_G0 + 3);
      double da0 = T.dn1[q][ta][0] * k00 + T.dn1[q][ta][1] * k10, da1 = T.dn1[q][ta][0] * k01 + T.dn1[q][ta][1] * k11;
      double db0 = T.dn1[q][tb][0] * k00 + T.dn1[q][tb][1] * k10, db1 = T.dn1[q][tb][0] * k01 + T.dn1[q][tb][1] * k11;
      TRows A = t_rows(T.n1[q][ta], da0, da1, E0, E1, E2, g0, g1), B = t_rows(T.n1[q][tb], db0, db1, E0, E1, E2, g0, g1);
      double cb11 = hb * cm11, cb12 = hb * cm12, cb33 = hb * cm33;
      outer_acc(K, A.k0, cb11 * B.k0 + cb12 * B.k1);
      outer_acc(K, A.k1, cb12 * B.k0 + cb11 * B.k1);
      outer_acc(K, A.k2, cb33 * B.k2);
      outer_acc(K, A.g0, cs * B.g0);
      outer_acc(K, A.g1, cs * B.g1);
      outer_acc(K, A.dr, cd * B.dr);
    } else {
      bool tr = (a >= 9);
      int ua = tr ? b : a, tb = (tr ? a : b) - 9;
      V3 g0 = ldv3(r + REC_G0), g1 = ldv3(r + REC_G0 + 3);
      double da0 = T.dphi2[q][ua][0] * k00 + T.dphi2[q][ua][1] * k10, da1 = T.dphi2[q][ua][0] * k01 + T.dphi2[q][ua][1] * k11;
      double db0 = T.dn1[q][tb][0] * k00 + T.dn1[q][tb][1] * k10, db1 = T.dn1[q][tb][0] * k01 + T.dn1[q][tb][1] * k11;
      URows A = u_rows(da0, da1, E0, E1, E2);
      TRows B = t_rows(T.n1[q][tb], db0, db1, E0, E1, E2, g0, g1);
      if (!tr) {
        outer_acc(K, A.g0, cs * B.g0); outer_acc(K, A.g1, cs * B.g1); outer_acc(K, A.dr, cd * B.dr);
      } else {
        outer_acc(K, cs * B.g0, A.g0); outer_acc(K, cs * B.g1, A.g1); outer_acc(K, cd * B.dr, A.dr);
      }
    }
  }
}

#define GATHER_THREADS 256

template <bool PLANAR>
__global__ void __launch_bounds__(GATHER_THREADS)
k_gather(const int32_t* __restrict__ centry, const int32_t* __restrict__ cdest, const int32_t* __restrict__ cptr,
         const double* __restrict__ rec, int stride, const Tables* __restrict__ tab, int nq, int64_t ncontrib,
         double* __restrict__ vals) {
  __shared__ SmemTab T;
  __shared__ double sK[GATHER_THREADS][9];
  {
    double* d = (double*)&T;
    const double* s0 = &tab->dphi2[0][0][0];
    for (int i = threadIdx.x; i < FEMO_MAX_NQ * 18; i += blockDim.x) d[i] = s0[i];
    double* d1 = &T.n1[0][0]; const double* s1 = &tab->n1[0][0];
    for (int i = threadIdx.x; i < FEMO_MAX_NQ * 4; i += blockDim.x) d1[i] = s1[i];
    double* d2 = &T.dn1[0][0][0]; const double* s2 = &tab->dn1[0][0][0];
    for (int i = threadIdx.x; i < FEMO_MAX_NQ * 8; i += blockDim.x) d2[i] = s2[i];
  }
  __syncthreads();
  const int64_t i = blockIdx.x * (int64_t)GATHER_THREADS + threadIdx.x;
  double K[9];
  int32_t dest = -1;
  if (i < ncontrib) {
    int32_t e = centry[i];
    dest = cdest[i];
    int32_t cell = e / 169; int pair = e - cell * 169;
    int a = pair / 13, b = pair - a * 13;
    const double* rc = rec + (int64_t)cell * stride;
    if (PLANAR) block_planar(rc, T, nq, a, b, K); else block_general(rc, T, nq, a, b, K);
#pragma unroll
    for (int k = 0; k < 9; ++k) sK[threadIdx.x][k] = K[k];
  }
  __syncthreads();
  if (i >= ncontrib) return;
  const int32_t first = cptr[dest];
  if ((int64_t)first != i) return;               // not the head of its segment
  const int32_t len = cptr[dest + 1] - first;
  for (int j = 1; j < len; ++j) {                 // fixed (cell-ascending) summation order
    int t = threadIdx.x + j;
    if (t < GATHER_THREADS) {
#pragma unroll
      for (int k = 0; k < 9; ++k) K[k] += sK[t][k];
    } else {                                      // segment spills into the next CTA: recompute
      int32_t e = centry[i + j];
      int32_t cell = e / 169; int pair = e - cell * 169;
      int a = pair / 13, b = pair - a * 13;
      double K2[9];
      const double* rc = rec + (int64_t)cell * stride;
      if (PLANAR) block_planar(rc, T, nq, a, b, K2); else block_general(rc, T, nq, a, b, K2);
#pragma unroll
      for (int k = 0; k < 9; ++k) K[k] += K2[k];
    }
  }
  double* o = vals + (int64_t)dest * 9;
#pragma unroll
  for (int k = 0; k < 9; ++k) o[k] = K[k];
}

// ------------------------------------------------------------------------------------------
// Warped (non-planar) cells, cell-centric ("warped cells").  k_gather<false> re-reads the 3.2 KB
// record of a cell once per contribution (169 times): L2-bound.  Here
//   k_warped_elem    one CTA per group of 32 cells: the group's records are staged in shared memory once, lane = cell,
//                    a warp takes one task = node a with WP_NB consecutive nodes b >= a of the same kind at a time
//                    (the rows of a, scaled by the CLT coefficients, are shared by the tile); the 91 blocks a <= b of the
//                    symmetric element matrix go to a scratch array  elem[group][pair][lane][10] (9 + 1 pad);
//   k_warped_gather  one thread per BSR block: sums its contributions (block or transposed block) from the scratch in
//                    list order and the CTA writes 256 blocks as one contiguous run.
// Every block of the matrix is written once, the summation order is fixed: bitwise reproducible, no atomics.
#define WP_PAIRS 91
#define WP_BS 10                      /* doubles per block in the scratch: 9 + 1 pad, so that a block is five 16-byte loads */
#define WP_THREADS 256
#define WP_MAX_TASKS 91

struct WpTasks { int n, n_uu; unsigned char a[WP_MAX_TASKS], b0[WP_MAX_TASKS], nb[WP_MAX_TASKS]; };

__host__ __device__ __forceinline__ int wp_pair(int a, int b) { return a * 13 - a * (a - 1) / 2 + (b - a); }   // a <= b

#define WP_RF(f) r[(f) * 32]
#define WP_LOAD_FRAME                                                                                                        \
  const double k00 = WP_RF(REC_K00), k01 = WP_RF(REC_K01), k10 = WP_RF(REC_K10), k11 = WP_RF(REC_K11);                       \
  const V3 E0 = v3(WP_RF(REC_E0), WP_RF(REC_E0 + 1), WP_RF(REC_E0 + 2)),                                                     \
           E1 = v3(WP_RF(REC_E0 + 3), WP_RF(REC_E0 + 4), WP_RF(REC_E0 + 5)),                                                 \
           E2 = v3(WP_RF(REC_E0 + 6), WP_RF(REC_E0 + 7), WP_RF(REC_E0 + 8));

// Displacement - displacement blocks.  With d = the two physical gradient components of a node in the local frame,
//   K_ab = sum_q  da0 db0 M00 + da1 db0 M10 + da0 db1 M10^T + da1 db1 M11
// where the 3x3 matrices depend on the point only (weighted CLT coefficients, c3p/m = cm33 +- cd/4):
//   M00 = cm11 E0 E0^T + c3p E1 E1^T + cs E2 E2^T,  M11 = cm11 E1 E1^T + c3p E0 E0^T + cs E2 E2^T,  M10 = cm12 E1 E0^T + c3m E0 E1^T
// (u_rows: e0 = d0 E0, e1 = d1 E1, e2 = d1 E0 + d0 E1, g0 = d0 E2, g1 = d1 E2, dr = (d1 E0 - d0 E1)/2, expanded).
// The "U" records hold k (4) and M00 (6), M11 (6), M10 (9): 25 fields like the frame records.
#define REC_U_M00 4
#define REC_U_M11 10
#define REC_U_M10 16
__host__ __device__ __forceinline__ void wp_m_tensors(V3 E0, V3 E1, V3 E2, double cm11, double cm12, double cm33, double cs,
                                                      double cd, double* m /* 21 */) {
  const double c3p = cm33 + 0.25 * cd, c3m = cm33 - 0.25 * cd;
  const double a0[3] = {E0.x, E0.y, E0.z}, a1[3] = {E1.x, E1.y, E1.z}, a2[3] = {E2.x, E2.y, E2.z};
  int n = 0;
  for (int i = 0; i < 3; ++i)
    for (int j = i; j < 3; ++j) {
      m[n] = cm11 * a0[i] * a0[j] + c3p * a1[i] * a1[j] + cs * a2[i] * a2[j];
      m[6 + n] = cm11 * a1[i] * a1[j] + c3p * a0[i] * a0[j] + cs * a2[i] * a2[j];
      ++n;
    }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) m[12 + 3 * i + j] = cm12 * a1[i] * a0[j] + c3m * a0[i] * a1[j];
}

// node a (rows) against nb <= NB nodes b0.. (columns): W0 = da0 M00 + da1 M10, W1 = da0 M10^T + da1 M11 once per point,
// then K_ab += db0 W0 + db1 W1 (18 FMAs per pair and point)
template <int NB>
__host__ __device__ __forceinline__ void warped_uu(const double* __restrict__ rs, const SmemTab& T, int nq, int a, int b0, int nb,
                                                   double (*K)[9]) {
#pragma unroll
  for (int j = 0; j < NB; ++j)
#pragma unroll
    for (int i = 0; i < 9; ++i) K[j][i] = 0.0;
  for (int q = 0; q < nq; ++q) {
    const double* r = rs + q * (REC_NQF_WARPED * 32);
    const double k00 = WP_RF(REC_K00), k01 = WP_RF(REC_K01), k10 = WP_RF(REC_K10), k11 = WP_RF(REC_K11);
    const double r0 = T.dphi2[q][a][0], r1 = T.dphi2[q][a][1];
    const double da0 = r0 * k00 + r1 * k10, da1 = r0 * k01 + r1 * k11;
    double W0[9], W1[9];
    {
      const double xx = WP_RF(REC_U_M00), xy = WP_RF(REC_U_M00 + 1), xz = WP_RF(REC_U_M00 + 2), yy = WP_RF(REC_U_M00 + 3),
                   yz = WP_RF(REC_U_M00 + 4), zz = WP_RF(REC_U_M00 + 5);
      W0[0] = da0 * xx; W0[1] = da0 * xy; W0[2] = da0 * xz; W0[3] = da0 * xy; W0[4] = da0 * yy; W0[5] = da0 * yz;
      W0[6] = da0 * xz; W0[7] = da0 * yz; W0[8] = da0 * zz;
    }
    {
      const double xx = WP_RF(REC_U_M11), xy = WP_RF(REC_U_M11 + 1), xz = WP_RF(REC_U_M11 + 2), yy = WP_RF(REC_U_M11 + 3),
                   yz = WP_RF(REC_U_M11 + 4), zz = WP_RF(REC_U_M11 + 5);
      W1[0] = da1 * xx; W1[1] = da1 * xy; W1[2] = da1 * xz; W1[3] = da1 * xy; W1[4] = da1 * yy; W1[5] = da1 * yz;
      W1[6] = da1 * xz; W1[7] = da1 * yz; W1[8] = da1 * zz;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const double m = WP_RF(REC_U_M10 + 3 * i + j);
        W0[3 * i + j] += da1 * m;
        W1[3 * j + i] += da0 * m;
      }
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      if (j < nb) {
        const double s0 = T.dphi2[q][b0 + j][0], s1 = T.dphi2[q][b0 + j][1];
        const double db0 = s0 * k00 + s1 * k10, db1 = s0 * k01 + s1 * k11;
#pragma unroll
        for (int i = 0; i < 9; ++i) K[j][i] += db0 * W0[i] + db1 * W1[i];
      }
    }
  }
}

// Displacement node a (rows) against the four rotation nodes (columns): only shear and drilling couple, and the column
// rows are N times a vector of the frame (t_rows: g0 = N E1, g1 = -N E0, dr = N E2):  K_ab = N_b (c3 E1^T - c4 E0^T + c5 E2^T)
__host__ __device__ __forceinline__ void warped_ut(const double* __restrict__ rs, const SmemTab& T, int nq, int a, double (*K)[9]) {
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int i = 0; i < 9; ++i) K[j][i] = 0.0;
  for (int q = 0; q < nq; ++q) {
    const double* r = rs + q * (REC_NQF_WARPED * 32);
    WP_LOAD_FRAME
    const double cs = WP_RF(REC_CS), cd = WP_RF(REC_CD);
    const double r0 = T.dphi2[q][a][0], r1 = T.dphi2[q][a][1];
    const URows A = u_rows(r0 * k00 + r1 * k10, r0 * k01 + r1 * k11, E0, E1, E2);
    double W[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    outer_acc(W, cs * A.g0, E1); outer_acc(W, (-cs) * A.g1, E0); outer_acc(W, cd * A.dr, E2);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const double N = T.n1[q][j];
#pragma unroll
      for (int i = 0; i < 9; ++i) K[j][i] += N * W[i];
    }
  }
}

// Rotation node a (rows) against nb <= 4 rotation nodes b0.. (columns).  The rows of a column node are linear in its value
// and its two gradient components (t_rows):  k0 = N X0 - d0 E1,  k1 = N X1 + d1 E0,  k2 = N X2 + d0 E0 - d1 E1,
// g0 = N E1, g1 = -N E0, dr = N E2  with  X0 = E0 x g_0, X1 = E1 x g_1, X2 = E0 x g_1 + E1 x g_0  (the dE2/dX terms), so
//   K_ab = N_b WN + d0_b W0 + d1_b W1,   WN = c0 X0^T + c1 X1^T + c2 X2^T + c3 E1^T - c4 E0^T + c5 E2^T,
//   W0 = c2 E0^T - c0 E1^T,   W1 = c1 E0^T - c2 E1^T     (c: the rows of a scaled by the bending / shear / drilling coefficients)
__host__ __device__ __forceinline__ void warped_tt(const double* __restrict__ rs, const SmemTab& T, int nq, int a, int b0, int nb,
                                                   double (*K)[9]) {
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int i = 0; i < 9; ++i) K[j][i] = 0.0;
  const int ta = a - 9;
  for (int q = 0; q < nq; ++q) {
    const double* r = rs + q * (REC_NQF_WARPED * 32);
    WP_LOAD_FRAME
    const V3 g0 = v3(WP_RF(REC_G0), WP_RF(REC_G0 + 1), WP_RF(REC_G0 + 2)), g1 = v3(WP_RF(REC_G0 + 3), WP_RF(REC_G0 + 4), WP_RF(REC_G0 + 5));
    const double hb = WP_RF(REC_HB), cs = WP_RF(REC_CS), cd = WP_RF(REC_CD);
    const double cb11 = hb * WP_RF(REC_CM11), cb12 = hb * WP_RF(REC_CM12), cb33 = hb * WP_RF(REC_CM33);
    const V3 X0 = cross(E0, g0), X1 = cross(E1, g1), X2 = cross(E0, g1) + cross(E1, g0);
    const double Na = T.n1[q][ta], r0 = T.dn1[q][ta][0], r1 = T.dn1[q][ta][1];
    const double d0 = r0 * k00 + r1 * k10, d1 = r0 * k01 + r1 * k11;
    const V3 ak0 = Na * X0 - d0 * E1, ak1 = Na * X1 + d1 * E0, ak2 = Na * X2 + d0 * E0 - d1 * E1;
    const V3 c0 = cb11 * ak0 + cb12 * ak1, c1 = cb12 * ak0 + cb11 * ak1, c2 = cb33 * ak2;
    const V3 c3 = (cs * Na) * E1, c4 = (-cs * Na) * E0, c5 = (cd * Na) * E2;
    double WN[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, W0[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, W1[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    outer_acc(WN, c0, X0); outer_acc(WN, c1, X1); outer_acc(WN, c2, X2);
    outer_acc(WN, c3, E1); outer_acc(WN, (-1.0) * c4, E0); outer_acc(WN, c5, E2);
    outer_acc(W0, c2, E0); outer_acc(W0, (-1.0) * c0, E1);
    outer_acc(W1, c1, E0); outer_acc(W1, (-1.0) * c2, E1);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (j < nb) {
        const int tb = b0 - 9 + j;
        const double s0 = T.dn1[q][tb][0], s1 = T.dn1[q][tb][1], Nb = T.n1[q][tb];
        const double e0 = s0 * k00 + s1 * k10, e1 = s0 * k01 + s1 * k11;
#pragma unroll
        for (int i = 0; i < 9; ++i) K[j][i] += Nb * WN[i] + e0 * W0[i] + e1 * W1[i];
      }
    }
  }
}
#undef WP_LOAD_FRAME
#undef WP_RF

__device__ __forceinline__ void wp_store_block(double* __restrict__ o, const double* K) {      // o is 16-byte aligned
  double2* o2 = reinterpret_cast<double2*>(o);
  o2[0] = make_double2(K[0], K[1]); o2[1] = make_double2(K[2], K[3]); o2[2] = make_double2(K[4], K[5]);
  o2[3] = make_double2(K[6], K[7]); o2[4] = make_double2(K[8], 0.0);
}

// quadrature record of (cell, q) in the group layout [q][field][lane]: role 0 the fields of k_cell_records' warped branch
// (frames, dE2/dX terms, coefficients), role 1 the "U" record (k and the three M matrices)
__device__ __noinline__ void warped_record(const double* __restrict__ xyz, const int32_t* __restrict__ cv,
                                           const double* __restrict__ uhat, const double* __restrict__ h,
                                           const double* __restrict__ E, const double* __restrict__ nu,
                                           const double* __restrict__ hE, const Tables* __restrict__ tab, int ewm, int role,
                                           int64_t cell, int q, double* __restrict__ o) {
  V3 xv[4], uh[4];
  int32_t vid[4];
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    vid[v] = cv[cell * 4 + v];
    xv[v] = ldv3(xyz + 3 * (int64_t)vid[v]);
    if (uhat) uh[v] = ldv3(uhat + 3 * (int64_t)vid[v]);
  }
  QpGeom g;
  qp_geometry(xv, uhat ? uh : nullptr, tab->X0[q], tab->X1[q], g);
  double hq, Eq, nuq;
  if (ewm) { hq = h[cell]; Eq = E[cell]; nuq = nu[cell]; }
  else {
    hq = Eq = nuq = 0;
#pragma unroll
    for (int v = 0; v < 4; ++v) { double N = tab->n1[q][v]; hq += N * h[vid[v]]; Eq += N * E[vid[v]]; nuq += N * nu[vid[v]]; }
  }
  const Coefs cf = clt_coefs(hq, Eq, nuq, hE[cell], g.Ju, 0);
  const double wd = tab->W[q] * g.detJ;
  const double k00 = dot(g.Kx0, g.E0), k01 = dot(g.Kx0, g.E1), k10 = dot(g.Kx1, g.E0), k11 = dot(g.Kx1, g.E1);
  o[REC_K00 * 32] = k00; o[REC_K01 * 32] = k01; o[REC_K10 * 32] = k10; o[REC_K11 * 32] = k11;
  if (role == 1) {
    double m[21];
    wp_m_tensors(g.E0, g.E1, g.E2, wd * cf.cm11, wd * cf.cm12, wd * cf.cm33, wd * cf.cs, wd * cf.cd, m);
#pragma unroll
    for (int i = 0; i < 21; ++i) o[(REC_U_M00 + i) * 32] = m[i];
    return;
  }
  o[REC_CM11 * 32] = wd * cf.cm11; o[REC_CM12 * 32] = wd * cf.cm12; o[REC_CM33 * 32] = wd * cf.cm33;
  o[REC_HB * 32] = hq * hq / 12.0;
  o[REC_CS * 32] = wd * cf.cs; o[REC_CD * 32] = wd * cf.cd;
  const V3 g0 = k00 * g.dE2_0 + k10 * g.dE2_1, g1 = k01 * g.dE2_0 + k11 * g.dE2_1;
  const double fr[15] = {g.E0.x, g.E0.y, g.E0.z, g.E1.x, g.E1.y, g.E1.z, g.E2.x, g.E2.y, g.E2.z, g0.x, g0.y, g0.z, g1.x, g1.y, g1.z};
#pragma unroll
  for (int i = 0; i < 15; ++i) o[(REC_E0 + i) * 32] = fr[i];
}

// Two CTAs per group of 32 cells: role 1 (odd blockIdx) the displacement-displacement tasks over the U records, role 0
// the displacement-rotation and rotation-rotation tasks over the frame records; each CTA builds its own records.
template <int NB>
__global__ void __launch_bounds__(WP_THREADS, 2)
k_warped_elem(const double* __restrict__ xyz, const int32_t* __restrict__ cv, const double* __restrict__ uhat,
              const double* __restrict__ h, const double* __restrict__ E, const double* __restrict__ nu,
              const double* __restrict__ hE, const Tables* __restrict__ tab, int ewm, int64_t nc, int nq, WpTasks tasks,
              double* __restrict__ elem) {
  extern __shared__ __align__(16) double wp_rec[];          // [nq][25][32]
  __shared__ SmemTab T;
  __shared__ int next_task;
  const int tid = threadIdx.x, lane = tid & 31;
  const int role = blockIdx.x & 1;
  const int64_t group = blockIdx.x >> 1;
  {
    double* d = (double*)&T;
    const double* s0 = &tab->dphi2[0][0][0];
    for (int i = tid; i < FEMO_MAX_NQ * 18; i += WP_THREADS) d[i] = s0[i];
    double* d1 = &T.n1[0][0]; const double* s1 = &tab->n1[0][0];
    for (int i = tid; i < FEMO_MAX_NQ * 4; i += WP_THREADS) d1[i] = s1[i];
    double* d2 = &T.dn1[0][0][0]; const double* s2 = &tab->dn1[0][0][0];
    for (int i = tid; i < FEMO_MAX_NQ * 8; i += WP_THREADS) d2[i] = s2[i];
    if (tid == 0) next_task = role ? 0 : tasks.n_uu;
    // the group's records, one (cell, point) per thread and pass: lane = cell, warp = point
    const int64_t cell = group * 32 + lane;
    for (int q = tid >> 5; q < nq; q += WP_THREADS / 32) {
      double* o = wp_rec + (size_t)q * (REC_NQF_WARPED * 32) + lane;
      if (cell < nc) warped_record(xyz, cv, uhat, h, E, nu, hE, tab, ewm, role, cell, q, o);
      else
        for (int f = 0; f < REC_NQF_WARPED; ++f) o[f * 32] = 0.0;
    }
  }
  __syncthreads();
  const int t_end = role ? tasks.n_uu : tasks.n;
  double* out = elem + group * (WP_PAIRS * 32 * WP_BS) + lane * WP_BS;
  for (;;) {
    int t = 0;
    if (lane == 0) t = atomicAdd(&next_task, 1);
    t = __shfl_sync(0xffffffffu, t, 0);
    if (t >= t_end) break;
    const int a = tasks.a[t], b0 = tasks.b0[t], nb = tasks.nb[t];
    if (role) {
      double K[NB][9];
      warped_uu<NB>(wp_rec + lane, T, nq, a, b0, nb, K);
#pragma unroll
      for (int j = 0; j < NB; ++j)
        if (j < nb) wp_store_block(out + (int64_t)wp_pair(a, b0 + j) * (32 * WP_BS), K[j]);
    } else {
      double K[4][9];
      if (a < 9) warped_ut(wp_rec + lane, T, nq, a, K);
      else warped_tt(wp_rec + lane, T, nq, a, b0, nb, K);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < nb) wp_store_block(out + (int64_t)wp_pair(a, b0 + j) * (32 * WP_BS), K[j]);
    }
  }
}

// One thread per BSR block: sums the contributions (block a <= b of the element matrix, or its transpose) in list order
// with five 16-byte loads each; the CTA writes its 256 blocks as one contiguous run.  Every block of the matrix is written
// exactly once.  (Summing only the upper triangle and writing the mirror blocks 72 bytes at a time halves the scratch
// reads but measured slower: 4.5 against 3.6 ms at config-3 size, profiles/r02s_warped_bench.json.)
// The runs of 256 blocks are taken in the order of the first cell contributing to them (`order`, built once): block (A, B)
// and its mirror (B, A) read the same element-matrix block, rows A and B are far apart in the matrix (vertex, edge, centre
// and rotation nodes are numbered one after the other) but close in this order, so the second read hits L2.
__global__ void __launch_bounds__(256)
k_warped_gather(const int32_t* __restrict__ centry, const int32_t* __restrict__ cptr, const int32_t* __restrict__ order,
                const double* __restrict__ elem, int64_t ndest, double* __restrict__ vals) {
  __shared__ double sK[256 * 9];
  const int64_t run = order ? order[blockIdx.x] : blockIdx.x;
  const int64_t d = run * 256 + threadIdx.x;
  double K[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  if (d < ndest) {
    const int32_t i1 = cptr[d + 1];
    for (int32_t i = cptr[d]; i < i1; ++i) {            // fixed (cell-ascending) summation order
      const int32_t e = centry[i];
      const int32_t cell = e / 169; const int pair = e - cell * 169;
      const int a = pair / 13, b = pair - a * 13;
      const bool tr = a > b;
      const int p = tr ? wp_pair(b, a) : wp_pair(a, b);
      const double2* s2 = reinterpret_cast<const double2*>(elem + (((int64_t)(cell >> 5) * WP_PAIRS + p) * 32 + (cell & 31)) * WP_BS);
      const double2 v0 = s2[0], v1 = s2[1], v2 = s2[2], v3 = s2[3], v4 = s2[4];
      if (!tr) {
        K[0] += v0.x; K[1] += v0.y; K[2] += v1.x; K[3] += v1.y; K[4] += v2.x; K[5] += v2.y; K[6] += v3.x; K[7] += v3.y; K[8] += v4.x;
      } else {
        K[0] += v0.x; K[3] += v0.y; K[6] += v1.x; K[1] += v1.y; K[4] += v2.x; K[7] += v2.y; K[2] += v3.x; K[5] += v3.y; K[8] += v4.x;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 9; ++k) sK[threadIdx.x * 9 + k] = K[k];
  __syncthreads();
  const int64_t base = run * (256 * 9);
  const int64_t n = ndest * 9 - base;
  for (int j = threadIdx.x; j < 256 * 9 && j < n; j += 256) vals[base + j] = sK[j];
}

__global__ void k_warped_run_keys(const int32_t* __restrict__ centry, const int32_t* __restrict__ cptr, int64_t ndest,
                                  int64_t nruns, int32_t* __restrict__ key) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= nruns) return;
  int32_t cell = 0x7fffffff;                      // first cell contributing to the run's first non-empty block
  for (int64_t d = r * 256; d < ndest && d < (r + 1) * 256; ++d)
    if (cptr[d + 1] > cptr[d]) { cell = centry[cptr[d]] / 169; break; }
  key[r] = cell;
}

// tasks [0, n_uu): displacement node a against up to nb_max displacement nodes b >= a, largest first (the CTA's warps draw
// tasks from a counter: longest-processing-time order keeps the tail short); then the rotation-rotation tasks (at most
// nbt columns each) and one task per displacement node against the four rotation nodes
static WpTasks wp_make_tasks(int nb_max, int nbt = 2) {
  WpTasks t; t.n = 0;
  auto add = [&](int a, int b, int nb) { t.a[t.n] = (unsigned char)a; t.b0[t.n] = (unsigned char)b; t.nb[t.n] = (unsigned char)nb; ++t.n; };
  for (int size = nb_max; size >= 1; --size)
    for (int a = 0; a < 9; ++a)
      for (int b = a; b < 9; b += nb_max)
        if (((9 - b) < nb_max ? (9 - b) : nb_max) == size) add(a, b, size);
  t.n_uu = t.n;
  for (int size = nbt; size >= 1; --size)
    for (int a = 9; a < 13; ++a)
      for (int b = a; b < 13; b += nbt)
        if (((13 - b) < nbt ? (13 - b) : nbt) == size) add(a, b, size);
  for (int a = 0; a < 9; ++a) add(a, 9, 4);
  return t;
}

// ------------------------------------------------------------------------------------------
// Chunked symmetric assembly for planar cells.
//   one CTA per chunk of C consecutive cells; destinations with row <= col owned by the chunk
//   stage : expand the quadrature records of own + halo cells into shared memory
//           (physical shape-function gradients in the cell frame + weighted CLT coefficients)
//   round : 256 contributions -> 3x3 blocks (type-uniform warps) -> shared exchange buffer ->
//           segment heads add their followers in list order -> 9 lanes per block write the block
//           and its transpose (72 contiguous bytes each)
#define ASM_THREADS 256
#define TBL_Q 36                      /* doubles per quadrature point in the staged table */
#define TW_CM11 26
#define TW_C3P 28
#define TW_CS 30
#define TW_CB11 32
#define SK_STRIDE (ASM_THREADS + 1)

struct AsmRef {                        // reference tables staged once per CTA
  double dphi2[FEMO_MAX_NQ][9][2];
  double dn1[FEMO_MAX_NQ][4][2];
  double n1[FEMO_MAX_NQ][4];
};

template <int NQ>
__device__ __forceinline__ void asm_block(const double* __restrict__ tb, const double* __restrict__ fr,
                                          const AsmRef& R, int nq_rt, int a, int b, int type, double* K) {
  const int nq = NQ > 0 ? NQ : nq_rt;
  const V3 E0 = ldv3(fr), E1 = ldv3(fr + 3), E2 = ldv3(fr + 6);
#pragma unroll
  for (int i = 0; i < 9; ++i) K[i] = 0.0;
  if (type == 0) {
    double m00 = 0, m01 = 0, m10 = 0, m11 = 0, s = 0;
    const double* pa = tb + 2 * a; const double* pb = tb + 2 * b;
#pragma unroll
    for (int q = 0; q < nq; ++q) {
      const double2 da = *reinterpret_cast<const double2*>(pa + q * TBL_Q);
      const double2 db = *reinterpret_cast<const double2*>(pb + q * TBL_Q);
      const double2 w0 = *reinterpret_cast<const double2*>(tb + q * TBL_Q + TW_CM11);   // cm11, cm12
      const double2 w1 = *reinterpret_cast<const double2*>(tb + q * TBL_Q + TW_C3P);    // cm33+cd/4, cm33-cd/4
      const double cs = tb[q * TBL_Q + TW_CS];
      const double P00 = da.x * db.x, P01 = da.x * db.y, P10 = da.y * db.x, P11 = da.y * db.y;
      m00 += w0.x * P00 + w1.x * P11;
      m01 += w0.y * P01 + w1.y * P10;
      m10 += w0.y * P10 + w1.y * P01;
      m11 += w0.x * P11 + w1.x * P00;
      s += cs * (P00 + P11);
    }
    outer_acc(K, m00 * E0 + m10 * E1, E0);
    outer_acc(K, m01 * E0 + m11 * E1, E1);
    outer_acc(K, s * E2, E2);
  } else if (type == 1) {
    const int ta = a - 9, tbn = b - 9;
    double t00 = 0, t01 = 0, t10 = 0, t11 = 0, t22 = 0;
    const double* pa = tb + 18 + 2 * ta; const double* pb = tb + 18 + 2 * tbn;
#pragma unroll
    for (int q = 0; q < nq; ++q) {
      const double2 da = *reinterpret_cast<const double2*>(pa + q * TBL_Q);
      const double2 db = *reinterpret_cast<const double2*>(pb + q * TBL_Q);
      const double2 wsd = *reinterpret_cast<const double2*>(tb + q * TBL_Q + TW_CS);     // cs, cd
      const double2 wb = *reinterpret_cast<const double2*>(tb + q * TBL_Q + TW_CB11);    // cb11, cb12
      const double cb33 = tb[q * TBL_Q + TW_CB11 + 2];
      const double NN = R.n1[q][ta] * R.n1[q][tbn];
      const double P00 = da.x * db.x, P01 = da.x * db.y, P10 = da.y * db.x, P11 = da.y * db.y;
      const double csn = wsd.x * NN;
      t00 += wb.x * P11 + cb33 * P00 + csn;
      t11 += wb.x * P00 + cb33 * P11 + csn;
      t01 -= wb.y * P10 + cb33 * P01;
      t10 -= wb.y * P01 + cb33 * P10;
      t22 += wsd.y * NN;
    }
    outer_acc(K, t00 * E0 + t10 * E1, E0);
    outer_acc(K, t01 * E0 + t11 * E1, E1);
    outer_acc(K, t22 * E2, E2);
  } else {
    // row = displacement node a, column = rotation node b (row < col always holds for this type)
    const int tbn = b - 9;
    double A0 = 0, A1 = 0, D0 = 0, D1 = 0;
    const double* pa = tb + 2 * a;
#pragma unroll
    for (int q = 0; q < nq; ++q) {
      const double2 da = *reinterpret_cast<const double2*>(pa + q * TBL_Q);
      const double2 wsd = *reinterpret_cast<const double2*>(tb + q * TBL_Q + TW_CS);
      const double N = R.n1[q][tbn];
      const double cs = wsd.x * N, cd = wsd.y * N;
      A0 += cs * da.x; A1 += cs * da.y; D0 += cd * da.x; D1 += cd * da.y;
    }
    outer_acc(K, E2, A0 * E1 - A1 * E0);
    outer_acc(K, 0.5 * (D1 * E0 - D0 * E1), E2);
  }
}

template <int NQ>
__global__ void __launch_bounds__(ASM_THREADS)
k_asm_planar(const uint32_t* __restrict__ it_pack, const int32_t* __restrict__ it_dest,
             const int32_t* __restrict__ chunk_ptr, const int32_t* __restrict__ halo_ptr,
             const int32_t* __restrict__ halo_cells, const int32_t* __restrict__ tpos,
             const double* __restrict__ rec, int rec_stride, const Tables* __restrict__ tab, int nq_rt, int C,
             int64_t nc, int tbl_stride, double* __restrict__ vals) {
  extern __shared__ __align__(16) double smem[];
  const int nq = NQ > 0 ? NQ : nq_rt;
  AsmRef& R = *reinterpret_cast<AsmRef*>(smem);
  double* sK = smem + sizeof(AsmRef) / sizeof(double);                 // 9 x SK_STRIDE
  int* sHeadTid = reinterpret_cast<int*>(sK + 9 * SK_STRIDE + 1);      // ASM_THREADS (keeps 16-byte alignment)
  int* sHeadPos = sHeadTid + ASM_THREADS;
  int* sHeadTp = sHeadPos + ASM_THREADS;
  int* sWarpCnt = sHeadTp + ASM_THREADS;                               // 8 (+ next_base at [8])
  double* sFrame = reinterpret_cast<double*>(sWarpCnt + 16);           // nslots x 9   (max 255 slots)
  const int64_t chunk = blockIdx.x;
  const int h0 = halo_ptr[chunk], nhalo = halo_ptr[chunk + 1] - h0;
  const int nslots = C + nhalo;
  double* sTbl = sFrame + ((nslots * 9 + 1) & ~1);                     // nslots x tbl_stride, 16-byte aligned
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  {
    double* d = &R.dphi2[0][0][0]; const double* s0 = &tab->dphi2[0][0][0];
    for (int i = tid; i < FEMO_MAX_NQ * 18; i += ASM_THREADS) d[i] = s0[i];
    double* d1 = &R.dn1[0][0][0]; const double* s1 = &tab->dn1[0][0][0];
    for (int i = tid; i < FEMO_MAX_NQ * 8; i += ASM_THREADS) d1[i] = s1[i];
    double* d2 = &R.n1[0][0]; const double* s2 = &tab->n1[0][0];
    for (int i = tid; i < FEMO_MAX_NQ * 4; i += ASM_THREADS) d2[i] = s2[i];
  }
  __syncthreads();
  // ---- stage own + halo cells ------------------------------------------------------------
  for (int s = wid; s < nslots; s += ASM_THREADS / 32) {
    int64_t cell = s < C ? chunk * C + s : (int64_t)halo_cells[h0 + s - C];
    if (cell >= nc) continue;                                         // tail chunk
    const double* rc = rec + cell * rec_stride;
    if (lane < 9) sFrame[s * 9 + lane] = rc[lane];
    double* T = sTbl + (int64_t)s * tbl_stride;
    for (int idx = lane; idx < 2 * nq; idx += 32) {
      const int q = idx % nq, half = idx / nq;
      const double* r = rc + REC_HEADER + q * REC_NQF_PLANAR;
      const double k00 = r[REC_K00], k01 = r[REC_K01], k10 = r[REC_K10], k11 = r[REC_K11];
      double* o = T + q * TBL_Q;
      if (half == 0) {
#pragma unroll
        for (int a = 0; a < 9; ++a) {
          const double g0 = R.dphi2[q][a][0], g1 = R.dphi2[q][a][1];
          o[2 * a] = g0 * k00 + g1 * k10; o[2 * a + 1] = g0 * k01 + g1 * k11;
        }
      } else {
#pragma unroll
        for (int a = 0; a < 4; ++a) {
          const double g0 = R.dn1[q][a][0], g1 = R.dn1[q][a][1];
          o[18 + 2 * a] = g0 * k00 + g1 * k10; o[18 + 2 * a + 1] = g0 * k01 + g1 * k11;
        }
        const double cm11 = r[REC_CM11], cm12 = r[REC_CM12], cm33 = r[REC_CM33], hb = r[REC_HB], cs = r[REC_CS],
                     cd = r[REC_CD];
        o[TW_CM11] = cm11; o[TW_CM11 + 1] = cm12;
        o[TW_C3P] = cm33 + 0.25 * cd; o[TW_C3P + 1] = cm33 - 0.25 * cd;
        o[TW_CS] = cs; o[TW_CS + 1] = cd;
        o[TW_CB11] = hb * cm11; o[TW_CB11 + 1] = hb * cm12; o[TW_CB11 + 2] = hb * cm33;
      }
    }
  }
  __syncthreads();
  // ---- rounds of ASM_THREADS contributions ---------------------------------------------------
  const int end = chunk_ptr[chunk + 1];
  int base = chunk_ptr[chunk];
  while (base < end) {
    const int j = base + tid;
    double K[9];
    uint32_t pk = 0;
    if (j < end) {
      pk = it_pack[j];
      const int slot = pk & 0xff, pair = (pk >> 8) & 0xff;
      const int a = pair / 13, b = pair - a * 13;
      const int type = (a < 9 && b < 9) ? 0 : ((a >= 9) ? 1 : 2);
      asm_block<NQ>(sTbl + (int64_t)slot * tbl_stride, sFrame + slot * 9, R, nq, a, b, type, K);
#pragma unroll
      for (int k = 0; k < 9; ++k) sK[k * SK_STRIDE + tid] = K[k];
    }
    if (tid == 0) sWarpCnt[8] = base + ASM_THREADS;
    __syncthreads();
    const bool head = (pk >> 16) & 1u;
    const int len = (int)(pk >> 17);
    bool emit = false;
    if (head) {
      if (tid + len <= ASM_THREADS) {
        for (int f = 1; f < len; ++f) {                               // fixed (cell-ascending) order
#pragma unroll
          for (int k = 0; k < 9; ++k) K[k] += sK[k * SK_STRIDE + tid + f];
        }
        emit = true;
      } else {
        sWarpCnt[8] = base + tid;                                     // segment continues past this round: redo it next round
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, emit);
    if (lane == 0) sWarpCnt[wid] = __popc(bal);
    __syncthreads();
    if (emit) {
      int off = __popc(bal & ((1u << lane) - 1u));
      for (int w = 0; w < wid; ++w) off += sWarpCnt[w];
      const int d = it_dest[j];
#pragma unroll
      for (int k = 0; k < 9; ++k) sK[k * SK_STRIDE + tid] = K[k];
      sHeadTid[off] = tid; sHeadPos[off] = d; sHeadTp[off] = tpos[d];
    }
    int nheads = 0;
#pragma unroll
    for (int w = 0; w < ASM_THREADS / 32; ++w) nheads += sWarpCnt[w];
    const int next_base = sWarpCnt[8];
    __syncthreads();
    for (int idx = tid; idx < nheads * 9; idx += ASM_THREADS) {
      const int h = idx / 9, k = idx - 9 * h;
      const double v = sK[k * SK_STRIDE + sHeadTid[h]];
      vals[(int64_t)sHeadPos[h] * 9 + k] = v;
      const int tp = sHeadTp[h];
      if (tp >= 0) vals[(int64_t)tp * 9 + (k % 3) * 3 + k / 3] = v;
    }
    __syncthreads();
    base = next_base;
  }
}

size_t asm_smem_bytes(int max_slots, int tbl_stride) {
  size_t b = sizeof(AsmRef) + sizeof(double) * (9 * SK_STRIDE + 1) + sizeof(int) * (3 * ASM_THREADS + 16);
  b += sizeof(double) * (size_t)((max_slots * 9 + 1) & ~1);
  b += sizeof(double) * (size_t)max_slots * tbl_stride;
  return b;
}

template <int NQ>
int launch_asm_planar(femo_ctx* c) {
  const int tbl_stride = TBL_Q * c->nq + 2;
  const size_t smem = asm_smem_bytes(c->plan_max_slots, tbl_stride);
  static size_t configured[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (configured[c->device & 7] < smem) {
    CUDA_TRY(cudaFuncSetAttribute(k_asm_planar<NQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[c->device & 7] = smem;
  }
  k_asm_planar<NQ><<<(unsigned)c->nchunks, ASM_THREADS, smem, c->stream>>>(
      c->it_pack, c->it_dest, c->chunk_ptr, c->halo_ptr, c->halo_cells, c->tpos, c->rec, c->rec_stride, c->tab, c->nq,
      c->plan_C, c->nc, tbl_stride, c->vals);
  return check_launch(c, "k_asm_planar");
}

// ------------------------------------------------------------------------------------------
// penalty BC: 5x5 trace matrix per (edge, cell), nodes (q2 va, q2 mid, q2 vb, q1 va, q1 vb)
__global__ void k_penalty_me(const double* __restrict__ xyz, const int32_t* __restrict__ cv,
                             const double* __restrict__ uhat, const double* __restrict__ hE,
                             const int32_t* __restrict__ pen_cell, int64_t n_pen, double* __restrict__ me) {
  int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= n_pen) return;
  int32_t pc = pen_cell[p];
  int64_t cell = pc >> 2; int le = pc & 3;
  const int LE[4][2] = {{0, 1}, {0, 2}, {1, 3}, {2, 3}};
  V3 xv[4], uh[4];
  for (int v = 0; v < 4; ++v) {
    int32_t id = cv[cell * 4 + v];
    xv[v] = ldv3(xyz + 3 * (int64_t)id);
    if (uhat) uh[v] = ldv3(uhat + 3 * (int64_t)id);
  }
  V3 t = xv[LE[le][1]] - xv[LE[le][0]];
  double L = sqrt(dot(t, t));
  t = (1.0 / L) * t;
  const double gx[3] = {0.5 - 0.5 * 0.7745966692414834, 0.5, 0.5 + 0.5 * 0.7745966692414834};
  const double gw[3] = {5.0 / 18.0, 8.0 / 18.0, 5.0 / 18.0};   // degree-4 facet rule (utils_dolfinx.py:556)
  double M[25];
  for (int i = 0; i < 25; ++i) M[i] = 0;
  for (int k = 0; k < 3; ++k) {
    double s = gx[k];
    double X0 = (le == 0 || le == 3) ? s : (le == 2 ? 1.0 : 0.0);
    double X1 = (le == 1 || le == 2) ? s : (le == 3 ? 1.0 : 0.0);
    QpGeom g;
    qp_geometry(xv, uhat ? uh : nullptr, X0, X1, g);
    V3 Nf = cross(t, g.E2);
    // Nanson: J F^-T N
    const double* Fi = g.Finv;
    V3 nn = v3(Fi[0] * Nf.x + Fi[3] * Nf.y + Fi[6] * Nf.z, Fi[1] * Nf.x + Fi[4] * Nf.y + Fi[7] * Nf.z,
               Fi[2] * Nf.x + Fi[5] * Nf.y + Fi[8] * Nf.z);
    double sc = g.Ju * sqrt(dot(nn, nn));
    double tr[5] = {2.0 * (s - 0.5) * (s - 1.0), 4.0 * s * (1.0 - s), 2.0 * s * (s - 0.5), 1.0 - s, s};
    double wgt = gw[k] * L * sc * FEMO_BETA / hE[cell];
    for (int i = 0; i < 5; ++i)
      for (int j = 0; j < 5; ++j)
        if ((i < 3) == (j < 3)) M[i * 5 + j] += wgt * tr[i] * tr[j];
  }
  for (int i = 0; i < 25; ++i) me[p * 25 + i] = M[i];
}

__global__ void k_penalty_add(const int32_t* __restrict__ dest_pos, const int32_t* __restrict__ dest_ptr,
                              const int32_t* __restrict__ src, const double* __restrict__ me, int64_t n_dest,
                              double* __restrict__ pen_val, double* __restrict__ vals) {
  int64_t d = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (d >= n_dest) return;
  double v = 0;
  for (int32_t k = dest_ptr[d]; k < dest_ptr[d + 1]; ++k) v += me[src[k]];
  pen_val[d] = v;
  double* o = vals + (int64_t)dest_pos[d] * 9;
  o[0] += v; o[4] += v; o[8] += v;
}

// ------------------------------------------------------------------------------------------
// load vector: F_I = sum_{cells containing Q2 node I} sum_q w detJ J phi2_a f(q)   (model:320)
// Two passes, both deterministic: one thread per cell integrates its 9 nodal contributions (27 doubles into
// `cellbuf`, stride 39), then one thread per Q2 node sums its <= 4 (more at junctions) cells in ascending order.
__global__ void __launch_bounds__(128)
k_load_cells(const double* __restrict__ xyz, const int32_t* __restrict__ cv, const double* __restrict__ uhat,
             const double* __restrict__ f, const Tables* __restrict__ tab, int nq, int ewp, int64_t nc,
             double* __restrict__ cellbuf) {
  const int64_t cell = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (cell >= nc) return;
  V3 xv[4], uh[4], fv[4];
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    const int32_t id = cv[cell * 4 + v];
    xv[v] = ldv3(xyz + 3 * (int64_t)id);
    if (uhat) uh[v] = ldv3(uhat + 3 * (int64_t)id);
    if (!ewp) fv[v] = ldv3(f + 3 * (int64_t)id);
  }
  const V3 fc = ewp ? ldv3(f + 3 * cell) : v3(0, 0, 0);
  V3 acc[9];
#pragma unroll
  for (int a = 0; a < 9; ++a) acc[a] = v3(0, 0, 0);
  for (int q = 0; q < nq; ++q) {
    QpGeom g;
    qp_geometry(xv, uhat ? uh : nullptr, tab->X0[q], tab->X1[q], g);
    V3 fq = fc;
    if (!ewp) fq = tab->n1[q][0] * fv[0] + tab->n1[q][1] * fv[1] + tab->n1[q][2] * fv[2] + tab->n1[q][3] * fv[3];
    const V3 sf = (tab->W[q] * g.detJ * g.Ju) * fq;
#pragma unroll
    for (int a = 0; a < 9; ++a) acc[a] = acc[a] + tab->phi2[q][a] * sf;
  }
  double* out = cellbuf + cell * 39;
#pragma unroll
  for (int a = 0; a < 9; ++a) { out[3 * a] = acc[a].x; out[3 * a + 1] = acc[a].y; out[3 * a + 2] = acc[a].z; }
}

__global__ void k_load_gather(const int32_t* __restrict__ nc_ptr, const int32_t* __restrict__ nc_adj,
                              const double* __restrict__ cellbuf, int64_t nq2, double* __restrict__ F) {
  const int64_t I = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (I >= nq2) return;
  double a0 = 0, a1 = 0, a2 = 0;
  for (int32_t k = nc_ptr[I]; k < nc_ptr[I + 1]; ++k) {
    const int32_t e = nc_adj[k];
    const int64_t cell = e / 13;
    const int a = e - (int32_t)cell * 13;
    const double* s = cellbuf + cell * 39 + 3 * a;
    a0 += s[0]; a1 += s[1]; a2 += s[2];
  }
  F[3 * I] = a0; F[3 * I + 1] = a1; F[3 * I + 2] = a2;
}

}  // namespace

// warped meshes: records + element matrices in one kernel (events: ev1 between the two kernels, so that last_timings
// reports element matrices / gather)
static int assemble_jacobian_warped(femo_ctx* c, const double* uh) {
  const size_t smem = sizeof(double) * c->nq * REC_NQF_WARPED * 32;
  const unsigned groups = (unsigned)cdiv(c->nc, 32);
  const WpTasks tasks = wp_make_tasks(c->warped_nb, c->warped_nbt);
#define WP_LAUNCH(NB)                                                                                            \
  {                                                                                                             \
    CUDA_TRY(cudaFuncSetAttribute(k_warped_elem<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));  \
    k_warped_elem<NB><<<2 * groups, WP_THREADS, smem, c->stream>>>(c->xyz, c->cv, uh, c->h, c->E, c->nu, c->hE, c->tab, c->ewm, \
                                                               c->nc, c->nq, tasks, c->elem);                   \
  }
  if (c->warped_nb == 1) WP_LAUNCH(1) else if (c->warped_nb == 2) WP_LAUNCH(2) else if (c->warped_nb == 4) WP_LAUNCH(4) else WP_LAUNCH(3)
#undef WP_LAUNCH
  TRY(check_launch(c, "k_warped_elem"));
  cudaEventRecord(c->ev1, c->stream);
  k_warped_gather<<<(unsigned)cdiv(c->nnzb, 256), 256, 0, c->stream>>>(c->centry, c->cptr, c->wp_order, c->elem, c->nnzb, c->vals);
  TRY(check_launch(c, "k_warped_gather"));
  return 0;
}

// scratch of the warped path: 91 padded blocks per cell, and the order in which k_warped_gather takes its runs of 256
// blocks (FEMO_WARPED_ORDER=0: matrix order).  Called once from femo_shell_setup.
int warped_setup(femo_ctx* c) {
  TRY(dev_alloc(&c->elem, (size_t)cdiv(c->nc, 32) * WP_PAIRS * 32 * WP_BS));
  dev_free(&c->wp_order);
  const char* eo = getenv("FEMO_WARPED_ORDER");
  if (eo && atoi(eo) == 0) return 0;
  const int64_t nruns = cdiv(c->nnzb, 256);
  TRY(dev_alloc(&c->wp_order, (size_t)nruns));
  k_warped_run_keys<<<(unsigned)cdiv(nruns, 128), 128, 0, c->stream>>>(c->centry, c->cptr, c->nnzb, nruns, c->wp_order);
  TRY(check_launch(c, "k_warped_run_keys"));
  std::vector<int32_t> key((size_t)nruns), ord((size_t)nruns);
  CUDA_TRY(cudaMemcpyAsync(key.data(), c->wp_order, sizeof(int32_t) * nruns, cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  for (int64_t i = 0; i < nruns; ++i) ord[i] = (int32_t)i;
  std::stable_sort(ord.begin(), ord.end(), [&](int32_t x, int32_t y) { return key[x] < key[y]; });
  CUDA_TRY(cudaMemcpyAsync(c->wp_order, ord.data(), sizeof(int32_t) * nruns, cudaMemcpyHostToDevice, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

// older kernels: FEMO_TILE=0, FEMO_WARPED=0, or a plan that does not fit shared memory
static int assemble_jacobian_legacy(femo_ctx* c, const double* uh) {
  k_cell_records<<<(unsigned)cdiv(c->nc * c->nq, 128), 128, 0, c->stream>>>(
      c->xyz, c->cv, uh, c->h, c->E, c->nu, c->hE, c->tab, c->nc, c->nq, c->ewm, c->rec_nqf, c->rec_stride, c->rec);
  TRY(check_launch(c, "k_cell_records"));
  cudaEventRecord(c->ev1, c->stream);
  unsigned grid = (unsigned)cdiv(c->ncontrib, GATHER_THREADS);
  const int tbl_stride = TBL_Q * c->nq + 2;
  if (c->all_planar && c->plan_ok && asm_smem_bytes(c->plan_max_slots, tbl_stride) <= 220 * 1024) {
    switch (c->nq) {
      case 16: TRY(launch_asm_planar<16>(c)); break;
      case 9: TRY(launch_asm_planar<9>(c)); break;
      case 4: TRY(launch_asm_planar<4>(c)); break;
      default: TRY(launch_asm_planar<0>(c)); break;
    }
  } else if (c->all_planar)
    k_gather<true><<<grid, GATHER_THREADS, 0, c->stream>>>(c->centry, c->cdest, c->cptr, c->rec, c->rec_stride, c->tab,
                                                           c->nq, c->ncontrib, c->vals);
  else
    k_gather<false><<<grid, GATHER_THREADS, 0, c->stream>>>(c->centry, c->cdest, c->cptr, c->rec, c->rec_stride, c->tab,
                                                            c->nq, c->ncontrib, c->vals);
  TRY(check_launch(c, "k_gather"));
  return 0;
}

int assemble_stiffness(femo_ctx* c) {
  if (!c->setup_done) return femo_fail("femo_assemble_jacobian: call femo_shell_setup first");
  const double* uh = c->uhat_zero ? nullptr : c->uhat;
  cudaEventRecord(c->ev0, c->stream);
  if (c->tile_ok) {
    TRY(assemble_jacobian_tiles(c, uh));
  } else if (c->warped_ok) {
    TRY(assemble_jacobian_warped(c, uh));
  } else {
    TRY(assemble_jacobian_legacy(c, uh));
  }
  cudaEvent_t ev2; cudaEventCreate(&ev2);
  cudaEventRecord(ev2, c->stream);
  c->launches += 2;
  if (c->n_pen > 0) {
    k_penalty_me<<<(unsigned)cdiv(c->n_pen, 128), 128, 0, c->stream>>>(c->xyz, c->cv, uh, c->hE, c->pen_cell, c->n_pen, c->pen_me);
    TRY(check_launch(c, "k_penalty_me"));
    k_penalty_add<<<(unsigned)cdiv(c->n_pen_dest, 128), 128, 0, c->stream>>>(c->pen_dest_pos, c->pen_dest_ptr, c->pen_src,
                                                                          c->pen_me, c->n_pen_dest, c->pen_val, c->vals);
    TRY(check_launch(c, "k_penalty_add"));
    c->launches += 2;
  }
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev0, c->ev1); c->t_rec = ms;
  cudaEventElapsedTime(&ms, c->ev1, ev2); c->t_asm = ms;
  cudaEventDestroy(ev2);
  c->jac_valid = true;
  c->prec_valid = false;
  return 0;
}

int assemble_load(femo_ctx* c, double* F) {
  if (c->kind == 1) return hex_load(c, F);
  const double* uh = c->uhat_zero ? nullptr : c->uhat;
  CUDA_TRY(cudaMemsetAsync(F, 0, sizeof(double) * c->ndof, c->stream));
  k_load_cells<<<(unsigned)cdiv(c->nc, 128), 128, 0, c->stream>>>(c->xyz, c->cv, uh, c->f, c->tab, c->nq, c->ewp, c->nc, c->cellbuf);
  k_load_gather<<<(unsigned)cdiv(c->nq2, 256), 256, 0, c->stream>>>(c->nc_ptr, c->nc_adj, c->cellbuf, c->nq2, F);
  TRY(check_launch(c, "k_load"));
  c->launches += 2;
  return 0;
}
