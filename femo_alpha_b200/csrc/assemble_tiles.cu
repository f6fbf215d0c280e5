// Tiled assembly of dR/du for planar-cell meshes (third generation).
//
// Replaces assembleMatrix(dR_du) (femo_alpha/fea/utils_dolfinx.py:200-221) for the forms of
// rm_shell/linear_shell_fenicsx/linear_shell_model.py:220-333, like assemble.cu, but organised so that the
// fp64 pipe and not shared memory is the limiter:
//
//   phase 1  one CTA per chunk of 32 cells (processing order of tile_plan.hpp).  lane = cell, warp = tile of
//            node pairs: 6 (3x3) Q2-Q2 tiles, 1 Q1-Q1 tile, 3 (3x4) Q2-Q1 tiles.  Every thread keeps its 45-50
//            local-frame moment accumulators in registers over the quadrature loop; reference gradients are
//            warp-uniform shared-memory broadcasts, the per-cell inverse Jacobians / CLT coefficients are
//            coalesced loads of the [chunk][q][field][lane] record.  Contributions of cells of OTHER chunks to
//            blocks this chunk owns ("halo items", a few % of the work) are computed one pair per thread.
//   phase 2  one thread per owned destination block (row <= col): gathers the moments of its <= 4 (more on
//            non-manifold junctions) contributions from shared memory in fixed order, rotates them with the
//            cell frames into the global 3x3 block, and the CTA writes block + transposed block, 72 contiguous
//            bytes each.  Every block of the matrix is written exactly once: no atomics, bitwise reproducible.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "femo_internal.cuh"
#include "tile_plan.hpp"

namespace {

using femo_tiles::C;
using femo_tiles::MOM_CELL;
using femo_tiles::MOM_HALO;
using femo_tiles::MOM_HALO_BASE;
using femo_tiles::MOM_TT;
using femo_tiles::MOM_UT;

#define TL_THREADS 192
// 10 fields per (cell, quadrature point).  Round 1 also stored cm33 +- cd/4; they are one FMA away from cm33 and cd, and
// dropping them shrinks the record stream by a sixth and makes room in shared memory for double-buffered halo records.
#define TL_NF 10
#define TF_K00 0
#define TF_K01 1
#define TF_K10 2
#define TF_K11 3
#define TF_CM11 4
#define TF_CM12 5
#define TF_CM33 6
#define TF_CD 7
#define TF_CS 8
#define TF_HB 9     /* h^2/12 */

struct TileTab {                       // reference tables staged once per CTA (6000 B)
  double dphi2[FEMO_MAX_NQ][9][2];
  double dn1[FEMO_MAX_NQ][4][2];
  double n1[FEMO_MAX_NQ][4];
};

// one thread per processing slot (cell): vertex data and the cell frame are set up once, then the quadrature loop
// writes the [chunk][q][field][lane] records (16 consecutive lanes = 128 contiguous bytes per field)
__global__ void k_cell_records_tiles(const double* __restrict__ xyz, const int32_t* __restrict__ cv,
                                     const double* __restrict__ uhat, const double* __restrict__ h,
                                     const double* __restrict__ E, const double* __restrict__ nu,
                                     const double* __restrict__ hE, const Tables* __restrict__ tab,
                                     const int32_t* __restrict__ cell_order, int64_t nchunks, int nq, int ewm,
                                     double* __restrict__ rec, double* __restrict__ frame) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= nchunks * C) return;
  const int lane = (int)(p % C);
  const int64_t chunk = p / C;
  const int32_t cell = cell_order[p];
  double* r0 = rec + chunk * ((int64_t)nq * TL_NF * C) + lane;
  double* fr = frame + p * 9;
  if (cell < 0) {                                       // padding slot of the last chunk
    for (int q = 0; q < nq; ++q)
#pragma unroll
      for (int f = 0; f < TL_NF; ++f) r0[(q * TL_NF + f) * C] = 0.0;
    for (int k = 0; k < 9; ++k) fr[k] = 0.0;
    return;
  }
  V3 xv[4], uh[4];
  int32_t vid[4];
  double hv[4], Ev[4], nv[4];
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    vid[v] = cv[(int64_t)cell * 4 + v];
    xv[v] = ldv3(xyz + 3 * (int64_t)vid[v]);
    if (uhat) uh[v] = ldv3(uhat + 3 * (int64_t)vid[v]);
    const int64_t mi = ewm ? (int64_t)cell : (int64_t)vid[v];
    hv[v] = h[mi]; Ev[v] = E[mi]; nv[v] = nu[mi];
  }
  // planar cell: one constant in-plane frame per cell, taken at the cell centre (isotropic CLT)
  QpGeom gc;
  qp_geometry(xv, nullptr, 0.5, 0.5, gc);
  fr[0] = gc.E0.x; fr[1] = gc.E0.y; fr[2] = gc.E0.z;
  fr[3] = gc.E1.x; fr[4] = gc.E1.y; fr[5] = gc.E1.z;
  fr[6] = gc.E2.x; fr[7] = gc.E2.y; fr[8] = gc.E2.z;
  const double hEc = hE[cell];
  for (int q = 0; q < nq; ++q) {
    QpGeom g;
    qp_geometry(xv, uhat ? uh : nullptr, tab->X0[q], tab->X1[q], g);
    double hq, Eq, nuq;
    if (ewm) { hq = hv[0]; Eq = Ev[0]; nuq = nv[0]; }
    else {
      hq = Eq = nuq = 0;
#pragma unroll
      for (int v = 0; v < 4; ++v) { const double N = tab->n1[q][v]; hq += N * hv[v]; Eq += N * Ev[v]; nuq += N * nv[v]; }
    }
    const Coefs cf = clt_coefs(hq, Eq, nuq, hEc, g.Ju, 0);
    const double wd = tab->W[q] * g.detJ;
    const double cm33 = wd * cf.cm33, cd = wd * cf.cd;
    double* r = r0 + q * (TL_NF * C);
    r[TF_K00 * C] = dot(g.Kx0, gc.E0); r[TF_K01 * C] = dot(g.Kx0, gc.E1);
    r[TF_K10 * C] = dot(g.Kx1, gc.E0); r[TF_K11 * C] = dot(g.Kx1, gc.E1);
    r[TF_CM11 * C] = wd * cf.cm11; r[TF_CM12 * C] = wd * cf.cm12;
    r[TF_CS * C] = wd * cf.cs; r[TF_CD * C] = cd;
    r[TF_CM33 * C] = cm33; r[TF_HB * C] = hq * hq / 12.0;
  }
}

// ---- phase 1: register-tiled local-frame moments -------------------------------------------------------------
// physical gradient of a reference shape function in the cell frame: (g0 k00 + g1 k10, g0 k01 + g1 k11)
#define PHYS(gx, gy, g) do { gx = (g).x * k00 + (g).y * k10; gy = (g).x * k01 + (g).y * k11; } while (0)

// Q2-Q2 tile (I, J): nodes 3I..3I+2 x 3J..3J+2, moments (m00, m01, m10, m11, s) per pair
__device__ __forceinline__ void tile_uu(const double* __restrict__ rc, const TileTab& R, int nq, int I, int J,
                                        double* __restrict__ mom) {
  double acc[3][3][5];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
      for (int m = 0; m < 5; ++m) acc[i][j][m] = 0.0;
#pragma unroll 1
  for (int q = 0; q < nq; ++q) {
    const double* r = rc + q * (TL_NF * C);
    const double k00 = r[TF_K00 * C], k01 = r[TF_K01 * C], k10 = r[TF_K10 * C], k11 = r[TF_K11 * C];
    const double w11 = r[TF_CM11 * C], w12 = r[TF_CM12 * C], ws = r[TF_CS * C];
    const double w33 = r[TF_CM33 * C], wd4 = r[TF_CD * C];
    const double w3p = fma(0.25, wd4, w33), w3m = fma(-0.25, wd4, w33);          // cm33 +- cd/4
    double bx[3], by[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const double2 g = *reinterpret_cast<const double2*>(&R.dphi2[q][3 * J + j][0]);
      PHYS(bx[j], by[j], g);
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double2 g = *reinterpret_cast<const double2*>(&R.dphi2[q][3 * I + i][0]);
      double ax, ay;
      PHYS(ax, ay, g);
      const double a11x = w11 * ax, a11y = w11 * ay, a3px = w3p * ax, a3py = w3p * ay;
      const double a12x = w12 * ax, a12y = w12 * ay, a3mx = w3m * ax, a3my = w3m * ay;
      const double asx = ws * ax, asy = ws * ay;
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        acc[i][j][0] = fma(a11x, bx[j], fma(a3py, by[j], acc[i][j][0]));     // cm11 ax bx + (cm33 + cd/4) ay by
        acc[i][j][1] = fma(a12x, by[j], fma(a3my, bx[j], acc[i][j][1]));     // cm12 ax by + (cm33 - cd/4) ay bx
        acc[i][j][2] = fma(a12y, bx[j], fma(a3mx, by[j], acc[i][j][2]));     // cm12 ay bx + (cm33 - cd/4) ax by
        acc[i][j][3] = fma(a11y, by[j], fma(a3px, bx[j], acc[i][j][3]));     // cm11 ay by + (cm33 + cd/4) ax bx
        acc[i][j][4] = fma(asx, bx[j], fma(asy, by[j], acc[i][j][4]));       // cs grad.grad   (u3-u3 shear)
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int a = 3 * I + i, b = 3 * J + j;
      if (a <= b) {
        double* o = mom + 5 * (a * 9 - a * (a - 1) / 2 + (b - a));
#pragma unroll
        for (int m = 0; m < 5; ++m) o[m] = acc[i][j][m];
      }
    }
}

// Q1-Q1 tile: 10 symmetric pairs, moments (t00, t01, t10, t11, t22)
// qb, qs: this thread visits quadrature points qb, qb+qs, ...; part: 0 = store the result, 1 = store a partial sum
// (odd points of a q-split half-warp pair), 2 = add the partner's partial sum, then store
__device__ __forceinline__ void tile_tt(const double* __restrict__ rc, const TileTab& R, int nq,
                                        double* __restrict__ mom, int qb = 0, int qs = 1, int part = 0) {
  double acc[10][5];
#pragma unroll
  for (int i = 0; i < 10; ++i)
#pragma unroll
    for (int m = 0; m < 5; ++m) acc[i][m] = 0.0;
#pragma unroll 1
  for (int q = qb; q < nq; q += qs) {
    const double* r = rc + q * (TL_NF * C);
    const double k00 = r[TF_K00 * C], k01 = r[TF_K01 * C], k10 = r[TF_K10 * C], k11 = r[TF_K11 * C];
    const double hb = r[TF_HB * C], cs = r[TF_CS * C], cd = r[TF_CD * C];
    const double cb11 = hb * r[TF_CM11 * C], cb12 = hb * r[TF_CM12 * C], cb33 = hb * r[TF_CM33 * C];
    double dx[4], dy[4], N[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const double2 g = *reinterpret_cast<const double2*>(&R.dn1[q][a][0]);
      PHYS(dx[a], dy[a], g);
      N[a] = R.n1[q][a];
    }
    int idx = 0;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const double a11x = cb11 * dx[a], a11y = cb11 * dy[a], a33x = cb33 * dx[a], a33y = cb33 * dy[a];
      const double a12x = cb12 * dx[a], a12y = cb12 * dy[a], ns = cs * N[a], nd = cd * N[a];
#pragma unroll
      for (int b = a; b < 4; ++b, ++idx) {
        acc[idx][0] = fma(a11y, dy[b], fma(a33x, dx[b], fma(ns, N[b], acc[idx][0])));   // cb11 P11 + cb33 P00 + cs NN
        acc[idx][1] = fma(-a12y, dx[b], fma(-a33x, dy[b], acc[idx][1]));               // -(cb12 P10 + cb33 P01)
        acc[idx][2] = fma(-a12x, dy[b], fma(-a33y, dx[b], acc[idx][2]));               // -(cb12 P01 + cb33 P10)
        acc[idx][3] = fma(a11x, dx[b], fma(a33y, dy[b], fma(ns, N[b], acc[idx][3])));   // cb11 P00 + cb33 P11 + cs NN
        acc[idx][4] = fma(nd, N[b], acc[idx][4]);                                       // drilling
      }
    }
  }
  if (part == 1) {
#pragma unroll
    for (int i = 0; i < 10; ++i)
#pragma unroll
      for (int m = 0; m < 5; ++m) mom[MOM_TT + 5 * i + m] = acc[i][m];
  }
  if (part != 0) __syncwarp();
  if (part != 1) {
#pragma unroll
    for (int i = 0; i < 10; ++i)
#pragma unroll
      for (int m = 0; m < 5; ++m) mom[MOM_TT + 5 * i + m] = part == 2 ? acc[i][m] + mom[MOM_TT + 5 * i + m] : acc[i][m];
  }
}

// Q2-Q1 tile I: displacement nodes 3I..3I+2 x rotation nodes 0..3, moments (A0, A1, D0, D1)
__device__ __forceinline__ void tile_ut(const double* __restrict__ rc, const TileTab& R, int nq, int I,
                                        double* __restrict__ mom, int qb = 0, int qs = 1, int part = 0) {
  double acc[3][4][4];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int m = 0; m < 4; ++m) acc[i][j][m] = 0.0;
#pragma unroll 1
  for (int q = qb; q < nq; q += qs) {
    const double* r = rc + q * (TL_NF * C);
    const double k00 = r[TF_K00 * C], k01 = r[TF_K01 * C], k10 = r[TF_K10 * C], k11 = r[TF_K11 * C];
    const double cs = r[TF_CS * C], cd = r[TF_CD * C];
    double ax[3], ay[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double2 g = *reinterpret_cast<const double2*>(&R.dphi2[q][3 * I + i][0]);
      PHYS(ax[i], ay[i], g);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const double N = R.n1[q][j];
      const double csn = cs * N, cdn = cd * N;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        acc[i][j][0] = fma(csn, ax[i], acc[i][j][0]); acc[i][j][1] = fma(csn, ay[i], acc[i][j][1]);
        acc[i][j][2] = fma(cdn, ax[i], acc[i][j][2]); acc[i][j][3] = fma(cdn, ay[i], acc[i][j][3]);
      }
    }
  }
  if (part == 1) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int m = 0; m < 4; ++m) mom[MOM_UT + 4 * ((3 * I + i) * 4 + j) + m] = acc[i][j][m];
  }
  if (part != 0) __syncwarp();
  if (part != 1) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          double* o = mom + MOM_UT + 4 * ((3 * I + i) * 4 + j) + m;
          *o = part == 2 ? acc[i][j][m] + *o : acc[i][j][m];
        }
  }
}


// one halo contribution (cell of another chunk, local pair lo <= hi): same moments, one pair per thread.  The records come
// from L2 in 8-byte pieces (one 32-byte sector each): the loop is unrolled so that HALO_UNROLL points' loads are in flight
// (r02f experiments: the halo items cost 0.69 of 3.69 ms, bound by L2 latency and sector traffic, not by arithmetic)
// HALO_UNROLL is a template parameter (a constant expression: the unroll pragma does not expand macros): 4 for records read
// from global memory (8 was slower there: 3.86 against 3.69 ms), 8 for the shared-memory copies
// rc: the cell's record; point q, field f at rc[q * qs + f * fs] (global [chunk][q][field][lane] layout: qs = TL_NF * C,
// fs = C; shared-memory copy of a halo cell: qs = TL_NF, fs = 1)
template <int HALO_UNROLL>
__device__ __forceinline__ void halo_item(const double* __restrict__ rc, int qs, int fs, const TileTab& R, int nq, int a, int b,
                                          double* __restrict__ o) {
  double m0 = 0, m1 = 0, m2 = 0, m3 = 0, m4 = 0;
  if (b < 9) {
#pragma unroll HALO_UNROLL
    for (int q = 0; q < nq; ++q) {
      const double* r = rc + q * qs;
      const double k00 = r[TF_K00 * fs], k01 = r[TF_K01 * fs], k10 = r[TF_K10 * fs], k11 = r[TF_K11 * fs];
      const double2 ga = *reinterpret_cast<const double2*>(&R.dphi2[q][a][0]);
      const double2 gb = *reinterpret_cast<const double2*>(&R.dphi2[q][b][0]);
      double ax, ay, bx, by;
      PHYS(ax, ay, ga); PHYS(bx, by, gb);
      const double w11 = r[TF_CM11 * fs], w12 = r[TF_CM12 * fs], ws = r[TF_CS * fs];
      const double w33 = r[TF_CM33 * fs], wd4 = r[TF_CD * fs];
      const double w3p = fma(0.25, wd4, w33), w3m = fma(-0.25, wd4, w33);
      m0 = fma(w11 * ax, bx, fma(w3p * ay, by, m0));
      m1 = fma(w12 * ax, by, fma(w3m * ay, bx, m1));
      m2 = fma(w12 * ay, bx, fma(w3m * ax, by, m2));
      m3 = fma(w11 * ay, by, fma(w3p * ax, bx, m3));
      m4 = fma(ws * ax, bx, fma(ws * ay, by, m4));
    }
  } else if (a >= 9) {
    const int ta = a - 9, tb = b - 9;
#pragma unroll HALO_UNROLL
    for (int q = 0; q < nq; ++q) {
      const double* r = rc + q * qs;
      const double k00 = r[TF_K00 * fs], k01 = r[TF_K01 * fs], k10 = r[TF_K10 * fs], k11 = r[TF_K11 * fs];
      const double2 ga = *reinterpret_cast<const double2*>(&R.dn1[q][ta][0]);
      const double2 gb = *reinterpret_cast<const double2*>(&R.dn1[q][tb][0]);
      double ax, ay, bx, by;
      PHYS(ax, ay, ga); PHYS(bx, by, gb);
      const double hb = r[TF_HB * fs], cs = r[TF_CS * fs], cd = r[TF_CD * fs];
      const double cb11 = hb * r[TF_CM11 * fs], cb12 = hb * r[TF_CM12 * fs], cb33 = hb * r[TF_CM33 * fs];
      const double Na = R.n1[q][ta], Nb = R.n1[q][tb];
      const double ns = cs * Na;
      m0 = fma(cb11 * ay, by, fma(cb33 * ax, bx, fma(ns, Nb, m0)));
      m1 = fma(-(cb12 * ay), bx, fma(-(cb33 * ax), by, m1));
      m2 = fma(-(cb12 * ax), by, fma(-(cb33 * ay), bx, m2));
      m3 = fma(cb11 * ax, bx, fma(cb33 * ay, by, fma(ns, Nb, m3)));
      m4 = fma(cd * Na, Nb, m4);
    }
  } else {
    const int tb = b - 9;
#pragma unroll HALO_UNROLL
    for (int q = 0; q < nq; ++q) {
      const double* r = rc + q * qs;
      const double k00 = r[TF_K00 * fs], k01 = r[TF_K01 * fs], k10 = r[TF_K10 * fs], k11 = r[TF_K11 * fs];
      const double2 ga = *reinterpret_cast<const double2*>(&R.dphi2[q][a][0]);
      double ax, ay;
      PHYS(ax, ay, ga);
      const double N = R.n1[q][tb];
      const double csn = r[TF_CS * fs] * N, cdn = r[TF_CD * fs] * N;
      m0 = fma(csn, ax, m0); m1 = fma(csn, ay, m1); m2 = fma(cdn, ax, m2); m3 = fma(cdn, ay, m3);
    }
  }
  o[0] = m0; o[1] = m1; o[2] = m2; o[3] = m3; o[4] = m4;
}

__device__ __forceinline__ void cp_async4(void* s, const void* g) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(s);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(a), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async8(void* s, const void* g) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(s);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(a), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async16(void* s, const void* g) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(s);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(a), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---- the kernel -------------------------------------------------------------------------------------------------
// 6 warps per CTA, 2 CTAs per SM (one CTA's phase 2 overlaps the other's phase 1).  A warp holds two half-warp
// tiles of the chunk's 16 cells:  w0-w2: the six Q2-Q2 tiles; w3: Q1-Q1 | idle; w4: Q2-Q1 tiles 0 | 1; w5: Q2-Q1 tile 2 | idle; the idle
// half of w3 and w4, w5 compute the halo items.  The chunk's records and plan arrive by cp.async.
struct TileSmem {                      // offsets in doubles
  int mom, frame, rec, plan, total;
};

__host__ __device__ inline TileSmem tile_smem_layout(int nq, int hitem_cap, int hcell_cap, int dest_cap, int ref_cap) {
  TileSmem L;
  L.mom = (int)(sizeof(TileTab) / sizeof(double));
  L.frame = L.mom + MOM_HALO_BASE + MOM_HALO * hitem_cap;
  L.rec = L.frame + (C + hcell_cap) * 9;
  L.rec = (L.rec + 1) & ~1;
  int rec_d = nq * TL_NF * C, sk_d = TL_THREADS * 9;          // records (phase 1) alias the output staging (phase 2)
  L.plan = L.rec + (rec_d > sk_d ? rec_d : sk_d);
  L.total = L.plan + (3 * dest_cap + ref_cap + 1) / 2 + 1;
  return L;
}

__global__ void __launch_bounds__(TL_THREADS, 2)
k_asm_tiles(const int32_t* __restrict__ dest_ptr, const int32_t* __restrict__ dest_pos,
            const int32_t* __restrict__ dest_tpos, const uint32_t* __restrict__ dest_info,
            const int32_t* __restrict__ ref_ptr, const uint32_t* __restrict__ refs,
            const int32_t* __restrict__ hitem_ptr, const uint32_t* __restrict__ hitems,
            const int32_t* __restrict__ hcell_ptr, const int32_t* __restrict__ hcells,
            const double* __restrict__ rec, const double* __restrict__ frame, const Tables* __restrict__ tab, int nq,
            int hitem_cap, int hcell_cap, int dest_cap, int ref_cap, double* __restrict__ vals) {
  extern __shared__ __align__(16) double smem[];
  const TileSmem L = tile_smem_layout(nq, hitem_cap, hcell_cap, dest_cap, ref_cap);
  TileTab& R = *reinterpret_cast<TileTab*>(smem);
  double* sMom = smem + L.mom;
  double* sFrame = smem + L.frame;
  double* sRec = smem + L.rec;
  double* sK = sRec;                                                  // reused after phase 1
  uint32_t* sInfo = reinterpret_cast<uint32_t*>(smem + L.plan);      // dest_cap
  int* sPos = reinterpret_cast<int*>(sInfo + dest_cap);               // dest_cap
  int* sTpos = sPos + dest_cap;                                       // dest_cap
  uint32_t* sRefs = reinterpret_cast<uint32_t*>(sTpos + dest_cap);   // ref_cap
  const int64_t chunk = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int slot = lane & (C - 1), half = lane >> 4;
  const int d0 = dest_ptr[chunk], nd = dest_ptr[chunk + 1] - d0;
  const int r0 = ref_ptr[chunk], nr = ref_ptr[chunk + 1] - r0;
  // ---- asynchronous copies: the chunk's quadrature records (group 0), its plan (group 1) ----
  {
    const double2* g = reinterpret_cast<const double2*>(rec + chunk * ((int64_t)nq * TL_NF * C));
    double2* sr = reinterpret_cast<double2*>(sRec);
    const int n2 = nq * TL_NF * C / 2;
    for (int i = tid; i < n2; i += TL_THREADS) cp_async16(sr + i, g + i);
    cp_async_commit();
    for (int i = tid; i < nd; i += TL_THREADS) {
      cp_async4(sInfo + i, dest_info + d0 + i); cp_async4(sPos + i, dest_pos + d0 + i); cp_async4(sTpos + i, dest_tpos + d0 + i);
    }
    for (int i = tid; i < nr; i += TL_THREADS) cp_async4(sRefs + i, refs + r0 + i);
    cp_async_commit();
  }
  // ---- reference tables and cell frames ----
  {
    double* d = &R.dphi2[0][0][0]; const double* s0 = &tab->dphi2[0][0][0];
    for (int i = tid; i < nq * 18; i += TL_THREADS) d[i] = s0[i];
    double* d1 = &R.dn1[0][0][0]; const double* s1 = &tab->dn1[0][0][0];
    for (int i = tid; i < nq * 8; i += TL_THREADS) d1[i] = s1[i];
    double* d2 = &R.n1[0][0]; const double* s2 = &tab->n1[0][0];
    for (int i = tid; i < nq * 4; i += TL_THREADS) d2[i] = s2[i];
    const double* f0 = frame + chunk * (C * 9);
    for (int i = tid; i < C * 9; i += TL_THREADS) sFrame[i] = f0[i];
    const int hc0 = hcell_ptr[chunk], nhc = hcell_ptr[chunk + 1] - hc0;
    for (int i = tid; i < nhc * 9; i += TL_THREADS) {
      const int s = i / 9, k = i - 9 * s;
      sFrame[(C + s) * 9 + k] = frame[(int64_t)hcells[hc0 + s] * 9 + k];
    }
  }
  cp_async_wait<1>();
  __syncthreads();
  // ---- phase 1 ----
  {
    const double* rc = sRec + slot;
    double* mom = sMom + slot * MOM_CELL;
    if (wid < 3) {
      const int t = 2 * wid + half;                  // tiles (0,0) (0,1) (0,2) (1,1) (1,2) (2,2)
      const int I = t < 3 ? 0 : (t < 5 ? 1 : 2);
      const int J = t < 3 ? t : (t < 5 ? t - 2 : 2);
      tile_uu(rc, R, nq, I, J, mom);
    } else if (wid == 3) {
      if (half == 0) tile_tt(rc, R, nq, mom);
    } else if (wid == 4) {
      tile_ut(rc, R, nq, half, mom);
    } else if (half == 0) {
      tile_ut(rc, R, nq, 2, mom);
    }
    // halo items, one pair per thread: the idle half of w3, then w4 and w5 after their light tiles
    const int hidx = wid == 3 ? (half ? lane - 16 : -1) : (wid >= 4 ? tid - 4 * 32 + 16 : -1);
    if (hidx >= 0) {
      const int h0 = hitem_ptr[chunk], nh = hitem_ptr[chunk + 1] - h0;
      for (int i = hidx; i < nh; i += 80) {
        const uint32_t it = hitems[h0 + i];
        const int64_t p = it >> 8;
        const int code = it & 0xff, a = code / 13, b = code - 13 * a;
        const double* rh = rec + (p / C) * ((int64_t)nq * TL_NF * C) + (p % C);
        halo_item<4>(rh, TL_NF * C, C, R, nq, a, b, sMom + MOM_HALO_BASE + MOM_HALO * i);
      }
    }
  }
  cp_async_wait<0>();
  __syncthreads();
  // ---- phase 2 ----
  for (int base = 0; base < nd; base += TL_THREADS) {
    const int di = base + tid;
    if (di < nd) {
      const uint32_t info = sInfo[di];
      const int start = info & 0xFFFFF, n = info >> 20;
      double K[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) K[k] = 0.0;
      for (int r = 0; r < n; ++r) {                                     // fixed (processing-order) summation
        const uint32_t w = sRefs[start + r];
        const double* m = sMom + (w & 0xFFFF);
        const int type = (w >> 16) & 3;
        const double* fr = sFrame + (w >> 19) * 9;
        const V3 E0 = ldv3(fr), E1 = ldv3(fr + 3), E2 = ldv3(fr + 6);
        if (type < 2) {
          const double m00 = m[0], m11 = m[3], s = m[4];
          double m01 = m[1], m10 = m[2];
          if ((w >> 18) & 1) { const double t = m01; m01 = m10; m10 = t; }
          outer_acc(K, m00 * E0 + m10 * E1, E0);
          outer_acc(K, m01 * E0 + m11 * E1, E1);
          outer_acc(K, s * E2, E2);
        } else {
          // K[u_i, th_j] = A0 E2 E1^T - A1 E2 E0^T + 0.5 D1 E0 E2^T - 0.5 D0 E1 E2^T
          const double A0 = m[0], A1 = m[1], D0 = m[2], D1 = m[3];
          outer_acc(K, E2, A0 * E1 - A1 * E0);
          outer_acc(K, 0.5 * (D1 * E0 - D0 * E1), E2);
        }
      }
#pragma unroll
      for (int k = 0; k < 9; ++k) sK[tid * 9 + k] = K[k];
    }
    __syncthreads();
    const int cnt = min(TL_THREADS, nd - base);
    if (tid < 189) {                                                    // 21 blocks x 9 entries per pass, k fixed per thread
      const int k = tid % 9, kt = (k % 3) * 3 + k / 3;
      for (int hh = tid / 9; hh < cnt; hh += 21) {
        const double v = sK[hh * 9 + k];
        vals[(int64_t)sPos[base + hh] * 9 + k] = v;
        const int tp = sTpos[base + hh];
        if (tp >= 0) vals[(int64_t)tp * 9 + kt] = v;
      }
    }
    __syncthreads();
  }
}

// ---- warp-specialised persistent variant ------------------------------------------------------------------------
// One CTA per SM, 12 warps.  Six COMPUTE warps run phase 1 of chunk k+1 while six WRITER warps run the halo items,
// phase 2 and the stores of chunk k: the fp64 pipe no longer idles during the memory phases.  The tile moments are
// double-buffered in shared memory and handed over with named barriers (producer: bar.arrive full / bar.sync empty;
// consumer: bar.sync full / bar.arrive empty).  Warp roles are laid out so that every SM sub-partition (warp id mod 4)
// carries one Q2-Q2 warp or the Q1-Q1 / Q2-Q1 warps, plus writers:
//   compute: w0 w1 w2 = Q2-Q2 tile pairs; w3 = Q1-Q1 tile, quadrature points split over the two half-warps;
//            w7 = Q2-Q1 tiles 0 | 1; w11 = Q2-Q1 tile 2, quadrature points split
//   writers: w4 w5 w6 w8 w9 w10
#define BAR_FULL 1      /* +buf */
#define BAR_EMPTY 3     /* +buf */
#define BAR_COMPUTE 5
#define BAR_WRITER 6
#define BAR_HALO 7

__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }

struct WsSmem { int mom0, mom_stride, hmom0, hmom_stride, frame, rec0, rec_stride, sk, plan, flags, hrec, total; };   // offsets in doubles

// sk_rows: 3x3 blocks staged per store round (writer threads x destinations per thread); hrec_cells: halo cells whose
// records are staged in shared memory (0: the halo items read them from global memory)
__host__ __device__ inline WsSmem ws_smem_layout(int nq, int hitem_cap, int hcell_cap, int dest_cap, int ref_cap, int sk_rows = 384,
                                                 int hrec_cells = 0) {
  WsSmem L;
  int o = (int)(sizeof(TileTab) / sizeof(double));
  L.mom0 = o; L.mom_stride = MOM_HALO_BASE + 1; o += 2 * L.mom_stride;
  L.hmom0 = o; L.hmom_stride = MOM_HALO * hitem_cap; o += 2 * L.hmom_stride;
  L.frame = o; o += (C + hcell_cap) * 9;
  o = (o + 1) & ~1;
  L.rec0 = o; L.rec_stride = nq * TL_NF * C; o += 2 * L.rec_stride;
  L.sk = o; o += sk_rows * 9;
  L.plan = o; o += (3 * dest_cap + ref_cap + 1) / 2 + 1;
  L.flags = o; o += 8;
  // records (per-cell stride nq * TL_NF + 1 doubles: an even stride maps every cell to the same banks and serialises the
  // loads of a warp whose items span several cells -- measured 4.49 ms against 3.76 without staging) + processing indices
  // two buffers (the records of chunk k+1 are copied while chunk k is processed) + three cell lists (k, k+1, k+2) of ints
  L.hrec = o; o += 2 * hrec_cells * (nq * TL_NF + 1) + (3 * hrec_cells + 1) / 2;
  L.total = o;
  return L;
}

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// MODE 0: 12 warps, 6 compute (w0-3, w7, w11) + 6 writers, all at the compiler's register count.
// MODE 1: 16 warps, 8 compute-class warps (two warpgroups: six tile warps + two halo-item warps, raised from 128 to 168
//         registers with setmaxnreg) + 8 writer warps (two warpgroups, lowered to 80).  The registers the writers give
//         back (256 x 48) must cover what the compute warps ask for (256 x 40): setmaxnreg.inc only draws from the CTA pool.
template <int MODE>
struct WsCfg {
  static constexpr int NT = MODE ? 512 : 384;       // threads per CTA
  static constexpr int CT = MODE ? 256 : 192;       // compute threads
  static constexpr int WT = MODE ? 256 : 192;       // writer threads
  static constexpr int DPT = MODE ? 1 : 2;          // destinations per writer thread and round
  static constexpr int HALO0 = MODE ? 4 : 3;        // first compute warp that takes halo items
  static constexpr int NCW = MODE ? 8 : 6;          // compute warps
};

template <int MODE>
__global__ void __launch_bounds__(WsCfg<MODE>::NT, 1)
k_asm_tiles_ws(const int32_t* __restrict__ dest_ptr, const int32_t* __restrict__ dest_pos,
               const int32_t* __restrict__ dest_tpos, const uint32_t* __restrict__ dest_info,
               const int32_t* __restrict__ ref_ptr, const uint32_t* __restrict__ refs,
               const int32_t* __restrict__ hitem_ptr, const uint32_t* __restrict__ hitems,
               const int32_t* __restrict__ hcell_ptr, const int32_t* __restrict__ hcells,
               const double* __restrict__ rec, const double* __restrict__ frame, const Tables* __restrict__ tab, int nq,
               int hitem_cap, int hcell_cap, int dest_cap, int ref_cap, int64_t nchunks, double* __restrict__ vals, int dbg,
               int hrec_cells, const int32_t* __restrict__ hfix) {
  // dbg (FEMO_TILE_DEBUG, timing experiments only -- results are wrong when set): 1 no global stores, 2 no tile arithmetic,
  // 4 no halo items, 8 no gather / rotation on the writer side
  using Cfg = WsCfg<MODE>;
  constexpr int WS_THREADS = Cfg::NT, CT = Cfg::CT, WT = Cfg::WT, DPT = Cfg::DPT;
  extern __shared__ __align__(16) double smem[];
  const WsSmem L = ws_smem_layout(nq, hitem_cap, hcell_cap, dest_cap, ref_cap, WT * DPT, hrec_cells);
  TileTab& R = *reinterpret_cast<TileTab*>(smem);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int slot = lane & (C - 1), half = lane >> 4;
  {
    double* d = &R.dphi2[0][0][0]; const double* s0 = &tab->dphi2[0][0][0];
    for (int i = tid; i < nq * 18; i += WS_THREADS) d[i] = s0[i];
    double* d1 = &R.dn1[0][0][0]; const double* s1 = &tab->dn1[0][0][0];
    for (int i = tid; i < nq * 8; i += WS_THREADS) d1[i] = s1[i];
    double* d2 = &R.n1[0][0]; const double* s2 = &tab->n1[0][0];
    for (int i = tid; i < nq * 4; i += WS_THREADS) d2[i] = s2[i];
  }
  __syncthreads();
  const int64_t G = gridDim.x;
  const int64_t first = blockIdx.x;
  const int n_it = first < nchunks ? (int)((nchunks - first + G - 1) / G) : 0;
  const bool compute = MODE ? wid < 8 : (wid < 4 || wid == 7 || wid == 11);
  const int64_t rec_chunk = (int64_t)nq * TL_NF * C;
  if (compute) {
    // ================================ COMPUTE WARPS: phase 1 ================================
    if (MODE) asm volatile("setmaxnreg.inc.sync.aligned.u32 168;");
    const int cw = MODE ? wid : (wid < 4 ? wid : (wid == 7 ? 4 : 5));
    const int ctid = cw * 32 + lane;
    const int n2 = (int)(rec_chunk / 2);
    if (n_it > 0) {
      const double2* g = reinterpret_cast<const double2*>(rec + first * rec_chunk);
      double2* sr = reinterpret_cast<double2*>(smem + L.rec0);
      for (int i = ctid; i < n2; i += CT) cp_async16(sr + i, g + i);
      if (hrec_cells > 0 && cw >= Cfg::HALO0) {        // halo pipeline prologue: lists of chunks 0 and 1, records of chunk 0
        double* hr = smem + L.hrec;
        int32_t* hpb = reinterpret_cast<int32_t*>(hr + 2 * hrec_cells * (nq * TL_NF + 1));
        const int htid = (cw - Cfg::HALO0) * 32 + lane, HT = (Cfg::NCW - Cfg::HALO0) * 32;
        const int per = nq * TL_NF;
        for (int i = htid; i < hrec_cells; i += HT) {
          cp_async4(hpb + i, hfix + first * hrec_cells + i);
          if (n_it > 1) cp_async4(hpb + hrec_cells + i, hfix + (first + G) * hrec_cells + i);
        }
        for (int i = htid; i < hrec_cells * per; i += HT) {
          const int sc = i / per, idx = i - sc * per;
          const int64_t pc = hfix[first * hrec_cells + sc];          // fixed-width rows, -1 padded
          if (pc >= 0) cp_async8(hr + sc * (per + 1) + idx, rec + (pc / C) * rec_chunk + (pc % C) + (int64_t)idx * C);
        }
      }
    }
    cp_async_commit();
    for (int it = 0; it < n_it; ++it) {
      const int64_t chunk = first + it * G;
      const int buf = it & 1;
      cp_async_wait<0>();
      bar_sync(BAR_COMPUTE, CT);               // rec[buf] complete; everybody left rec[buf ^ 1] (iteration it-1)
      // halo warps: the records of a chunk's (few) halo cells are staged in shared memory -- a chunk's ~115 halo items come
      // from ~9 cells, and read from global memory per item they cost 13x the L2 sectors (r02f: 0.69 of 3.69 ms).  The
      // copies of chunk it+1 are issued now and land during this iteration (two record buffers); the cell lists run two
      // chunks ahead (three small buffers), so no global load sits between the barrier and the copies.
      int nhc_s = 0;
      if (hrec_cells > 0 && cw >= Cfg::HALO0) {
        double* hr = smem + L.hrec;
        const int per = nq * TL_NF;
        int32_t* hpb = reinterpret_cast<int32_t*>(hr + 2 * hrec_cells * (per + 1));
        const int32_t* hp = hpb + (it % 3) * hrec_cells;
        while (nhc_s < hrec_cells && hp[nhc_s] >= 0) ++nhc_s;
        const int htid = (cw - Cfg::HALO0) * 32 + lane, HT = (Cfg::NCW - Cfg::HALO0) * 32;
        if (it + 1 < n_it) {
          const int32_t* hpn = hpb + ((it + 1) % 3) * hrec_cells;
          double* hrn = hr + (buf ^ 1) * hrec_cells * (per + 1);
          for (int i = htid; i < hrec_cells * per; i += HT) {
            const int sc = i / per, idx = i - sc * per;
            const int64_t pc = hpn[sc];
            if (pc >= 0) cp_async8(hrn + sc * (per + 1) + idx, rec + (pc / C) * rec_chunk + (pc % C) + (int64_t)idx * C);
          }
        }
        if (it + 2 < n_it)
          for (int i = htid; i < hrec_cells; i += HT) cp_async4(hpb + ((it + 2) % 3) * hrec_cells + i, hfix + (chunk + 2 * G) * hrec_cells + i);
        cp_async_commit();
      }
      if (it + 1 < n_it) {
        const double2* g = reinterpret_cast<const double2*>(rec + (chunk + G) * rec_chunk);
        double2* sr = reinterpret_cast<double2*>(smem + L.rec0 + (buf ^ 1) * L.rec_stride);
        for (int i = ctid; i < n2; i += CT) cp_async16(sr + i, g + i);
      }
      cp_async_commit();
      if (it >= 2) bar_sync(BAR_EMPTY + buf, WS_THREADS);   // writers are done with mom[buf] (chunk it-2)
      const double* rc = smem + L.rec0 + buf * L.rec_stride + slot;
      double* mom = smem + L.mom0 + buf * L.mom_stride + slot * MOM_CELL;
      if (dbg & 2) {
      } else if (cw < 3) {
        const int t = 2 * cw + half;                 // tiles (0,0) (0,1) (0,2) (1,1) (1,2) (2,2)
        const int I = t < 3 ? 0 : (t < 5 ? 1 : 2);
        const int J = t < 3 ? t : (t < 5 ? t - 2 : 2);
        tile_uu(rc, R, nq, I, J, mom);
      } else if (cw == 3) {
        tile_tt(rc, R, nq, mom, half, 2, half ? 1 : 2);
      } else if (cw == 4) {
        tile_ut(rc, R, nq, half, mom);
      } else if (cw == 5) {
        tile_ut(rc, R, nq, 2, mom, half, 2, half ? 1 : 2);
      }
      if (cw >= Cfg::HALO0 && !(dbg & 4)) {                          // halo items on the lighter warps: one pair per thread
        const int h0 = hitem_ptr[chunk], nh = hitem_ptr[chunk + 1] - h0;
        double* hm = smem + L.hmom0 + buf * L.hmom_stride;
        // records of this chunk's halo cells: copied during the previous iteration, complete since the loop-top wait + barrier
        const double* hr = smem + L.hrec + buf * hrec_cells * (nq * TL_NF + 1);
        const int32_t* hp = reinterpret_cast<const int32_t*>(smem + L.hrec + 2 * hrec_cells * (nq * TL_NF + 1)) + (it % 3) * hrec_cells;
        for (int i = (cw - Cfg::HALO0) * 32 + lane; i < nh; i += (Cfg::NCW - Cfg::HALO0) * 32) {
          const uint32_t item = hitems[h0 + i];
          const int64_t p = item >> 8;
          const int code = item & 0xff, a = code / 13, b = code - 13 * a;
          if (hrec_cells > 0) {
            int sc = 0;
            while (sc < nhc_s - 1 && hp[sc] != (int32_t)p) ++sc;     // <= hrec_cells entries; p is always among them
            halo_item<8>(hr + sc * (nq * TL_NF + 1), TL_NF, 1, R, nq, a, b, hm + MOM_HALO * i);
          } else {
            const double* rh = rec + (p / C) * rec_chunk + (p % C);
            halo_item<4>(rh, TL_NF * C, C, R, nq, a, b, hm + MOM_HALO * i);
          }
        }
      }
      __threadfence_block();
      bar_arrive(BAR_FULL + buf, WS_THREADS);
    }
    cp_async_wait<0>();
  } else {
    // ================================ WRITER WARPS: halo items, phase 2, stores ================================
    if (MODE) asm volatile("setmaxnreg.dec.sync.aligned.u32 80;");
    const int ww = MODE ? wid - 8 : (wid < 7 ? wid - 4 : wid - 5);
    const int wtid = ww * 32 + lane;
    double* sFrame = smem + L.frame;
    double* sK = smem + L.sk;
    uint32_t* sInfo = reinterpret_cast<uint32_t*>(smem + L.plan);
    int* sPos = reinterpret_cast<int*>(sInfo + dest_cap);
    int* sTpos = sPos + dest_cap;
    uint32_t* sRefs = reinterpret_cast<uint32_t*>(sTpos + dest_cap);
    int* sFlag = reinterpret_cast<int*>(smem + L.flags);
    for (int it = 0; it < n_it; ++it) {
      const int64_t chunk = first + it * G;
      const int buf = it & 1;
      const int d0 = dest_ptr[chunk], nd = dest_ptr[chunk + 1] - d0;
      const int r0 = ref_ptr[chunk], nr = ref_ptr[chunk + 1] - r0;
      const int hc0 = hcell_ptr[chunk], nhc = hcell_ptr[chunk + 1] - hc0;
      bar_sync(BAR_WRITER, WT);                 // previous chunk's stores issued: plan, frames, sK reusable
      for (int i = wtid; i < nd; i += WT) {
        cp_async4(sInfo + i, dest_info + d0 + i); cp_async4(sPos + i, dest_pos + d0 + i); cp_async4(sTpos + i, dest_tpos + d0 + i);
      }
      for (int i = wtid; i < nr; i += WT) cp_async4(sRefs + i, refs + r0 + i);
      cp_async_commit();
      {
        const double* f0 = frame + chunk * (C * 9);
        for (int i = wtid; i < C * 9; i += WT) sFrame[i] = f0[i];
        for (int i = wtid; i < nhc * 9; i += WT) {
          const int s = i / 9, k = i - 9 * s;
          sFrame[(C + s) * 9 + k] = frame[(int64_t)hcells[hc0 + s] * 9 + k];
        }
      }
      if (it + 1 < n_it) {                           // pull the next chunk's plan towards L2 while this one is processed
        const int64_t nx = chunk + G;
        const int e0 = dest_ptr[nx], ne = dest_ptr[nx + 1] - e0, q0 = ref_ptr[nx], nqr = ref_ptr[nx + 1] - q0;
        const int l1 = (ne + 31) / 32, l2 = (nqr + 31) / 32;      // 128-byte lines per array
        for (int i = wtid; i < 3 * l1 + l2 + 10; i += WT) {
          const void* a;
          if (i < l1) a = dest_info + e0 + 32 * i;
          else if (i < 2 * l1) a = dest_pos + e0 + 32 * (i - l1);
          else if (i < 3 * l1) a = dest_tpos + e0 + 32 * (i - 2 * l1);
          else if (i < 3 * l1 + l2) a = refs + q0 + 32 * (i - 3 * l1);
          else a = frame + nx * (C * 9) + 16 * (i - 3 * l1 - l2 < 9 ? i - 3 * l1 - l2 : 8);
          prefetch_l2(a);
        }
      }
      cp_async_wait<0>();
      bar_sync(BAR_FULL + buf, WS_THREADS);          // moments of this chunk are in mom[buf] / hmom[buf]; all writers are here
      {                                              // do all cells of this chunk (and its halo) share one frame?
        const int nf = (C + nhc) * 9;
        bool same = true;
        for (int i = wtid; i < nf; i += WT)
          same = same && (__double_as_longlong(sFrame[i]) == __double_as_longlong(sFrame[i % 9]));
        same = __all_sync(0xffffffffu, same);
        if (lane == 0) sFlag[ww] = same ? 1 : 0;
      }
      bar_sync(BAR_WRITER, WT);
      bool uniform = true;
#pragma unroll
      for (int w = 0; w < WT / 32; ++w) uniform = uniform && sFlag[w] != 0;
      const V3 F0 = ldv3(sFrame), F1 = ldv3(sFrame + 3), F2 = ldv3(sFrame + 6);
      const double* sMom = smem + L.mom0 + buf * L.mom_stride;
      const double* sHm = smem + L.hmom0 + buf * L.hmom_stride;
      // Rounds of 32 destinations per WARP: gather + rotate into the warp's own slice of sK, then the same warp stores its
      // 32 blocks and their mirror images (lane j of pass i holds value 32 i + j of the slice: 3.6 blocks of 72 contiguous
      // bytes per store instruction, as before) -- no CTA-wide barrier inside a chunk (round 2; the two barriers per round
      // of 256 destinations cost the writer warps 17 % of their time, profiles/r02n_ncu_full_k_asm_tiles_ws_details.csv)
      double* myK = sK + ww * (32 * 9);
      for (int base = ww * 32; base < nd; base += WT) {
        const int di = base + lane;
        double K[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) K[k] = 0.0;
        if (di < nd && !(dbg & 8)) {
          const uint32_t info = sInfo[di];
          const int start = info & 0xFFFFF, n = info >> 20;
          if (uniform) {
            // flat patch: add the local-frame moments of all contributions first (fixed order), rotate once
            double a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0;
            int type = 0;
            for (int r = 0; r < n; ++r) {
              const uint32_t w = sRefs[start + r];
              const int off = w & 0xFFFF;
              const double* m = off < MOM_HALO_BASE ? sMom + off : sHm + (off - MOM_HALO_BASE);
              type = (w >> 16) & 3;
              const bool tr = (w >> 18) & 1;
              a0 += m[0]; a3 += m[3];
              a1 += tr ? m[2] : m[1]; a2 += tr ? m[1] : m[2];
              if (type < 2) a4 += m[4];
            }
            if (type < 2) {
              outer_acc(K, a0 * F0 + a2 * F1, F0);
              outer_acc(K, a1 * F0 + a3 * F1, F1);
              outer_acc(K, a4 * F2, F2);
            } else {
              outer_acc(K, F2, a0 * F1 - a1 * F0);
              outer_acc(K, 0.5 * (a3 * F0 - a2 * F1), F2);
            }
          } else {
            for (int r = 0; r < n; ++r) {                                 // fixed (processing-order) summation
              const uint32_t w = sRefs[start + r];
              const int off = w & 0xFFFF;
              const double* m = off < MOM_HALO_BASE ? sMom + off : sHm + (off - MOM_HALO_BASE);
              const int type = (w >> 16) & 3;
              const double* fr = sFrame + (w >> 19) * 9;
              const V3 E0 = ldv3(fr), E1 = ldv3(fr + 3), E2 = ldv3(fr + 6);
              if (type < 2) {
                const double m00 = m[0], m11 = m[3], sv = m[4];
                double m01 = m[1], m10 = m[2];
                if ((w >> 18) & 1) { const double t = m01; m01 = m10; m10 = t; }
                outer_acc(K, m00 * E0 + m10 * E1, E0);
                outer_acc(K, m01 * E0 + m11 * E1, E1);
                outer_acc(K, sv * E2, E2);
              } else {
                const double A0 = m[0], A1 = m[1], D0 = m[2], D1 = m[3];
                outer_acc(K, E2, A0 * E1 - A1 * E0);
                outer_acc(K, 0.5 * (D1 * E0 - D0 * E1), E2);
              }
            }
          }
        }
        __syncwarp();                                  // the previous round's stores have read the slice
#pragma unroll
        for (int k = 0; k < 9; ++k) myK[lane * 9 + k] = K[k];
        __syncwarp();
        const int cnt = min(32, nd - base);
        if (!(dbg & 1)) {
          int k = lane % 9, hh = lane / 9;
#pragma unroll 1
          for (int i = 0; i < 9; ++i) {
            if (hh < cnt) {
              const double v = myK[32 * i + lane];
              vals[(int64_t)sPos[base + hh] * 9 + k] = v;
              const int tp = sTpos[base + hh];
              if (tp >= 0) vals[(int64_t)tp * 9 + (k % 3) * 3 + k / 3] = v;
            }
            k += 5; hh += 3;                           // value index + 32 = 3 blocks + 5 entries
            if (k >= 9) { k -= 9; hh += 1; }
          }
        }
      }
      if (it + 2 < n_it) {
        __threadfence_block();
        bar_arrive(BAR_EMPTY + buf, WS_THREADS);       // mom[buf] may be overwritten (chunk it+2)
      }
    }
  }
}

template <typename T>
int upload(T** dst, const std::vector<T>& v) {
  TRY(dev_alloc(dst, v.size()));
  if (!v.empty()) CUDA_TRY(cudaMemcpy(*dst, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice));
  return 0;
}

template <typename T>
int download(std::vector<T>& v, const T* src, size_t n) {
  v.resize(n);
  if (n) CUDA_TRY(cudaMemcpy(v.data(), src, sizeof(T) * n, cudaMemcpyDeviceToHost));
  return 0;
}

}  // namespace

void tile_plan_free(femo_ctx* c) {
  c->tile_ok = false;
  dev_free(&c->tl_cell_order); dev_free(&c->tl_dest_ptr); dev_free(&c->tl_dest_pos); dev_free(&c->tl_dest_tpos);
  dev_free(&c->tl_dest_info); dev_free(&c->tl_refs); dev_free(&c->tl_hitems); dev_free(&c->tl_ref_ptr);
  dev_free(&c->tl_hitem_ptr); dev_free(&c->tl_hcell_ptr); dev_free(&c->tl_hcells); dev_free(&c->tl_rec);
  dev_free(&c->tl_frame); dev_free(&c->tl_hfix);
}

// FEMO_TILE=0 disables the tiled kernel; FEMO_TILE_ORDER=identity keeps the caller's cell order (default: Morton)
int tile_plan_setup(femo_ctx* c) {
  tile_plan_free(c);
  const char* en = getenv("FEMO_TILE");
  if (en && atoi(en) == 0) return 0;
  if (!c->all_planar) return 0;
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  std::vector<int32_t> rowptr, colidx, nc_ptr, nc_adj, cv, order;
  TRY(download(rowptr, c->rowptr, (size_t)c->nb + 1));
  TRY(download(colidx, c->colidx, (size_t)c->nnzb));
  TRY(download(nc_ptr, c->nc_ptr, (size_t)c->nb + 1));
  TRY(download(nc_adj, c->nc_adj, (size_t)c->nc * 13));
  const char* eo = getenv("FEMO_TILE_ORDER");
  if (!(eo && strcmp(eo, "identity") == 0)) {
    std::vector<double> xyz, hE;
    TRY(download(xyz, c->xyz, (size_t)c->nn * 3));
    TRY(download(cv, c->cv, (size_t)c->nc * 4));
    TRY(download(hE, c->hE, (size_t)c->nc));
    double s = 0;
    for (double v : hE) s += v;
    femo_tiles::morton_order(c->nc, xyz.data(), cv.data(), 0.70710678 * s / (double)c->nc, order);
  }
  femo_tiles::Plan P;
  std::string err = femo_tiles::build(c->nc, c->nb, rowptr.data(), colidx.data(), nc_ptr.data(), nc_adj.data(),
                                      order.empty() ? nullptr : order.data(), P,
                                      getenv("FEMO_TILE_SORT") ? atoi(getenv("FEMO_TILE_SORT")) : 2);   // 0 position, 1 (type, position), 2 (type, n, position)
  if (!err.empty()) return femo_fail(err);
  const int hitem_cap = (P.max_hitems + 1) & ~1, hcell_cap = P.max_hcells;
  const int dest_cap = P.max_dests, ref_cap = P.max_refs;
  size_t smem = sizeof(double) * (size_t)tile_smem_layout(c->nq, hitem_cap, hcell_cap, dest_cap, ref_cap).total;
  if (smem > 227 * 1024) return 0;                      // halo too large for one CTA: older kernels take over
  const char* ews = getenv("FEMO_TILE_WS");             // FEMO_TILE_WS=0: two independent CTAs per SM instead of warp roles
  const int ws_req = ews ? atoi(ews) : 2;                 // 1: 6+6 warps, 2: 8+8 warps with setmaxnreg
  const int sk_rows = ws_req == 2 ? WsCfg<1>::WT * WsCfg<1>::DPT : WsCfg<0>::WT * WsCfg<0>::DPT;
  size_t smem_ws = sizeof(double) * (size_t)ws_smem_layout(c->nq, hitem_cap, hcell_cap, dest_cap, ref_cap, sk_rows, 0).total;
  c->tl_ws = smem_ws <= 227 * 1024 ? ws_req : 0;
  c->tl_hrec = 0;
  if (c->tl_ws) {
    // FEMO_TILE_HREC=1: stage the halo cells' records in shared memory (when they fit) instead of reading them from global
    // memory per halo item.  OFF by default: measured at cfg 3 (profiles/r02l_asm_experiments*.log) the items get 2.7x
    // cheaper, but the ~1 400 scattered 8-byte cp.async per chunk cost more LSU time than they save (3.55 against 3.39 ms).
    const char* eh = getenv("FEMO_TILE_HREC");
    const size_t with_h = sizeof(double) * (size_t)ws_smem_layout(c->nq, hitem_cap, hcell_cap, dest_cap, ref_cap, sk_rows, hcell_cap).total;
    if (eh && atoi(eh) == 1 && hcell_cap > 0 && with_h <= 227 * 1024) { c->tl_hrec = hcell_cap; smem_ws = with_h; }
    smem = smem_ws;
  }
  if (c->tl_hrec) {                                      // fixed-width halo-cell lists: row `chunk`, -1 padded
    std::vector<int32_t> fixed((size_t)P.nchunks * hcell_cap, -1);
    for (int64_t ch = 0; ch < P.nchunks; ++ch)
      for (int32_t k = P.hcell_ptr[ch]; k < P.hcell_ptr[ch + 1]; ++k) fixed[(size_t)ch * hcell_cap + (k - P.hcell_ptr[ch])] = P.hcells[k];
    TRY(upload(&c->tl_hfix, fixed));
  }
  TRY(upload(&c->tl_cell_order, P.cell_order)); TRY(upload(&c->tl_dest_ptr, P.dest_ptr));
  TRY(upload(&c->tl_dest_pos, P.dest_pos)); TRY(upload(&c->tl_dest_tpos, P.dest_tpos));
  TRY(upload(&c->tl_dest_info, P.dest_info)); TRY(upload(&c->tl_ref_ptr, P.ref_ptr)); TRY(upload(&c->tl_refs, P.refs));
  TRY(upload(&c->tl_hitem_ptr, P.hitem_ptr)); TRY(upload(&c->tl_hitems, P.hitems));
  TRY(upload(&c->tl_hcell_ptr, P.hcell_ptr)); TRY(upload(&c->tl_hcells, P.hcells));
  c->tl_nchunks = P.nchunks; c->tl_ndest = (int64_t)P.dest_pos.size(); c->tl_nref = (int64_t)P.refs.size();
  c->tl_nhitem = (int64_t)P.hitems.size();
  c->tl_hitem_cap = hitem_cap; c->tl_hcell_cap = hcell_cap; c->tl_dest_cap = dest_cap; c->tl_ref_cap = ref_cap;
  c->tl_smem = smem;
  TRY(dev_alloc(&c->tl_rec, (size_t)P.nchunks * c->nq * TL_NF * C));
  TRY(dev_alloc(&c->tl_frame, (size_t)P.nchunks * C * 9));
  CUDA_TRY(cudaFuncSetAttribute(k_asm_tiles, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  CUDA_TRY(cudaFuncSetAttribute(k_asm_tiles_ws<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  CUDA_TRY(cudaFuncSetAttribute(k_asm_tiles_ws<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  {
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, c->device));
    c->tl_sms = prop.multiProcessorCount;
  }
  if (getenv("FEMO_VERBOSE"))
    fprintf(stderr, "[femo] tile plan: %lld chunks of %d cells, %lld dests, %lld refs, %lld halo items (max/chunk %d) from <= %d halo cells (%s), smem %zu B\n",
            (long long)P.nchunks, C, (long long)c->tl_ndest, (long long)c->tl_nref, (long long)c->tl_nhitem, P.max_hitems, P.max_hcells,
            c->tl_hrec ? "records staged in shared memory" : "records read from global memory", smem);
  c->tile_ok = true;
  return 0;
}

// records + tiled assembly on c->stream; c->ev0 / c->ev1 bracket the records kernel (timed by the caller)
int assemble_jacobian_tiles(femo_ctx* c, const double* uh) {
  const int64_t nthreads = c->tl_nchunks * C;
  k_cell_records_tiles<<<(unsigned)cdiv(nthreads, 64), 64, 0, c->stream>>>(
      c->xyz, c->cv, uh, c->h, c->E, c->nu, c->hE, c->tab, c->tl_cell_order, c->tl_nchunks, c->nq, c->ewm, c->tl_rec,
      c->tl_frame);
  TRY(check_launch(c, "k_cell_records_tiles"));
  cudaEventRecord(c->ev1, c->stream);
  if (c->tl_ws) {
    const unsigned grid = (unsigned)std::min<int64_t>(c->tl_nchunks, c->tl_sms);
    const char* ed = getenv("FEMO_TILE_DEBUG");
    const int dbg = ed ? atoi(ed) : 0;
    if (c->tl_ws == 2)
      k_asm_tiles_ws<1><<<grid, WsCfg<1>::NT, c->tl_smem, c->stream>>>(
        c->tl_dest_ptr, c->tl_dest_pos, c->tl_dest_tpos, c->tl_dest_info, c->tl_ref_ptr, c->tl_refs, c->tl_hitem_ptr,
        c->tl_hitems, c->tl_hcell_ptr, c->tl_hcells, c->tl_rec, c->tl_frame, c->tab, c->nq, c->tl_hitem_cap,
        c->tl_hcell_cap, c->tl_dest_cap, c->tl_ref_cap, c->tl_nchunks, c->vals, dbg, c->tl_hrec, c->tl_hfix);
    else
      k_asm_tiles_ws<0><<<grid, WsCfg<0>::NT, c->tl_smem, c->stream>>>(
        c->tl_dest_ptr, c->tl_dest_pos, c->tl_dest_tpos, c->tl_dest_info, c->tl_ref_ptr, c->tl_refs, c->tl_hitem_ptr,
        c->tl_hitems, c->tl_hcell_ptr, c->tl_hcells, c->tl_rec, c->tl_frame, c->tab, c->nq, c->tl_hitem_cap,
        c->tl_hcell_cap, c->tl_dest_cap, c->tl_ref_cap, c->tl_nchunks, c->vals, dbg, c->tl_hrec, c->tl_hfix);
    return check_launch(c, "k_asm_tiles_ws");
  }
  k_asm_tiles<<<(unsigned)c->tl_nchunks, TL_THREADS, c->tl_smem, c->stream>>>(
      c->tl_dest_ptr, c->tl_dest_pos, c->tl_dest_tpos, c->tl_dest_info, c->tl_ref_ptr, c->tl_refs, c->tl_hitem_ptr,
      c->tl_hitems, c->tl_hcell_ptr, c->tl_hcells, c->tl_rec, c->tl_frame, c->tab, c->nq, c->tl_hitem_cap, c->tl_hcell_cap,
      c->tl_dest_cap, c->tl_ref_cap, c->vals);
  return check_launch(c, "k_asm_tiles");
}
