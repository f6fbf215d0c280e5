// BSR(3x3) SpMV, block-Jacobi preconditioner and PCG for the state and adjoint solves.
//
// Replaces PETSc KSP preonly+LU(MUMPS) (femo_alpha/fea/utils_dolfinx.py:438-468,514-531;
// fea/fea_dolfinx.py:173-203) and MatMult (utils_dolfinx.py:275-306).  A = dR/du with the
// dolfinx BC convention (rows+cols of clamped dofs zeroed, unit diagonal) is applied on the
// fly through `bcmask`, so only one matrix is stored.
//
// All reductions are two-stage with fixed grids and fixed summation order -> bitwise
// reproducible iterations (no floating-point atomics anywhere).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include "femo_internal.cuh"

#define RED_BLOCKS 1184          /* 148 SMs x 8 */
#define RED_THREADS 256

namespace {

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}

// deterministic CTA reduction of up to 3 values; result valid in thread 0
__device__ __forceinline__ void block_sum3(double& a, double& b, double& c) {
  __shared__ double sa[32], sb[32], sc[32];
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
  if (lane == 0) { sa[wid] = a; sb[wid] = b; sc[wid] = c; }
  __syncthreads();
  if (wid == 0) {
    int nw = (blockDim.x + 31) >> 5;
    a = lane < nw ? sa[lane] : 0.0; b = lane < nw ? sb[lane] : 0.0; c = lane < nw ? sc[lane] : 0.0;
    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
  }
  __syncthreads();
}

// every CTA sums the same partial array in the same order -> identical scalar everywhere
__device__ __forceinline__ double sum_partials(const double* __restrict__ part, int n) {
  __shared__ double s[32];
  __shared__ double result;
  double v = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) v += part[i];
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = warp_sum(v);
  if (lane == 0) s[wid] = v;
  __syncthreads();
  if (wid == 0) {
    int nw = (blockDim.x + 31) >> 5;
    v = lane < nw ? s[lane] : 0.0;
    v = warp_sum(v);
    if (lane == 0) result = v;
  }
  __syncthreads();
  double r = result;
  __syncthreads();
  return r;
}

// Column words: low 29 bits = block column, top 3 bits = clamp flags of its dofs (femo_ctx::colbc); with plain
// colidx (K x, no BC) the flags are 0.  `rbits` (per block row clamp flags) is null for K x.
#define COL_OF(w) ((int64_t)((w) & 0x1fffffffu))

// y = K x or y = A x.  LANES lanes cooperate on one block row, one block per lane and pass.
// Optionally accumulates the partial dot x.y per CTA (deterministic) into dot_part[blockIdx.x].
// VT: storage type of the block values -- double (the operator) or float (the copy the multigrid smoothers use: half the
// bytes of a bandwidth-bound kernel; products and sums are fp64 either way)
template <int LANES, int MINB, typename VT = double>
__global__ void __launch_bounds__(256, MINB)
k_spmv(const int32_t* __restrict__ rowptr, const uint32_t* __restrict__ colw, const VT* __restrict__ vals,
       const double* __restrict__ x, const uint8_t* __restrict__ rbits, int64_t nb, double* __restrict__ y,
       double* __restrict__ dot_part, const double* __restrict__ own, const int32_t* __restrict__ rows, int64_t row_base = 0) {
  const int lane = threadIdx.x % LANES;
  constexpr int GPW = 32 / LANES;                       // row groups per warp
  const int sub = (threadIdx.x & 31) / LANES;
  const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarp = (gridDim.x * (int64_t)blockDim.x) >> 5;
  double dacc = 0.0;
  // warp-uniform trip count: every lane takes part in the full-mask shuffles below
  for (int64_t rb = warp * GPW; rb < nb; rb += nwarp * GPW) {
    // `rows` (multi-GPU): the nb listed block rows (interior or boundary rows of the owned part); null: rows 0 .. nb-1
    const bool valid = rb + sub < nb;
    // row_base (rows == null): the nb rows row_base .. row_base + nb - 1 (row ranges of the pipelined residual, api.cu)
    const int64_t row = rows ? (valid ? (int64_t)rows[rb + sub] : 0) : row_base + rb + sub;
    const int32_t s = valid ? rowptr[row] : 0, e = valid ? rowptr[row + 1] : 0;
    double y0 = 0, y1 = 0, y2 = 0;
    for (int32_t k = s + lane; k < e; k += LANES) {
      const uint32_t w = colw[k];
      const VT* ap = vals + (int64_t)k * 9;
      double a[9];
#pragma unroll
      for (int i = 0; i < 9; ++i) a[i] = (double)ap[i];
      const double* xp = x + 3 * COL_OF(w);
      const double x0 = (w & 0x20000000u) ? 0.0 : xp[0], x1 = (w & 0x40000000u) ? 0.0 : xp[1], x2 = (w & 0x80000000u) ? 0.0 : xp[2];
      y0 += a[0] * x0 + a[1] * x1 + a[2] * x2;
      y1 += a[3] * x0 + a[4] * x1 + a[5] * x2;
      y2 += a[6] * x0 + a[7] * x1 + a[8] * x2;
    }
#pragma unroll
    for (int o = LANES / 2; o > 0; o >>= 1) {
      y0 += __shfl_down_sync(0xffffffffu, y0, o, LANES);
      y1 += __shfl_down_sync(0xffffffffu, y1, o, LANES);
      y2 += __shfl_down_sync(0xffffffffu, y2, o, LANES);
    }
    if (lane == 0 && valid) {
      const double xr0 = x[3 * row], xr1 = x[3 * row + 1], xr2 = x[3 * row + 2];
      if (rbits) {
        const unsigned b = rbits[row];
        if (b & 1u) y0 = xr0;
        if (b & 2u) y1 = xr1;
        if (b & 4u) y2 = xr2;
      }
      y[3 * row] = y0; y[3 * row + 1] = y1; y[3 * row + 2] = y2;
      const double ow = own ? own[row] : 1.0;
      dacc += ow * (xr0 * y0 + xr1 * y1 + xr2 * y2);
    }
  }
  if (dot_part) {
    double b = 0, c2 = 0;
    block_sum3(dacc, b, c2);
    if (threadIdx.x == 0) dot_part[blockIdx.x] = dacc;
  }
}

// y[row] += sign * sum_d pen_val[d] * x[col(d)]  (penalty part of dR/du, identity 3x3 blocks)
__global__ void k_penalty_apply(const int32_t* __restrict__ prow, const int32_t* __restrict__ prow_ptr,
                                const int32_t* __restrict__ dcol, const double* __restrict__ pval, int64_t nrows,
                                const double* __restrict__ x, double sign, double* __restrict__ y) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= nrows) return;
  double a0 = 0, a1 = 0, a2 = 0;
  for (int32_t d = prow_ptr[i]; d < prow_ptr[i + 1]; ++d) {
    int64_t cidx = dcol[d]; double v = pval[d];
    a0 += v * x[3 * cidx]; a1 += v * x[3 * cidx + 1]; a2 += v * x[3 * cidx + 2];
  }
  int64_t r = prow[i];
  y[3 * r] += sign * a0; y[3 * r + 1] += sign * a1; y[3 * r + 2] += sign * a2;
}

// in-place Gauss-Jordan inverse of an SPD n x n matrix (n = 3 or 6), no pivoting needed
template <int N>
__device__ __forceinline__ void spd_inverse(double* A) {
#pragma unroll
  for (int k = 0; k < N; ++k) {
    double ip = 1.0 / A[k * N + k];
    A[k * N + k] = 1.0;
#pragma unroll
    for (int j = 0; j < N; ++j) A[k * N + j] *= ip;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      if (i == k) continue;
      double f = A[i * N + k];
      A[i * N + k] = 0.0;
#pragma unroll
      for (int j = 0; j < N; ++j) A[i * N + j] -= f * A[k * N + j];
    }
  }
}

// block-Jacobi set-up: 6x6 (u_v, theta_v) at vertices, 3x3 at edge / centre nodes, of A (with BCs)
__global__ void k_precond_setup(const double* __restrict__ vals, const int32_t* __restrict__ diag_pos,
                                const int32_t* __restrict__ ut_pos, const int32_t* __restrict__ tu_pos,
                                const double* __restrict__ mask, const double* __restrict__ own, int64_t nn,
                                int64_t nq2, double* __restrict__ dinv6, double* __restrict__ dinv3) {
  int64_t I = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (I >= nq2) return;
  const double ow = own ? own[I] : 1.0;           // ghost nodes: rows are incomplete -> identity block
  if (I < nn) {
    double A[36];
    const double* uu = vals + (int64_t)diag_pos[I] * 9;
    const double* tt = vals + (int64_t)diag_pos[nq2 + I] * 9;
    const double* ut = vals + (int64_t)ut_pos[I] * 9;
    const double* tu = vals + (int64_t)tu_pos[I] * 9;
    double m[6];
    for (int i = 0; i < 3; ++i) { m[i] = ow * mask[3 * I + i]; m[3 + i] = ow * mask[3 * (nq2 + I) + i]; }
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        A[i * 6 + j] = uu[3 * i + j]; A[i * 6 + 3 + j] = ut[3 * i + j];
        A[(3 + i) * 6 + j] = tu[3 * i + j]; A[(3 + i) * 6 + 3 + j] = tt[3 * i + j];
      }
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 6; ++j) A[i * 6 + j] = m[i] * m[j] * A[i * 6 + j] + ((i == j) ? (1.0 - m[i]) : 0.0);
    spd_inverse<6>(A);
    for (int i = 0; i < 36; ++i) dinv6[I * 36 + i] = A[i];
  } else {
    double A[9];
    const double* uu = vals + (int64_t)diag_pos[I] * 9;
    double m[3] = {ow * mask[3 * I], ow * mask[3 * I + 1], ow * mask[3 * I + 2]};
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) A[3 * i + j] = m[i] * m[j] * uu[3 * i + j] + ((i == j) ? (1.0 - m[i]) : 0.0);
    spd_inverse<3>(A);
    for (int i = 0; i < 9; ++i) dinv3[(I - nn) * 9 + i] = A[i];
  }
}

// z = M^-1 r for the node owned by this thread
__device__ __forceinline__ void precond_node(int64_t I, int64_t nn, int64_t nq2, const double* __restrict__ dinv6,
                                             const double* __restrict__ dinv3, const double* rl, double* zl) {
  if (I < nn) {
    const double* D = dinv6 + I * 36;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      double s = 0;
#pragma unroll
      for (int j = 0; j < 6; ++j) s += D[i * 6 + j] * rl[j];
      zl[i] = s;
    }
  } else {
    const double* D = dinv3 + (I - nn) * 9;
#pragma unroll
    for (int i = 0; i < 3; ++i) zl[i] = D[3 * i] * rl[0] + D[3 * i + 1] * rl[1] + D[3 * i + 2] * rl[2];
  }
}

// PCG initialisation: r = b - Ax (Ax given), z = M^-1 r, p = z; partials of r.z, r.r, b.b
__global__ void __launch_bounds__(RED_THREADS)
k_pcg_init(const double* __restrict__ b, const double* __restrict__ Ax, const double* __restrict__ dinv6,
           const double* __restrict__ dinv3, int64_t nn, int64_t nq2, double* __restrict__ r, double* __restrict__ z,
           double* __restrict__ p, double* __restrict__ part_rz, double* __restrict__ part_rr,
           double* __restrict__ part_bb, const double* __restrict__ own) {
  double rz = 0, rr = 0, bb = 0;
  for (int64_t I = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; I < nq2; I += gridDim.x * (int64_t)blockDim.x) {
    int nd = I < nn ? 6 : 3;
    const double ow = own ? own[I] : 1.0;
    double rl[6], zl[6];
    for (int i = 0; i < nd; ++i) {
      int64_t d = (i < 3) ? 3 * I + i : 3 * (nq2 + I) + (i - 3);
      double bv = b[d];
      rl[i] = bv - Ax[d];
      bb += ow * bv * bv;
    }
    precond_node(I, nn, nq2, dinv6, dinv3, rl, zl);
    for (int i = 0; i < nd; ++i) {
      int64_t d = (i < 3) ? 3 * I + i : 3 * (nq2 + I) + (i - 3);
      r[d] = rl[i]; z[d] = zl[i]; p[d] = zl[i];
      rz += ow * rl[i] * zl[i]; rr += ow * rl[i] * rl[i];
    }
  }
  block_sum3(rz, rr, bb);
  if (threadIdx.x == 0) { part_rz[blockIdx.x] = rz; part_rr[blockIdx.x] = rr; part_bb[blockIdx.x] = bb; }
}

// alpha = rz_old / pAp; x += alpha p; r -= alpha Ap; z = M^-1 r; partials of r.z and r.r
__global__ void __launch_bounds__(RED_THREADS)
k_pcg_update(const double* __restrict__ part_pAp, int n_pAp, const double* __restrict__ part_rz_old, int n_rz_old,
             const double* __restrict__ own, const double* __restrict__ p, const double* __restrict__ Ap, const double* __restrict__ dinv6,
             const double* __restrict__ dinv3, int64_t nn, int64_t nq2, double* __restrict__ x,
             double* __restrict__ r, double* __restrict__ z, double* __restrict__ part_rz_new,
             double* __restrict__ part_rr) {
  const double pAp = sum_partials(part_pAp, n_pAp);
  const double rz_old = sum_partials(part_rz_old, n_rz_old);
  const double alpha = (pAp != 0.0) ? rz_old / pAp : 0.0;
  double rz = 0, rr = 0, dummy = 0;
  for (int64_t I = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; I < nq2; I += gridDim.x * (int64_t)blockDim.x) {
    int nd = I < nn ? 6 : 3;
    const double ow = own ? own[I] : 1.0;
    double rl[6], zl[6];
    for (int i = 0; i < nd; ++i) {
      int64_t d = (i < 3) ? 3 * I + i : 3 * (nq2 + I) + (i - 3);
      x[d] += alpha * p[d];
      rl[i] = r[d] - alpha * Ap[d];
    }
    precond_node(I, nn, nq2, dinv6, dinv3, rl, zl);
    for (int i = 0; i < nd; ++i) {
      int64_t d = (i < 3) ? 3 * I + i : 3 * (nq2 + I) + (i - 3);
      r[d] = rl[i]; z[d] = zl[i];
      rz += ow * rl[i] * zl[i]; rr += ow * rl[i] * rl[i];
    }
  }
  block_sum3(rz, rr, dummy);
  if (threadIdx.x == 0) { part_rz_new[blockIdx.x] = rz; part_rr[blockIdx.x] = rr; }
}

// beta = rz_new / rz_old; p = z + beta p; CTA 0 publishes r.r
__global__ void __launch_bounds__(RED_THREADS)
k_pcg_dir(const double* __restrict__ part_rz_new, const double* __restrict__ part_rz_old,
          const double* __restrict__ part_rr, int n_part, const double* __restrict__ z, int64_t n,
          double* __restrict__ p, double* __restrict__ rr_out) {
  const double rzn = sum_partials(part_rz_new, n_part);
  const double rzo = sum_partials(part_rz_old, n_part);
  const double beta = (rzo != 0.0) ? rzn / rzo : 0.0;
  if (blockIdx.x == 0) {
    double rr = sum_partials(part_rr, n_part);
    if (threadIdx.x == 0) *rr_out = rr;
  }
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += gridDim.x * (int64_t)blockDim.x)
    p[i] = z[i] + beta * p[i];
}

__global__ void __launch_bounds__(RED_THREADS)
k_dot(const double* __restrict__ a, const double* __restrict__ b, int64_t n, double* __restrict__ part) {
  double s = 0, u = 0, v = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += gridDim.x * (int64_t)blockDim.x) s += a[i] * b[i];
  block_sum3(s, u, v);
  if (threadIdx.x == 0) part[blockIdx.x] = s;
}

// owner-masked dot (multi-GPU): own is per block node
__global__ void __launch_bounds__(RED_THREADS)
k_dot_own(const double* __restrict__ a, const double* __restrict__ b, const double* __restrict__ own, int64_t nb,
          double* __restrict__ part) {
  double s = 0, u = 0, v = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nb; i += gridDim.x * (int64_t)blockDim.x)
    s += own[i] * (a[3 * i] * b[3 * i] + a[3 * i + 1] * b[3 * i + 1] + a[3 * i + 2] * b[3 * i + 2]);
  block_sum3(s, u, v);
  if (threadIdx.x == 0) part[blockIdx.x] = s;
}

// CTA k reduces partial array k (up to 3 arrays) into out[k]
__global__ void k_sum_partials3(const double* __restrict__ p0, const double* __restrict__ p1,
                                const double* __restrict__ p2, int n, double* __restrict__ out) {
  const double* part = blockIdx.x == 0 ? p0 : (blockIdx.x == 1 ? p1 : p2);
  double v = sum_partials(part, n);
  if (threadIdx.x == 0) out[blockIdx.x] = v;
}

__global__ void k_sum_partials(const double* __restrict__ part, int n, double* __restrict__ out) {
  double v = sum_partials(part, n);
  if (threadIdx.x == 0) *out = v;
}

// ---- single-reduction PCG (Chronopoulos & Gear) for multi-GPU contexts ----------------------------------------------------
// With u = M^-1 r, w = A u, gamma = (r, u), delta = (w, u):  beta = gamma / gamma_old,
// alpha = gamma / (delta - beta gamma / alpha_old),  p = u + beta p,  s = w + beta s,  x += alpha p,  r -= alpha s.
// Both inner products are available right after the SpMV, so ONE all-reduce per iteration carries (gamma, delta, r.r).
// S = (gamma, delta, rr) all-reduced; Oin = (gamma_old, alpha_old) of the previous iteration, Oout receives this one's.
__global__ void __launch_bounds__(RED_THREADS)
k_cg_update(const double* __restrict__ S, const double* __restrict__ Oin, double* __restrict__ Oout, int first,
            const double* __restrict__ own, const double* __restrict__ w, const double* __restrict__ dinv6,
            const double* __restrict__ dinv3, int64_t nn, int64_t nq2, double* __restrict__ u, double* __restrict__ p,
            double* __restrict__ s, double* __restrict__ x, double* __restrict__ r, double* __restrict__ part_ru,
            double* __restrict__ part_rr) {
  const double gamma = S[0], delta = S[1];
  double beta = 0.0, alpha;
  if (first || Oin[0] == 0.0 || Oin[1] == 0.0) alpha = (delta != 0.0) ? gamma / delta : 0.0;
  else {
    beta = gamma / Oin[0];
    const double den = delta - beta * gamma / Oin[1];
    alpha = (den != 0.0) ? gamma / den : 0.0;
  }
  const bool use_old = beta != 0.0;
  if (blockIdx.x == 0 && threadIdx.x == 0) { Oout[0] = gamma; Oout[1] = alpha; }
  double ru = 0, rr = 0, dummy = 0;
  for (int64_t I = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; I < nq2; I += gridDim.x * (int64_t)blockDim.x) {
    const int nd = I < nn ? 6 : 3;
    const double ow = own ? own[I] : 1.0;
    double rl[6], zl[6];
    for (int i = 0; i < nd; ++i) {
      const int64_t d = (i < 3) ? 3 * I + i : 3 * (nq2 + I) + (i - 3);
      const double pn = use_old ? u[d] + beta * p[d] : u[d];
      const double sn = use_old ? w[d] + beta * s[d] : w[d];
      p[d] = pn; s[d] = sn;
      x[d] += alpha * pn;
      rl[i] = r[d] - alpha * sn;
    }
    precond_node(I, nn, nq2, dinv6, dinv3, rl, zl);
    for (int i = 0; i < nd; ++i) {
      const int64_t d = (i < 3) ? 3 * I + i : 3 * (nq2 + I) + (i - 3);
      r[d] = rl[i]; u[d] = zl[i];
      ru += ow * rl[i] * zl[i]; rr += ow * rl[i] * rl[i];
    }
  }
  block_sum3(ru, rr, dummy);
  if (threadIdx.x == 0) { part_ru[blockIdx.x] = ru; part_rr[blockIdx.x] = rr; }
}

// CTA 0: out[0] = sum part_ru; CTA 1: out[1] = sum dot_int + sum dot_bnd; CTA 2: out[2] = sum part_rr; CTA 3: out[3] = sum part_bb
__global__ void k_cg_sum(const double* __restrict__ part_ru, const double* __restrict__ part_rr, const double* __restrict__ part_bb,
                         int n_red, const double* __restrict__ dot_int, int n_int, const double* __restrict__ dot_bnd, int n_bnd,
                         double* __restrict__ out) {
  double v;
  if (blockIdx.x == 0) v = sum_partials(part_ru, n_red);
  else if (blockIdx.x == 1) { v = sum_partials(dot_int, n_int); if (n_bnd > 0) v += sum_partials(dot_bnd, n_bnd); }
  else if (blockIdx.x == 2) v = sum_partials(part_rr, n_red);
  else v = sum_partials(part_bb, n_red);
  if (threadIdx.x == 0) out[blockIdx.x] = v;
}

// ---- roofline denominators -------------------------------------------------------------
__global__ void k_fma_peak(double* out, int iters) {
  double a0 = threadIdx.x * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double m = 0.999999, c = 1e-6;
  for (int i = 0; i < iters; ++i) {
    a0 = a0 * m + c; a1 = a1 * m + c; a2 = a2 * m + c; a3 = a3 * m + c;
    a4 = a4 * m + c; a5 = a5 * m + c; a6 = a6 * m + c; a7 = a7 * m + c;
  }
  out[blockIdx.x * (int64_t)blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
// fp64 tensor-core peak: 4 independent mma.sync m8n8k4 f64 chains per warp (512 flop per mma)
__global__ void k_dmma_peak(double* out, int iters) {
  double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  double c0[2] = {0, 0}, c1[2] = {0, 0}, c2[2] = {0, 0}, c3[2] = {0, 0};
  for (int i = 0; i < iters; ++i) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0[0]), "+d"(c0[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c1[0]), "+d"(c1[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c2[0]), "+d"(c2[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c3[0]), "+d"(c3[1]) : "d"(a), "d"(b));
  }
  out[blockIdx.x * (int64_t)blockDim.x + threadIdx.x] = c0[0] + c0[1] + c1[0] + c1[1] + c2[0] + c2[1] + c3[0] + c3[1];
}
__global__ void k_copy(const double4* __restrict__ a, double4* __restrict__ b, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += gridDim.x * (int64_t)blockDim.x) b[i] = a[i];
}

}  // namespace

#define SPMV_BLOCKS (148 * 8)   /* stride of the partial arrays in `red` (>= spmv_grid(), RED_BLOCKS) */
// 8 lanes per block row, 6 CTAs of 256 threads per SM in one wave (40 registers).  Measured at cfg 3 (ncu, one launch):
// 16 lanes + fp64 mask gathers 1.76 ms -> packed column/flag words, single wave 1.38 ms -> 8 lanes ~1.2 ms;
// a 27-lane coalesced layout (2.27 ms) and a cp.async-staged pipeline (2.3 - 3.1 ms) lose to the memory-level
// parallelism of one block (12 independent loads) per lane.
static int spmv_grid() { return 148 * 6; }
static int launch_spmv(femo_ctx* c, const double* x, bool with_bc, double* y, double* dot_part, const double* own) {
  const uint32_t* colw = with_bc ? c->colbc : reinterpret_cast<const uint32_t*>(c->colidx);
  const uint8_t* rbits = with_bc ? c->bcbits : nullptr;
  k_spmv<8, 6><<<spmv_grid(), 256, 0, c->stream>>>(c->rowptr, colw, c->vals, x, rbits, c->nb, y, dot_part, own, nullptr);
  return 0;
}
// the listed rows only (A with BCs); grid sized to the list (boundary lists are short)
static int launch_spmv_rows(femo_ctx* c, const double* x, double* y, const int32_t* rows, int64_t nrows, double* dot_part, int* grid_out) {
  int grid = (int)std::min<int64_t>(spmv_grid(), std::max<int64_t>(1, cdiv(nrows, 32)));     // 32 block rows per CTA and pass
  if (grid_out) *grid_out = grid;
  if (nrows > 0 || dot_part)
    k_spmv<8, 6><<<grid, 256, 0, c->stream>>>(c->rowptr, c->colbc, c->vals, x, c->bcbits, nrows, y, dot_part, c->own, rows);
  return 0;
}

// y = A x on a multi-GPU context: the halo exchange of x (pack -> ncclSend/Recv -> unpack on the context's communication
// stream) runs while the interior rows -- owned rows without a ghost column -- are multiplied; the boundary rows follow
// once the ghosts have arrived.  Ghost rows are not computed (their y entries are never used: every consumer is
// owner-masked).  dot_int / dot_bnd (optional): per-CTA partials of the owner-masked x.y of the two launches; the
// numbers of partials written are returned in n_int / n_bnd.
int dist_spmv(femo_ctx* c, double* x, double* y, double* dot_int, double* dot_bnd, int* n_int, int* n_bnd) {
  if (!c->jac_valid) return femo_fail("Jacobian not assembled");
  if (c->world <= 1 || !c->rows_int) {     // single GPU, or no row split (overlap switched off): exchange, then all rows
    if (c->world > 1) TRY(halo_exchange(c, x));
    TRY(launch_spmv(c, x, true, y, dot_int, c->world > 1 ? c->own : nullptr));
    if (n_int) *n_int = spmv_grid();
    if (n_bnd) *n_bnd = 0;
    c->launches += 1;
    return check_launch(c, "k_spmv");
  }
  // trace (FEMO_TRACE_OVERLAP=1): [0] start, [1] halo done (communication stream), [2] interior rows done, [3] boundary rows
  // done (context's stream); read back by femo_dist_overlap_info
  const bool tr = c->trace_overlap != 0;
  if (tr) cudaEventRecord(c->ev_trace[0], c->stream);
  TRY(halo_exchange_begin(c, x));
  if (tr) cudaEventRecord(c->ev_trace[1], c->comm_stream);
  int g0 = 0, g1 = 0;
  TRY(launch_spmv_rows(c, x, y, c->rows_int, c->n_rows_int, dot_int, &g0));
  if (tr) cudaEventRecord(c->ev_trace[2], c->stream);
  TRY(halo_exchange_end(c));
  TRY(launch_spmv_rows(c, x, y, c->rows_bnd, c->n_rows_bnd, dot_bnd, &g1));
  if (tr) cudaEventRecord(c->ev_trace[3], c->stream);
  if (n_int) *n_int = g0;
  if (n_bnd) *n_bnd = g1;
  c->launches += 2;
  return check_launch(c, "k_spmv (interior / boundary rows)");
}
int bsr_spmv(femo_ctx* c, const double* x, double* y, int with_bc) {
  if (!c->jac_valid) return femo_fail("Jacobian not assembled");
  TRY(launch_spmv(c, x, with_bc != 0, y, nullptr, nullptr));
  TRY(check_launch(c, "k_spmv"));
  c->launches += 1;
  return 0;
}

// y = (vals) x with the pattern of this context (K while the dynamic operator A_eff occupies c->vals)
int bsr_spmv_vals(femo_ctx* c, const double* vals, const double* x, double* y, int with_bc) {
  const uint32_t* colw = with_bc ? c->colbc : reinterpret_cast<const uint32_t*>(c->colidx);
  const uint8_t* rbits = with_bc ? c->bcbits : nullptr;
  k_spmv<8, 6><<<spmv_grid(), 256, 0, c->stream>>>(c->rowptr, colw, vals, x, rbits, c->nb, y, nullptr, nullptr, nullptr);
  TRY(check_launch(c, "k_spmv"));
  c->launches += 1;
  return 0;
}

// ---- fp32 copy of the block values (multigrid smoothers, mg.cu) ---------------------------------------------------------------
__global__ void k_vals_to_float(const double* __restrict__ v, int64_t n, float* __restrict__ o) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += gridDim.x * (int64_t)blockDim.x) o[i] = (float)v[i];
}
int vals_to_float(femo_ctx* c, float* out) {
  k_vals_to_float<<<148 * 8, 256, 0, c->stream>>>(c->vals, c->nnzb * 9, out);
  c->launches += 1;
  return check_launch(c, "k_vals_to_float");
}
// y = A32 x (BCs applied) with the context's pattern; multi-GPU contexts: halo exchange overlapped with the interior rows
// exactly like dist_spmv
int bsr_spmv_f32(femo_ctx* c, const float* vals32, double* x, double* y) {
  if (c->world <= 1 || !c->rows_int) {
    if (c->world > 1) TRY(halo_exchange(c, x));
    k_spmv<8, 6, float><<<spmv_grid(), 256, 0, c->stream>>>(c->rowptr, c->colbc, vals32, x, c->bcbits, c->nb, y, nullptr, nullptr, nullptr);
    c->launches += 1;
    return check_launch(c, "k_spmv<float>");
  }
  TRY(halo_exchange_begin(c, x));
  const int gi = (int)std::min<int64_t>(spmv_grid(), std::max<int64_t>(1, cdiv(c->n_rows_int, 32)));
  const int gb = (int)std::min<int64_t>(spmv_grid(), std::max<int64_t>(1, cdiv(c->n_rows_bnd, 32)));
  if (c->n_rows_int > 0)
    k_spmv<8, 6, float><<<gi, 256, 0, c->stream>>>(c->rowptr, c->colbc, vals32, x, c->bcbits, c->n_rows_int, y, nullptr, c->own, c->rows_int);
  TRY(halo_exchange_end(c));
  if (c->n_rows_bnd > 0)
    k_spmv<8, 6, float><<<gb, 256, 0, c->stream>>>(c->rowptr, c->colbc, vals32, x, c->bcbits, c->n_rows_bnd, y, nullptr, c->own, c->rows_bnd);
  c->launches += 2;
  return check_launch(c, "k_spmv<float> (interior / boundary rows)");
}

// y[rows r0 .. r1) = K x (no BCs), the other rows of y untouched: row ranges of the pipelined residual (femo_evaluate_residual)
int bsr_spmv_range(femo_ctx* c, const double* x, double* y, int64_t r0, int64_t r1) {
  if (!c->jac_valid) return femo_fail("Jacobian not assembled");
  if (r1 <= r0) return 0;
  const int grid = (int)std::min<int64_t>(spmv_grid(), std::max<int64_t>(1, cdiv(r1 - r0, 32)));
  k_spmv<8, 6><<<grid, 256, 0, c->stream>>>(c->rowptr, reinterpret_cast<const uint32_t*>(c->colidx), c->vals, x, nullptr, r1 - r0, y, nullptr,
                                            nullptr, nullptr, r0);
  c->launches += 1;
  return check_launch(c, "k_spmv (row range)");
}

int penalty_apply(femo_ctx* c, const double* x, double* y, double sign) {
  if (c->n_pen_rows == 0) return 0;
  k_penalty_apply<<<(unsigned)cdiv(c->n_pen_rows, 128), 128, 0, c->stream>>>(c->pen_row, c->pen_row_ptr, c->pen_dest_col,
                                                                           c->pen_val, c->n_pen_rows, x, sign, y);
  TRY(check_launch(c, "k_penalty_apply"));
  c->launches += 1;
  return 0;
}

int precond_setup(femo_ctx* c) {
  if (!c->jac_valid) return femo_fail("Jacobian not assembled");
  k_precond_setup<<<(unsigned)cdiv(c->nq2, 128), 128, 0, c->stream>>>(c->vals, c->diag_pos, c->ut_pos, c->tu_pos, c->bcmask,
                                                                     c->own, c->nn, c->nq2, c->dinv6, c->dinv3);
  TRY(check_launch(c, "k_precond_setup"));
  c->launches += 1;
  c->prec_valid = true;
  return 0;
}

int dev_dot(femo_ctx* c, const double* a, const double* b, int64_t n, double* out) {
  double* part = c->red + 6 * RED_BLOCKS;
  const bool dist = c->world > 1 && c->own && n == c->ndof;     // state-sized vectors: owner-masked, summed over ranks
  if (dist) k_dot_own<<<RED_BLOCKS, RED_THREADS, 0, c->stream>>>(a, b, c->own, c->nb, part);
  else k_dot<<<RED_BLOCKS, RED_THREADS, 0, c->stream>>>(a, b, n, part);
  k_sum_partials<<<1, RED_THREADS, 0, c->stream>>>(part, RED_BLOCKS, c->scal + 8);
  TRY(check_launch(c, "k_dot"));
  c->launches += 2;
  if (dist) TRY(dist_allreduce_sum(c, c->scal + 8, 1));
  CUDA_TRY(cudaMemcpyAsync(out, c->scal + 8, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

// out_dev[0] = a . b without host synchronisation (single GPU; capturable in a CUDA graph)
int dev_dot_async(femo_ctx* c, const double* a, const double* b, int64_t n, double* out_dev) {
  double* part = c->red + 6 * RED_BLOCKS;
  k_dot<<<RED_BLOCKS, RED_THREADS, 0, c->stream>>>(a, b, n, part);
  k_sum_partials<<<1, RED_THREADS, 0, c->stream>>>(part, RED_BLOCKS, out_dev);
  TRY(check_launch(c, "k_dot"));
  c->launches += 2;
  return 0;
}

// Multi-GPU PCG: Chronopoulos-Gear recurrences (one all-reduce of three doubles per iteration) over the overlapped SpMV
// (halo exchange hidden behind the interior rows).  Per iteration: k_cg_update, [pack, send/recv, unpack || interior
// rows], boundary rows, k_cg_sum, ncclAllReduce.  FEMO_PCG_CLASSIC=1 selects the two-reduction loop of round 1.
static int pcg_solve_dist(femo_ctx* c, const double* b, double* x, double rtol, int max_it, int* iters, double* relres,
                          bool fixed_iters) {
  double* D_int = c->red;
  double* D_bnd = c->red + 5 * SPMV_BLOCKS;
  double* P_ru = c->red + 1 * SPMV_BLOCKS;
  double* P_rr = c->red + 3 * SPMV_BLOCKS;
  double* P_bb = c->red + 4 * SPMV_BLOCKS;
  double* S = c->scal + 0;                       // gamma, delta, rr, bb
  double* O[2] = {c->scal + 24, c->scal + 26};   // (gamma_old, alpha_old) ping-pong
  double* u = c->z; double* w = c->Ap; double* p = c->p; double* sv = c->tmp;
  const int64_t n = c->ndof;
  cudaEventRecord(c->ev0, c->stream);
  int nI = 0, nB = 0;
  TRY(dist_spmv(c, x, c->Ap, nullptr, nullptr, nullptr, nullptr));                 // r0 = b - A x0
  k_pcg_init<<<RED_BLOCKS, RED_THREADS, 0, c->stream>>>(b, c->Ap, c->dinv6, c->dinv3, c->nn, c->nq2, c->r, u, p, P_ru, P_rr, P_bb, c->own);
  TRY(dist_spmv(c, u, w, D_int, D_bnd, &nI, &nB));
  k_cg_sum<<<4, RED_THREADS, 0, c->stream>>>(P_ru, P_rr, P_bb, RED_BLOCKS, D_int, nI, D_bnd, nB, S);
  TRY(check_launch(c, "pcg init"));
  c->launches += 2;
  TRY(dist_allreduce_sum(c, S, 4));
  double hs[4];
  CUDA_TRY(cudaMemcpyAsync(hs, S, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  const double bb = hs[3];
  double rr = hs[2];
  const double bnorm = sqrt(bb);
  int it = 0;
  double rel = bnorm > 0 ? sqrt(rr) / bnorm : 0.0;
  if (bnorm == 0.0 && !fixed_iters) {
    CUDA_TRY(cudaMemsetAsync(x, 0, sizeof(double) * n, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    *iters = 0; *relres = 0.0; c->t_solve = 0; c->t_iter = 0;
    return 0;
  }
  const int check_every = 25;
  bool done = (!fixed_iters && rel <= rtol);
  while (!done && it < max_it) {
    const int nloop = std::min(check_every, max_it - it);
    for (int k = 0; k < nloop; ++k, ++it) {
      k_cg_update<<<RED_BLOCKS, RED_THREADS, 0, c->stream>>>(S, O[it & 1], O[(it + 1) & 1], it == 0 ? 1 : 0, c->own, w, c->dinv6,
                                                             c->dinv3, c->nn, c->nq2, u, p, sv, x, c->r, P_ru, P_rr);
      TRY(dist_spmv(c, u, w, D_int, D_bnd, &nI, &nB));
      k_cg_sum<<<3, RED_THREADS, 0, c->stream>>>(P_ru, P_rr, P_bb, RED_BLOCKS, D_int, nI, D_bnd, nB, S);
      TRY(dist_allreduce_sum(c, S, 3));
    }
    c->launches += 2 * (int64_t)nloop;
    TRY(check_launch(c, "pcg iteration"));
    CUDA_TRY(cudaMemcpyAsync(&rr, S + 2, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    rel = sqrt(rr) / bnorm;
    if (!(rel == rel)) return femo_fail("pcg: NaN residual (matrix not SPD?)");
    if (!fixed_iters && rel <= rtol) done = true;
  }
  TRY(halo_exchange(c, x));                  // ghost entries of the solution follow their owners
  cudaEventRecord(c->ev1, c->stream);
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev0, c->ev1);
  c->t_solve = ms; c->t_iter = it > 0 ? ms / it : 0.0;
  *iters = it; *relres = rel;
  return 0;
}

// red layout: [0]=pAp partials (SPMV_BLOCKS) | [1],[2] = rz ping-pong | [3] = rr | [4] = bb | [5] spare | [6] dot
// scal layout: [0]=bb [1]=rr (host reads) [2]=pAp | [10..11],[12..13] = (rz, rr) ping-pong pairs (multi-GPU)
//
// Multi-GPU (world > 1): ghost entries of the operand are refreshed before every SpMV (halo exchange),
// the dot-product partials are owner-masked, reduced to scalars on the device and summed over ranks with one
// ncclAllReduce per reduction point (pAp; (rz, rr)); the consumer kernels then read the scalar (n = 1).
int pcg_solve(femo_ctx* c, const double* b, double* x, double rtol, int max_it, int* iters, double* relres,
              bool fixed_iters) {
  if (!c->jac_valid) return femo_fail("pcg: Jacobian not assembled");
  if (!c->prec_valid) TRY(precond_setup(c));
  const bool dist = c->world > 1;
  if (dist && !c->nccl_comm) return femo_fail("pcg: multi-GPU context without communicator (femo_dist_init)");
  if (dist) {
    const char* env = getenv("FEMO_PCG_CLASSIC");
    if (!(env && env[0] == '1')) return pcg_solve_dist(c, b, x, rtol, max_it, iters, relres, fixed_iters);
  }
  double* P_pAp = c->red;
  double* P_rz[2] = {c->red + 1 * SPMV_BLOCKS, c->red + 2 * SPMV_BLOCKS};
  double* P_rr = c->red + 3 * SPMV_BLOCKS;
  double* P_bb = c->red + 4 * SPMV_BLOCKS;
  double* S_pair[2] = {c->scal + 10, c->scal + 12};
  const int64_t n = c->ndof;
  cudaEventRecord(c->ev0, c->stream);
  // r0 = b - A x0
  if (dist) TRY(halo_exchange(c, x));
  TRY(launch_spmv(c, x, true, c->Ap, nullptr, nullptr));
  k_pcg_init<<<RED_BLOCKS, RED_THREADS, 0, c->stream>>>(b, c->Ap, c->dinv6, c->dinv3, c->nn, c->nq2, c->r, c->z, c->p,
                                                        P_rz[0], P_rr, P_bb, c->own);
  if (dist) {
    // scal[0..2] <- (bb, rr, rz) summed over ranks; pair[0] = (rz, rr)
    k_sum_partials3<<<3, RED_THREADS, 0, c->stream>>>(P_bb, P_rr, P_rz[0], RED_BLOCKS, c->scal + 0);
    TRY(dist_allreduce_sum(c, c->scal + 0, 3));
    CUDA_TRY(cudaMemcpyAsync(S_pair[0], c->scal + 2, sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
    CUDA_TRY(cudaMemcpyAsync(S_pair[0] + 1, c->scal + 1, sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  } else {
    k_sum_partials<<<1, RED_THREADS, 0, c->stream>>>(P_bb, RED_BLOCKS, c->scal + 0);
    k_sum_partials<<<1, RED_THREADS, 0, c->stream>>>(P_rr, RED_BLOCKS, c->scal + 1);
  }
  TRY(check_launch(c, "pcg init"));
  c->launches += 4;
  double hs[2];
  CUDA_TRY(cudaMemcpyAsync(hs, c->scal, 2 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  const double bb = hs[0];
  double rr = hs[1];
  const double bnorm = sqrt(bb);
  int it = 0;
  double rel = bnorm > 0 ? sqrt(rr) / bnorm : 0.0;
  if (bnorm == 0.0 && !fixed_iters) {
    CUDA_TRY(cudaMemsetAsync(x, 0, sizeof(double) * n, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    *iters = 0; *relres = 0.0; c->t_solve = 0; c->t_iter = 0;
    return 0;
  }
  const int check_every = 25;
  bool done = (!fixed_iters && rel <= rtol);
  while (!done && it < max_it) {
    int nloop = std::min(check_every, max_it - it);
    for (int k = 0; k < nloop; ++k, ++it) {
      int cur = it & 1, nxt = cur ^ 1;
      if (dist) {
        TRY(halo_exchange(c, c->p));
        TRY(launch_spmv(c, c->p, true, c->Ap, P_pAp, c->own));
        k_sum_partials<<<1, RED_THREADS, 0, c->stream>>>(P_pAp, spmv_grid(), c->scal + 2);
        TRY(dist_allreduce_sum(c, c->scal + 2, 1));
        k_pcg_update<<<RED_BLOCKS, RED_THREADS, 0, c->stream>>>(c->scal + 2, 1, S_pair[cur], 1, c->own, c->p, c->Ap, c->dinv6,
                                                                c->dinv3, c->nn, c->nq2, x, c->r, c->z, P_rz[nxt], P_rr);
        k_sum_partials3<<<2, RED_THREADS, 0, c->stream>>>(P_rz[nxt], P_rr, P_rr, RED_BLOCKS, S_pair[nxt]);
        TRY(dist_allreduce_sum(c, S_pair[nxt], 2));
        k_pcg_dir<<<RED_BLOCKS, RED_THREADS, 0, c->stream>>>(S_pair[nxt], S_pair[cur], S_pair[nxt] + 1, 1, c->z, n, c->p,
                                                             c->scal + 1);
        c->launches += 2;
      } else {
        TRY(launch_spmv(c, c->p, true, c->Ap, P_pAp, nullptr));
        k_pcg_update<<<RED_BLOCKS, RED_THREADS, 0, c->stream>>>(P_pAp, spmv_grid(), P_rz[cur], RED_BLOCKS, nullptr, c->p, c->Ap,
                                                                c->dinv6, c->dinv3, c->nn, c->nq2, x, c->r, c->z, P_rz[nxt], P_rr);
        k_pcg_dir<<<RED_BLOCKS, RED_THREADS, 0, c->stream>>>(P_rz[nxt], P_rz[cur], P_rr, RED_BLOCKS, c->z, n, c->p, c->scal + 1);
      }
    }
    c->launches += 3 * (int64_t)nloop;
    TRY(check_launch(c, "pcg iteration"));
    CUDA_TRY(cudaMemcpyAsync(&rr, c->scal + 1, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    rel = sqrt(rr) / bnorm;
    if (!(rel == rel)) return femo_fail("pcg: NaN residual (matrix not SPD?)");
    if (!fixed_iters && rel <= rtol) done = true;
  }
  if (dist) TRY(halo_exchange(c, x));      // ghost entries of the solution follow their owners
  cudaEventRecord(c->ev1, c->stream);
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev0, c->ev1);
  c->t_solve = ms; c->t_iter = it > 0 ? ms / it : 0.0;
  *iters = it; *relres = rel;
  return 0;
}

int measure_peaks(femo_ctx* c, double* tf, double* gbs) {
  // fp64 FMA chain: 8 independent chains/thread, 1024 threads x 148*4 CTAs
  const int iters = 1 << 14, blocks = 148 * 8, threads = 512;
  double* out = nullptr;
  TRY(dev_alloc(&out, (size_t)blocks * threads));
  k_fma_peak<<<blocks, threads, 0, c->stream>>>(out, 64);
  cudaEventRecord(c->ev0, c->stream);
  k_fma_peak<<<blocks, threads, 0, c->stream>>>(out, iters);
  cudaEventRecord(c->ev1, c->stream);
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  float ms = 0; cudaEventElapsedTime(&ms, c->ev0, c->ev1);
  *tf = 2.0 * 8.0 * iters * (double)blocks * threads / (ms * 1e-3) / 1e12;
  if (c->dmma_tf >= 0) {   // also measure the fp64 tensor-core (DMMA) rate
    k_dmma_peak<<<blocks, threads, 0, c->stream>>>(out, 64);
    cudaEventRecord(c->ev0, c->stream);
    k_dmma_peak<<<blocks, threads, 0, c->stream>>>(out, 1 << 12);
    cudaEventRecord(c->ev1, c->stream);
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    c->dmma_tf = 4.0 * 512.0 * (1 << 12) * (double)blocks * (threads / 32) / (ms * 1e-3) / 1e12;
  }
  dev_free(&out);
  const int64_t nbytes = (int64_t)1 << 30;
  double4 *a = nullptr, *b = nullptr;
  TRY(dev_alloc(&a, nbytes / 32)); TRY(dev_alloc(&b, nbytes / 32));
  CUDA_TRY(cudaMemsetAsync(a, 0, nbytes, c->stream));
  k_copy<<<148 * 16, 256, 0, c->stream>>>(a, b, nbytes / 32);
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(c->ev0, c->stream);
    k_copy<<<148 * 16, 256, 0, c->stream>>>(a, b, nbytes / 32);
    cudaEventRecord(c->ev1, c->stream);
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    best = ms < best ? ms : best;
  }
  *gbs = 2.0 * nbytes / (best * 1e-3) / 1e9;
  dev_free(&a); dev_free(&b);
  c->launches += 8;
  return 0;
}
