// Sparse direct solver of libfemo_b200.so:  A x = b  with A = dR/du (BCs applied), by block-tridiagonal factorisation.
//
// Replaces the MUMPS LU the reference puts behind every solve (femo_alpha/fea/utils_dolfinx.py:438-468 NewtonSolver
// "preonly + lu", :514-531 setUpKSP_MUMPS; fea/fea_dolfinx.py:159-203) where a direct solve is the right tool on the
// GPU: the coarsest level of the V-cycle (mg.cu), meshes small enough to be solved outright (the reference's wing box
// as shipped: 32 226 dofs), and the time stepper's constant effective operator (dynamic.cu).  Round 1 used cuSOLVER
// potrf + potri on the DENSE matrix (12.8 GB and 3.1 s at 39 939 dofs); this file needs no library.
//
//   ordering     breadth-first level sets of the block-node graph (the BSR pattern) from a pseudo-peripheral node: level
//                k couples to levels k-1, k, k+1 only, so A is block tridiagonal with blocks D_k (m_k x m_k), O_k
//                (m_k x m_{k-1}); on a shell mesh the fronts hold O(sqrt(n)) dofs
//   factor       S_0 = D_0,  G_k = O_k S_{k-1}^-1,  S_k = D_k - G_k O_k^T   (block LDL^T);  S_k^-1, G_k and G_k^T are
//                stored as dense matrices: a solve is then 2 matrix-vector sweeps, one small kernel per level
//   S^-1         blocked right-looking Cholesky (64 x 64 tiles factorised and inverted in shared memory), recursive
//                inversion of the triangular factor, S^-1 = L^-T L^-1: everything but the tile kernel is a GEMM.
//                (A recursive Schur-complement inversion loses ~5 digits per level on thin-shell blocks and the sweeps
//                amplify it: scripts/blocktri_proto.py, the NumPy prototype of this file, shows both.)
//   small n      n <= 8192: the sweeps are replaced by ONE matrix-vector product with the dense inverse (permuted order),
//                built from the factor with GEMMs
// fp64 throughout, no atomics, fixed summation orders: repeated solves are bitwise identical.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <string>
#include <vector>
#include "femo_internal.cuh"

#define CT 64                 /* Cholesky tile */
#define DENSE_LIMIT 8192      /* up to this many dofs a solve is one GEMV with the explicit inverse */

struct CoarseSolver {
  int64_t n = 0, nb = 0, mmax = 0;
  int nl = 0;
  std::vector<int64_t> lp;            // dof offset of every level (nl + 1)
  std::vector<size_t> oS, oG;         // offsets of S_k^-1 and of G_k / G_k^T in their pools (doubles)
  int32_t *perm = nullptr;            // n: position in level order -> dof
  int32_t *order = nullptr;           // nb: block nodes in level order
  int32_t *node_lev = nullptr, *node_pos = nullptr;   // nb: level of a block node, its position inside the level
  double *Sinv = nullptr, *G = nullptr, *GT = nullptr;
  double *z = nullptr, *xp = nullptr;
  double *dense = nullptr;            // n x n inverse in level order (n <= DENSE_LIMIT)
  int *flag = nullptr;                // device: set when a Cholesky pivot is not positive
  double bytes = 0;
  bool factored = false;
  // the sweeps are 2 small kernels per level set (hundreds of launches, launch-bound): captured once per (b, x) pointer
  // pair into a CUDA graph and replayed; the factor arrays never move, so a re-factorisation keeps the graphs valid
  struct SweepGraph { const double* b = nullptr; double* x = nullptr; cudaGraphExec_t exec = nullptr; };
  SweepGraph graphs[4];
  cudaStream_t gstream = nullptr;     // capture needs a non-legacy stream
};

namespace {

// ---- C = alpha op(A) op(B) + beta C, row-major with leading dimensions; 64 x 64 tile per CTA, 4 x 4 per thread -----------
template <bool TA, bool TB>
__global__ void __launch_bounds__(256)
k_gemm(int M, int N, int K, double alpha, const double* __restrict__ A, int lda, const double* __restrict__ B, int ldb,
       double beta, double* __restrict__ C, int ldc) {
  __shared__ double As[16][65], Bs[16][65];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int row0 = blockIdx.y * 64, col0 = blockIdx.x * 64;
  double acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += 16) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int e = tid + 256 * q;
      {  // A tile: As[kk][r] = op(A)[row0 + r][k0 + kk]
        const int r = TA ? (e & 63) : (e >> 4), kk = TA ? (e >> 6) : (e & 15);
        const int gr = row0 + r, gk = k0 + kk;
        double v = 0.0;
        if (gr < M && gk < K) v = TA ? A[(size_t)gk * lda + gr] : A[(size_t)gr * lda + gk];
        As[kk][r] = v;
      }
      {  // B tile: Bs[kk][cc] = op(B)[k0 + kk][col0 + cc]
        const int cc = TB ? (e >> 4) : (e & 63), kk = TB ? (e & 15) : (e >> 6);
        const int gc = col0 + cc, gk = k0 + kk;
        double v = 0.0;
        if (gc < N && gk < K) v = TB ? B[(size_t)gc * ldb + gk] : B[(size_t)gk * ldb + gc];
        Bs[kk][cc] = v;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[kk][ty + 16 * i]; b[i] = Bs[kk][tx + 16 * i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = row0 + ty + 16 * i;
    if (r >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int cidx = col0 + tx + 16 * j;
      if (cidx >= N) continue;
      double* p = C + (size_t)r * ldc + cidx;
      *p = alpha * acc[i][j] + (beta != 0.0 ? beta * *p : 0.0);
    }
  }
}

int gemm(femo_ctx* c, bool ta, bool tb, int64_t M, int64_t N, int64_t K, double alpha, const double* A, int64_t lda,
         const double* B, int64_t ldb, double beta, double* C, int64_t ldc) {
  if (M <= 0 || N <= 0) return 0;
  dim3 grid((unsigned)cdiv(N, 64), (unsigned)cdiv(M, 64));
  cudaStream_t s = c->stream;
  if (!ta && !tb) k_gemm<false, false><<<grid, 256, 0, s>>>((int)M, (int)N, (int)K, alpha, A, (int)lda, B, (int)ldb, beta, C, (int)ldc);
  else if (!ta && tb) k_gemm<false, true><<<grid, 256, 0, s>>>((int)M, (int)N, (int)K, alpha, A, (int)lda, B, (int)ldb, beta, C, (int)ldc);
  else if (ta && !tb) k_gemm<true, false><<<grid, 256, 0, s>>>((int)M, (int)N, (int)K, alpha, A, (int)lda, B, (int)ldb, beta, C, (int)ldc);
  else k_gemm<true, true><<<grid, 256, 0, s>>>((int)M, (int)N, (int)K, alpha, A, (int)lda, B, (int)ldb, beta, C, (int)ldc);
  c->launches += 1;
  return 0;
}

// ---- one CT x CT diagonal tile: L = chol(A_tile) (lower), T = L^-1; both written as full kb x kb tiles (zeros above the
// diagonal).  One CTA; the tile lives in shared memory.
__global__ void __launch_bounds__(256)
k_chol_tile(const double* __restrict__ A, int lda, int kb, double* __restrict__ Lout, int ldl, double* __restrict__ Tout,
            int ldt, int* __restrict__ flag) {
  extern __shared__ double tile_sm[];                 // 2 x CT x (CT + 1) doubles (65 KB: dynamic, opted in by the host)
  double (*a)[CT + 1] = reinterpret_cast<double (*)[CT + 1]>(tile_sm);
  double (*t)[CT + 1] = reinterpret_cast<double (*)[CT + 1]>(tile_sm + CT * (CT + 1));
  const int tid = threadIdx.x;
  for (int e = tid; e < kb * kb; e += 256) { const int i = e / kb, j = e - i * kb; a[i][j] = (j <= i) ? A[(size_t)i * lda + j] : 0.0; }
  __syncthreads();
  for (int j = 0; j < kb; ++j) {
    const double d = a[j][j];
    if (!(d > 0.0)) { if (tid == 0) atomicExch(flag, 1); }
    const double sd = sqrt(d > 0.0 ? d : 1.0);
    __syncthreads();
    for (int i = j + tid; i < kb; i += 256) a[i][j] = (i == j) ? sd : a[i][j] / sd;
    __syncthreads();
    // trailing update of the lower triangle: a[i][k] -= a[i][j] a[k][j],  j < k <= i
    const int w = kb - j - 1;
    for (int e = tid; e < w * w; e += 256) {
      const int i = j + 1 + e / w, k = j + 1 + e % w;
      if (k <= i) a[i][k] -= a[i][j] * a[k][j];
    }
    __syncthreads();
  }
  // T = L^-1: thread col solves L t = e_col by forward substitution
  for (int col = tid; col < kb; col += 256) {
    for (int i = 0; i < kb; ++i) {
      double s = (i == col) ? 1.0 : 0.0;
      for (int k = col; k < i; ++k) s -= a[i][k] * t[k][col];
      t[i][col] = (i < col) ? 0.0 : s / a[i][i];
    }
  }
  __syncthreads();
  for (int e = tid; e < kb * kb; e += 256) {
    const int i = e / kb, j = e - i * kb;
    Lout[(size_t)i * ldl + j] = a[i][j];
    Tout[(size_t)i * ldt + j] = t[i][j];
  }
}

__global__ void k_copy2d(const double* __restrict__ src, int lds, double* __restrict__ dst, int ldd, int rows, int cols) {
  const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (e >= (int64_t)rows * cols) return;
  const int i = (int)(e / cols), j = (int)(e - (int64_t)i * cols);
  dst[(size_t)i * ldd + j] = src[(size_t)i * lds + j];
}

// ---- dense diagonal / sub-diagonal blocks of the BC-applied operator for one level -------------------------------------------
// one thread per (block node of the level, BSR entry slot): D[3 p + a][3 q + b] for columns of the same level,
// O[3 p + a][3 q + b] for columns of the previous one; clamped dofs: zero row and column, unit diagonal
__global__ void k_level_blocks(const int32_t* __restrict__ order, int64_t first, int64_t count, const int32_t* __restrict__ rowptr,
                               const int32_t* __restrict__ colidx, const double* __restrict__ vals,
                               const double* __restrict__ bcmask, const int32_t* __restrict__ node_lev,
                               const int32_t* __restrict__ node_pos, int lev, int64_t m, int64_t mprev, double* __restrict__ D,
                               double* __restrict__ O) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= count) return;
  const int64_t I = order[first + p];
  for (int32_t k = rowptr[I]; k < rowptr[I + 1]; ++k) {
    const int64_t J = colidx[k];
    const int lj = node_lev[J];
    if (lj != lev && lj != lev - 1) continue;
    const int64_t q = node_pos[J];
    double* dst = (lj == lev) ? D + (3 * p) * m + 3 * q : O + (3 * p) * mprev + 3 * q;
    const int64_t ld = (lj == lev) ? m : mprev;
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        const double mr = bcmask[3 * I + a], mc = bcmask[3 * J + b];
        double v = vals[(int64_t)k * 9 + 3 * a + b] * mr * mc;
        if (I == J && a == b && mr == 0.0) v = 1.0;
        dst[a * ld + b] = v;
      }
  }
}

// ---- sweeps: one warp per row -------------------------------------------------------------------------------------------------
// z_k[r] = b[perm_k[r]] - sum_j G_k[r][j] z_{k-1}[j]
__global__ void __launch_bounds__(256)
k_sweep_fwd(const double* __restrict__ Gk, int m, int mprev, const double* __restrict__ zprev, const double* __restrict__ b,
            const int32_t* __restrict__ perm_k, double* __restrict__ zk) {
  const int row = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (row >= m) return;
  double s = 0.0;
  const double* g = Gk + (size_t)row * mprev;
  for (int j = lane; j < mprev; j += 32) s = fma(g[j], zprev[j], s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) zk[row] = b[perm_k[row]] - s;
}

// x_k[r] = sum_j Sinv_k[r][j] z_k[j] - sum_j GT_{k+1}[r][j] x_{k+1}[j];  also scattered to the caller's ordering
__global__ void __launch_bounds__(256)
k_sweep_bwd(const double* __restrict__ Sk, int m, const double* __restrict__ zk, const double* __restrict__ GTn, int mnext,
            const double* __restrict__ xnext, const int32_t* __restrict__ perm_k, double* __restrict__ xk, double* __restrict__ out) {
  const int row = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (row >= m) return;
  double s = 0.0, u = 0.0;
  const double* a = Sk + (size_t)row * m;
  for (int j = lane; j < m; j += 32) s = fma(a[j], zk[j], s);
  if (mnext > 0) {
    const double* g = GTn + (size_t)row * mnext;
    for (int j = lane; j < mnext; j += 32) u = fma(g[j], xnext[j], u);
  }
  s -= u;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) { xk[row] = s; out[perm_k[row]] = s; }
}

__global__ void k_gather_perm(const double* __restrict__ b, const int32_t* __restrict__ perm, int64_t n, double* __restrict__ z) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) z[i] = b[perm[i]];
}

// out[perm[row]] = sum_j X[row][j] z[j]   (dense inverse in level order)
__global__ void __launch_bounds__(256)
k_dense_apply(const double* __restrict__ X, const double* __restrict__ z, int64_t n, const int32_t* __restrict__ perm,
              double* __restrict__ out) {
  const int64_t row = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= n) return;
  const double* a = X + row * n;
  double s = 0.0;
  for (int64_t j = lane; j < n; j += 32) s = fma(a[j], z[j], s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) out[perm[row]] = s;
}

// r = b - A x with every product and sum carried in double-double (error-free TwoProd through fma, TwoSum), one thread per
// block row.  Iterative refinement needs the residual to more digits than the solution has: in plain fp64 it stalls at
// eps |A| |x| / |b| (~1e-8 on the wing box, kappa ~ 1e12).  A with the BC convention (clamped rows / columns -> identity).
__device__ __forceinline__ void dd_add_prod(double a, double b, double& hi, double& lo) {
  const double p = __dmul_rn(a, b), pe = __fma_rn(a, b, -p);    // a b = p + pe exactly (intrinsics: never contracted)
  const double s = __dadd_rn(hi, p), bb = s - hi;               // TwoSum(hi, p)
  const double se = (hi - (s - bb)) + (p - bb);
  hi = s; lo += se + pe;
}
__global__ void k_residual_dd(const int32_t* __restrict__ rowptr, const uint32_t* __restrict__ colw, const double* __restrict__ vals,
                              const uint8_t* __restrict__ rbits, const double* __restrict__ x, const double* __restrict__ b,
                              int64_t nb, double* __restrict__ r) {
  const int64_t I = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (I >= nb) return;
  double hi[3] = {0, 0, 0}, lo[3] = {0, 0, 0};
  for (int32_t k = rowptr[I]; k < rowptr[I + 1]; ++k) {
    const uint32_t w = colw[k];
    const double* a = vals + (int64_t)k * 9;
    const double* xp = x + 3 * (int64_t)(w & 0x1fffffffu);
    const double xc[3] = {(w & 0x20000000u) ? 0.0 : xp[0], (w & 0x40000000u) ? 0.0 : xp[1], (w & 0x80000000u) ? 0.0 : xp[2]};
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) dd_add_prod(a[3 * i + j], xc[j], hi[i], lo[i]);
  }
  const unsigned bits = rbits[I];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double bi = b[3 * I + i];
    if (bits & (1u << i)) r[3 * I + i] = bi - x[3 * I + i];                 // identity row
    else r[3 * I + i] = (bi - hi[i]) - lo[i];                               // bi - hi is exact to one rounding when they nearly cancel
  }
}

__global__ void k_set_diag(double* A, int64_t n) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) A[i * n + i] = 1.0;
}

// ---- host: level sets -----------------------------------------------------------------------------------------------------------
// breadth-first levels from `start` over the nodes with lev < 0; returns the last level number used
int bfs(const std::vector<int32_t>& rp, const std::vector<int32_t>& ci, int32_t start, int first_level, std::vector<int32_t>& lev,
        std::vector<int32_t>* visited) {
  std::vector<int32_t> front{start}, next;
  lev[start] = first_level;
  if (visited) visited->push_back(start);
  int l = first_level;
  while (true) {
    next.clear();
    for (int32_t i : front)
      for (int32_t k = rp[i]; k < rp[i + 1]; ++k) {
        const int32_t j = ci[k];
        if (lev[j] < 0) { lev[j] = l + 1; next.push_back(j); if (visited) visited->push_back(j); }
      }
    if (next.empty()) break;
    ++l;
    front.swap(next);
  }
  return l;
}

void level_sets(const std::vector<int32_t>& rp, const std::vector<int32_t>& ci, std::vector<int32_t>& lev, int* n_levels) {
  const int64_t nb = (int64_t)rp.size() - 1;
  lev.assign(nb, -1);
  int next_level = 0;
  for (int32_t seed = 0; seed < nb; ++seed) {
    if (lev[seed] >= 0) continue;
    // pseudo-peripheral node of this component: repeat the search from a minimum-degree node of the last level while
    // the eccentricity grows
    int32_t s = seed;
    int ecc = -1;
    std::vector<int32_t> visited;
    for (int rep = 0; rep < 6; ++rep) {
      visited.clear();
      const int last = bfs(rp, ci, s, 0, lev, &visited);
      int32_t best = -1;
      for (int32_t v : visited)
        if (lev[v] == last && (best < 0 || rp[v + 1] - rp[v] < rp[best + 1] - rp[best])) best = v;
      for (int32_t v : visited) lev[v] = -1;
      if (last <= ecc) break;
      ecc = last;
      s = best;
    }
    next_level = bfs(rp, ci, s, next_level, lev, nullptr) + 1;
  }
  *n_levels = next_level;
}

// S (m x m, destroyed) -> Sinv (m x m); L, W: m x m scratch, T: m x CT scratch
int spd_inverse_dev(femo_ctx* c, CoarseSolver* cs, double* S, double* L, double* T, double* W, int64_t m, double* Sinv) {
  const size_t tile_bytes = sizeof(double) * 2 * CT * (CT + 1);
  static bool opted_in = false;
  if (!opted_in) {
    CUDA_TRY(cudaFuncSetAttribute(k_chol_tile, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_bytes));
    opted_in = true;
  }
  for (int64_t k = 0; k < m; k += CT) {
    const int64_t kb = std::min<int64_t>(CT, m - k), e = k + kb;
    k_chol_tile<<<1, 256, tile_bytes, c->stream>>>(S + k * m + k, (int)m, (int)kb, L + k * m + k, (int)m, T + k * CT, CT, cs->flag);
    c->launches += 1;
    if (e < m) {
      TRY(gemm(c, false, true, m - e, kb, kb, 1.0, S + e * m + k, m, T + k * CT, CT, 0.0, L + e * m + k, m));        // panel
      TRY(gemm(c, false, true, m - e, m - e, kb, -1.0, L + e * m + k, m, L + e * m + k, m, 1.0, S + e * m + e, m));   // trailing
    }
  }
  // Linv (lower triangular, dense storage) into S
  CUDA_TRY(cudaMemsetAsync(S, 0, sizeof(double) * m * m, c->stream));
  struct Rec {
    femo_ctx* c; double *S, *L, *T, *W; int64_t m;
    int run(int64_t r0, int64_t mm) {
      if (mm <= CT) {
        k_copy2d<<<(unsigned)cdiv(mm * mm, 256), 256, 0, c->stream>>>(T + r0 * CT, CT, S + r0 * m + r0, (int)m, (int)mm, (int)mm);
        c->launches += 1;
        return 0;
      }
      const int64_t nblk = cdiv(mm, CT), m1 = (nblk / 2) * CT, m2 = mm - m1;
      TRY(run(r0, m1));
      TRY(run(r0 + m1, m2));
      TRY(gemm(c, false, false, m2, m1, m1, 1.0, L + (r0 + m1) * m + r0, m, S + r0 * m + r0, m, 0.0, W, m1));
      TRY(gemm(c, false, false, m2, m1, m2, -1.0, S + (r0 + m1) * m + (r0 + m1), m, W, m1, 0.0, S + (r0 + m1) * m + r0, m));
      return 0;
    }
  } rec{c, S, L, T, W, m};
  TRY(rec.run(0, m));
  TRY(gemm(c, true, false, m, m, m, 1.0, S, m, S, m, 0.0, Sinv, m));
  return check_launch(c, "coarse: dense SPD inverse");
}

}  // namespace

void coarse_free(CoarseSolver* cs) {
  if (!cs) return;
  for (auto& g : cs->graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
  if (cs->gstream) cudaStreamDestroy(cs->gstream);
  dev_free(&cs->perm); dev_free(&cs->order); dev_free(&cs->node_lev); dev_free(&cs->node_pos);
  dev_free(&cs->Sinv); dev_free(&cs->G); dev_free(&cs->GT); dev_free(&cs->z); dev_free(&cs->xp); dev_free(&cs->dense);
  dev_free(&cs->flag);
  delete cs;
}

// symbolic phase: level sets of the BSR pattern, permutation, storage.  Once per pattern.
int coarse_analyse(femo_ctx* c, CoarseSolver** out) {
  if (c->world > 1) return femo_fail("direct solver: single-GPU contexts only (replicated coarse levels, small meshes)");
  CoarseSolver* cs = new CoarseSolver();
  cs->n = c->ndof; cs->nb = c->nb;
  std::vector<int32_t> rp((size_t)c->nb + 1), ci((size_t)c->nnzb);
  cudaMemcpy(rp.data(), c->rowptr, sizeof(int32_t) * (c->nb + 1), cudaMemcpyDeviceToHost);
  cudaMemcpy(ci.data(), c->colidx, sizeof(int32_t) * c->nnzb, cudaMemcpyDeviceToHost);
  std::vector<int32_t> lev;
  level_sets(rp, ci, lev, &cs->nl);
  std::vector<int64_t> cnt((size_t)cs->nl, 0);
  for (int64_t i = 0; i < c->nb; ++i) cnt[lev[i]] += 1;
  std::vector<int64_t> nlp((size_t)cs->nl + 1, 0);
  for (int l = 0; l < cs->nl; ++l) nlp[l + 1] = nlp[l] + cnt[l];
  std::vector<int32_t> order((size_t)c->nb), pos((size_t)c->nb), perm((size_t)c->ndof);
  std::vector<int64_t> fill(nlp.begin(), nlp.end() - 1);
  for (int64_t i = 0; i < c->nb; ++i) { const int64_t p = fill[lev[i]]++; order[p] = (int32_t)i; pos[i] = (int32_t)(p - nlp[lev[i]]); }
  for (int64_t p = 0; p < c->nb; ++p)
    for (int d = 0; d < 3; ++d) perm[3 * p + d] = 3 * order[p] + d;
  cs->lp.resize(cs->nl + 1);
  for (int l = 0; l <= cs->nl; ++l) cs->lp[l] = 3 * nlp[l];
  cs->oS.resize(cs->nl); cs->oG.resize(cs->nl);
  size_t nS = 0, nG = 0;
  for (int l = 0; l < cs->nl; ++l) {
    const int64_t m = cs->lp[l + 1] - cs->lp[l], mp = l ? cs->lp[l] - cs->lp[l - 1] : 0;
    cs->mmax = std::max(cs->mmax, m);
    cs->oS[l] = nS; nS += (size_t)m * m;
    cs->oG[l] = nG; nG += (size_t)m * mp;
  }
  cs->bytes = 8.0 * ((double)nS + 2.0 * (double)nG);
  const char* env = getenv("FEMO_DIRECT_MAX_GB");
  const double cap = (env ? atof(env) : 64.0) * 1e9;
  if (cs->bytes + 8.0 * 4.0 * (double)cs->mmax * cs->mmax > cap) {
    const std::string msg = "direct solver: the block-tridiagonal factor needs " + std::to_string(cs->bytes / 1e9) + " GB (" +
                            std::to_string(cs->nl) + " level sets, largest front " + std::to_string(cs->mmax) +
                            " dofs); raise FEMO_DIRECT_MAX_GB or use the multigrid / PCG solvers";
    coarse_free(cs);
    return femo_fail(msg);
  }
  int r = 0;
  if (r == 0) r = dev_alloc(&cs->perm, (size_t)c->ndof);
  if (r == 0) r = dev_alloc(&cs->order, (size_t)c->nb);
  if (r == 0) r = dev_alloc(&cs->node_lev, (size_t)c->nb);
  if (r == 0) r = dev_alloc(&cs->node_pos, (size_t)c->nb);
  if (r == 0) r = dev_alloc(&cs->Sinv, nS);
  if (r == 0) r = dev_alloc(&cs->G, nG);
  if (r == 0) r = dev_alloc(&cs->GT, nG);
  if (r == 0) r = dev_alloc(&cs->z, (size_t)c->ndof);
  if (r == 0) r = dev_alloc(&cs->xp, (size_t)c->ndof);
  if (r == 0) r = dev_alloc(&cs->flag, 1);
  if (r != 0) { coarse_free(cs); return r; }
  cudaMemcpy(cs->perm, perm.data(), sizeof(int32_t) * c->ndof, cudaMemcpyHostToDevice);
  cudaMemcpy(cs->order, order.data(), sizeof(int32_t) * c->nb, cudaMemcpyHostToDevice);
  cudaMemcpy(cs->node_lev, lev.data(), sizeof(int32_t) * c->nb, cudaMemcpyHostToDevice);
  cudaMemcpy(cs->node_pos, pos.data(), sizeof(int32_t) * c->nb, cudaMemcpyHostToDevice);
  if (getenv("FEMO_VERBOSE"))
    fprintf(stderr, "[femo] direct solver: %lld dofs, %d level sets, largest front %lld dofs, factor %.2f GB\n", (long long)cs->n,
            cs->nl, (long long)cs->mmax, cs->bytes / 1e9);
  *out = cs;
  return 0;
}

// numeric phase from the current c->vals / c->bcmask
int coarse_factor(femo_ctx* c, CoarseSolver* cs) {
  if (!c->jac_valid) return femo_fail("direct solver: Jacobian not assembled");
  const int64_t mm = cs->mmax;
  double *S = nullptr, *L = nullptr, *T = nullptr, *W = nullptr, *O = nullptr;
  struct Scratch { double **a, **b, **cc, **d, **e; ~Scratch() { dev_free(a); dev_free(b); dev_free(cc); dev_free(d); dev_free(e); } } guard{&S, &L, &T, &W, &O};
  TRY(dev_alloc(&S, (size_t)mm * mm)); TRY(dev_alloc(&L, (size_t)mm * mm)); TRY(dev_alloc(&T, (size_t)mm * CT));
  TRY(dev_alloc(&W, (size_t)mm * mm)); TRY(dev_alloc(&O, (size_t)mm * mm));
  CUDA_TRY(cudaMemsetAsync(cs->flag, 0, sizeof(int), c->stream));
  for (int l = 0; l < cs->nl; ++l) {
    const int64_t m = cs->lp[l + 1] - cs->lp[l], mp = l ? cs->lp[l] - cs->lp[l - 1] : 0;
    CUDA_TRY(cudaMemsetAsync(S, 0, sizeof(double) * m * m, c->stream));
    if (mp) CUDA_TRY(cudaMemsetAsync(O, 0, sizeof(double) * m * mp, c->stream));
    k_level_blocks<<<(unsigned)cdiv(m / 3, 128), 128, 0, c->stream>>>(cs->order, cs->lp[l] / 3, m / 3, c->rowptr, c->colidx, c->vals,
                                                                     c->bcmask, cs->node_lev, cs->node_pos, l, m, mp, S, O);
    c->launches += 1;
    double* Sinv = cs->Sinv + cs->oS[l];
    if (l > 0) {
      const double* Sp = cs->Sinv + cs->oS[l - 1];
      double *G = cs->G + cs->oG[l], *GT = cs->GT + cs->oG[l];
      TRY(gemm(c, false, false, m, mp, mp, 1.0, O, mp, Sp, mp, 0.0, G, mp));        // G_l = O_l S_{l-1}^-1
      TRY(gemm(c, false, true, mp, m, mp, 1.0, Sp, mp, O, mp, 0.0, GT, m));         // G_l^T
      TRY(gemm(c, false, true, m, m, mp, -1.0, G, mp, O, mp, 1.0, S, m));           // S_l = D_l - G_l O_l^T
    }
    TRY(spd_inverse_dev(c, cs, S, L, T, W, m, Sinv));
  }
  int bad = 0;
  CUDA_TRY(cudaMemcpyAsync(&bad, cs->flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  if (bad) return femo_fail("direct solver: the operator is not positive definite (a Cholesky pivot is <= 0; unconstrained rigid-body modes?)");
  // small systems: explicit inverse (level order), X = blocktri_solve(I)
  if (cs->n <= DENSE_LIMIT) {
    const int64_t n = cs->n;
    if (!cs->dense) TRY(dev_alloc(&cs->dense, (size_t)n * n));
    double* Z = cs->dense;
    double* P = nullptr;                       // row panel
    struct G1 { double** p; ~G1() { dev_free(p); } } g1{&P};
    TRY(dev_alloc(&P, (size_t)mm * n));
    CUDA_TRY(cudaMemsetAsync(Z, 0, sizeof(double) * n * n, c->stream));
    k_set_diag<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(Z, n);
    for (int l = 1; l < cs->nl; ++l) {
      const int64_t m = cs->lp[l + 1] - cs->lp[l], mp = cs->lp[l] - cs->lp[l - 1];
      // rows of level l-1 are non-zero in the columns [0, lp[l]) only
      TRY(gemm(c, false, false, m, cs->lp[l], mp, -1.0, cs->G + cs->oG[l], mp, Z + cs->lp[l - 1] * n, n, 1.0, Z + cs->lp[l] * n, n));
    }
    for (int l = cs->nl - 1; l >= 0; --l) {
      const int64_t m = cs->lp[l + 1] - cs->lp[l];
      TRY(gemm(c, false, false, m, n, m, 1.0, cs->Sinv + cs->oS[l], m, Z + cs->lp[l] * n, n, 0.0, P, n));
      if (l + 1 < cs->nl) {
        const int64_t mn = cs->lp[l + 2] - cs->lp[l + 1];
        TRY(gemm(c, false, false, m, n, mn, -1.0, cs->GT + cs->oG[l + 1], mn, Z + cs->lp[l + 1] * n, n, 1.0, P, n));
      }
      CUDA_TRY(cudaMemcpyAsync(Z + cs->lp[l] * n, P, sizeof(double) * m * n, cudaMemcpyDeviceToDevice, c->stream));
    }
    TRY(check_launch(c, "coarse: explicit inverse"));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
  }
  cs->factored = true;
  return 0;
}

static int coarse_sweeps(femo_ctx* c, CoarseSolver* cs, const double* b, double* x) {
  for (int l = 0; l < cs->nl; ++l) {
    const int64_t m = cs->lp[l + 1] - cs->lp[l], mp = l ? cs->lp[l] - cs->lp[l - 1] : 0;
    k_sweep_fwd<<<(unsigned)cdiv(m * 32, 256), 256, 0, c->stream>>>(cs->G + cs->oG[l], (int)m, (int)mp, l ? cs->z + cs->lp[l - 1] : nullptr, b,
                                                                   cs->perm + cs->lp[l], cs->z + cs->lp[l]);
  }
  for (int l = cs->nl - 1; l >= 0; --l) {
    const int64_t m = cs->lp[l + 1] - cs->lp[l], mn = (l + 1 < cs->nl) ? cs->lp[l + 2] - cs->lp[l + 1] : 0;
    k_sweep_bwd<<<(unsigned)cdiv(m * 32, 256), 256, 0, c->stream>>>(cs->Sinv + cs->oS[l], (int)m, cs->z + cs->lp[l],
                                                                   mn ? cs->GT + cs->oG[l + 1] : nullptr, (int)mn,
                                                                   mn ? cs->xp + cs->lp[l + 1] : nullptr, cs->perm + cs->lp[l],
                                                                   cs->xp + cs->lp[l], x);
  }
  c->launches += 2 * (int64_t)cs->nl;
  return check_launch(c, "coarse: sweeps");
}

// x = A^-1 b (device vectors, caller's ordering; b and x may alias).  Kernels only: capturable in a CUDA graph.
int coarse_apply(femo_ctx* c, CoarseSolver* cs, const double* b, double* x) {
  if (!cs || !cs->factored) return femo_fail("direct solver: not factorised");
  const int64_t n = cs->n;
  if (cs->dense) {
    k_gather_perm<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(b, cs->perm, n, cs->z);
    k_dense_apply<<<(unsigned)cdiv(n * 32, 256), 256, 0, c->stream>>>(cs->dense, cs->z, n, cs->perm, x);
    c->launches += 2;
    return check_launch(c, "coarse: dense apply");
  }
  // graph replay of the sweeps (FEMO_COARSE_GRAPH=0: plain launches).  Not while the caller itself is capturing.
  static const bool use_graph = !(getenv("FEMO_COARSE_GRAPH") && getenv("FEMO_COARSE_GRAPH")[0] == '0');
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (use_graph) cudaStreamIsCapturing(c->stream, &st);
  if (!use_graph || st != cudaStreamCaptureStatusNone || cs->nl < 8) return coarse_sweeps(c, cs, b, x);
  CoarseSolver::SweepGraph* g = nullptr;
  for (auto& e : cs->graphs) if (e.exec && e.b == b && e.x == x) g = &e;
  if (!g) {
    for (auto& e : cs->graphs) if (!e.exec) { g = &e; break; }
    if (!g) return coarse_sweeps(c, cs, b, x);                       // more pointer pairs than slots: plain launches
    if (!cs->gstream) CUDA_TRY(cudaStreamCreateWithFlags(&cs->gstream, cudaStreamNonBlocking));
    cudaStream_t keep = c->stream;
    const int64_t keep_launches = c->launches;
    c->stream = cs->gstream;
    cudaGraph_t graph = nullptr;
    int r = 0;
    if (cudaStreamBeginCapture(cs->gstream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) r = 1;
    if (r == 0) r = coarse_sweeps(c, cs, b, x);
    const cudaError_t ee = cudaStreamEndCapture(cs->gstream, &graph);
    c->stream = keep; c->launches = keep_launches;
    if (r != 0 || ee != cudaSuccess || !graph || cudaGraphInstantiate(&g->exec, graph, 0) != cudaSuccess) {
      if (graph) cudaGraphDestroy(graph);
      g->exec = nullptr;
      cudaGetLastError();
      return coarse_sweeps(c, cs, b, x);                             // capture unavailable: plain launches
    }
    cudaGraphDestroy(graph);
    g->b = b; g->x = x;
  }
  CUDA_TRY(cudaGraphLaunch(g->exec, c->stream));
  c->launches += 2 * (int64_t)cs->nl;
  return 0;
}

int coarse_num_levels(const CoarseSolver* cs) { return cs ? cs->nl : 0; }
int64_t coarse_max_front(const CoarseSolver* cs) { return cs ? cs->mmax : 0; }
double coarse_bytes(const CoarseSolver* cs) { return cs ? cs->bytes + (cs->dense ? 8.0 * (double)cs->n * cs->n : 0.0) : 0.0; }
bool coarse_is_dense(const CoarseSolver* cs) { return cs && cs->dense; }

// ---- C ABI: the direct solver as a solver of its own (small / medium meshes without a hierarchy) -----------------------------
extern "C" {

// analyse (first call) and factorise the current operator (assembled on demand)
int femo_direct_setup(femo_ctx* c) {
  if (!c || !c->setup_done) return femo_fail("femo_direct_setup: context not set up");
  int prev = 0; cudaGetDevice(&prev); cudaSetDevice(c->device);
  int r = 0;
  if (c->jac_valid && c->direct && c->direct->factored && c->direct_epoch == c->mat_epoch) { cudaSetDevice(prev); return 0; }   // up to date
  if (!c->jac_valid) r = assemble_jacobian(c);
  if (r == 0 && !c->direct) r = coarse_analyse(c, &c->direct);
  if (r == 0) r = coarse_factor(c, c->direct);
  if (r == 0) c->direct_epoch = c->mat_epoch;
  cudaSetDevice(prev);
  return r;
}

}  // extern "C"

// r = b - A x over all local block rows, accumulated in double-double (A with the BC rows / columns replaced by identity, as
// every SpMV of the library); multi-GPU callers exchange the halo of x before and of r after
int residual_dd(femo_ctx* c, const double* x, const double* b, double* r) {
  k_residual_dd<<<(unsigned)cdiv(c->nb, 128), 128, 0, c->stream>>>(c->rowptr, c->colbc, c->vals, c->bcbits, x, b, c->nb, r);
  c->launches += 1;
  return check_launch(c, "k_residual_dd");
}

extern "C" {

// A x = rhs with `refine` steps of iterative refinement on a residual accumulated in double-double (k_residual_dd);
// relres = |rhs - A x| / |rhs| of the result
int femo_direct_solve(femo_ctx* c, const double* rhs, double* sol, int refine, double* relres, int on_device, int zero_bc) {
  if (!c || !c->setup_done) return femo_fail("femo_direct_solve: context not set up");
  if (!c->direct || !c->direct->factored || c->direct_epoch != c->mat_epoch || !c->jac_valid)
    return femo_fail("femo_direct_solve: call femo_direct_setup first (and again after the inputs or the BCs change)");
  int prev = 0; cudaGetDevice(&prev); cudaSetDevice(c->device);
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
  const int64_t n = c->ndof;
  CUDA_TRY(cudaMemcpyAsync(c->b, rhs, sizeof(double) * n, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
  TRY(vec_mul(c, c->b, c->bcmask, n, c->b));
  TRY(coarse_apply(c, c->direct, c->b, c->tmp2));
  double bb = 0, rr = 0;
  TRY(dev_dot(c, c->b, c->b, n, &bb));
  for (int it = 0; it <= refine; ++it) {
    k_residual_dd<<<(unsigned)cdiv(c->nb, 128), 128, 0, c->stream>>>(c->rowptr, c->colbc, c->vals, c->bcbits, c->tmp2, c->b, c->nb, c->r);
    c->launches += 1;
    TRY(check_launch(c, "k_residual_dd"));
    TRY(dev_dot(c, c->r, c->r, n, &rr));
    if (it == refine || !(rr > 0)) break;
    TRY(coarse_apply(c, c->direct, c->r, c->z));
    TRY(vec_axpby(c, 1.0, c->tmp2, 1.0, c->z, n, c->tmp2));
  }
  if (relres) *relres = bb > 0 ? sqrt(rr / bb) : 0.0;
  if (zero_bc) TRY(vec_mul(c, c->tmp2, c->bcmask, n, c->tmp2));
  CUDA_TRY(cudaMemcpyAsync(sol, c->tmp2, sizeof(double) * n, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

// info[0] = level sets, [1] = largest front (dofs), [2] = bytes of the factor, [3] = 1 when solves use the explicit inverse
int femo_direct_info(femo_ctx* c, double* info4) {
  if (!c || !info4) return femo_fail("femo_direct_info: null argument");
  if (!c->direct) return femo_fail("femo_direct_info: call femo_direct_setup first");
  info4[0] = coarse_num_levels(c->direct); info4[1] = (double)coarse_max_front(c->direct);
  info4[2] = coarse_bytes(c->direct); info4[3] = coarse_is_dense(c->direct) ? 1.0 : 0.0;
  return 0;
}

}  // extern "C"
