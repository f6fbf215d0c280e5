// Geometric multigrid preconditioner for the state / adjoint solves  A x = b  (A = dR/du with BCs).
//
// Replaces the MUMPS LU behind FEA.solve / solveLinearBwd (femo_alpha/fea/utils_dolfinx.py:438-468,514-531;
// fea_dolfinx.py:159-203) when the mesh comes with a hierarchy of uniformly refined quad meshes
// (femo_alpha_b200/multigrid.py): block-Jacobi PCG needs O(1e5) iterations at 1e7 DOF on thin shells, a V-cycle
// preconditioned PCG ~1e2, mesh-independently (scripts/mg_study.py).
//
//   level l        a femo_ctx holding the Q2xQ1 discretisation of mesh l (re-discretised, not Galerkin), BCs set
//   smoother       Chebyshev polynomial of degree `deg` in  D^-1 A  (D = the 6x6 / 3x3 block-Jacobi of solver.cu),
//                  on [lmax/10, 1.1 lmax], lmax by power iteration: linear and A-self-adjoint -> the V-cycle is SPD
//   transfer       nested Q2xQ1 spaces: P = coarse basis evaluated at the fine nodes (block-scalar CSR, rows = fine
//                  block nodes), R = P^T stored as its own CSR so that both directions are gathers (no atomics)
//   coarsest       direct solve: block-tridiagonal factorisation on level sets (coarse.cu; no library)
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "femo_internal.cuh"

struct MgLevel {
  femo_ctx* c = nullptr;
  int64_t nbf = 0, nbc = 0;
  int32_t *P_ptr = nullptr, *P_col = nullptr, *R_ptr = nullptr, *R_col = nullptr;
  double *P_w = nullptr, *R_w = nullptr;
  double lmax = 0;
  double *x = nullptr, *b = nullptr, *r = nullptr, *d = nullptr, *t = nullptr;
  float* vals32 = nullptr;     // fp32 copy of the level's block values for the smoother / residual inside the V-cycle
  double *ev = nullptr;        // last iterate of the power iteration (warm start of the next set-up)
  int power_its = 0;
};

struct femo_mg {
  std::vector<MgLevel> L;
  int deg = 3;
  CoarseSolver* cs = nullptr;  // direct solver of the coarsest operator (coarse.cu)
  bool ready = false;
  double *pr = nullptr, *pz = nullptr, *pp = nullptr, *pAp = nullptr, *px = nullptr, *pb = nullptr;   // PCG vectors (level 0)
  // iterative refinement around the PCG (femo_mg_set_refine): x += PCG(b - A x), the residual accumulated in double-double
  int refine = 0;
  double refine_rtol = 1e-6;
  double *pxa = nullptr, *pb0 = nullptr;   // accumulated solution, right-hand side of the outer problem
  int64_t vcycles = 0;
};

namespace {

__device__ __forceinline__ void mg_precond_node(int64_t I, int64_t nn, const double* __restrict__ dinv6,
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

// one Chebyshev step, fused per Jacobi group (vertex: u_v and theta_v, 6 dofs; edge / centre node: 3 dofs):
//   r <- r0 - sr * Ad        (r0 = b when first, else r)
//   d <- cd * d + cz * D^-1 r
//   x <- (first ? 0 : x) + d
__global__ void k_cheb_step(const double* r0, const double* __restrict__ Ad, double sr, double cd, double cz,
                            int zero_x, const double* __restrict__ dinv6, const double* __restrict__ dinv3, int64_t nn,
                            int64_t nq2, double* r, double* d, double* x) {
  const int64_t I = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (I >= nq2) return;
  const int nd = I < nn ? 6 : 3;
  double rl[6], zl[6];
  for (int i = 0; i < nd; ++i) {
    const int64_t k = (i < 3) ? 3 * I + i : 3 * (nq2 + I) + (i - 3);
    rl[i] = r0[k] - (Ad ? sr * Ad[k] : 0.0);
  }
  mg_precond_node(I, nn, dinv6, dinv3, rl, zl);
  for (int i = 0; i < nd; ++i) {
    const int64_t k = (i < 3) ? 3 * I + i : 3 * (nq2 + I) + (i - 3);
    const double dn = (cd != 0.0 ? cd * d[k] : 0.0) + cz * zl[i];
    r[k] = rl[i]; d[k] = dn;
    x[k] = (zero_x ? 0.0 : x[k]) + dn;
  }
}

__global__ void k_axpby(int64_t n, double a, const double* x, double b, const double* y, double* o) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) o[i] = a * x[i] + (b != 0.0 ? b * y[i] : 0.0);
}

// y[3 i + c] (+)= mask * sum_k w_k x[3 col_k + c]      (block-scalar CSR, one thread per (row, component))
__global__ void k_transfer(const int32_t* __restrict__ ptr, const int32_t* __restrict__ col, const double* __restrict__ w,
                           const double* __restrict__ x, const double* __restrict__ mask, int64_t nrows, int add,
                           double* __restrict__ y) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= 3 * nrows) return;
  const int64_t i = t / 3; const int cmp = (int)(t - 3 * i);
  double s = 0;
  for (int32_t k = ptr[i]; k < ptr[i + 1]; ++k) {
    const double wk = w[k];
    if (wk != 0.0) s += wk * x[3 * (int64_t)col[k] + cmp];      // zero weights: ghost rows of a distributed fine level
  }
  s *= mask[t];
  y[t] = add ? y[t] + s : s;
}

__global__ void k_mg_mask(int64_t n, const double* __restrict__ mask, double* __restrict__ x) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) x[i] *= mask[i];
}

// y = A x of a level; on a distributed level the ghosts of x are refreshed first (overlapped with the interior rows).
// v32: the level's fp32 copy -- used inside the preconditioner only (smoothing, residual before restriction, spectral
// bound), never for the operator of the outer PCG
int mg_spmv(femo_ctx* c, double* x, double* y, const float* v32 = nullptr) {
  if (v32) return bsr_spmv_f32(c, v32, x, y);
  if (c->world > 1) return dist_spmv(c, x, y, nullptr, nullptr, nullptr, nullptr);
  return bsr_spmv(c, x, y, 1);
}

int mg_smooth(femo_mg* m, int l, bool zero_guess) {
  MgLevel& Lv = m->L[l];
  femo_ctx* c = Lv.c;
  const double lo = 0.1 * Lv.lmax, hi = 1.1 * Lv.lmax;
  const double theta = 0.5 * (hi + lo), delta = 0.5 * (hi - lo), sigma = theta / delta;
  double rho = 1.0 / sigma;
  const unsigned grid = (unsigned)cdiv(c->nq2, 128);
  for (int k = 0; k < m->deg; ++k) {
    if (k == 0) {
      if (zero_guess) {
        k_cheb_step<<<grid, 128, 0, c->stream>>>(Lv.b, nullptr, 0.0, 0.0, 1.0 / theta, 1, c->dinv6, c->dinv3, c->nn, c->nq2,
                                                 Lv.r, Lv.d, Lv.x);
      } else {
        TRY(mg_spmv(c, Lv.x, Lv.t, Lv.vals32));             // distributed level: halo exchange hidden behind the interior rows
        k_cheb_step<<<grid, 128, 0, c->stream>>>(Lv.b, Lv.t, 1.0, 0.0, 1.0 / theta, 0, c->dinv6, c->dinv3, c->nn, c->nq2,
                                                 Lv.r, Lv.d, Lv.x);
      }
    } else {
      TRY(mg_spmv(c, Lv.d, Lv.t, Lv.vals32));
      const double rho_new = 1.0 / (2.0 * sigma - rho);
      k_cheb_step<<<grid, 128, 0, c->stream>>>(Lv.r, Lv.t, 1.0, rho_new * rho, 2.0 * rho_new / delta, 0, c->dinv6, c->dinv3,
                                               c->nn, c->nq2, Lv.r, Lv.d, Lv.x);
      rho = rho_new;
    }
    c->launches += 1;
  }
  return check_launch(c, "k_cheb_step");
}

// L[l].b -> L[l].x
int mg_vcycle(femo_mg* m, int l) {
  MgLevel& Lv = m->L[l];
  femo_ctx* c = Lv.c;
  const int64_t n = c->ndof;
  if (l == (int)m->L.size() - 1) {
    return coarse_apply(c, m->cs, Lv.b, Lv.x);
  }
  MgLevel& Lc = m->L[l + 1];
  TRY(mg_smooth(m, l, true));
  TRY(mg_spmv(c, Lv.x, Lv.t, Lv.vals32));
  k_axpby<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(n, 1.0, Lv.b, -1.0, Lv.t, Lv.r);
  k_transfer<<<(unsigned)cdiv(3 * Lv.nbc, 256), 256, 0, c->stream>>>(Lv.R_ptr, Lv.R_col, Lv.R_w, Lv.r, Lc.c->bcmask, Lv.nbc, 0, Lc.b);
  TRY(check_launch(c, "mg restrict"));
  c->launches += 2;
  // distributed fine level (dist_multigrid.py): R carries zero weights on ghost rows, so every rank holds the
  // contribution of the fine nodes it owns.  Distributed coarse level (nested partition): those contributions land on
  // owned AND ghost coarse nodes -> reverse halo exchange (owners += ghost copies).  Replicated coarse level: the sum
  // over ranks is the coarse residual, on every rank.
  if (Lc.c->world > 1) TRY(halo_accumulate(Lc.c, Lc.b));
  else if (c->world > 1) TRY(dist_allreduce_sum(c, Lc.b, (int)Lc.c->ndof));
  TRY(mg_vcycle(m, l + 1));
  if (Lc.c->world > 1) TRY(halo_exchange(Lc.c, Lc.x));      // the prolongation reads ghost coarse nodes
  k_transfer<<<(unsigned)cdiv(3 * Lv.nbf, 256), 256, 0, c->stream>>>(Lv.P_ptr, Lv.P_col, Lv.P_w, Lc.x, c->bcmask, Lv.nbf, 1, Lv.x);
  TRY(check_launch(c, "mg prolong"));
  c->launches += 1;
  return mg_smooth(m, l, false);
}

// start vector of the power iteration: a fixed function of the node POSITION, so that the spectral bound (and with it the
// smoother and the iteration counts) does not depend on the numbering, the partition or the number of ranks
__global__ void k_mg_seed(const double* __restrict__ xyz, const int32_t* __restrict__ edges, const int32_t* __restrict__ cv,
                          int64_t nn, int64_t ne, int64_t nc, double* __restrict__ v) {
  const int64_t I = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t nq2 = nn + ne + nc;
  if (I >= nq2 + nn) return;
  const V3 p = block_coord(xyz, edges, cv, nn, ne, nc, I);
  const double ph = 12.9898 * p.x + 78.233 * p.y + 37.719 * p.z + (I >= nq2 ? 0.61 : 0.0);
#pragma unroll
  for (int k = 0; k < 3; ++k) v[3 * I + k] = sin(ph + 1.7 * k + 0.11) + 0.5 * cos(3.1 * ph - 0.9 * k);
}

// largest eigenvalue of D^-1 A by power iteration.  First set-up: from the position-seeded vector, at least 8 steps, until
// two successive estimates agree to 1 %, at most 24.  Later set-ups (new inputs, same mesh) restart from the previous
// iterate kept in Lv.ev and need 2 - 6 steps: the dominant eigenvector hardly moves between optimisation iterates.
int mg_lambda_max(femo_mg* m, int l) {
  MgLevel& Lv = m->L[l];
  femo_ctx* c = Lv.c;
  const int64_t n = c->ndof;
  const bool warm = Lv.ev != nullptr;
  if (!warm) {
    TRY(dev_alloc(&Lv.ev, (size_t)n));
    k_mg_seed<<<(unsigned)cdiv(c->nb, 256), 256, 0, c->stream>>>(c->xyz, c->edges, c->cv, c->nn, c->ne, c->nc, Lv.ev);
    TRY(check_launch(c, "k_mg_seed"));
  }
  CUDA_TRY(cudaMemcpyAsync(Lv.x, Lv.ev, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
  k_mg_mask<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(n, c->bcmask, Lv.x);
  double lam = 1.0, lam_prev = 0.0;
  const int min_it = warm ? 2 : 8, max_it = warm ? 6 : 24;
  int its_done = 0;
  for (int it = 0; it <= max_it; ++it) {
    double nrm2 = 0;
    its_done = it;
    TRY(dev_dot(c, Lv.x, Lv.x, n, &nrm2));
    if (!(nrm2 > 0)) return femo_fail("multigrid: zero vector in the power iteration");
    if (it > 0) { lam_prev = lam; lam = sqrt(nrm2); }
    if (it == max_it || (it >= min_it && fabs(lam - lam_prev) <= 1e-2 * lam)) break;
    k_axpby<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(n, 1.0 / sqrt(nrm2), Lv.x, 0.0, Lv.x, Lv.b);
    TRY(mg_spmv(c, Lv.b, Lv.t, Lv.vals32));
    k_cheb_step<<<(unsigned)cdiv(c->nq2, 128), 128, 0, c->stream>>>(Lv.t, nullptr, 0.0, 0.0, 1.0, 1, c->dinv6, c->dinv3, c->nn,
                                                                    c->nq2, Lv.r, Lv.d, Lv.x);
    TRY(check_launch(c, "mg power iteration"));
  }
  CUDA_TRY(cudaMemcpyAsync(Lv.ev, Lv.b, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));   // last normalised iterate
  Lv.lmax = lam;
  Lv.power_its = its_done;
  return 0;
}

int mg_coarse_factor(femo_mg* m) {
  femo_ctx* c = m->L.back().c;
  if (!m->cs) TRY(coarse_analyse(c, &m->cs));
  return coarse_factor(c, m->cs);
}

}  // namespace

int mg_num_levels(femo_mg* m) { return m ? (int)m->L.size() : 0; }
femo_ctx* mg_level_ctx(femo_mg* m, int l) { return m->L[l].c; }
void mg_invalidate(femo_mg* m) { if (m) m->ready = false; }
extern "C" {

int femo_mg_create(int n_levels, femo_ctx** ctxs, femo_mg** out) {
  if (n_levels < 2 || !ctxs || !out) return femo_fail("femo_mg_create: need at least two levels");
  femo_mg* m = new femo_mg();
  m->L.resize(n_levels);
  for (int l = 0; l < n_levels; ++l) {
    if (!ctxs[l] || !ctxs[l]->setup_done) { delete m; return femo_fail("femo_mg_create: level context not set up"); }
    if (ctxs[l]->kind != 0) { delete m; return femo_fail("femo_mg_create: RM-shell contexts only"); }
    // multi-GPU hierarchies: the finest k levels are distributed (nested partitions, dist_multigrid.py), the rest is
    // replicated on every rank; the coarsest level is solved directly and must be replicated
    if (ctxs[l]->world > 1 && l > 0 && ctxs[l - 1]->world <= 1) { delete m; return femo_fail("femo_mg_create: a distributed level below a replicated one"); }
    if (ctxs[l]->world > 1 && l == n_levels - 1) { delete m; return femo_fail("femo_mg_create: the coarsest level must be replicated (single-GPU context)"); }
    if (ctxs[l]->device != ctxs[0]->device) { delete m; return femo_fail("femo_mg_create: levels must share one device"); }
    m->L[l].c = ctxs[l];
    ctxs[l]->stream = ctxs[0]->stream;
  }
  *out = m;
  return 0;
}

static void mg_free_level(MgLevel& Lv) {
  dev_free(&Lv.P_ptr); dev_free(&Lv.P_col); dev_free(&Lv.P_w); dev_free(&Lv.R_ptr); dev_free(&Lv.R_col); dev_free(&Lv.R_w);
  dev_free(&Lv.x); dev_free(&Lv.b); dev_free(&Lv.r); dev_free(&Lv.d); dev_free(&Lv.t); dev_free(&Lv.ev); dev_free(&Lv.vals32);
}

int femo_mg_destroy(femo_mg* m) {
  if (!m) return 0;
  cudaSetDevice(m->L[0].c->device);
  for (auto& Lv : m->L) mg_free_level(Lv);
  coarse_free(m->cs); m->cs = nullptr; dev_free(&m->pr); dev_free(&m->pz); dev_free(&m->pp); dev_free(&m->pAp); dev_free(&m->px); dev_free(&m->pb); dev_free(&m->pxa); dev_free(&m->pb0);
  delete m;
  return 0;
}

// transfer between level l (fine) and l+1 (coarse): P has n_fine_blocks rows, R = P^T has n_coarse_blocks rows
int femo_mg_set_transfer(femo_mg* m, int l, const int32_t* P_ptr, const int32_t* P_col, const double* P_w,
                         const int32_t* R_ptr, const int32_t* R_col, const double* R_w) {
  if (!m || l < 0 || l + 1 >= (int)m->L.size()) return femo_fail("femo_mg_set_transfer: bad level");
  MgLevel& Lv = m->L[l];
  cudaSetDevice(Lv.c->device);
  Lv.nbf = Lv.c->nb; Lv.nbc = m->L[l + 1].c->nb;
  const int64_t nnz = P_ptr[Lv.nbf];
  if (R_ptr[Lv.nbc] != nnz) return femo_fail("femo_mg_set_transfer: P and R differ in size");
  TRY(dev_alloc(&Lv.P_ptr, (size_t)Lv.nbf + 1)); TRY(dev_alloc(&Lv.P_col, (size_t)nnz)); TRY(dev_alloc(&Lv.P_w, (size_t)nnz));
  TRY(dev_alloc(&Lv.R_ptr, (size_t)Lv.nbc + 1)); TRY(dev_alloc(&Lv.R_col, (size_t)nnz)); TRY(dev_alloc(&Lv.R_w, (size_t)nnz));
  CUDA_TRY(cudaMemcpy(Lv.P_ptr, P_ptr, sizeof(int32_t) * (Lv.nbf + 1), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(Lv.P_col, P_col, sizeof(int32_t) * nnz, cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(Lv.P_w, P_w, sizeof(double) * nnz, cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(Lv.R_ptr, R_ptr, sizeof(int32_t) * (Lv.nbc + 1), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(Lv.R_col, R_col, sizeof(int32_t) * nnz, cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(Lv.R_w, R_w, sizeof(double) * nnz, cudaMemcpyHostToDevice));
  m->ready = false;
  return 0;
}

// (re)build the preconditioner from the levels' current inputs: assemble every level, block-Jacobi inverses,
// spectral bounds, factorisation of the coarsest operator
int femo_mg_setup(femo_mg* m, int cheb_degree, int reuse_coarse_inverse) {
  if (!m) return femo_fail("null multigrid handle");
  cudaSetDevice(m->L[0].c->device);
  m->deg = cheb_degree > 0 ? cheb_degree : 3;
  const bool verbose = getenv("FEMO_VERBOSE") != nullptr;
  struct Events {                      // released on every return path
    cudaEvent_t e[4] = {nullptr, nullptr, nullptr, nullptr};
    Events() { for (auto& x : e) cudaEventCreate(&x); }
    ~Events() { for (auto& x : e) if (x) cudaEventDestroy(x); }
  } evs;
  cudaEvent_t e0 = evs.e[0], e1 = evs.e[1], e2 = evs.e[2], e3 = evs.e[3];
  cudaStream_t st = m->L[0].c->stream;
  cudaEventRecord(e0, st);
  for (size_t l = 0; l < m->L.size(); ++l) {
    femo_ctx* c = m->L[l].c;
    if (!c->jac_valid) TRY(assemble_jacobian(c));
    TRY(precond_setup(c));
    // fp32 copy for the smoothers of every level but the coarsest (FEMO_MG_FP32=0: smooth with the fp64 values)
    const char* e32 = getenv("FEMO_MG_FP32");
    const bool want32 = !(e32 && e32[0] == '0') && l + 1 < m->L.size();
    if (want32) {
      if (!m->L[l].vals32) TRY(dev_alloc(&m->L[l].vals32, (size_t)c->nnzb * 9));
      TRY(vals_to_float(c, m->L[l].vals32));
    } else dev_free(&m->L[l].vals32);
  }
  cudaEventRecord(e1, st);
  for (size_t l = 0; l < m->L.size(); ++l) {
    MgLevel& Lv = m->L[l];
    femo_ctx* c = Lv.c;
    if (l + 1 < m->L.size() && !Lv.P_ptr) return femo_fail("femo_mg_setup: transfer operators missing");
    const size_t n = (size_t)c->ndof;
    if (!Lv.x) {
      // zero-initialised: on a distributed level the ghost rows of an operator product are never written (dist_spmv),
      // and the ownership-weighted restriction multiplies those entries by zero -- they must stay finite
      double** vs[] = {&Lv.x, &Lv.b, &Lv.r, &Lv.d, &Lv.t};
      for (double** v : vs) { TRY(dev_alloc(v, n)); CUDA_TRY(cudaMemsetAsync(*v, 0, sizeof(double) * n, c->stream)); }
    }
    if (l + 1 < m->L.size()) {
      cudaStreamSynchronize(st);
      const auto t0 = std::chrono::steady_clock::now();
      TRY(mg_lambda_max(m, (int)l));
      cudaStreamSynchronize(st);
      if (verbose)
        fprintf(stderr, "[femo] mg level %zu (%lld dofs%s): lambda_max %.4f after %d power iterations, %.1f ms\n", l, (long long)c->ndof,
                c->world > 1 ? ", distributed" : "", Lv.lmax, Lv.power_its,
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    }
  }
  cudaEventRecord(e2, st);
  // a lagged coarse inverse (from the previous inputs) is still SPD: a valid, slightly weaker preconditioner
  if (!(reuse_coarse_inverse && m->cs)) TRY(mg_coarse_factor(m));
  cudaEventRecord(e3, st);
  cudaEventSynchronize(e3);
  if (verbose) {
    float a = 0, b = 0, d = 0;
    cudaEventElapsedTime(&a, e0, e1); cudaEventElapsedTime(&b, e1, e2); cudaEventElapsedTime(&d, e2, e3);
    fprintf(stderr, "[femo] mg setup: assemble + block-Jacobi %.1f ms, spectral bounds %.1f ms, coarse factorisation (n = %lld) %.1f ms\n",
            a, b, (long long)m->L.back().c->ndof, d);
  }
  if (!m->pr) {
    const size_t n = (size_t)m->L[0].c->ndof;
    double** vs[] = {&m->pr, &m->pz, &m->pp, &m->pAp, &m->px, &m->pb};
    for (double** v : vs) { TRY(dev_alloc(v, n)); CUDA_TRY(cudaMemsetAsync(*v, 0, sizeof(double) * n, m->L[0].c->stream)); }
  }
  m->ready = true;
  return 0;
}

int femo_mg_get_lambda(femo_mg* m, int l, double* lam) {
  if (!m || l < 0 || l >= (int)m->L.size()) return femo_fail("bad level");
  *lam = m->L[l].lmax;
  return 0;
}

// PCG on the finest level with one V-cycle as preconditioner.  rhs / sol: n_dof doubles (host or device);
// clamped entries of rhs are ignored (homogeneous Dirichlet), x0 = 0; zero_bc zeroes the clamped entries of sol
// (state_operation.py:217-218).
// Iterative refinement around femo_mg_solve's PCG: `steps` > 0 adds that many corrections  x += PCG(b - A x)  with the
// residual accumulated in double-double, each solved to inner_rtol relative to its own right-hand side.  relres of
// femo_mg_solve is then the TRUE residual of the result (CG's recurrence residual drifts from it by ~eps |A| |x|, and the
// true one cannot go below that either: x is stored in fp64).
int femo_mg_set_refine(femo_mg* m, int steps, double inner_rtol) {
  if (!m) return femo_fail("femo_mg_set_refine: null hierarchy");
  if (steps < 0 || steps > 8 || !(inner_rtol > 0 && inner_rtol < 1)) return femo_fail("femo_mg_set_refine: steps in [0, 8], 0 < inner_rtol < 1");
  m->refine = steps; m->refine_rtol = inner_rtol;
  return 0;
}

int femo_mg_solve(femo_mg* m, const double* rhs, double* sol, double rtol, int max_it, int* iters, double* relres,
                  int on_device, int zero_bc) {
  if (!m || !m->ready) return femo_fail("femo_mg_solve: call femo_mg_setup first (and again after the inputs change)");
  femo_ctx* c = m->L[0].c;
  cudaSetDevice(c->device);
  if (!c->jac_valid) return femo_fail("femo_mg_solve: inputs changed since femo_mg_setup");
  const int64_t n = c->ndof;
  MgLevel& L0 = m->L[0];
  CUDA_TRY(cudaMemcpyAsync(m->pb, rhs, sizeof(double) * n, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
  // b <- mask * b
  k_mg_mask<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(n, c->bcmask, m->pb);
  cudaEventRecord(c->ev0, c->stream);
  double bb = 0;
  TRY(dev_dot(c, m->pb, m->pb, n, &bb));
  const double bnorm0 = sqrt(bb);
  int it = 0;
  double rel = 0;
  // PCG on A x = pb from x = 0 until |r| <= stop (recurrence residual); returns |r| / |pb|
  auto pcg = [&](double stop, double* rel_out) -> int {
    CUDA_TRY(cudaMemsetAsync(m->px, 0, sizeof(double) * n, c->stream));
    CUDA_TRY(cudaMemcpyAsync(m->pr, m->pb, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
    double pbb = 0, rz = 0, rr = 0;
    TRY(dev_dot(c, m->pb, m->pb, n, &pbb));
    *rel_out = 0;
    if (!(pbb > 0) || sqrt(pbb) <= stop) return 0;
    bool first = true;
    while (it < max_it) {
      CUDA_TRY(cudaMemcpyAsync(L0.b, m->pr, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
      TRY(mg_vcycle(m, 0));
      m->vcycles += 1;
      double rz_new = 0;
      TRY(dev_dot(c, m->pr, L0.x, n, &rz_new));
      if (!(rz_new > 0)) return femo_fail("femo_mg_solve: preconditioner is not positive definite (r.z = " + std::to_string(rz_new) + ")");
      if (first) CUDA_TRY(cudaMemcpyAsync(m->pp, L0.x, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
      else k_axpby<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(n, 1.0, L0.x, rz_new / rz, m->pp, m->pp);
      first = false;
      rz = rz_new;
      TRY(mg_spmv(c, m->pp, m->pAp));
      double pAp = 0;
      TRY(dev_dot(c, m->pp, m->pAp, n, &pAp));
      if (!(pAp > 0)) return femo_fail("femo_mg_solve: operator is not positive definite");
      const double alpha = rz / pAp;
      k_axpby<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(n, 1.0, m->px, alpha, m->pp, m->px);
      k_axpby<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(n, 1.0, m->pr, -alpha, m->pAp, m->pr);
      c->launches += 3;
      ++it;
      TRY(dev_dot(c, m->pr, m->pr, n, &rr));
      *rel_out = sqrt(rr / pbb);
      if (!(*rel_out == *rel_out)) return femo_fail("femo_mg_solve: NaN residual");
      if (sqrt(rr) <= stop) break;
    }
    return 0;
  };
  if (bnorm0 > 0) {
    TRY(pcg(rtol * bnorm0, &rel));
    if (m->refine > 0) {
      if (!m->pxa) { TRY(dev_alloc(&m->pxa, (size_t)n)); TRY(dev_alloc(&m->pb0, (size_t)n)); }
      CUDA_TRY(cudaMemcpyAsync(m->pb0, m->pb, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
      CUDA_TRY(cudaMemcpyAsync(m->pxa, m->px, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
      for (int s = 0; s <= m->refine; ++s) {
        if (c->world > 1) TRY(halo_exchange(c, m->pxa));
        TRY(residual_dd(c, m->pxa, m->pb0, m->pb));          // the correction's right-hand side
        if (c->world > 1) TRY(halo_exchange(c, m->pb));
        double rr = 0;
        TRY(dev_dot(c, m->pb, m->pb, n, &rr));
        rel = sqrt(rr) / bnorm0;                               // true residual of the accumulated solution
        if (s == m->refine || !(rr > 0) || it >= max_it) break;
        double rel_in = 0;
        TRY(pcg(m->refine_rtol * sqrt(rr), &rel_in));
        k_axpby<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(n, 1.0, m->pxa, 1.0, m->px, m->pxa);
        c->launches += 1;
      }
      CUDA_TRY(cudaMemcpyAsync(m->px, m->pxa, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
    }
  } else {
    CUDA_TRY(cudaMemsetAsync(m->px, 0, sizeof(double) * n, c->stream));
  }
  if (zero_bc) k_mg_mask<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(n, c->bcmask, m->px);
  if (c->world > 1) TRY(halo_exchange(c, m->px));          // ghost entries of the solution follow their owners
  cudaEventRecord(c->ev1, c->stream);
  CUDA_TRY(cudaMemcpyAsync(sol, m->px, sizeof(double) * n, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev0, c->ev1);
  c->t_solve = ms; c->t_iter = it > 0 ? ms / it : 0.0;
  if (iters) *iters = it;
  if (relres) *relres = rel;
  return 0;
}

}  // extern "C"
