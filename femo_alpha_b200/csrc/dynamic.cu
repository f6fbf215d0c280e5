// Time-stepped (dynamic) Reissner-Mindlin shell: SURVEY 8 row f1, BASELINE config 4.
//
// Replaces femo_alpha/dynamic_rm_shell: PlateSim.solve_dynamic_problem (plate_sim.py:281-361) with the residual
// of PlateSim.SVK_residual (:190-215) under the implicit midpoint rule (:135-140)
//     w_mid = (w + w_old)/2,   wdot = 2/dt (w - w_old) - wdot_old,   wddot = (wdot - wdot_old)/dt,
//     R_i   = M wddot_i + K w_mid - F(f_i) - W,
// and the forward / reverse sweeps of StateOperation.apply_inverse_jacobian / compute_jacvec_product
// (state_operation_dynamic.py:140-432,447-702), TotalStrainEnergyOperation and VolumeOperation.
//
// With c = 2/dt^2:   dR/dw_cur = A = c M + K/2,   dR/dw_old = B = -c M + K/2,   dR/dwdot_old = C = -(2/dt) M.
// The reference re-assembles and LU-factorises A in every step (solveNonlinear_mod, nonlinear_utils.py:210-233)
// and stores N^2 vectors in the reverse sweep (state_operation_dynamic.py:613).  Here:
//   * K comes from the static assembly kernels, M is block-scalar (one double per BSR block: rho h int phi_a phi_b on
//     the u-u blocks, rho h hE^2 int N_a N_b on the theta-theta blocks; linear_shell_model.py:335-348) gathered in
//     the fixed cell order of the contribution list, and A is formed once per design into the buffer every solver
//     reads (`vals`), K moving to `vals_K`;
//   * a step is: load vector, one M product, one solve, one K product (shared by the strain energy of this level
//     and the right-hand side of the next);
//   * the state history (levels x n_dof) stays resident in HBM;
//   * the reverse sweep is the recurrence of the two-field system (w_i, wdot_i) with adjoints (lam_i, mu_i):
//        mu_i  = -C^T lam_{i+1} - mu_{i+1},    A^T lam_i = d_i - B^T lam_{i+1} + (2/dt)(mu_i - mu_{i+1}),
//     algebraically the same adjoints as the reference's O(N^2) store (tests/test_oracle_dynamic.py).
// Solvers for A x = b: dense inverse (cuSOLVER potrf/potri once, one GEMV per step) up to 46 000 dofs -- BASELINE
// config 4's meshes have 420 .. 5 238 dofs -- else block-Jacobi PCG warm-started from the previous level, or V-cycle
// PCG when a multigrid hierarchy is attached (its levels carry their own A).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>
#include "femo_internal.cuh"

extern "C" int femo_mg_setup(femo_mg* mg, int cheb_degree, int reuse_coarse_inverse);
extern "C" int femo_mg_solve(femo_mg* mg, const double* rhs, double* sol, double rtol, int max_it, int* iters, double* relres,
                             int on_device, int zero_bc);

struct femo_dyn {
  femo_ctx* c = nullptr;
  femo_mg* mg = nullptr;
  double dt = 0, g_acc = 0;
  int n_steps = 0, levels = 1, self_weight = 0;
  int solver_req = 0, solver = 0;
  bool ready = false, energies_valid = false;
  double* hist = nullptr;                     // levels * ndof, level-major
  double *io1 = nullptr, *io2 = nullptr;      // staging of host-side histories
  CoarseSolver* cs = nullptr;      // direct solver of the effective operator (coarse.cu)
  double *v = nullptr, *kw = nullptr, *t1 = nullptr, *t2 = nullptr, *rhs = nullptr, *W = nullptr, *mu = nullptr, *lm = nullptr;
  double* dh = nullptr;
  double *fh = nullptr, *en_dev = nullptr;    // load history (levels x n_force) and per-level w.Kw on the device
  // the forward loop of the dense solver is launch-bound (13 small launches per step): captured once into a CUDA
  // graph and replayed by every later femo_dyn_solve of the same shape (FEMO_DYN_GRAPH=0 disables)
  cudaStream_t gstream = nullptr;             // capture needs a non-default stream: the context runs on it for the call
  cudaGraphExec_t fwd_graph = nullptr;
  int64_t fwd_graph_rows = -1;
  int fwd_graph_nodes = 0;
  std::vector<double> energies;
  double ms_fwd = 0, ms_adj = 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
};

namespace {

struct DevGuard {
  int prev = 0;
  explicit DevGuard(int dev) { cudaGetDevice(&prev); cudaSetDevice(dev); }
  ~DevGuard() { cudaSetDevice(prev); }
};

// run the context on another stream for the lifetime of this object (both streams drained at the switches)
struct StreamSwap {
  femo_ctx* c = nullptr;
  cudaStream_t old = nullptr;
  StreamSwap(femo_ctx* ctx, cudaStream_t s) : c(ctx), old(ctx->stream) { cudaStreamSynchronize(old); c->stream = s; }
  ~StreamSwap() { cudaStreamSynchronize(c->stream); c->stream = old; }
};

// mvals[d] = sum over the contributions (cell, a, b) of destination block d, ascending cell order, of the scalar
// element mass entry (u-u or theta-theta pairs; u-theta blocks carry no mass)
__global__ void k_mass_gather(const int32_t* __restrict__ cptr, const int32_t* __restrict__ centry,
                              const double* __restrict__ me, int64_t nnzb, double* __restrict__ mvals) {
  const int64_t d = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (d >= nnzb) return;
  double m = 0.0;
  for (int32_t k = cptr[d]; k < cptr[d + 1]; ++k) {
    const int32_t e = centry[k];
    const int64_t cell = e / 169;
    const int ab = e - (int32_t)cell * 169, a = ab / 13, b = ab - a * 13;
    if (a < 9 && b < 9) m += me[cell * 97 + 9 * a + b];
    else if (a >= 9 && b >= 9) m += me[cell * 97 + 81 + 4 * (a - 9) + (b - 9)];
  }
  mvals[d] = m;
}

// A = K/2 + c M  (M on the diagonals of the 3x3 blocks)
__global__ void k_form_eff(const double* __restrict__ K, const double* __restrict__ mvals, double cm, int64_t nnzb,
                           double* __restrict__ A) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= nnzb * 9) return;
  const int64_t d = t / 9; const int k = (int)(t - d * 9);
  A[t] = 0.5 * K[t] + ((k == 0 || k == 4 || k == 8) ? cm * mvals[d] : 0.0);
}

// y = M x, block-scalar BSR, 8 lanes per block row (fixed order: bitwise reproducible)
__global__ void __launch_bounds__(256) k_mass_spmv(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
                                                   const double* __restrict__ mvals, const double* __restrict__ x, int64_t nb,
                                                   double* __restrict__ y) {
  const int lane = threadIdx.x & 7;
  const int64_t row = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 3;
  const bool valid = row < nb;
  const int32_t s = valid ? rowptr[row] : 0, e = valid ? rowptr[row + 1] : 0;
  double y0 = 0, y1 = 0, y2 = 0;
  for (int32_t k = s + lane; k < e; k += 8) {
    const double m = mvals[k];
    const double* xp = x + 3 * (int64_t)colidx[k];
    y0 += m * xp[0]; y1 += m * xp[1]; y2 += m * xp[2];
  }
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) {
    y0 += __shfl_down_sync(0xffffffffu, y0, o, 8);
    y1 += __shfl_down_sync(0xffffffffu, y1, o, 8);
    y2 += __shfl_down_sync(0xffffffffu, y2, o, 8);
  }
  if (lane == 0 && valid) { y[3 * row] = y0; y[3 * row + 1] = y1; y[3 * row + 2] = y2; }
}

// out = m .* (a x + b y + c z + d u)   (null pointers skipped; m null = no mask)
__global__ void k_lin4(int64_t n, double a, const double* x, double b, const double* y, double c, const double* z, double d,
                       const double* u, const double* __restrict__ m, double* out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = 0.0;
  if (x) s += a * x[i];
  if (y) s += b * y[i];
  if (z) s += c * z[i];
  if (u) s += d * u[i];
  out[i] = m ? m[i] * s : s;
}
// out = m .* x + (1 - m) .* y
__global__ void k_blend(int64_t n, const double* __restrict__ m, const double* x, const double* y, double* out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) out[i] = m[i] * x[i] + (1.0 - m[i]) * y[i];
}

int lin4(femo_ctx* c, double a, const double* x, double b, const double* y, double cc, const double* z, double d,
         const double* u, const double* m, double* out) {
  k_lin4<<<(unsigned)cdiv(c->ndof, 256), 256, 0, c->stream>>>(c->ndof, a, x, b, y, cc, z, d, u, m, out);
  c->launches += 1;
  return check_launch(c, "k_lin4");
}

int mass_spmv(femo_ctx* c, const double* x, double* y) {
  k_mass_spmv<<<(unsigned)cdiv(c->nb * 8, 256), 256, 0, c->stream>>>(c->rowptr, c->colidx, c->mvals, x, c->nb, y);
  c->launches += 1;
  return check_launch(c, "k_mass_spmv");
}

int dyn_enable(femo_ctx* c, double dt) {
  // multi-GPU contexts (femo_dist_setup): every operand of an M / K product is ghost-consistent by construction
  // (solutions leave pcg_solve with their halo exchanged, and the recurrences combine them entry by entry); products
  // are complete on owned rows, which is all the owner-masked PCG and dot products read
  if (c->n_pen > 0) return femo_fail("dynamic path: strong Dirichlet BCs only (femo_set_strong_bc)");
  if (!c->vals_K) TRY(dev_alloc(&c->vals_K, (size_t)c->nnzb * 9));
  if (!c->mvals) TRY(dev_alloc(&c->mvals, (size_t)c->nnzb));
  if (!c->dyn_on || c->dyn_dt != dt) { c->jac_valid = false; c->prec_valid = false; }
  c->dyn_on = 1; c->dyn_dt = dt;
  return 0;
}

void dyn_disable(femo_ctx* c) {
  if (c->dyn_swapped) { std::swap(c->vals, c->vals_K); c->dyn_swapped = false; }
  c->dyn_on = 0;
  c->jac_valid = false; c->prec_valid = false;
}

// K -> vals_K, M -> mvals, A -> vals
int dyn_assemble(femo_ctx* c) {
  if (c->dyn_swapped) { std::swap(c->vals, c->vals_K); c->dyn_swapped = false; }
  TRY(assemble_stiffness(c));
  c->jac_valid = false;
  TRY(dev_alloc(&c->mass_me, (size_t)c->nc * 97));
  TRY(mass_elements(c, c->mass_me));
  k_mass_gather<<<(unsigned)cdiv(c->nnzb, 256), 256, 0, c->stream>>>(c->cptr, c->centry, c->mass_me, c->nnzb, c->mvals);
  TRY(check_launch(c, "k_mass_gather"));
  k_form_eff<<<(unsigned)cdiv(c->nnzb * 9, 256), 256, 0, c->stream>>>(c->vals, c->mvals, 2.0 / (c->dyn_dt * c->dyn_dt), c->nnzb,
                                                                    c->vals_K);
  TRY(check_launch(c, "k_form_eff"));
  c->launches += 2;
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  dev_free(&c->mass_me);
  std::swap(c->vals, c->vals_K);
  c->dyn_swapped = true;
  c->jac_valid = true; c->prec_valid = false;
  return 0;
}

int ensure_io(femo_dyn* d, double** buf) {
  if (!*buf) TRY(dev_alloc(buf, (size_t)d->levels * d->c->ndof));
  return 0;
}

// A x = b (b masked).  x holds the initial guess of the PCG on entry.
int dyn_linear_solve(femo_dyn* d, const double* b, double* x, double rtol, int max_it, int64_t* iters) {
  femo_ctx* c = d->c;
  int it = 0; double rr = 0;
  if (d->solver == 1) {
    TRY(coarse_apply(c, d->cs, b, x));
    it = 1;
  } else if (d->solver == 3) {
    // warm start (femo_mg_solve starts from zero): solve A delta = b - A x0 to the tolerance rtol |b| / |b - A x0|
    const int64_t n = c->ndof;
    double bb = 0, r0 = 0;
    TRY(dev_dot(c, b, b, n, &bb));
    TRY(bsr_spmv(c, x, d->t1, 1));
    TRY(lin4(c, 1.0, b, -1.0, d->t1, 0, nullptr, 0, nullptr, c->bcmask, d->t1));
    TRY(dev_dot(c, d->t1, d->t1, n, &r0));
    if (bb > 0 && r0 > rtol * rtol * bb) {
      const double rt = std::min(0.5, rtol * sqrt(bb / r0));
      TRY(femo_mg_solve(d->mg, d->t1, d->t2, rt, max_it, &it, &rr, 1, 1));
      TRY(lin4(c, 1.0, x, 1.0, d->t2, 0, nullptr, 0, nullptr, c->bcmask, x));
    } else if (!(bb > 0)) {
      CUDA_TRY(cudaMemsetAsync(x, 0, sizeof(double) * n, c->stream));
    }
  } else {
    TRY(pcg_solve(c, b, x, rtol, max_it, &it, &rr, false));
  }
  if (iters) *iters += it;
  return 0;
}

int dyn_ready(femo_dyn* d) {
  if (!d->ready || !d->c->jac_valid || !d->c->dyn_swapped) return femo_dyn_setup(d, d->solver_req);
  return 0;
}

int dyn_energies(femo_dyn* d) {
  femo_ctx* c = d->c;
  const int64_t n = c->ndof;
  d->energies.assign(d->levels, 0.0);
  for (int i = 1; i < d->levels; ++i) {
    const double* wi = d->hist + (size_t)i * n;
    TRY(bsr_spmv_vals(c, c->vals_K, wi, d->kw, 0));
    double e = 0;
    TRY(dev_dot(c, wi, d->kw, n, &e));
    d->energies[i] = 0.5 * e;
  }
  d->energies_valid = true;
  return 0;
}

}  // namespace

// assembly entry point of every solver: K, or A while the context is in dynamic mode
int assemble_jacobian(femo_ctx* c) {
  c->mat_epoch += 1;
  if (c->kind == 1) return hex_assemble(c);
  if (c->dyn_on) return dyn_assemble(c);
  return assemble_stiffness(c);
}

extern "C" {

int femo_dyn_create(femo_ctx* c, double dt, int n_steps, int add_self_weight, double g_factor, femo_dyn** out) {
  if (!c || !c->setup_done || !out) return femo_fail("femo_dyn_create: context not set up");
  if (c->kind != 0) return femo_fail("femo_dyn_create: RM-shell contexts only");
  if (!(dt > 0) || n_steps < 1) return femo_fail("femo_dyn_create: need dt > 0 and n_steps >= 1");
  DevGuard g(c->device);
  TRY(dyn_enable(c, dt));
  femo_dyn* d = new femo_dyn();
  d->c = c; d->dt = dt; d->n_steps = n_steps; d->levels = n_steps + 1;
  d->self_weight = add_self_weight ? 1 : 0;
  d->g_acc = add_self_weight ? g_factor * 9.81 : 0.0;      // plate_sim.py:205-209
  const size_t n = (size_t)c->ndof;
  int r = dev_alloc(&d->hist, n * d->levels);
  double** vecs[] = {&d->v, &d->kw, &d->t1, &d->t2, &d->rhs, &d->W, &d->mu, &d->lm};
  for (double** v : vecs) if (r == 0) r = dev_alloc(v, n);
  if (r == 0) r = dev_alloc(&d->dh, (size_t)c->n_mat);
  if (r != 0) { femo_dyn_destroy(d); return r; }
  cudaMemsetAsync(d->hist, 0, sizeof(double) * n * d->levels, c->stream);
  cudaEventCreate(&d->e0); cudaEventCreate(&d->e1);
  d->energies.assign(d->levels, 0.0);
  *out = d;
  return 0;
}

int femo_dyn_destroy(femo_dyn* d) {
  if (!d) return 0;
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  cudaStreamSynchronize(c->stream);
  if (d->mg) { for (int l = 0; l < mg_num_levels(d->mg); ++l) dyn_disable(mg_level_ctx(d->mg, l)); mg_invalidate(d->mg); }
  dyn_disable(c);
  double** vecs[] = {&d->hist, &d->io1, &d->io2, &d->v, &d->kw, &d->t1, &d->t2, &d->rhs, &d->W, &d->mu, &d->lm, &d->dh,
                     &d->fh, &d->en_dev};
  for (double** v : vecs) dev_free(v);
  coarse_free(d->cs); d->cs = nullptr;
  if (d->fwd_graph) cudaGraphExecDestroy(d->fwd_graph);
  if (d->gstream) cudaStreamDestroy(d->gstream);
  if (d->e0) cudaEventDestroy(d->e0);
  if (d->e1) cudaEventDestroy(d->e1);
  delete d;
  return 0;
}

int femo_dyn_set_mg(femo_dyn* d, femo_mg* mg) {
  if (!d) return femo_fail("null dynamic handle");
  DevGuard g(d->c->device);
  if (mg) {
    if (mg_num_levels(mg) < 2 || mg_level_ctx(mg, 0) != d->c) return femo_fail("femo_dyn_set_mg: level 0 of the hierarchy must be this context");
    for (int l = 0; l < mg_num_levels(mg); ++l) TRY(dyn_enable(mg_level_ctx(mg, l), d->dt));
    mg_invalidate(mg);
  }
  d->mg = mg; d->ready = false;
  return 0;
}

int femo_dyn_setup(femo_dyn* d, int solver) {
  if (!d) return femo_fail("null dynamic handle");
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  if (!c->dyn_on) return femo_fail("femo_dyn_setup: context left dynamic mode");
  d->solver_req = solver;
  int s = solver;
  if (c->world > 1) {
    if (s == 1 || s == 3) return femo_fail("femo_dyn_setup: multi-GPU contexts step with the distributed block-Jacobi PCG (solver 0 or 2)");
    s = 2;
  }
  if (s == 0) s = (c->ndof <= 46000) ? 1 : (d->mg ? 3 : 2);
  if (s == 3 && !d->mg) return femo_fail("femo_dyn_setup: no multigrid hierarchy attached (femo_dyn_set_mg)");
  if (s < 1 || s > 3) return femo_fail("femo_dyn_setup: solver must be 0 (auto), 1 (dense), 2 (PCG) or 3 (multigrid)");
  if (s == 3) {
    for (int l = 0; l < mg_num_levels(d->mg); ++l) mg_level_ctx(d->mg, l)->jac_valid = false;
    TRY(femo_mg_setup(d->mg, 3, 0));          // assembles A on every level through assemble_jacobian()
  } else {
    TRY(dyn_assemble(c));
    if (s == 1) { if (!d->cs) TRY(coarse_analyse(c, &d->cs)); TRY(coarse_factor(c, d->cs)); }
    else TRY(precond_setup(c));
  }
  if (d->self_weight) TRY(self_weight_load(c, d->g_acc, d->W));
  else CUDA_TRY(cudaMemsetAsync(d->W, 0, sizeof(double) * c->ndof, c->stream));
  d->solver = s; d->ready = true; d->energies_valid = false;
  if (d->fwd_graph) { cudaGraphExecDestroy(d->fwd_graph); d->fwd_graph = nullptr; }     // buffers may have moved
  return 0;
}

int femo_dyn_solve(femo_dyn* d, const double* f_history, int64_t n_rows, double rtol, int max_it, int64_t* total_iters,
                   int on_device) {
  if (!d || !f_history || n_rows < 1) return femo_fail("femo_dyn_solve: bad arguments");
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  TRY(dyn_ready(d));
  const int64_t n = c->ndof;
  const double cm = 2.0 / (d->dt * d->dt), c2 = 2.0 / d->dt;
  int64_t iters = 0;
  const char* genv = getenv("FEMO_DYN_GRAPH");
  const bool use_graph = d->solver == 1 && !(genv && genv[0] == '0');
  if (use_graph && !d->gstream) CUDA_TRY(cudaStreamCreateWithFlags(&d->gstream, cudaStreamNonBlocking));
  StreamSwap sw(c, use_graph ? d->gstream : c->stream);
  // the rows of the load history that are used (row min(i, n_rows - 1) at level i, plate_sim.py:314-319) go to the
  // device in one copy; per-level strain energies are reduced on the device and read back once
  const int64_t rows = std::min<int64_t>(n_rows, d->levels);
  if (!d->fh) TRY(dev_alloc(&d->fh, (size_t)d->levels * c->n_f));
  if (!d->en_dev) TRY(dev_alloc(&d->en_dev, (size_t)d->levels));
  CUDA_TRY(cudaMemcpyAsync(d->fh, f_history, sizeof(double) * rows * c->n_f,
                           on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
  CUDA_TRY(cudaMemsetAsync(d->en_dev, 0, sizeof(double) * d->levels, c->stream));
  d->energies.assign(d->levels, 0.0);

  auto issue_steps = [&]() -> int {
    CUDA_TRY(cudaMemsetAsync(d->hist, 0, sizeof(double) * n, c->stream));    // zero initial conditions (plate_sim.py:305-309)
    CUDA_TRY(cudaMemsetAsync(d->v, 0, sizeof(double) * n, c->stream));
    CUDA_TRY(cudaMemsetAsync(d->kw, 0, sizeof(double) * n, c->stream));
    for (int i = 1; i < d->levels; ++i) {
      const int64_t row = std::min<int64_t>(i, rows - 1);
      CUDA_TRY(cudaMemcpyAsync(c->f, d->fh + row * c->n_f, sizeof(double) * c->n_f, cudaMemcpyDeviceToDevice, c->stream));
      TRY(assemble_load(c, c->b));
      const double* wo = d->hist + (size_t)(i - 1) * n;
      double* wi = d->hist + (size_t)i * n;
      // rhs = F + W - B w_old - C v_old = F + W - K w_old / 2 + M (c w_old + 2/dt v_old)
      TRY(lin4(c, cm, wo, c2, d->v, 0, nullptr, 0, nullptr, nullptr, d->t1));
      TRY(mass_spmv(c, d->t1, d->t2));
      TRY(lin4(c, 1.0, c->b, 1.0, d->W, -0.5, d->kw, 1.0, d->t2, c->bcmask, d->rhs));
      if (d->solver != 1)                                                    // warm start of the iterative solvers
        CUDA_TRY(cudaMemcpyAsync(wi, wo, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
      TRY(dyn_linear_solve(d, d->rhs, wi, rtol, max_it, &iters));
      TRY(lin4(c, c2, wi, -c2, wo, -1.0, d->v, 0, nullptr, nullptr, d->v));   // wdot_i (plate_sim.py:243-244)
      TRY(bsr_spmv_vals(c, c->vals_K, wi, d->kw, 0));
      if (c->world > 1) {                                                     // owner-masked, summed over ranks
        double e = 0;
        TRY(dev_dot(c, wi, d->kw, n, &e));
        d->energies[i] = e;
      } else {
        TRY(dev_dot_async(c, wi, d->kw, n, d->en_dev + i));                   // 2 x strain energy (plate_sim.py:337)
      }
    }
    return 0;
  };

  cudaEventRecord(d->e0, c->stream);
  if (use_graph) {
    if (!d->fwd_graph || d->fwd_graph_rows != rows) {
      if (d->fwd_graph) { cudaGraphExecDestroy(d->fwd_graph); d->fwd_graph = nullptr; }
      const int64_t l0 = c->launches;
      cudaGraph_t graph = nullptr;
      CUDA_TRY(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
      const int rc = issue_steps();
      const cudaError_t ce = cudaStreamEndCapture(c->stream, &graph);
      if (rc != 0) { if (graph) cudaGraphDestroy(graph); return rc; }
      if (ce != cudaSuccess) return femo_fail(std::string("femo_dyn_solve: stream capture failed: ") + cudaGetErrorString(ce));
      const cudaError_t ie = cudaGraphInstantiate(&d->fwd_graph, graph, 0);
      cudaGraphDestroy(graph);
      if (ie != cudaSuccess) { d->fwd_graph = nullptr; return femo_fail(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(ie)); }
      d->fwd_graph_rows = rows;
      d->fwd_graph_nodes = (int)(c->launches - l0);
      c->launches = l0;
    }
    CUDA_TRY(cudaGraphLaunch(d->fwd_graph, c->stream));
    c->launches += d->fwd_graph_nodes;
    iters = d->n_steps;
  } else {
    TRY(issue_steps());
  }
  cudaEventRecord(d->e1, c->stream);
  if (c->world == 1) {
    d->energies.assign(d->levels, 0.0);
    CUDA_TRY(cudaMemcpyAsync(d->energies.data(), d->en_dev, sizeof(double) * d->levels, cudaMemcpyDeviceToHost, c->stream));
  }
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  for (double& e : d->energies) e *= 0.5;
  float ms = 0; cudaEventElapsedTime(&ms, d->e0, d->e1); d->ms_fwd = ms;
  d->energies_valid = true;
  if (total_iters) *total_iters = iters;
  return 0;
}

int femo_dyn_get_history(femo_dyn* d, double* w_hist, int on_device) {
  if (!d || !w_hist) return femo_fail("femo_dyn_get_history: bad arguments");
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  CUDA_TRY(cudaMemcpyAsync(w_hist, d->hist, sizeof(double) * c->ndof * d->levels,
                           on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

int femo_dyn_set_history(femo_dyn* d, const double* w_hist, int on_device) {
  if (!d || !w_hist) return femo_fail("femo_dyn_set_history: bad arguments");
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  CUDA_TRY(cudaMemcpyAsync(d->hist, w_hist, sizeof(double) * c->ndof * d->levels,
                           on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  d->energies_valid = false;
  return 0;
}

int femo_dyn_get_energies(femo_dyn* d, double* energies) {
  if (!d || !energies) return femo_fail("femo_dyn_get_energies: bad arguments");
  DevGuard g(d->c->device);
  TRY(dyn_ready(d));
  if (!d->energies_valid) TRY(dyn_energies(d));
  for (int i = 0; i < d->levels; ++i) energies[i] = d->energies[i];
  return 0;
}

int femo_dyn_solve_adjoint(femo_dyn* d, const double* d_out_hist, double* lam_hist, double rtol, int max_it,
                           int64_t* total_iters, int on_device) {
  if (!d || !d_out_hist || !lam_hist) return femo_fail("femo_dyn_solve_adjoint: bad arguments");
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  TRY(dyn_ready(d));
  const int64_t n = c->ndof;
  const double cm = 2.0 / (d->dt * d->dt), c2 = 2.0 / d->dt;
  const double* D = d_out_hist;
  double* L = lam_hist;
  if (!on_device) {
    TRY(ensure_io(d, &d->io1)); TRY(ensure_io(d, &d->io2));
    CUDA_TRY(cudaMemcpyAsync(d->io1, d_out_hist, sizeof(double) * n * d->levels, cudaMemcpyHostToDevice, c->stream));
    D = d->io1; L = d->io2;
  }
  int64_t iters = 0;
  cudaEventRecord(d->e0, c->stream);
  CUDA_TRY(cudaMemsetAsync(L, 0, sizeof(double) * n, c->stream));            // level 0: state_operation_dynamic.py:600-601
  CUDA_TRY(cudaMemsetAsync(d->mu, 0, sizeof(double) * n, c->stream));
  for (int i = d->levels - 1; i >= 1; --i) {
    const double* di = D + (size_t)i * n;
    double* li = L + (size_t)i * n;
    if (i == d->levels - 1) {
      TRY(lin4(c, 1.0, di, 0, nullptr, 0, nullptr, 0, nullptr, c->bcmask, d->rhs));
    } else {
      const double* ln = L + (size_t)(i + 1) * n;                             // clamped entries already zero
      TRY(mass_spmv(c, ln, d->t1));                                           // M lam_{i+1}
      TRY(bsr_spmv_vals(c, c->vals_K, ln, d->t2, 0));                         // K lam_{i+1}
      // mu_i = -C^T lam_{i+1} - mu_{i+1} = 2/dt M lam_{i+1} - mu_{i+1}
      TRY(lin4(c, c2, d->t1, -1.0, d->mu, 0, nullptr, 0, nullptr, nullptr, d->lm));
      // rhs = d_i - B^T lam_{i+1} + 2/dt (mu_i - mu_{i+1}),  B^T lam = K lam / 2 - c M lam
      TRY(lin4(c, -0.5, d->t2, cm, d->t1, c2, d->lm, -c2, d->mu, nullptr, d->t2));
      TRY(lin4(c, 1.0, di, 1.0, d->t2, 0, nullptr, 0, nullptr, c->bcmask, d->rhs));
      std::swap(d->mu, d->lm);
    }
    // initial guess of the iterative solvers: the adjoint of the level above (smooth in time)
    if (i == d->levels - 1) CUDA_TRY(cudaMemsetAsync(li, 0, sizeof(double) * n, c->stream));
    else CUDA_TRY(cudaMemcpyAsync(li, L + (size_t)(i + 1) * n, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
    TRY(dyn_linear_solve(d, d->rhs, li, rtol, max_it, &iters));
    TRY(lin4(c, 1.0, li, 0, nullptr, 0, nullptr, 0, nullptr, c->bcmask, li));
  }
  cudaEventRecord(d->e1, c->stream);
  if (!on_device) CUDA_TRY(cudaMemcpyAsync(lam_hist, L, sizeof(double) * n * d->levels, cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  float ms = 0; cudaEventElapsedTime(&ms, d->e0, d->e1); d->ms_adj = ms;
  if (total_iters) *total_iters = iters;
  return 0;
}

int femo_dyn_solve_forward(femo_dyn* d, const double* d_res_hist, double* x_hist, double rtol, int max_it,
                           int64_t* total_iters, int on_device) {
  if (!d || !d_res_hist || !x_hist) return femo_fail("femo_dyn_solve_forward: bad arguments");
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  TRY(dyn_ready(d));
  const int64_t n = c->ndof;
  const double cm = 2.0 / (d->dt * d->dt), c2 = 2.0 / d->dt;
  const double* D = d_res_hist;
  double* X = x_hist;
  if (!on_device) {
    TRY(ensure_io(d, &d->io1)); TRY(ensure_io(d, &d->io2));
    CUDA_TRY(cudaMemcpyAsync(d->io1, d_res_hist, sizeof(double) * n * d->levels, cudaMemcpyHostToDevice, c->stream));
    D = d->io1; X = d->io2;
  }
  int64_t iters = 0;
  CUDA_TRY(cudaMemcpyAsync(X, D, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));   // level 0 is the identity
  CUDA_TRY(cudaMemsetAsync(d->mu, 0, sizeof(double) * n, c->stream));                         // d wdot
  for (int i = 1; i < d->levels; ++i) {
    const double* xo = X + (size_t)(i - 1) * n;
    const double* di = D + (size_t)i * n;
    double* xi = X + (size_t)i * n;
    TRY(lin4(c, cm, xo, c2, d->mu, 0, nullptr, 0, nullptr, nullptr, d->t1));
    TRY(mass_spmv(c, d->t1, d->t2));
    TRY(bsr_spmv_vals(c, c->vals_K, xo, d->t1, 0));
    TRY(lin4(c, 1.0, di, 1.0, d->t2, -0.5, d->t1, 0, nullptr, c->bcmask, d->rhs));
    CUDA_TRY(cudaMemsetAsync(d->lm, 0, sizeof(double) * n, c->stream));
    TRY(dyn_linear_solve(d, d->rhs, d->lm, rtol, max_it, &iters));
    k_blend<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(n, c->bcmask, d->lm, di, xi);
    TRY(check_launch(c, "k_blend"));
    c->launches += 1;
    TRY(lin4(c, c2, xi, -c2, xo, -1.0, d->mu, 0, nullptr, nullptr, d->mu));
  }
  if (!on_device) CUDA_TRY(cudaMemcpyAsync(x_hist, X, sizeof(double) * n * d->levels, cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  if (total_iters) *total_iters = iters;
  return 0;
}

int femo_dyn_dRdw_apply(femo_dyn* d, const double* x_hist, double* y_hist, int transpose, int on_device) {
  if (!d || !x_hist || !y_hist) return femo_fail("femo_dyn_dRdw_apply: bad arguments");
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  TRY(dyn_ready(d));
  const int64_t n = c->ndof;
  const double cm = 2.0 / (d->dt * d->dt), c2 = 2.0 / d->dt;
  const double* X = x_hist;
  double* Y = y_hist;
  if (!on_device) {
    TRY(ensure_io(d, &d->io1)); TRY(ensure_io(d, &d->io2));
    CUDA_TRY(cudaMemcpyAsync(d->io1, x_hist, sizeof(double) * n * d->levels, cudaMemcpyHostToDevice, c->stream));
    X = d->io1; Y = d->io2;
  }
  CUDA_TRY(cudaMemsetAsync(d->mu, 0, sizeof(double) * n, c->stream));
  if (!transpose) {
    CUDA_TRY(cudaMemcpyAsync(Y, X, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
    for (int i = 1; i < d->levels; ++i) {
      const double* xo = X + (size_t)(i - 1) * n;
      const double* xi = X + (size_t)i * n;
      double* yi = Y + (size_t)i * n;
      // y_i = A_bc x_i + mask (B x_{i-1} + C dwdot_{i-1})
      TRY(lin4(c, cm, xo, c2, d->mu, 0, nullptr, 0, nullptr, nullptr, d->t1));
      TRY(mass_spmv(c, d->t1, d->t2));
      TRY(bsr_spmv_vals(c, c->vals_K, xo, d->t1, 0));
      TRY(bsr_spmv(c, xi, d->rhs, 1));
      TRY(lin4(c, 0.5, d->t1, -1.0, d->t2, 0, nullptr, 0, nullptr, c->bcmask, d->lm));
      TRY(lin4(c, 1.0, d->rhs, 1.0, d->lm, 0, nullptr, 0, nullptr, nullptr, yi));
      TRY(lin4(c, c2, xi, -c2, xo, -1.0, d->mu, 0, nullptr, nullptr, d->mu));
    }
  } else {
    // y_i = A_bc x_i + B^T (mask x_{i+1}) + 2/dt (nu_{i+1} - nu_i),  nu_i = 2/dt M (mask x_{i+1}) - nu_{i+1};  level 0: identity
    for (int i = d->levels - 1; i >= 0; --i) {
      const double* xi = X + (size_t)i * n;
      double* yi = Y + (size_t)i * n;
      if (i > 0) TRY(bsr_spmv(c, xi, d->rhs, 1));
      else CUDA_TRY(cudaMemcpyAsync(d->rhs, xi, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
      if (i == d->levels - 1) {
        CUDA_TRY(cudaMemcpyAsync(yi, d->rhs, sizeof(double) * n, cudaMemcpyDeviceToDevice, c->stream));
        continue;
      }
      TRY(lin4(c, 1.0, X + (size_t)(i + 1) * n, 0, nullptr, 0, nullptr, 0, nullptr, c->bcmask, d->lm));
      TRY(mass_spmv(c, d->lm, d->t1));
      TRY(bsr_spmv_vals(c, c->vals_K, d->lm, d->t2, 0));
      // nu_i (level 0 has no S_0 row: only the + 2/dt nu_1 term remains)
      if (i > 0) {
        TRY(lin4(c, c2, d->t1, -1.0, d->mu, 0, nullptr, 0, nullptr, nullptr, d->lm));        // nu_i
        TRY(lin4(c, 0.5, d->t2, -cm, d->t1, c2, d->mu, -c2, d->lm, nullptr, d->t2));
        TRY(lin4(c, 1.0, d->rhs, 1.0, d->t2, 0, nullptr, 0, nullptr, nullptr, yi));
        std::swap(d->mu, d->lm);
      } else {
        TRY(lin4(c, 0.5, d->t2, -cm, d->t1, c2, d->mu, 0, nullptr, nullptr, d->t2));
        TRY(lin4(c, 1.0, d->rhs, 1.0, d->t2, 0, nullptr, 0, nullptr, nullptr, yi));
      }
    }
  }
  if (!on_device) CUDA_TRY(cudaMemcpyAsync(y_hist, Y, sizeof(double) * n * d->levels, cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

int femo_dyn_dRdh_T_apply(femo_dyn* d, const double* lam_hist, double* out, int on_device) {
  if (!d || !lam_hist || !out) return femo_fail("femo_dyn_dRdh_T_apply: bad arguments");
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  TRY(dyn_ready(d));
  const int64_t n = c->ndof;
  const double cm = 2.0 / (d->dt * d->dt), c2 = 2.0 / d->dt;
  const double* L = lam_hist;
  if (!on_device) {
    TRY(ensure_io(d, &d->io1));
    CUDA_TRY(cudaMemcpyAsync(d->io1, lam_hist, sizeof(double) * n * d->levels, cudaMemcpyHostToDevice, c->stream));
    L = d->io1;
  }
  CUDA_TRY(cudaMemsetAsync(d->mu, 0, sizeof(double) * n, c->stream));         // wdot_{i-1}
  for (int i = 1; i < d->levels; ++i) {
    const double* wo = d->hist + (size_t)(i - 1) * n;
    const double* wi = d->hist + (size_t)i * n;
    // dR_i/dh at (w_i, w_{i-1}, wdot_{i-1}): dK/dh w_mid + dM/dh wddot_i - dW/dh, BC rows zeroed
    TRY(lin4(c, 0.5, wi, 0.5, wo, 0, nullptr, 0, nullptr, nullptr, d->t1));
    TRY(lin4(c, cm, wi, -cm, wo, -c2, d->mu, 0, nullptr, nullptr, d->t2));
    TRY(lin4(c, 1.0, L + (size_t)i * n, 0, nullptr, 0, nullptr, 0, nullptr, c->bcmask, d->lm));
    TRY(stiff_param_T(c, d->t1, d->lm, i > 1));
    TRY(mass_param_T(c, d->t2, d->lm, d->g_acc, true));
    TRY(lin4(c, c2, wi, -c2, wo, -1.0, d->mu, 0, nullptr, nullptr, d->mu));
  }
  TRY(cellbuf_to_material(c, d->dh));
  CUDA_TRY(cudaMemcpyAsync(out, d->dh, sizeof(double) * c->n_mat, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost,
                           c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

int femo_dyn_dRdh_apply(femo_dyn* d, const double* dh, double* out_hist, int on_device) {
  if (!d || !dh || !out_hist) return femo_fail("femo_dyn_dRdh_apply: bad arguments");
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  TRY(dyn_ready(d));
  const int64_t n = c->ndof;
  const double cm = 2.0 / (d->dt * d->dt), c2 = 2.0 / d->dt;
  double* Y = out_hist;
  if (!on_device) { TRY(ensure_io(d, &d->io2)); Y = d->io2; }
  CUDA_TRY(cudaMemcpyAsync(d->dh, dh, sizeof(double) * c->n_mat, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                           c->stream));
  CUDA_TRY(cudaMemsetAsync(Y, 0, sizeof(double) * n, c->stream));
  CUDA_TRY(cudaMemsetAsync(d->mu, 0, sizeof(double) * n, c->stream));
  for (int i = 1; i < d->levels; ++i) {
    const double* wo = d->hist + (size_t)(i - 1) * n;
    const double* wi = d->hist + (size_t)i * n;
    TRY(lin4(c, 0.5, wi, 0.5, wo, 0, nullptr, 0, nullptr, nullptr, d->t1));
    TRY(lin4(c, cm, wi, -cm, wo, -c2, d->mu, 0, nullptr, nullptr, d->t2));
    TRY(stiff_param_fwd(c, d->t1, d->dh, d->rhs));
    TRY(mass_param_fwd(c, d->t2, d->dh, d->g_acc, d->lm));
    TRY(lin4(c, 1.0, d->rhs, 1.0, d->lm, 0, nullptr, 0, nullptr, c->bcmask, Y + (size_t)i * n));
    TRY(lin4(c, c2, wi, -c2, wo, -1.0, d->mu, 0, nullptr, nullptr, d->mu));
  }
  if (!on_device) CUDA_TRY(cudaMemcpyAsync(out_hist, Y, sizeof(double) * n * d->levels, cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

int femo_dyn_total_strain_energy(femo_dyn* d, double* value) {
  if (!d || !value) return femo_fail("femo_dyn_total_strain_energy: bad arguments");
  DevGuard g(d->c->device);
  TRY(dyn_ready(d));
  if (!d->energies_valid) TRY(dyn_energies(d));
  double t = 0;
  for (int i = 0; i < d->levels; ++i) t += d->energies[i];      // total_strain_energy_operation.py:68-71
  *value = t;
  return 0;
}

int femo_dyn_total_strain_energy_grad(femo_dyn* d, int arg, double* gout, int on_device) {
  if (!d || !gout) return femo_fail("femo_dyn_total_strain_energy_grad: bad arguments");
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  TRY(dyn_ready(d));
  const int64_t n = c->ndof;
  if (arg == FEMO_ARG_STATE) {           // dE/dw_i = K w_i   (total_strain_energy_operation.py:96-99)
    double* G = gout;
    if (!on_device) { TRY(ensure_io(d, &d->io2)); G = d->io2; }
    for (int i = 0; i < d->levels; ++i) TRY(bsr_spmv_vals(c, c->vals_K, d->hist + (size_t)i * n, G + (size_t)i * n, 0));
    if (!on_device) CUDA_TRY(cudaMemcpyAsync(gout, G, sizeof(double) * n * d->levels, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return 0;
  }
  if (arg != FEMO_THICKNESS) return femo_fail("femo_dyn_total_strain_energy_grad: arg must be FEMO_THICKNESS or FEMO_ARG_STATE");
  // dE/dh = sum_i 1/2 w_i^T dK/dh w_i
  for (int i = 1; i < d->levels; ++i) {
    const double* wi = d->hist + (size_t)i * n;
    TRY(lin4(c, 0.5, wi, 0, nullptr, 0, nullptr, 0, nullptr, nullptr, d->t1));
    TRY(stiff_param_T(c, d->t1, wi, i > 1));
  }
  TRY(cellbuf_to_material(c, d->dh));
  CUDA_TRY(cudaMemcpyAsync(gout, d->dh, sizeof(double) * c->n_mat, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost,
                           c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

int femo_dyn_mass_apply(femo_dyn* d, const double* x, double* y, int on_device) {
  if (!d || !x || !y) return femo_fail("femo_dyn_mass_apply: bad arguments");
  femo_ctx* c = d->c;
  DevGuard g(c->device);
  TRY(dyn_ready(d));
  CUDA_TRY(cudaMemcpyAsync(d->t1, x, sizeof(double) * c->ndof, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
  TRY(mass_spmv(c, d->t1, d->t2));
  CUDA_TRY(cudaMemcpyAsync(y, d->t2, sizeof(double) * c->ndof, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

int femo_dyn_info(femo_dyn* d, double* ms_forward, double* ms_adjoint, int* solver) {
  if (!d) return femo_fail("null dynamic handle");
  if (ms_forward) *ms_forward = d->ms_fwd;
  if (ms_adjoint) *ms_adjoint = d->ms_adj;
  if (solver) *solver = d->solver;
  return 0;
}

}  // extern "C"
