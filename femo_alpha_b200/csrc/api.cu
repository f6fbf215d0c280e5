// extern "C" surface of libfemo_b200.so (see include/femo_b200.h) + context management.
#include <cmath>
#include <cstring>
#include "femo_internal.cuh"
#include "tables.hpp"

static thread_local std::string g_err;
void femo_set_error(const std::string& s) { g_err = s; }
int femo_fail(const std::string& s) { g_err = s; return 1; }

int check_launch(femo_ctx* c, const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return femo_fail(std::string(what) + ": " + cudaGetErrorString(e));
  (void)c;
  return 0;
}

namespace {

struct DevGuard {
  int prev = 0;
  explicit DevGuard(int dev) { cudaGetDevice(&prev); cudaSetDevice(dev); }
  ~DevGuard() { cudaSetDevice(prev); }
};


__global__ void k_fill(double* p, int64_t n, double v) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
__global__ void k_mask_clear(double* mask, const int32_t* dofs, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) mask[dofs[i]] = 0.0;
}
__global__ void k_bc_bits(const double* __restrict__ mask, int64_t nb, uint8_t* __restrict__ bits) {
  int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (b < nb) bits[b] = (uint8_t)((mask[3 * b] == 0.0 ? 1 : 0) | (mask[3 * b + 1] == 0.0 ? 2 : 0) | (mask[3 * b + 2] == 0.0 ? 4 : 0));
}
__global__ void k_col_bc(const int32_t* __restrict__ colidx, const uint8_t* __restrict__ bits, int64_t nnzb,
                         uint32_t* __restrict__ colbc) {
  int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k < nnzb) { const uint32_t col = (uint32_t)colidx[k]; colbc[k] = col | ((uint32_t)bits[col] << 29); }
}
__global__ void k_pick_blocks(const double* __restrict__ vals, const int64_t* __restrict__ pos, int64_t n, double* __restrict__ out) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t < 9 * n) out[t] = vals[pos[t / 9] * 9 + t % 9];
}
__global__ void k_mul(const double* a, const double* m, int64_t n, double* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) out[i] = a[i] * m[i];
}
__global__ void k_axpby(double a, const double* x, double b, const double* y, int64_t n, double* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) out[i] = a * x[i] + b * y[i];
}
__global__ void k_block_coords(const double* xyz, const int32_t* edges, const int32_t* cv, int64_t nn, int64_t ne,
                               int64_t nc, double* out) {
  int64_t I = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (I >= 2 * nn + ne + nc) return;
  const V3 p = block_coord(xyz, edges, cv, nn, ne, nc, I);
  out[3 * I] = p.x; out[3 * I + 1] = p.y; out[3 * I + 2] = p.z;
}
__global__ void k_bsr_apply_bc(const int32_t* rowptr, const int32_t* colidx, const double* mask, int64_t nb, double* vals) {
  int64_t row = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (row >= nb) return;
  for (int32_t k = rowptr[row]; k < rowptr[row + 1]; ++k) {
    int64_t col = colidx[k];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        double mi = mask[3 * row + i], mj = mask[3 * col + j];
        double v = mi * mj * vals[(int64_t)k * 9 + 3 * i + j];
        if (row == col && i == j) v += 1.0 - mi;
        vals[(int64_t)k * 9 + 3 * i + j] = v;
      }
  }
}

int fill(femo_ctx* c, double* p, int64_t n, double v) {
  k_fill<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(p, n, v);
  c->launches += 1;
  return check_launch(c, "k_fill");
}

double** input_slot(femo_ctx* c, int which, int64_t* n) {
  switch (which) {
    case FEMO_THICKNESS: *n = c->n_mat; return &c->h;
    case FEMO_E: *n = c->n_mat; return &c->E;
    case FEMO_NU: *n = c->n_mat; return &c->nu;
    case FEMO_DENSITY: *n = c->n_mat; return &c->rho;
    case FEMO_F_SOLID: *n = c->n_f; return &c->f;
    case FEMO_UHAT: *n = 3 * c->nn; return &c->uhat;
  }
  return nullptr;
}

// copy a caller vector (host or device) into a device buffer / back
int copy_in(femo_ctx* c, double* dst, const double* src, int64_t n, int on_device) {
  CUDA_TRY(cudaMemcpyAsync(dst, src, sizeof(double) * n, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
  return 0;
}
int copy_out(femo_ctx* c, double* dst, const double* src, int64_t n, int on_device) {
  CUDA_TRY(cudaMemcpyAsync(dst, src, sizeof(double) * n, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

}  // namespace

int vec_mul(femo_ctx* c, const double* a, const double* m, int64_t n, double* out) {
  k_mul<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(a, m, n, out);
  c->launches += 1;
  return check_launch(c, "k_mul");
}
int vec_axpby(femo_ctx* c, double a, const double* x, double b, const double* y, int64_t n, double* out) {
  k_axpby<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(a, x, b, y, n, out);
  c->launches += 1;
  return check_launch(c, "k_axpby");
}

namespace {

#define NEED_SETUP(c, name) \
  if (!(c) || !(c)->setup_done) return femo_fail(std::string(name) + ": context not set up (femo_mesh_set + femo_shell_setup)")

// entry points that read shell data refuse the solid contexts of hex.cu
#define NEED_SHELL(c, name) \
  NEED_SETUP(c, name);      \
  if ((c)->kind != 0) return femo_fail(std::string(name) + ": RM-shell entry point called on a hexahedral-solid context (use femo_hex_*)")

}  // namespace

extern "C" {

const char* femo_last_error(void) { return g_err.c_str(); }
int femo_version(void) { return 100; }

int femo_ctx_create(int device, femo_ctx** out) {
  if (!out) return femo_fail("femo_ctx_create: null output pointer");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return femo_fail(std::string("femo_ctx_create: no CUDA device available (") + cudaGetErrorString(e) +
                     "); libfemo_b200 has no CPU fallback");
  if (device < 0 || device >= ndev) return femo_fail("femo_ctx_create: bad device index");
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) return femo_fail("femo_ctx_create: libfemo_b200 is built for sm_100a (Blackwell) only");
  femo_ctx* c = new femo_ctx();
  c->device = device;
  DevGuard g(device);
  CUDA_TRY(cudaEventCreate(&c->ev0));
  CUDA_TRY(cudaEventCreate(&c->ev1));
  *out = c;
  return 0;
}

int femo_ctx_destroy(femo_ctx* c) {
  if (!c) return 0;
  DevGuard g(c->device);
  cudaStreamSynchronize(c->stream);
  dist_free(c);
  tile_plan_free(c);
  coarse_free(c->direct); c->direct = nullptr;
  void* ptrs[] = {c->xyz, c->cv, c->ce, c->edges, c->cell_blocks, c->hE, c->tab, c->h, c->E, c->nu, c->rho, c->f, c->uhat,
                  c->rowptr, c->colidx, c->vals, c->centry, c->cdest, c->cptr, c->diag_pos, c->ut_pos, c->tu_pos, c->nc_ptr,
                  c->nc_adj, c->rec, c->elem, c->wp_order, c->bcmask, c->bcbits, c->colbc, c->pen_cell, c->pen_dest_pos, c->pen_dest_ptr, c->pen_src, c->pen_me,
                  c->pen_val, c->pen_dest_col, c->tab4, c->cell_tags, c->pen_row, c->pen_row_ptr, c->w, c->b, c->r, c->z, c->p, c->Ap, c->tmp,
                  c->tmp2, c->dinv6, c->dinv3, c->red, c->scal, c->cellbuf, c->it_pack, c->it_dest, c->chunk_ptr, c->halo_ptr,
                  c->halo_cells, c->tpos, c->tabm, c->vals_K, c->mvals, c->mass_me, c->hx_cells, c->hx_geo, c->hx_geo2, c->hx_rho, c->pen_vkey, c->pen_vptr, c->pen_vadj, c->pen_bkey, c->pen_bptr,
                  c->pen_badj, c->pen_buf,
                  c->hx_scale, c->hx_vol, c->hx_f, c->hx_cellvec};
  for (void* p : ptrs) if (p) cudaFree(p);
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  for (auto& e : c->ev_copy) if (e) cudaEventDestroy(e);
  delete c;
  return 0;
}

int femo_set_stream(femo_ctx* c, void* s) {
  if (!c) return femo_fail("null context");
  c->stream = (cudaStream_t)s;
  return 0;
}

int femo_mesh_set(femo_ctx* c, int64_t n_nodes, const double* xyz, int64_t n_cells, int npc, const int32_t* conn) {
  if (!c) return femo_fail("null context");
  if (c->kind != 0) return femo_fail("femo_mesh_set: this context holds a hexahedral solid (femo_hex_setup)");
  if (npc == 3) return femo_fail("femo_mesh_set: triangles (P2xP1) are not implemented; quadrilaterals only");
  if (npc != 4) return femo_fail("femo_mesh_set: nodes_per_cell must be 4");
  if (n_nodes <= 0 || n_cells <= 0 || !xyz || !conn) return femo_fail("femo_mesh_set: empty mesh");
  DevGuard g(c->device);
  c->nn = n_nodes; c->nc = n_cells; c->setup_done = false; c->jac_valid = false;
  TRY(dev_alloc(&c->xyz, (size_t)n_nodes * 3));
  CUDA_TRY(cudaMemcpy(c->xyz, xyz, sizeof(double) * n_nodes * 3, cudaMemcpyHostToDevice));
  return topo_build(c, conn);
}

int femo_shell_setup(femo_ctx* c, const femo_shell_opts* o) {
  if (!c || c->nc == 0) return femo_fail("femo_shell_setup: set the mesh first");
  DevGuard g(c->device);
  femo_shell_opts d = {0, 0, 6, 0};
  if (o) d = *o;
  int deg = d.quadrature_degree > 0 ? d.quadrature_degree : 6;
  int m = (deg + 2) / 2;
  if (m < 1 || m > FEMO_MAX_Q1D) return femo_fail("femo_shell_setup: quadrature_degree must be in [1, 9]");
  c->ewm = d.element_wise_material ? 1 : 0;
  c->ewp = d.elementwise_pressure ? 1 : 0;
  c->m1d = m; c->nq = m * m;
  build_tables(c->htab, m);
  TRY(dev_alloc(&c->tab, 1));
  CUDA_TRY(cudaMemcpy(c->tab, &c->htab, sizeof(Tables), cudaMemcpyHostToDevice));
  Tables t4;
  build_tables(t4, 3);
  TRY(dev_alloc(&c->tab4, 1));
  CUDA_TRY(cudaMemcpy(c->tab4, &t4, sizeof(Tables), cudaMemcpyHostToDevice));
  c->nq4 = 9;
  Tables tm;
  build_tables(tm, 4);            // 4x4 rule of the mass form (dynamic.cu)
  TRY(dev_alloc(&c->tabm, 1));
  CUDA_TRY(cudaMemcpy(c->tabm, &tm, sizeof(Tables), cudaMemcpyHostToDevice));
  c->area_cache.clear();
  TRY(pattern_build(c));
  if (c->all_planar) TRY(tile_plan_setup(c));
  if (c->all_planar && !c->tile_ok) TRY(asm_plan_build(c));
  c->n_mat = c->ewm ? c->nc : c->nn;
  c->n_f = 3 * (c->ewp ? c->nc : c->nn);
  TRY(dev_alloc(&c->h, c->n_mat)); TRY(dev_alloc(&c->E, c->n_mat)); TRY(dev_alloc(&c->nu, c->n_mat));
  TRY(dev_alloc(&c->rho, c->n_mat)); TRY(dev_alloc(&c->f, c->n_f)); TRY(dev_alloc(&c->uhat, 3 * c->nn));
  // initial values of fea.add_input (rm_shell_model.py:209-214); nu=1 there is singular, use 0.3
  TRY(fill(c, c->h, c->n_mat, 1e-3)); TRY(fill(c, c->E, c->n_mat, 1.0)); TRY(fill(c, c->nu, c->n_mat, 0.3));
  TRY(fill(c, c->rho, c->n_mat, 1.0)); TRY(fill(c, c->f, c->n_f, 1.0)); TRY(fill(c, c->uhat, 3 * c->nn, 0.0));
  c->uhat_zero = 1;
  c->rec_nqf = c->all_planar ? REC_NQF_PLANAR : REC_NQF_WARPED;
  c->rec_stride = REC_HEADER + c->nq * c->rec_nqf;
  c->warped_ok = 0;
  if (!c->all_planar) {
    // cell-centric kernels for warped cells (FEMO_WARPED=0: the round-1 contribution-gather kernel); the records of a
    // group of 32 cells (nq * 25 * 32 doubles) must fit a CTA's shared memory.  FEMO_WARPED_NB: node pairs per task (1..4)
    const char* ev = getenv("FEMO_WARPED");
    const char* en = getenv("FEMO_WARPED_NB");
    c->warped_nb = en ? atoi(en) : 3;
    if (c->warped_nb < 1 || c->warped_nb > 4) c->warped_nb = 3;
    const char* et = getenv("FEMO_WARPED_NBT");      // columns per rotation-rotation task (1..4)
    c->warped_nbt = et ? atoi(et) : 2;
    if (c->warped_nbt < 1 || c->warped_nbt > 4) c->warped_nbt = 2;
    if (!(ev && atoi(ev) == 0) && (size_t)c->nq * REC_NQF_WARPED * 32 * sizeof(double) <= 200 * 1024) c->warped_ok = 1;
  }
  if (c->warped_ok) {
    TRY(warped_setup(c));
  } else if (!c->tile_ok) TRY(dev_alloc(&c->rec, (size_t)c->nc * c->rec_stride));
  const int64_t n = c->ndof;
  TRY(dev_alloc(&c->bcmask, n)); TRY(fill(c, c->bcmask, n, 1.0));
  TRY(dev_alloc(&c->bcbits, (size_t)c->nb)); CUDA_TRY(cudaMemsetAsync(c->bcbits, 0, (size_t)c->nb, c->stream));
  if (c->nb >= (int64_t)1 << 29) return femo_fail("femo_shell_setup: more than 2^29 block nodes");
  TRY(dev_alloc(&c->colbc, (size_t)c->nnzb));
  k_col_bc<<<(unsigned)cdiv(c->nnzb, 256), 256, 0, c->stream>>>(c->colidx, c->bcbits, c->nnzb, c->colbc);
  TRY(check_launch(c, "k_col_bc"));
  c->n_bc = 0;
  double** vecs[] = {&c->w, &c->b, &c->r, &c->z, &c->p, &c->Ap, &c->tmp, &c->tmp2};
  for (double** v : vecs) { TRY(dev_alloc(v, n)); CUDA_TRY(cudaMemsetAsync(*v, 0, sizeof(double) * n, c->stream)); }
  TRY(dev_alloc(&c->dinv6, (size_t)c->nn * 36)); TRY(dev_alloc(&c->dinv3, (size_t)(c->ne + c->nc) * 9));
  TRY(dev_alloc(&c->red, 8 * 1184)); TRY(dev_alloc(&c->scal, 32));
  TRY(dev_alloc(&c->cellbuf, (size_t)c->nc * 39));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  noalloc_prepare(c->device);     // femo_dRdf_T_apply_noalloc: resolve its static scratch at set-up (never fails the set-up)
  c->setup_done = true; c->jac_valid = false; c->prec_valid = false;
  return 0;
}

int femo_get_sizes(femo_ctx* c, femo_sizes* s) {
  if (!c || !s) return femo_fail("null argument");
  s->n_nodes = c->nn; s->n_edges = c->ne; s->n_cells = c->nc; s->n_blocks = c->nb; s->n_dof = c->ndof;
  s->nnz_blocks = c->nnzb; s->n_contrib = c->ncontrib; s->n_material = c->n_mat; s->n_force = c->n_f;
  s->n_qp = c->nq; s->all_planar = c->all_planar;
  return 0;
}

int femo_get_block_coords(femo_ctx* c, double* out) {
  if (!c || c->nb == 0) return femo_fail("femo_get_block_coords: set the mesh first");
  DevGuard g(c->device);
  double* d = nullptr;
  TRY(dev_alloc(&d, (size_t)c->nb * 3));
  k_block_coords<<<(unsigned)cdiv(c->nb, 256), 256, 0, c->stream>>>(c->xyz, c->edges, c->cv, c->nn, c->ne, c->nc, d);
  TRY(check_launch(c, "k_block_coords"));
  c->launches += 1;
  int r = copy_out(c, out, d, c->nb * 3, 0);
  dev_free(&d);
  return r;
}

int femo_get_cell_blocks(femo_ctx* c, int32_t* out) {
  if (!c || !c->cell_blocks) return femo_fail("femo_get_cell_blocks: set the mesh first");
  DevGuard g(c->device);
  CUDA_TRY(cudaMemcpy(out, c->cell_blocks, sizeof(int32_t) * c->nc * 13, cudaMemcpyDeviceToHost));
  return 0;
}

int femo_set_strong_bc(femo_ctx* c, int64_t n, const int32_t* dofs) {
  NEED_SETUP(c, "femo_set_strong_bc");
  DevGuard g(c->device);
  TRY(fill(c, c->bcmask, c->ndof, 1.0));
  c->n_bc = n;
  if (n > 0) {
    for (int64_t i = 0; i < n; ++i)
      if (dofs[i] < 0 || dofs[i] >= c->ndof) return femo_fail("femo_set_strong_bc: dof index out of range");
    int32_t* d = nullptr;
    TRY(dev_alloc(&d, (size_t)n));
    CUDA_TRY(cudaMemcpyAsync(d, dofs, sizeof(int32_t) * n, cudaMemcpyHostToDevice, c->stream));
    k_mask_clear<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(c->bcmask, d, n);
    TRY(check_launch(c, "k_mask_clear"));
    c->launches += 1;
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    dev_free(&d);
  }
  k_bc_bits<<<(unsigned)cdiv(c->nb, 256), 256, 0, c->stream>>>(c->bcmask, c->nb, c->bcbits);
  TRY(check_launch(c, "k_bc_bits"));
  k_col_bc<<<(unsigned)cdiv(c->nnzb, 256), 256, 0, c->stream>>>(c->colidx, c->bcbits, c->nnzb, c->colbc);
  TRY(check_launch(c, "k_col_bc"));
  c->launches += 2;
  c->prec_valid = false;
  c->mat_epoch += 1;
  return 0;
}

int femo_set_penalty_bc(femo_ctx* c, int64_t n_nodes, const uint8_t* von) {
  NEED_SHELL(c, "femo_set_penalty_bc");
  if (n_nodes != c->nn) return femo_fail("femo_set_penalty_bc: vertex mask must have n_nodes entries");
  DevGuard g(c->device);
  return penalty_build(c, von);
}

int femo_set_input(femo_ctx* c, int which, const double* data, int64_t n, int on_device) {
  NEED_SHELL(c, "femo_set_input");
  DevGuard g(c->device);
  int64_t want = 0;
  double** slot = input_slot(c, which, &want);
  if (!slot) return femo_fail("femo_set_input: unknown input id");
  if (n == 1) {  // utils_dolfinx.update: len(v_values) == 1 -> v.vector.set(value)
    double v;
    if (on_device) CUDA_TRY(cudaMemcpy(&v, data, sizeof(double), cudaMemcpyDeviceToHost)); else v = data[0];
    TRY(fill(c, *slot, want, v));
    if (which == FEMO_UHAT) c->uhat_zero = (v == 0.0);
  } else {
    if (n != want) return femo_fail("femo_set_input: size mismatch (got " + std::to_string(n) + ", expected " + std::to_string(want) + ")");
    TRY(copy_in(c, *slot, data, n, on_device));
    if (which == FEMO_UHAT) {
      double nrm = 0;
      TRY(dev_dot(c, c->uhat, c->uhat, n, &nrm));
      c->uhat_zero = (nrm == 0.0);
    }
  }
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  // the dynamic operator A_eff = (2/dt^2) M + K/2 also depends on the density
  if ((which != FEMO_DENSITY || c->dyn_on) && which != FEMO_F_SOLID) { c->jac_valid = false; c->prec_valid = false; }
  return 0;
}

int femo_get_input(femo_ctx* c, int which, double* out, int64_t n, int on_device) {
  NEED_SHELL(c, "femo_get_input");
  DevGuard g(c->device);
  int64_t want = 0;
  double** slot = input_slot(c, which, &want);
  if (!slot || n != want) return femo_fail("femo_get_input: bad input id or size");
  return copy_out(c, out, *slot, n, on_device);
}

int femo_set_state(femo_ctx* c, const double* w, int on_device) {
  NEED_SETUP(c, "femo_set_state");
  DevGuard g(c->device);
  TRY(copy_in(c, c->w, w, c->ndof, on_device));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}
int femo_get_state(femo_ctx* c, double* w, int on_device) {
  NEED_SETUP(c, "femo_get_state");
  DevGuard g(c->device);
  return copy_out(c, w, c->w, c->ndof, on_device);
}

int femo_assemble_jacobian(femo_ctx* c) {
  NEED_SETUP(c, "femo_assemble_jacobian");
  DevGuard g(c->device);
  return assemble_jacobian(c);
}

int femo_assemble_load(femo_ctx* c, double* F, int on_device) {
  NEED_SETUP(c, "femo_assemble_load");
  DevGuard g(c->device);
  TRY(assemble_load(c, c->tmp));
  return copy_out(c, F, c->tmp, c->ndof, on_device);
}

int femo_assemble_residual(femo_ctx* c, const double* w, double* R, int on_device) {
  NEED_SETUP(c, "femo_assemble_residual");
  DevGuard g(c->device);
  if (!c->jac_valid) TRY(assemble_jacobian(c));
  const double* wd = c->w;
  if (w) { TRY(copy_in(c, c->tmp2, w, c->ndof, on_device)); wd = c->tmp2; }
  TRY(bsr_spmv(c, wd, c->tmp, 0));
  TRY(assemble_load(c, c->b));
  k_axpby<<<(unsigned)cdiv(c->ndof, 256), 256, 0, c->stream>>>(1.0, c->tmp, -1.0, c->b, c->ndof, c->tmp);
  TRY(check_launch(c, "k_axpby"));
  c->launches += 1;
  return copy_out(c, R, c->tmp, c->ndof, on_device);
}

// One evaluation of the residual through host buffers with the copies overlapped (bench.py's end-to-end figure; what
// StateOperation.compute_residual does per call: update every input, then assembleVector(R), state_operation.py:86-131):
//   * inputs the Jacobian depends on (thickness, E, nu, uhat) go up first on the context's stream, dR/du is assembled;
//   * the other inputs (F_solid, density) and the state go up on a second stream WHILE the assembly runs;
//   * R = K w - F is produced in row ranges and every range starts its way back to the host as soon as it is complete,
//     while the SpMV still works on the next one.
// which[i] / data[i] / sizes[i]: input ids and HOST arrays (pinned memory makes the copies asynchronous; pageable memory
// works, serialised by the driver); w_host: state (n_dof) or null for the resident one; R_host: n_dof doubles.
// On a rank of a multi-GPU context everything is local (local inputs, local state with its ghost entries, rows of the
// rank's blocks; owned rows are complete) exactly as in femo_set_input + femo_assemble_residual: no collective.
int femo_evaluate_residual(femo_ctx* c, int n_in, const int* which, const double* const* data, const int64_t* sizes,
                           const double* w_host, double* R_host) {
  NEED_SHELL(c, "femo_evaluate_residual");
  if (n_in < 0 || (n_in > 0 && (!which || !data || !sizes)) || !R_host) return femo_fail("femo_evaluate_residual: bad arguments");
  DevGuard g(c->device);
  if (!c->copy_stream) {
    CUDA_TRY(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    for (auto& e : c->ev_copy) CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  const int64_t n = c->ndof;
  bool matrix_changed = false, uhat_given = false;
  for (int pass = 0; pass < 2; ++pass)                     // pass 0: Jacobian inputs on the main stream; pass 1: the rest on the copy stream
    for (int i = 0; i < n_in; ++i) {
      int64_t want = 0;
      double** slot = input_slot(c, which[i], &want);
      if (!slot) return femo_fail("femo_evaluate_residual: unknown input id");
      if (sizes[i] != want && sizes[i] != 1) return femo_fail("femo_evaluate_residual: size mismatch for input " + std::to_string(which[i]));
      const bool for_matrix = which[i] != FEMO_F_SOLID && (which[i] != FEMO_DENSITY || c->dyn_on);
      if (for_matrix != (pass == 0)) continue;
      cudaStream_t st = pass == 0 ? c->stream : c->copy_stream;
      if (sizes[i] == 1) {                                  // utils_dolfinx.update: a single value fills the function
        cudaStream_t keep = c->stream; c->stream = st;
        const int r = fill(c, *slot, want, data[i][0]);
        c->stream = keep;
        TRY(r);
        if (which[i] == FEMO_UHAT) c->uhat_zero = (data[i][0] == 0.0);
      } else {
        CUDA_TRY(cudaMemcpyAsync(*slot, data[i], sizeof(double) * want, cudaMemcpyHostToDevice, st));
        if (which[i] == FEMO_UHAT) uhat_given = true;
      }
      if (for_matrix) matrix_changed = true;
    }
  if (w_host) CUDA_TRY(cudaMemcpyAsync(c->tmp2, w_host, sizeof(double) * n, cudaMemcpyHostToDevice, c->copy_stream));
  CUDA_TRY(cudaEventRecord(c->ev_copy[0], c->copy_stream));
  if (uhat_given) {                                         // which kernels run depends on uhat == 0 (as in femo_set_input)
    double nrm = 0;
    TRY(dev_dot(c, c->uhat, c->uhat, 3 * c->nn, &nrm));
    c->uhat_zero = (nrm == 0.0);
  }
  if (matrix_changed) { c->jac_valid = false; c->prec_valid = false; }
  if (!c->jac_valid) TRY(assemble_jacobian(c));             // runs while F_solid / density / w are still on their way up
  CUDA_TRY(cudaStreamWaitEvent(c->stream, c->ev_copy[0], 0));
  TRY(assemble_load(c, c->b));
  const double* wd = w_host ? c->tmp2 : c->w;
  const int NR = 4;                                         // row ranges; the last ones hide the first ones' way back
  for (int j = 0; j < NR; ++j) {
    const int64_t r0 = c->nb * j / NR, r1 = c->nb * (j + 1) / NR;
    TRY(bsr_spmv_range(c, wd, c->tmp, r0, r1));
    k_axpby<<<(unsigned)cdiv(3 * (r1 - r0), 256), 256, 0, c->stream>>>(1.0, c->tmp + 3 * r0, -1.0, c->b + 3 * r0, 3 * (r1 - r0), c->tmp + 3 * r0);
    c->launches += 1;
    CUDA_TRY(cudaEventRecord(c->ev_copy[1 + j], c->stream));
    CUDA_TRY(cudaStreamWaitEvent(c->copy_stream, c->ev_copy[1 + j], 0));
    CUDA_TRY(cudaMemcpyAsync(R_host + 3 * r0, c->tmp + 3 * r0, sizeof(double) * 3 * (r1 - r0), cudaMemcpyDeviceToHost, c->copy_stream));
  }
  TRY(check_launch(c, "femo_evaluate_residual"));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->copy_stream));
  return 0;
}

int femo_jacobian_apply(femo_ctx* c, const double* x, double* y, int with_bc, int on_device) {
  NEED_SETUP(c, "femo_jacobian_apply");
  DevGuard g(c->device);
  if (!c->jac_valid) TRY(assemble_jacobian(c));
  TRY(copy_in(c, c->tmp2, x, c->ndof, on_device));
  TRY(bsr_spmv(c, c->tmp2, c->tmp, with_bc));
  return copy_out(c, y, c->tmp, c->ndof, on_device);
}

int femo_get_bsr(femo_ctx* c, int32_t* rowptr, int32_t* colidx, double* vals, int with_bc) {
  NEED_SETUP(c, "femo_get_bsr");
  DevGuard g(c->device);
  if (rowptr) CUDA_TRY(cudaMemcpy(rowptr, c->rowptr, sizeof(int32_t) * (c->nb + 1), cudaMemcpyDeviceToHost));
  if (colidx) CUDA_TRY(cudaMemcpy(colidx, c->colidx, sizeof(int32_t) * c->nnzb, cudaMemcpyDeviceToHost));
  if (vals) {
    if (!c->jac_valid) TRY(assemble_jacobian(c));
    if (!with_bc) CUDA_TRY(cudaMemcpy(vals, c->vals, sizeof(double) * c->nnzb * 9, cudaMemcpyDeviceToHost));
    else {
      double* t = nullptr;
      TRY(dev_alloc(&t, (size_t)c->nnzb * 9));
      CUDA_TRY(cudaMemcpyAsync(t, c->vals, sizeof(double) * c->nnzb * 9, cudaMemcpyDeviceToDevice, c->stream));
      k_bsr_apply_bc<<<(unsigned)cdiv(c->nb, 128), 128, 0, c->stream>>>(c->rowptr, c->colidx, c->bcmask, c->nb, t);
      TRY(check_launch(c, "k_bsr_apply_bc"));
      c->launches += 1;
      int r = copy_out(c, vals, t, c->nnzb * 9, 0);
      dev_free(&t);
      return r;
    }
  }
  return 0;
}

// selected 3x3 blocks of dR/du by BSR position (diagnostics at sizes where femo_get_bsr would move gigabytes)
int femo_get_blocks(femo_ctx* c, int64_t n, const int64_t* positions, double* out) {
  NEED_SETUP(c, "femo_get_blocks");
  if (n < 0 || (n > 0 && (!positions || !out))) return femo_fail("femo_get_blocks: bad arguments");
  for (int64_t i = 0; i < n; ++i)
    if (positions[i] < 0 || positions[i] >= c->nnzb) return femo_fail("femo_get_blocks: block position out of range");
  DevGuard g(c->device);
  if (!c->jac_valid) TRY(assemble_jacobian(c));
  if (n == 0) return 0;
  int64_t* dp = nullptr; double* dv = nullptr;
  int r = dev_alloc(&dp, (size_t)n);
  if (r == 0) r = dev_alloc(&dv, (size_t)9 * n);
  if (r == 0 && cudaMemcpyAsync(dp, positions, sizeof(int64_t) * n, cudaMemcpyHostToDevice, c->stream) != cudaSuccess) r = femo_fail("femo_get_blocks: copy failed");
  if (r == 0) {
    k_pick_blocks<<<(unsigned)cdiv(9 * n, 256), 256, 0, c->stream>>>(c->vals, dp, n, dv);
    c->launches += 1;
    r = check_launch(c, "k_pick_blocks");
  }
  if (r == 0) r = copy_out(c, out, dv, 9 * n, 0);
  dev_free(&dp); dev_free(&dv);
  return r;
}

int femo_solve_state(femo_ctx* c, double rtol, int max_it, int* iters, double* relres) {
  NEED_SETUP(c, "femo_solve_state");
  DevGuard g(c->device);
  if (!c->jac_valid) TRY(assemble_jacobian(c));
  TRY(assemble_load(c, c->b));
  // homogeneous Dirichlet values (ubc = 0, rm_shell_model.py:174-176)
  k_mul<<<(unsigned)cdiv(c->ndof, 256), 256, 0, c->stream>>>(c->b, c->bcmask, c->ndof, c->b);
  k_mul<<<(unsigned)cdiv(c->ndof, 256), 256, 0, c->stream>>>(c->w, c->bcmask, c->ndof, c->w);
  TRY(check_launch(c, "k_mul"));
  c->launches += 2;
  int it = 0; double rr = 0;
  TRY(pcg_solve(c, c->b, c->w, rtol, max_it, &it, &rr, false));
  if (iters) *iters = it;
  if (relres) *relres = rr;
  return 0;
}

static int solve_generic(femo_ctx* c, const double* rhs, double* sol, double rtol, int max_it, int* iters, double* relres,
                         int on_device, bool zero_bc) {
  if (!c->jac_valid) TRY(assemble_jacobian(c));
  TRY(copy_in(c, c->b, rhs, c->ndof, on_device));
  CUDA_TRY(cudaMemsetAsync(c->tmp2, 0, sizeof(double) * c->ndof, c->stream));
  int it = 0; double rr = 0;
  TRY(pcg_solve(c, c->b, c->tmp2, rtol, max_it, &it, &rr, false));
  if (zero_bc) {  // state_operation.py:217-218
    k_mul<<<(unsigned)cdiv(c->ndof, 256), 256, 0, c->stream>>>(c->tmp2, c->bcmask, c->ndof, c->tmp2);
    TRY(check_launch(c, "k_mul"));
    c->launches += 1;
  }
  if (iters) *iters = it;
  if (relres) *relres = rr;
  return copy_out(c, sol, c->tmp2, c->ndof, on_device);
}

int femo_solve_adjoint(femo_ctx* c, const double* rhs, double* lam, double rtol, int max_it, int* iters, double* relres,
                       int on_device) {
  NEED_SETUP(c, "femo_solve_adjoint");
  DevGuard g(c->device);
  return solve_generic(c, rhs, lam, rtol, max_it, iters, relres, on_device, true);
}

int femo_solve_forward(femo_ctx* c, const double* rhs, double* du, double rtol, int max_it, int* iters, double* relres,
                       int on_device) {
  NEED_SETUP(c, "femo_solve_forward");
  DevGuard g(c->device);
  return solve_generic(c, rhs, du, rtol, max_it, iters, relres, on_device, false);
}

int femo_dRdf_T_apply(femo_ctx* c, int arg, const double* lam, double* out, int on_device) {
  NEED_SHELL(c, "femo_dRdf_T_apply");
  DevGuard g(c->device);
  int64_t n = 0;
  if (!input_slot(c, arg, &n)) return femo_fail("femo_dRdf_T_apply: unknown argument");
  TRY(copy_in(c, c->tmp2, lam, c->ndof, on_device));
  TRY(dRdf_T_apply(c, arg, c->tmp2, c->tmp));
  return copy_out(c, out, c->tmp, n, on_device);
}

int femo_dRdf_apply(femo_ctx* c, int arg, const double* df, double* out, int on_device) {
  NEED_SHELL(c, "femo_dRdf_apply");
  DevGuard g(c->device);
  int64_t n = 0;
  if (!input_slot(c, arg, &n)) return femo_fail("femo_dRdf_apply: unknown argument");
  TRY(copy_in(c, c->tmp2, df, n, on_device));
  TRY(dRdf_apply(c, arg, c->tmp2, c->tmp));
  return copy_out(c, out, c->tmp, c->ndof, on_device);
}

int femo_output(femo_ctx* c, int which, double* value) {
  NEED_SHELL(c, "femo_output");
  DevGuard g(c->device);
  return out_scalar(c, which, value);
}

int femo_output_grad(femo_ctx* c, int which, int arg, double* gout, int on_device) {
  NEED_SHELL(c, "femo_output_grad");
  DevGuard g(c->device);
  int64_t n = c->ndof;
  if (arg != FEMO_ARG_STATE && !input_slot(c, arg, &n)) return femo_fail("femo_output_grad: unknown argument");
  TRY(out_grad(c, which, arg, c->tmp));
  return copy_out(c, gout, c->tmp, n, on_device);
}

int femo_set_stress_params(femo_ctx* c, double m, double rho) {
  if (!c) return femo_fail("null context");
  c->stress_m = m; c->stress_rho = rho;
  return 0;
}

int femo_set_cell_tags(femo_ctx* c, int64_t n, const int32_t* tags) {
  NEED_SHELL(c, "femo_set_cell_tags");
  if (n != c->nc) return femo_fail("femo_set_cell_tags: need one tag per cell");
  DevGuard g(c->device);
  TRY(dev_alloc(&c->cell_tags, (size_t)n));
  CUDA_TRY(cudaMemcpy(c->cell_tags, tags, sizeof(int32_t) * n, cudaMemcpyHostToDevice));
  c->area_cache.clear();
  return 0;
}

int femo_pnorm_stress(femo_ctx* c, int tag, double* value) {
  NEED_SHELL(c, "femo_pnorm_stress");
  DevGuard g(c->device);
  return pnorm_value(c, tag, value);
}

int femo_pnorm_stress_grad(femo_ctx* c, int tag, int arg, double* gout, int on_device) {
  NEED_SHELL(c, "femo_pnorm_stress_grad");
  DevGuard g(c->device);
  int64_t n = c->ndof;
  if (arg != FEMO_ARG_STATE && !input_slot(c, arg, &n)) return femo_fail("femo_pnorm_stress_grad: unknown argument");
  TRY(pnorm_grad(c, tag, arg, c->tmp));
  return copy_out(c, gout, c->tmp, n, on_device);
}

int femo_stress_field(femo_ctx* c, int surface, double* out, int on_device) {
  NEED_SHELL(c, "femo_stress_field");
  if (surface < 0 || surface > 2) return femo_fail("femo_stress_field: surface must be 0 (Top), 1 (Mid) or 2 (Bot)");
  DevGuard g(c->device);
  double* t = nullptr;
  TRY(dev_alloc(&t, (size_t)c->nc * 4));
  int r = stress_field(c, surface, t);
  if (r == 0) r = copy_out(c, out, t, c->nc * 4, on_device);
  dev_free(&t);
  return r;
}

int femo_launch_count(femo_ctx* c, int64_t* n) {
  if (!c || !n) return femo_fail("null argument");
  *n = c->launches;
  return 0;
}

int femo_last_timings(femo_ctx* c, double* ms4) {
  if (!c || !ms4) return femo_fail("null argument");
  ms4[0] = c->t_rec; ms4[1] = c->t_asm; ms4[2] = c->t_solve; ms4[3] = c->t_iter;
  return 0;
}

int femo_last_dmma_tflops(femo_ctx* c, double* tf) {
  if (!c || !tf) return femo_fail("null argument");
  *tf = c->dmma_tf;
  return 0;
}

int femo_pcg_fixed(femo_ctx* c, int n_iter, double* ms_total) {
  NEED_SETUP(c, "femo_pcg_fixed");
  DevGuard g(c->device);
  if (!c->jac_valid) TRY(assemble_jacobian(c));
  TRY(assemble_load(c, c->b));
  k_mul<<<(unsigned)cdiv(c->ndof, 256), 256, 0, c->stream>>>(c->b, c->bcmask, c->ndof, c->b);
  c->launches += 1;
  CUDA_TRY(cudaMemsetAsync(c->tmp2, 0, sizeof(double) * c->ndof, c->stream));
  int it = 0; double rr = 0;
  TRY(pcg_solve(c, c->b, c->tmp2, 0.0, n_iter, &it, &rr, true));
  if (ms_total) *ms_total = c->t_solve;
  return 0;
}

int femo_measure_peaks(femo_ctx* c, double* tf, double* gbs) {
  if (!c) return femo_fail("null context");
  DevGuard g(c->device);
  return measure_peaks(c, tf, gbs);
}

}  // extern "C"
