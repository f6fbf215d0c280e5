// Generic 3-D forms on trilinear hexahedra behind the same context / solver machinery as the RM shell
// (BASELINE config 5: unit cube, 160^3 cells, cell-wise SIMP density).  The reference's generic FEA takes
// any UFL form (femo_alpha/fea/fea_dolfinx.py:71-116) and has no in-tree form for this configuration; here a small set
// of named forms (hex_element.cuh) is assembled into the BSR 3x3 matrix of a `kind = 1` femo_ctx, so that
// femo_set_strong_bc, femo_assemble_jacobian, femo_jacobian_apply, femo_get_bsr, femo_assemble_residual,
// femo_solve_state / _adjoint / _forward, femo_pcg_fixed, femo_set_state / femo_get_state work on it unchanged
// (k_spmv<8,6> and the fused block-Jacobi PCG of solver.cu) -- and so does femo_dist_setup / femo_dist_init: with the
// owner mask and halo lists of femo_alpha_b200/partition.py (build_local_part_hex) the solves run over NCCL like the
// shell's (owner-computes assembly with one ghost-cell layer, no collective in assembly).
//
// Assembly is destination-driven: one thread per BSR block (32 lanes per row node) sums the contributions of the cells
// around the row node that also hold the column node, in cell-ascending order -- no atomics, no intermediate element
// matrices, bitwise reproducible.  Algorithmic bytes per cell (structured interior): 27 blocks x 72 B written per node
// ~ 1944 B, reads 8 x 80 B geometry + 8 B density + 32 B connectivity; the geometry records are re-read once per
// block (warp-uniform addresses, served by L1/L2).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>

#include "femo_internal.cuh"
#include "hex_element.cuh"
#include "hex_pattern.hpp"

namespace {

using namespace femo_hex;

struct DevGuardH {
  int prev = 0;
  explicit DevGuardH(int dev) { cudaGetDevice(&prev); cudaSetDevice(dev); }
  ~DevGuardH() { cudaSetDevice(prev); }
};

__global__ void k_hex_geom(const double* __restrict__ xyz, const int32_t* __restrict__ cells, int64_t nc,
                           double* __restrict__ geo, double* __restrict__ vol, int* __restrict__ bad) {
  const int64_t cell = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (cell >= nc) return;
  double v = 0.0;
  for (int q = 0; q < 8; ++q) {
    const double det = geom_qp(xyz, cells + 8 * cell, q, geo + (cell * 8 + q) * GEO);
    if (!(det > 0.0)) atomicOr(bad, 1);
    v += det;
  }
  vol[cell] = v;
}

__global__ void k_hex_scale(const double* __restrict__ rho, double p, int64_t nc, double* __restrict__ scale) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < nc) scale[i] = pow(rho[i], p);
}

// One warp per row node i.  Per pass of up to 8 cells around i (HEX_PASS = 64 work items, hex_element.cuh):
//   1. the cells' geometry records (8 x 80 doubles) are copied to shared memory with cp.async (row stride 82
//      doubles: 16-byte aligned, and the 8 cells of a request fall into different banks);
//   2. lane l evaluates the item pair (2 l, 2 l + 1) = two columns of cell l / 4 (item_pair_body);
//   3. the blocks are staged component-major in the same shared region, and the lanes that own the destination blocks of
//      the row collect the items of their column in cell-ascending order (row_collect);
//   4. the row's blocks leave through a shared transpose as contiguous 256-byte stores.
// Deterministic, no atomics.  Rows with more than 8 cells take further passes (read-modify-write of their own blocks),
// rows with more than 32 blocks further destination chunks.
static_assert(HEX_PASS == 64, "k_hex_assemble stages one item pair per lane");
constexpr int HEX_WARPS = 4;
constexpr int HEX_GSTRIDE = 8 * GEO + 2;                 // 16-byte aligned rows, 4 banks apart
constexpr int HEX_REGION = 8 * HEX_GSTRIDE;              // >= 9 * HEX_PASS
template <int MINB>     // resident CTAs per SM the compiler plans for (5: 96 registers, 11.9 ms at 160^3; 7: 72 registers, spills, 21.2 ms -- r02h)
__global__ void __launch_bounds__(32 * HEX_WARPS, MINB)
k_hex_assemble(int form, double c1, double c2, const double* __restrict__ geo, const int32_t* __restrict__ cells,
               const double* __restrict__ scale, const int32_t* __restrict__ nc_ptr, const int32_t* __restrict__ nc_adj,
               const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx, int64_t nb,
               double* __restrict__ vals) {
  __shared__ __align__(16) double s_reg[HEX_WARPS][HEX_REGION];        // geometry records, then the staged blocks
  __shared__ double s_out[HEX_WARPS][32 * 9];
  __shared__ __align__(16) int32_t s_col[HEX_WARPS][HEX_PASS];
  const unsigned FULL = 0xffffffffu;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t i = blockIdx.x * (int64_t)HEX_WARPS + warp;
  if (i >= nb) return;                                   // whole warps leave together
  double* reg = s_reg[warp];
  const int32_t e0 = nc_ptr[i], n_cells_i = nc_ptr[i + 1] - e0;
  const int32_t r0 = rowptr[i], n_row = rowptr[i + 1] - r0;
  for (int c0 = 0; c0 < n_cells_i; c0 += 8) {
    const int n_pass = min(8, n_cells_i - c0);           // cells of this pass
    const int32_t my_ca = (lane < n_pass) ? nc_adj[e0 + c0 + lane] : 0;
    // 1. geometry records -> shared memory: 8 cells x 40 chunks of 16 bytes, ten asynchronous copies per lane in flight
    //    at once (a load -> store loop exposes the full memory latency ten times: 37 % of the stall samples in
    //    profiles/r01t_hex_asm_*.csv)
#pragma unroll
    for (int j = 0; j < (8 * 8 * GEO / 2) / 32; ++j) {
      const int idx = j * 32 + lane;
      const int cl = idx / (8 * GEO / 2), ch = idx - cl * (8 * GEO / 2);
      const int32_t ca = __shfl_sync(FULL, my_ca, cl);
      if (cl < n_pass) {
        const unsigned dst = (unsigned)__cvta_generic_to_shared(reg + cl * HEX_GSTRIDE + 2 * ch);
        const double* src = geo + (int64_t)(ca >> 3) * (8 * GEO) + 2 * ch;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    // 2. item pair of this lane
    const int cl = lane >> 2, b = 2 * (lane & 3);
    const int32_t ca = __shfl_sync(FULL, my_ca, cl);
    double B0[9], B1[9];
    int32_t col0 = -1, col1 = -1;
    if (cl < n_pass) {
      const int64_t cell = ca >> 3;
      const int2 cc = *reinterpret_cast<const int2*>(cells + 8 * cell + b);
      col0 = cc.x; col1 = cc.y;
      item_pair_body(form, c1, c2, reg + cl * HEX_GSTRIDE, scale[cell], ca & 7, b, B0, B1);
    }
    __syncwarp();                                        // every lane is done with the records
    // 3. stage and collect
    const int n_staged = 8 * n_pass;
    if (cl < n_pass) {
      *reinterpret_cast<int2*>(&s_col[warp][2 * lane]) = make_int2(col0, col1);
#pragma unroll
      for (int k = 0; k < 9; ++k) *reinterpret_cast<double2*>(&reg[k * HEX_PASS + 2 * lane]) = make_double2(B0[k], B1[k]);
    }
    __syncwarp();
    for (int k0 = 0; k0 < n_row; k0 += 32) {
      const int n_dst = min(32, n_row - k0);
      if (lane < n_dst) {
        double acc[9];
        row_collect(reg, s_col[warp], n_staged, colidx[r0 + k0 + lane], acc);
#pragma unroll
        for (int e = 0; e < 9; ++e) s_out[warp][9 * lane + e] = acc[e];
      }
      __syncwarp();
      // 4. contiguous stores of the row's blocks
      double* out = vals + (int64_t)(r0 + k0) * 9;
      if (c0 == 0) for (int idx = lane; idx < 9 * n_dst; idx += 32) out[idx] = s_out[warp][idx];
      else for (int idx = lane; idx < 9 * n_dst; idx += 32) out[idx] += s_out[warp][idx];
      __syncwarp();
    }
  }
}

// ---- gradient-record variant (FEMO_HEX_GRAD=1 at femo_hex_setup).  Correct on the B200 and bitwise equal to the default
// kernel, but not faster (0.382 vs 0.354 ms at 48^3, profiles/r01v_hex_grad_worker.log): off by default ----
__global__ void k_hex_geom2(const double* __restrict__ geo, int64_t n_rec, double* __restrict__ geo2) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;       // r = cell * 8 + q
  if (r < n_rec) geom2_qp(geo + r * GEO, (int)(r & 7), geo2 + r * GEO2);
}

// Same row algorithm as k_hex_assemble with one item per lane and sub-passes of 4 cells: the 4 x 208 doubles of gradient
// records are staged with cp.async (13 chunks of 16 bytes per lane), lane l evaluates item (cell l / 8, column l % 8), the
// blocks of up to two sub-passes (64 items) are staged, then collected and stored as in k_hex_assemble.
constexpr int HEXG_WARPS = 3;
constexpr int HEXG_CSTRIDE = 8 * GEO2 + 2;               // 210 doubles: 16-byte aligned, cells 4 banks apart
__global__ void __launch_bounds__(32 * HEXG_WARPS)
k_hex_assemble_g(int form, double c1, double c2, const double* __restrict__ geo2, const int32_t* __restrict__ cells,
                 const double* __restrict__ scale, const int32_t* __restrict__ nc_ptr, const int32_t* __restrict__ nc_adj,
                 const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx, int64_t nb,
                 double* __restrict__ vals) {
  __shared__ __align__(16) double s_geo[HEXG_WARPS][4 * HEXG_CSTRIDE];
  __shared__ __align__(16) double s_stg[HEXG_WARPS][9 * HEX_PASS];
  __shared__ double s_out[HEXG_WARPS][32 * 9];
  __shared__ __align__(16) int32_t s_col[HEXG_WARPS][HEX_PASS];
  const unsigned FULL = 0xffffffffu;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t i = blockIdx.x * (int64_t)HEXG_WARPS + warp;
  if (i >= nb) return;
  double* reg = s_geo[warp];
  double* stg = s_stg[warp];
  const int32_t e0 = nc_ptr[i], n_cells_i = nc_ptr[i + 1] - e0;
  const int32_t r0 = rowptr[i], n_row = rowptr[i + 1] - r0;
  constexpr int CH = 8 * GEO2 / 2;                       // 16-byte chunks per cell
  for (int c0 = 0; c0 < n_cells_i; c0 += 8) {
    const int n_chunk = min(8, n_cells_i - c0);          // cells of this chunk (two sub-passes of 4)
    for (int h = 0; h < 2 && 4 * h < n_chunk; ++h) {
      const int n_pass = min(4, n_chunk - 4 * h);
      const int32_t my_ca = (lane < n_pass) ? nc_adj[e0 + c0 + 4 * h + lane] : 0;
#pragma unroll
      for (int j = 0; j < (4 * CH) / 32; ++j) {
        const int idx = j * 32 + lane;
        const int cl = idx / CH, ch = idx - cl * CH;
        const int32_t ca = __shfl_sync(FULL, my_ca, cl);
        if (cl < n_pass) {
          const unsigned dst = (unsigned)__cvta_generic_to_shared(reg + cl * HEXG_CSTRIDE + 2 * ch);
          const double* src = geo2 + (int64_t)(ca >> 3) * (8 * GEO2) + 2 * ch;
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncwarp();
      const int cl = lane >> 3, b = lane & 7;
      const int32_t ca = __shfl_sync(FULL, my_ca, cl);
      if (cl < n_pass) {
        const int64_t cell = ca >> 3;
        double B[9];
        item_body_g(form, c1, c2, reg + cl * HEXG_CSTRIDE, scale[cell], ca & 7, b, B);
        const int t = 32 * h + lane;                     // = 8 * (cell position in the chunk) + b
        s_col[warp][t] = cells[8 * cell + b];
#pragma unroll
        for (int k = 0; k < 9; ++k) stg[k * HEX_PASS + t] = B[k];
      }
      __syncwarp();                                      // records free for the next sub-pass, staging visible
    }
    const int n_staged = 8 * n_chunk;
    for (int k0 = 0; k0 < n_row; k0 += 32) {
      const int n_dst = min(32, n_row - k0);
      if (lane < n_dst) {
        double acc[9];
        row_collect(stg, s_col[warp], n_staged, colidx[r0 + k0 + lane], acc);
#pragma unroll
        for (int e = 0; e < 9; ++e) s_out[warp][9 * lane + e] = acc[e];
      }
      __syncwarp();
      double* out = vals + (int64_t)(r0 + k0) * 9;
      if (c0 == 0) for (int idx = lane; idx < 9 * n_dst; idx += 32) out[idx] = s_out[warp][idx];
      else for (int idx = lane; idx < 9 * n_dst; idx += 32) out[idx] += s_out[warp][idx];
      __syncwarp();
    }
  }
}

// mode 0: consistent mass product coef * M_e x_e;  mode 1: cellcoef[e] * coef * K0_e x_e
__global__ void k_hex_cellvec(int mode, int form, double c1, double c2, const double* __restrict__ geo,
                              const int32_t* __restrict__ cells, const double* __restrict__ x,
                              const double* __restrict__ cellcoef, double coef, int64_t nc, double* __restrict__ cellvec) {
  const int64_t cell = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (cell >= nc) return;
  double ye[24];
  const double cf = coef * (cellcoef ? cellcoef[cell] : 1.0);
  if (mode == 0) cell_mass_apply(geo, cells, cell, x, cf, ye);
  else cell_apply(form, c1, c2, geo, cells, cell, x, cf, ye);
  double* out = cellvec + cell * 24;
  for (int k = 0; k < 24; ++k) out[k] = ye[k];
}

__global__ void k_hex_gather(const int32_t* __restrict__ nc_ptr, const int32_t* __restrict__ nc_adj,
                             const double* __restrict__ cellvec, int64_t nb, double* __restrict__ out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= nb) return;
  double v[3];
  node_gather(nc_ptr, nc_adj, cellvec, i, v);
  out[3 * i] = v[0]; out[3 * i + 1] = v[1]; out[3 * i + 2] = v[2];
}

// out[e] = coef[e] * a0_e(lam, u)
__global__ void k_hex_energy(int form, double c1, double c2, const double* __restrict__ geo, const int32_t* __restrict__ cells,
                             const double* __restrict__ u, const double* __restrict__ lm, const double* __restrict__ rho,
                             double p, int64_t nc, double* __restrict__ out) {
  const int64_t cell = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (cell >= nc) return;
  out[cell] = p * pow(rho[cell], p - 1.0) * cell_energy(form, c1, c2, geo, cells, cell, u, lm);
}

__global__ void k_hex_dscale(const double* __restrict__ rho, const double* drho, double p, int64_t nc, double* out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < nc) out[i] = p * pow(rho[i], p - 1.0) * drho[i];
}

__global__ void k_hex_fill(double* p, int64_t n, double v) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

__global__ void k_hex_colbc(const int32_t* __restrict__ colidx, int64_t nnzb, uint32_t* __restrict__ colbc) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < nnzb) colbc[i] = (uint32_t)colidx[i];          // no clamped dof yet
}

template <typename T>
int upload(T** dst, const std::vector<T>& src, cudaStream_t s) {
  TRY(dev_alloc(dst, src.size()));
  if (!src.empty()) CUDA_TRY(cudaMemcpyAsync(*dst, src.data(), sizeof(T) * src.size(), cudaMemcpyHostToDevice, s));
  return 0;
}

int hex_cellvec_gather(femo_ctx* c, int mode, const double* x, const double* cellcoef, double coef, double* out) {
  k_hex_cellvec<<<(unsigned)cdiv(c->nc, 128), 128, 0, c->stream>>>(mode, c->hx_form, c->hx_c1, c->hx_c2, c->hx_geo, c->hx_cells, x,
                                                                 cellcoef, coef, c->nc, c->hx_cellvec);
  k_hex_gather<<<(unsigned)cdiv(c->nb, 256), 256, 0, c->stream>>>(c->nc_ptr, c->nc_adj, c->hx_cellvec, c->nb, out);
  TRY(check_launch(c, "k_hex_cellvec/k_hex_gather"));
  c->launches += 2;
  return 0;
}

}  // namespace

// ---- internal entry points (dispatched from assemble_jacobian / assemble_load for kind == 1) ---------------------------
int hex_assemble(femo_ctx* c) {
  cudaEventRecord(c->ev0, c->stream);
  k_hex_scale<<<(unsigned)cdiv(c->nc, 256), 256, 0, c->stream>>>(c->hx_rho, c->hx_p, c->nc, c->hx_scale);
  TRY(check_launch(c, "k_hex_scale"));
  if (c->hx_geo2)
    k_hex_assemble_g<<<(unsigned)cdiv(c->nb, HEXG_WARPS), 32 * HEXG_WARPS, 0, c->stream>>>(
        c->hx_form, c->hx_c1, c->hx_c2, c->hx_geo2, c->hx_cells, c->hx_scale, c->nc_ptr, c->nc_adj, c->rowptr, c->colidx, c->nb,
        c->vals);
  else if (getenv("FEMO_HEX_MINB") && atoi(getenv("FEMO_HEX_MINB")) == 7)      // occupancy experiment
    k_hex_assemble<7><<<(unsigned)cdiv(c->nb, HEX_WARPS), 32 * HEX_WARPS, 0, c->stream>>>(
        c->hx_form, c->hx_c1, c->hx_c2, c->hx_geo, c->hx_cells, c->hx_scale, c->nc_ptr, c->nc_adj, c->rowptr, c->colidx, c->nb,
        c->vals);
  else
    k_hex_assemble<5><<<(unsigned)cdiv(c->nb, HEX_WARPS), 32 * HEX_WARPS, 0, c->stream>>>(
        c->hx_form, c->hx_c1, c->hx_c2, c->hx_geo, c->hx_cells, c->hx_scale, c->nc_ptr, c->nc_adj, c->rowptr, c->colidx, c->nb,
        c->vals);
  TRY(check_launch(c, "k_hex_assemble"));
  cudaEventRecord(c->ev1, c->stream);
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  float ms = 0; cudaEventElapsedTime(&ms, c->ev0, c->ev1);
  c->t_rec = 0; c->t_asm = ms;
  c->launches += 2;
  c->jac_valid = true; c->prec_valid = false;
  return 0;
}

int hex_load(femo_ctx* c, double* F) { return hex_cellvec_gather(c, 0, c->hx_f, nullptr, 1.0, F); }

namespace {

// node -> (cell, local index) adjacency and the BSR pattern of the node graph, built on the host (hex_pattern.hpp)
int hex_pattern(femo_ctx* c, const std::vector<int32_t>& cells) {
  Pattern P;
  const int64_t r = build_pattern(c->nb, c->nc, cells.data(), P);
  if (r > 0) return femo_fail("femo_hex_setup: node " + std::to_string(r - 1) + " belongs to no cell");
  if (r < 0) return femo_fail("femo_hex_setup: more than 2e9 matrix blocks");
  c->nnzb = (int64_t)P.colidx.size();
  c->ncontrib = c->nc * 64;
  TRY(upload(&c->nc_ptr, P.nc_ptr, c->stream));
  TRY(upload(&c->nc_adj, P.nc_adj, c->stream));
  TRY(upload(&c->rowptr, P.rowptr, c->stream));
  TRY(upload(&c->colidx, P.colidx, c->stream));
  TRY(upload(&c->diag_pos, P.diag, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));       // the host vectors go out of scope
  return 0;
}

#define NEED_HEX(c, name) \
  if (!(c) || !(c)->setup_done || (c)->kind != 1) return femo_fail(std::string(name) + ": not a hexahedral-solid context (femo_hex_setup)")

int hex_copy_in(femo_ctx* c, double* dst, const double* src, int64_t n, int on_device) {
  CUDA_TRY(cudaMemcpyAsync(dst, src, sizeof(double) * n, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
  return 0;
}
int hex_copy_out(femo_ctx* c, double* dst, const double* src, int64_t n, int on_device) {
  CUDA_TRY(cudaMemcpyAsync(dst, src, sizeof(double) * n, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

}  // namespace

extern "C" {

int femo_hex_setup(femo_ctx* c, int64_t n_nodes, const double* xyz, int64_t n_cells, const int32_t* conn, int form, double E0,
                   double nu, double simp_p) {
  if (!c) return femo_fail("femo_hex_setup: null context");
  if (c->setup_done || c->nc != 0) return femo_fail("femo_hex_setup: the context already holds a mesh (one problem per context)");
  if (n_nodes <= 0 || n_cells <= 0 || !xyz || !conn) return femo_fail("femo_hex_setup: empty mesh");
  if (form != 0 && form != 1) return femo_fail("femo_hex_setup: form must be 0 (vector Poisson) or 1 (linear elasticity)");
  if (!(E0 > 0) || (form == 1 && !(nu > -1.0 && nu < 0.5))) return femo_fail("femo_hex_setup: need E0 > 0 and -1 < nu < 0.5");
  if (n_nodes >= ((int64_t)1 << 29) || n_cells >= ((int64_t)1 << 28)) return femo_fail("femo_hex_setup: mesh too large for 32-bit indices");
  std::vector<int32_t> cells(conn, conn + n_cells * 8);
  for (int32_t v : cells) if (v < 0 || v >= n_nodes) return femo_fail("femo_hex_setup: connectivity entry out of range");
  DevGuardH g(c->device);
  c->kind = 1; c->hx_form = form; c->hx_p = simp_p;
  if (form == 1) { c->hx_c1 = E0 * nu / ((1 + nu) * (1 - 2 * nu)); c->hx_c2 = E0 / (2 * (1 + nu)); }
  else { c->hx_c1 = E0; c->hx_c2 = 0; }
  c->hx_nn = n_nodes; c->nn = 0; c->ne = 0; c->nc = n_cells; c->nq2 = n_nodes; c->nb = n_nodes; c->ndof = 3 * n_nodes;
  c->n_mat = n_cells; c->n_f = 3 * n_nodes;
  int r = hex_pattern(c, cells);
  if (r != 0) { c->nc = 0; return r; }
  const int64_t n = c->ndof;
  TRY(dev_alloc(&c->xyz, (size_t)n_nodes * 3));
  CUDA_TRY(cudaMemcpyAsync(c->xyz, xyz, sizeof(double) * n_nodes * 3, cudaMemcpyHostToDevice, c->stream));
  TRY(upload(&c->hx_cells, cells, c->stream));
  TRY(dev_alloc(&c->hx_geo, (size_t)n_cells * 8 * GEO));
  TRY(dev_alloc(&c->hx_rho, (size_t)n_cells)); TRY(dev_alloc(&c->hx_scale, (size_t)n_cells)); TRY(dev_alloc(&c->hx_vol, (size_t)n_cells));
  TRY(dev_alloc(&c->hx_f, (size_t)n)); TRY(dev_alloc(&c->hx_cellvec, (size_t)n_cells * 24));
  TRY(dev_alloc(&c->vals, (size_t)c->nnzb * 9));
  TRY(dev_alloc(&c->bcmask, (size_t)n)); TRY(dev_alloc(&c->bcbits, (size_t)c->nb)); TRY(dev_alloc(&c->colbc, (size_t)c->nnzb));
  double** vecs[] = {&c->w, &c->b, &c->r, &c->z, &c->p, &c->Ap, &c->tmp, &c->tmp2};
  for (double** v : vecs) { TRY(dev_alloc(v, (size_t)n)); CUDA_TRY(cudaMemsetAsync(*v, 0, sizeof(double) * n, c->stream)); }
  TRY(dev_alloc(&c->dinv6, (size_t)1)); TRY(dev_alloc(&c->dinv3, (size_t)c->nb * 9));
  TRY(dev_alloc(&c->red, 8 * 1184)); TRY(dev_alloc(&c->scal, 32));
  TRY(dev_alloc(&c->cellbuf, (size_t)n_cells));              // per-cell scalars (density partials)
  CUDA_TRY(cudaMemsetAsync(c->hx_f, 0, sizeof(double) * n, c->stream));
  CUDA_TRY(cudaMemsetAsync(c->bcbits, 0, (size_t)c->nb, c->stream));
  k_hex_fill<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(c->bcmask, n, 1.0);
  k_hex_fill<<<(unsigned)cdiv(n_cells, 256), 256, 0, c->stream>>>(c->hx_rho, n_cells, 1.0);
  k_hex_colbc<<<(unsigned)cdiv(c->nnzb, 256), 256, 0, c->stream>>>(c->colidx, c->nnzb, c->colbc);
  int* bad = nullptr;
  TRY(dev_alloc(&bad, (size_t)1));
  CUDA_TRY(cudaMemsetAsync(bad, 0, sizeof(int), c->stream));
  k_hex_geom<<<(unsigned)cdiv(n_cells, 128), 128, 0, c->stream>>>(c->xyz, c->hx_cells, n_cells, c->hx_geo, c->hx_vol, bad);
  TRY(check_launch(c, "k_hex_geom"));
  c->launches += 4;
  const char* genv = getenv("FEMO_HEX_GRAD");
  if (genv && genv[0] == '1') {                            // gradient records for k_hex_assemble_g
    TRY(dev_alloc(&c->hx_geo2, (size_t)n_cells * 8 * GEO2));
    k_hex_geom2<<<(unsigned)cdiv(n_cells * 8, 256), 256, 0, c->stream>>>(c->hx_geo, n_cells * 8, c->hx_geo2);
    TRY(check_launch(c, "k_hex_geom2"));
    c->launches += 1;
  }
  int hbad = 0;
  CUDA_TRY(cudaMemcpyAsync(&hbad, bad, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  dev_free(&bad);
  if (hbad) { c->nc = 0; return femo_fail("femo_hex_setup: a cell has a non-positive Jacobian (VTK_HEXAHEDRON node order expected)"); }
  c->n_bc = 0; c->n_pen = 0;
  c->setup_done = true; c->jac_valid = false; c->prec_valid = false;
  return 0;
}

/* which: 0 cell density (n_cells), 1 nodal body force (3 n_nodes) */
int femo_hex_set_input(femo_ctx* c, int which, const double* data, int64_t n, int on_device) {
  NEED_HEX(c, "femo_hex_set_input");
  if (!data) return femo_fail("femo_hex_set_input: null data");
  DevGuardH g(c->device);
  if (which == 0) {
    if (n != c->nc && n != 1) return femo_fail("femo_hex_set_input: density needs n_cells values (or 1)");
    if (n == 1) {
      if (on_device) return femo_fail("femo_hex_set_input: scalar broadcast takes a host value");
      k_hex_fill<<<(unsigned)cdiv(c->nc, 256), 256, 0, c->stream>>>(c->hx_rho, c->nc, data[0]);
      TRY(check_launch(c, "k_hex_fill")); c->launches += 1;
    } else TRY(hex_copy_in(c, c->hx_rho, data, n, on_device));
    c->jac_valid = false; c->prec_valid = false;
  } else if (which == 1) {
    if (n != c->ndof) return femo_fail("femo_hex_set_input: body force needs 3 n_nodes values");
    TRY(hex_copy_in(c, c->hx_f, data, n, on_device));
  } else return femo_fail("femo_hex_set_input: unknown input");
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

int femo_hex_get_input(femo_ctx* c, int which, double* out, int64_t n, int on_device) {
  NEED_HEX(c, "femo_hex_get_input");
  DevGuardH g(c->device);
  if (which == 0 && n == c->nc) return hex_copy_out(c, out, c->hx_rho, n, on_device);
  if (which == 1 && n == c->ndof) return hex_copy_out(c, out, c->hx_f, n, on_device);
  return femo_fail("femo_hex_get_input: unknown input or wrong size");
}

/* (dR/d arg)^T lam with R = K(rho) w - M f at the context's state w.  arg 0: p rho^(p-1) a0_e(lam, w) per cell; arg 1: -M lam */
int femo_hex_dRdf_T_apply(femo_ctx* c, int arg, const double* lam, double* out, int on_device) {
  NEED_HEX(c, "femo_hex_dRdf_T_apply");
  if (!lam || !out) return femo_fail("femo_hex_dRdf_T_apply: null argument");
  DevGuardH g(c->device);
  TRY(hex_copy_in(c, c->tmp2, lam, c->ndof, on_device));
  if (arg == 0) {
    k_hex_energy<<<(unsigned)cdiv(c->nc, 128), 128, 0, c->stream>>>(c->hx_form, c->hx_c1, c->hx_c2, c->hx_geo, c->hx_cells, c->w,
                                                                  c->tmp2, c->hx_rho, c->hx_p, c->nc, c->cellbuf);
    TRY(check_launch(c, "k_hex_energy")); c->launches += 1;
    return hex_copy_out(c, out, c->cellbuf, c->nc, on_device);
  }
  if (arg == 1) {
    TRY(hex_cellvec_gather(c, 0, c->tmp2, nullptr, -1.0, c->tmp));
    return hex_copy_out(c, out, c->tmp, c->ndof, on_device);
  }
  return femo_fail("femo_hex_dRdf_T_apply: unknown argument");
}

/* (dR/d arg) d arg.  arg 0: sum_e d rho_e p rho_e^(p-1) K0_e w_e; arg 1: -M df */
int femo_hex_dRdf_apply(femo_ctx* c, int arg, const double* df, double* out, int on_device) {
  NEED_HEX(c, "femo_hex_dRdf_apply");
  if (!df || !out) return femo_fail("femo_hex_dRdf_apply: null argument");
  DevGuardH g(c->device);
  if (arg == 0) {
    TRY(hex_copy_in(c, c->cellbuf, df, c->nc, on_device));
    k_hex_dscale<<<(unsigned)cdiv(c->nc, 256), 256, 0, c->stream>>>(c->hx_rho, c->cellbuf, c->hx_p, c->nc, c->cellbuf);
    TRY(check_launch(c, "k_hex_dscale")); c->launches += 1;
    TRY(hex_cellvec_gather(c, 1, c->w, c->cellbuf, 1.0, c->tmp));
  } else if (arg == 1) {
    TRY(hex_copy_in(c, c->tmp2, df, c->ndof, on_device));
    TRY(hex_cellvec_gather(c, 0, c->tmp2, nullptr, -1.0, c->tmp));
  } else return femo_fail("femo_hex_dRdf_apply: unknown argument");
  return hex_copy_out(c, out, c->tmp, c->ndof, on_device);
}

/* which 0: compliance int f . w dx = (M f) . w;  1: volume sum_e rho_e |e| */
int femo_hex_output(femo_ctx* c, int which, double* val) {
  NEED_HEX(c, "femo_hex_output");
  if (!val) return femo_fail("femo_hex_output: null output");
  DevGuardH g(c->device);
  if (which == 0) {
    TRY(hex_load(c, c->tmp));
    return dev_dot(c, c->tmp, c->w, c->ndof, val);
  }
  if (which == 1) {
    if (c->world == 1) return dev_dot(c, c->hx_rho, c->hx_vol, c->nc, val);
    // multi-GPU: every cell is counted by the rank that owns it (owned cells come first), then summed over ranks
    TRY(dev_dot_async(c, c->hx_rho, c->hx_vol, c->nc_owned, c->scal + 9));
    TRY(dist_allreduce_sum(c, c->scal + 9, 1));
    CUDA_TRY(cudaMemcpyAsync(val, c->scal + 9, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return 0;
  }
  return femo_fail("femo_hex_output: unknown output");
}

/* gradient of output `which` w.r.t. arg 0 (density, n_cells), 1 (body force, 3 n_nodes) or 100 (state, 3 n_nodes) */
int femo_hex_output_grad(femo_ctx* c, int which, int arg, double* gout, int on_device) {
  NEED_HEX(c, "femo_hex_output_grad");
  if (!gout) return femo_fail("femo_hex_output_grad: null output");
  if (which != 0 && which != 1) return femo_fail("femo_hex_output_grad: unknown output");
  if (arg != 0 && arg != 1 && arg != 100) return femo_fail("femo_hex_output_grad: unknown argument");
  DevGuardH g(c->device);
  const int64_t n = (arg == 0) ? c->nc : c->ndof;
  if (which == 0 && arg == 100) { TRY(hex_load(c, c->tmp)); return hex_copy_out(c, gout, c->tmp, n, on_device); }
  if (which == 0 && arg == 1) { TRY(hex_cellvec_gather(c, 0, c->w, nullptr, 1.0, c->tmp)); return hex_copy_out(c, gout, c->tmp, n, on_device); }
  if (which == 1 && arg == 0) return hex_copy_out(c, gout, c->hx_vol, n, on_device);
  double* z = (arg == 0) ? c->cellbuf : c->tmp;
  CUDA_TRY(cudaMemsetAsync(z, 0, sizeof(double) * n, c->stream));
  return hex_copy_out(c, gout, z, n, on_device);
}

/* node coordinates and connectivity back (tests, post-processing) */
int femo_hex_get_mesh(femo_ctx* c, double* xyz, int32_t* conn) {
  NEED_HEX(c, "femo_hex_get_mesh");
  DevGuardH g(c->device);
  if (xyz) CUDA_TRY(cudaMemcpy(xyz, c->xyz, sizeof(double) * c->nb * 3, cudaMemcpyDeviceToHost));
  if (conn) CUDA_TRY(cudaMemcpy(conn, c->hx_cells, sizeof(int32_t) * c->nc * 8, cudaMemcpyDeviceToHost));
  return 0;
}

}  // extern "C"
