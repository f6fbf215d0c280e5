// Internal declarations shared by the translation units of libfemo_b200.so.
// Element maths restates SURVEY.md Appendix A (reference UFL:
// femo_alpha/rm_shell/linear_shell_fenicsx/linear_shell_model.py:136-157,220-296,
// kinematics.py:12-106).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include <map>
#include "../../include/femo_b200.h"

#define FEMO_MAX_Q1D 5
#define FEMO_MAX_NQ 25
#define FEMO_SHEAR_K 0.833       /* linear_shell_model.py:149 */
#define FEMO_BETA 1.0e15         /* linear_shell_model.py:324 */
#define FEMO_REG_ALPHA1 1.0e-2   /* rm_shell_pde.py:67 */

// per-quadrature-point record written by k_cell_records and read by the gather assembly
#define REC_K00 0
#define REC_K01 1
#define REC_K10 2
#define REC_K11 3
#define REC_CM11 4
#define REC_CM12 5
#define REC_CM33 6
#define REC_HB 7
#define REC_CS 8
#define REC_CD 9
#define REC_NQF_PLANAR 10
#define REC_E0 10   /* 9 doubles: E0,E1,E2 */
#define REC_G0 19   /* 6 doubles: g_0, g_1 */
#define REC_NQF_WARPED 25
#define REC_HEADER 9  /* per-cell constant frame (planar path) */

void femo_set_error(const std::string& s);
int femo_fail(const std::string& s);

#define CUDA_TRY(expr)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess)                                                               \
      return femo_fail(std::string(#expr) + ": " + cudaGetErrorString(_e) + " at " +     \
                       __FILE__ + ":" + std::to_string(__LINE__));                       \
  } while (0)

#define TRY(expr)            \
  do {                       \
    int _r = (expr);         \
    if (_r != 0) return _r;  \
  } while (0)

// reference-element tables (device global memory; hot kernels stage them in shared memory)
struct Tables {
  int m, nq;
  double X0[FEMO_MAX_NQ], X1[FEMO_MAX_NQ], W[FEMO_MAX_NQ];
  double phi2[FEMO_MAX_NQ][9];
  double dphi2[FEMO_MAX_NQ][9][2];
  double n1[FEMO_MAX_NQ][4];
  double dn1[FEMO_MAX_NQ][4][2];
};

struct CoarseSolver;   // coarse.cu: block-tridiagonal direct solver

struct femo_ctx {
  int device = 0;
  cudaStream_t stream = 0;
  int64_t launches = 0;
  // mesh
  int64_t nn = 0, nc = 0, ne = 0, nq2 = 0, nb = 0, ndof = 0;
  double* xyz = nullptr;        // nn*3
  int32_t* cv = nullptr;        // nc*4 tensor-ordered vertices
  int32_t* ce = nullptr;        // nc*4 edge ids
  int32_t* edges = nullptr;     // ne*2
  int32_t* cell_blocks = nullptr;  // nc*13
  double* hE = nullptr;         // nc  CellDiameter
  int all_planar = 0;
  bool setup_done = false;
  // options
  int ewm = 0, ewp = 0, m1d = 4, nq = 16;
  Tables* tab = nullptr;        // device
  Tables* tab4 = nullptr;       // device: degree-4 rule (3x3) of the stress measure
  int nq4 = 9;
  double stress_m = 1e-6, stress_rho = 100.0;
  int32_t* cell_tags = nullptr; // nc (optional sub-domain ids)
  std::map<int, double> area_cache;
  Tables htab;                  // host copy
  // inputs (device)
  int64_t n_mat = 0, n_f = 0;
  double *h = nullptr, *E = nullptr, *nu = nullptr, *rho = nullptr, *f = nullptr, *uhat = nullptr;
  int uhat_zero = 1;
  // BSR pattern + deterministic contribution permutation
  int64_t nnzb = 0, ncontrib = 0;
  int32_t *rowptr = nullptr, *colidx = nullptr;
  double* vals = nullptr;
  int32_t *centry = nullptr;    // contribution -> cell*169 + a*13 + b, sorted by destination
  int32_t *cdest = nullptr;     // contribution -> destination BSR block
  int32_t *cptr = nullptr;      // destination -> first contribution (nnzb+1)
  int32_t *diag_pos = nullptr;  // nb
  // chunked symmetric assembly plan (planar meshes): see asm_plan_build in topology.cu
  bool plan_ok = false;
  int plan_C = 8;               // own cells per chunk (one CTA per chunk)
  int plan_max_slots = 0;       // max over chunks of own + halo cells
  int64_t n_items = 0, nchunks = 0, n_halo = 0;
  uint32_t *it_pack = nullptr;  // n_items: slot | pair<<8 | head<<16 | seglen<<17
  int32_t *it_dest = nullptr;   // n_items: destination BSR block (row <= col)
  int32_t *chunk_ptr = nullptr; // nchunks+1 -> items
  int32_t *halo_ptr = nullptr;  // nchunks+1 -> halo_cells
  int32_t *halo_cells = nullptr;
  int32_t *tpos = nullptr;      // nnzb: position of the transposed block, -1 on the diagonal
  int32_t *ut_pos = nullptr, *tu_pos = nullptr;   // nn: positions of (v, nq2+v) and (nq2+v, v)
  // tiled planar assembly (assemble_tiles.cu, tile_plan.hpp): chunks of 32 cells in processing order
  bool tile_ok = false;
  int tl_ws = 0;
  int tl_hrec = 0;              // halo cells whose records the tiled kernel stages in shared memory (0: read from global)
  int tl_sms = 148;
  int64_t tl_nchunks = 0, tl_ndest = 0, tl_nref = 0, tl_nhitem = 0;
  int tl_hitem_cap = 0, tl_hcell_cap = 0, tl_dest_cap = 0, tl_ref_cap = 0;
  size_t tl_smem = 0;
  int32_t *tl_cell_order = nullptr, *tl_dest_ptr = nullptr, *tl_dest_pos = nullptr, *tl_dest_tpos = nullptr;
  uint32_t *tl_dest_info = nullptr, *tl_refs = nullptr, *tl_hitems = nullptr;
  int32_t *tl_ref_ptr = nullptr, *tl_hitem_ptr = nullptr, *tl_hcell_ptr = nullptr, *tl_hcells = nullptr;
  int32_t *tl_hfix = nullptr;   // [chunk][tl_hrec] halo cells of a chunk, -1 padded (staged halo records)
  double *tl_rec = nullptr, *tl_frame = nullptr;  // [chunk][q][12 fields][32 lanes], [slot][9]
  // block node -> (cell, local index) adjacency, cell-ascending
  int32_t *nc_ptr = nullptr;    // nb+1
  int32_t *nc_adj = nullptr;    // nc*13 entries = cell*13 + a
  // quadrature records
  double* rec = nullptr;
  int rec_nqf = 0, rec_stride = 0;
  // warped meshes (assemble.cu: k_warped_elem + k_warped_gather): records grouped by 32 cells, scratch for the 91 blocks
  // a <= b of every element matrix
  int warped_ok = 0, warped_nb = 1, warped_nbt = 2;
  double* elem = nullptr;
  int32_t* wp_order = nullptr;  // runs of 256 blocks in the order of their first contributing cell (k_warped_gather)
  // boundary conditions
  double* bcmask = nullptr;     // ndof: 1 free, 0 clamped
  uint8_t* bcbits = nullptr;    // nb: bit d set = dof 3 b + d clamped (same information, 1 byte per block node)
  uint32_t* colbc = nullptr;    // nnzb: colidx | bcbits[colidx] << 29 (SpMV reads column and clamp flags in one word)
  int64_t n_bc = 0;
  // penalty (edge, cell) adjacencies
  int64_t n_pen = 0;            // number of (edge,cell) pairs
  int32_t* pen_cell = nullptr;  // n_pen: cell*4 + local edge
  int64_t n_pen_dest = 0;       // unique (row,col) blocks hit by the penalty
  int32_t* pen_dest_pos = nullptr;   // n_pen_dest: BSR position
  int32_t* pen_dest_ptr = nullptr;   // n_pen_dest+1 -> into pen_src
  int32_t* pen_src = nullptr;        // entries: pair*25 + (i*5+j)
  double* pen_me = nullptr;          // n_pen*25 trace mass matrices (scaled)
  double* pen_val = nullptr;         // n_pen_dest summed penalty coefficient (times I3)
  int32_t* pen_dest_col = nullptr;   // n_pen_dest: column block
  int64_t n_pen_rows = 0;
  int32_t* pen_row = nullptr;        // n_pen_rows: row block
  int32_t* pen_row_ptr = nullptr;    // n_pen_rows+1 -> into destinations
  // shape derivative of the penalty terms (penalty_shape.cu): pairs grouped by vertex and by trace block node
  int64_t n_pen_vkey = 0, n_pen_bkey = 0;
  int32_t *pen_vkey = nullptr, *pen_vptr = nullptr, *pen_vadj = nullptr;   // vertex -> entries pair*4 + local vertex
  int32_t *pen_bkey = nullptr, *pen_bptr = nullptr, *pen_badj = nullptr;   // block node -> entries pair*5 + trace index
  double* pen_buf = nullptr;         // n_pen*15 per-pair results
  // multi-GPU (one process per GPU): owner mask, halo lists, NCCL communicator (dist.cu)
  int world = 1, rank = 0;
  void* nccl_comm = nullptr;
  bool comm_owner = false;      // false: borrowed from another context of this rank (femo_dist_share_comm)
  int n_nbr = 0;
  std::vector<int> nbr_rank, send_off, recv_off;   // neighbour ranks and block offsets (n_nbr+1)
  int64_t n_send = 0, n_recv = 0;                 // blocks
  int32_t *send_blocks = nullptr, *recv_blocks = nullptr;
  double *send_buf = nullptr, *recv_buf = nullptr;
  double* own = nullptr;        // nb: 1 = block node owned by this rank, 0 = ghost (null when world == 1)
  int64_t nc_owned = 0;         // owned cells come first in the local cell order
  int64_t halo_count = 0, allreduce_count = 0;
  // halo / SpMV overlap: owned block rows without (interior) and with (boundary) a ghost column; the exchange runs on
  // its own stream between two events
  int32_t *rows_int = nullptr, *rows_bnd = nullptr;
  int64_t n_rows_int = 0, n_rows_bnd = 0;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_ready = nullptr, ev_halo = nullptr;
  cudaEvent_t ev_trace[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // FEMO_TRACE_OVERLAP=1: timed events of the last dist_spmv
  int trace_overlap = 0;
  // solver state
  double *w = nullptr, *b = nullptr, *r = nullptr, *z = nullptr, *p = nullptr, *Ap = nullptr, *tmp = nullptr;
  double *tmp2 = nullptr;
  double *dinv6 = nullptr;      // nn*36   inverse vertex blocks (u_v, theta_v)
  double *dinv3 = nullptr;      // (ne+nc)*9
  double *red = nullptr;        // reduction scratch
  double *scal = nullptr;       // device scalars
  double *cellbuf = nullptr;    // nc*39 element vectors
  bool jac_valid = false, prec_valid = false;
  int64_t mat_epoch = 0;        // bumped whenever c->vals or the BC masks change (consumers cache factorisations against it)
  CoarseSolver* direct = nullptr;   // femo_direct_* (coarse.cu)
  int64_t direct_epoch = -1;
  double t_rec = 0, t_asm = 0, t_solve = 0, t_iter = 0;
  double dmma_tf = 0;           // last measured fp64 mma.sync rate (TFLOP/s)
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaStream_t copy_stream = nullptr;   // femo_evaluate_residual: host <-> device copies overlapped with the kernels
  cudaEvent_t ev_copy[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  // dynamic (time-stepped) operator, dynamic.cu: while dyn_on, `vals` holds A_eff = (2/dt^2) M + K/2 (the matrix every
  // solver / preconditioner / multigrid level works on) and `vals_K` the stiffness K; `mvals` one scalar per BSR block
  // (the mass matrix is block-scalar: rho h int phi_a phi_b on u-u blocks, rho h hE^2 int N_a N_b on theta-theta blocks)
  int dyn_on = 0;
  bool dyn_swapped = false;
  double dyn_dt = 0.0;
  double *vals_K = nullptr, *mvals = nullptr, *mass_me = nullptr;
  Tables* tabm = nullptr;       // 4x4 rule of the mass form
  // generic 3-D forms on trilinear hexahedra (hex.cu): kind 1 contexts carry no shell data;
  // the BSR pattern, BC masks, Krylov vectors and the block-Jacobi PCG above are shared (nn = 0: every node is a
  // 3x3 block node of the `dinv3` kind)
  int kind = 0;                 // 0 RM shell, 1 hexahedral solid
  int hx_form = 1;              // 1 linear elasticity, 0 vector Poisson
  double hx_c1 = 0, hx_c2 = 0, hx_p = 3.0;   // (lam, mu) or (kappa, -); SIMP exponent
  int64_t hx_nn = 0;            // mesh nodes (= nb)
  int32_t* hx_cells = nullptr;  // nc*8, VTK_HEXAHEDRON order
  double *hx_geo = nullptr;     // nc*8*10 inverse Jacobians and weights
  double *hx_geo2 = nullptr;    // nc*8*26 gradient records (only with FEMO_HEX_GRAD=1: k_hex_assemble_g)
  double *hx_rho = nullptr, *hx_scale = nullptr, *hx_vol = nullptr;   // nc: density, rho^p, cell volume
  double *hx_f = nullptr;       // nn*3 nodal body force
  double *hx_cellvec = nullptr; // nc*24 element vectors
};

// ---- host-side helpers (context.cu) ---------------------------------------
template <typename T>
int dev_alloc(T** p, size_t n) {
  if (*p) { cudaFree(*p); *p = nullptr; }
  if (n == 0) n = 1;
  cudaError_t e = cudaMalloc((void**)p, n * sizeof(T));
  if (e != cudaSuccess) return femo_fail(std::string("cudaMalloc failed: ") + cudaGetErrorString(e));
  return 0;
}
template <typename T>
void dev_free(T** p) {
  if (*p) cudaFree(*p);
  *p = nullptr;
}
inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
int check_launch(femo_ctx* c, const char* what);

// ---- phase entry points ------------------------------------------------------
int topo_build(femo_ctx* c, const int32_t* conn_vtk_host);                // topology.cu
int pattern_build(femo_ctx* c);                                           // topology.cu
int penalty_build(femo_ctx* c, const uint8_t* vertex_on_host);            // topology.cu
int asm_plan_build(femo_ctx* c);                                          // topology.cu
int warped_setup(femo_ctx* c);                                            // assemble.cu
int tile_plan_setup(femo_ctx* c);                                         // assemble_tiles.cu
void tile_plan_free(femo_ctx* c);                                         // assemble_tiles.cu
int assemble_jacobian_tiles(femo_ctx* c, const double* uhat);             // assemble_tiles.cu (records + assembly)
int penalty_shape_T(femo_ctx* c, const double* lam, double* out_nn3);      // penalty_shape.cu: out += (dR_pen/duhat)^T lam
int penalty_shape_fwd(femo_ctx* c, const double* duhat, double* out_ndof); // penalty_shape.cu: out += (dR_pen/duhat) duhat
int hex_assemble(femo_ctx* c);                                            // hex.cu: K(rho) -> c->vals
int hex_load(femo_ctx* c, double* F);                                     // hex.cu: F = M f
int assemble_stiffness(femo_ctx* c);                                      // assemble.cu: K -> c->vals
int assemble_jacobian(femo_ctx* c);                                       // dynamic.cu: K, or A_eff while c->dyn_on
int assemble_load(femo_ctx* c, double* F_dev);                            // assemble.cu
int bsr_spmv(femo_ctx* c, const double* x, double* y, int with_bc);       // solver.cu
int bsr_spmv_vals(femo_ctx* c, const double* vals, const double* x, double* y, int with_bc);   // solver.cu
int bsr_spmv_range(femo_ctx* c, const double* x, double* y, int64_t r0, int64_t r1);           // solver.cu: block rows [r0, r1) of K x
int penalty_apply(femo_ctx* c, const double* x, double* y, double sign);  // solver.cu  y += sign * P x
int precond_setup(femo_ctx* c);                                           // solver.cu
int pcg_solve(femo_ctx* c, const double* b, double* x, double rtol, int max_it, int* iters,
              double* relres, bool fixed_iters);                          // solver.cu
int dev_dot(femo_ctx* c, const double* a, const double* b, int64_t n, double* out_host);   // solver.cu
int residual_dd(femo_ctx* c, const double* x, const double* b, double* r);   // coarse.cu: b - A x in double-double
int dev_dot_async(femo_ctx* c, const double* a, const double* b, int64_t n, double* out_dev);   // solver.cu (no sync)
int out_scalar(femo_ctx* c, int which, double* val);                      // outputs.cu
int out_grad(femo_ctx* c, int which, int arg, double* g_dev);             // outputs.cu
int dRdf_T_apply(femo_ctx* c, int arg, const double* lam, double* out);   // outputs.cu
int dRdf_apply(femo_ctx* c, int arg, const double* df, double* out);      // outputs.cu
void noalloc_prepare(int device);                                         // outputs.cu: address of the static scratch array
int measure_peaks(femo_ctx* c, double* tf, double* gbs);                  // solver.cu
int halo_exchange(femo_ctx* c, double* x);                                // dist.cu: ghost entries <- owners
int halo_exchange_begin(femo_ctx* c, double* x);                          // dist.cu: the same on the communication stream ...
int halo_exchange_end(femo_ctx* c);                                       // ... the context's stream waits for it here
int dist_spmv(femo_ctx* c, double* x, double* y, double* dot_int, double* dot_bnd, int* n_int, int* n_bnd);   // solver.cu
int vals_to_float(femo_ctx* c, float* out);                               // solver.cu: fp32 copy of c->vals (nnzb * 9)
int bsr_spmv_f32(femo_ctx* c, const float* vals32, double* x, double* y); // solver.cu: y = A32 x (BCs; halo overlapped when distributed)
int halo_accumulate(femo_ctx* c, double* x);                              // dist.cu: owners += ghost copies (reverse halo)
int dist_allreduce_sum(femo_ctx* c, double* dev_vals, int n);             // dist.cu: in-place on the stream
int dist_allreduce_host(femo_ctx* c, double* host_vals, int n);           // dist.cu
void dist_free(femo_ctx* c);                                              // dist.cu
int pnorm_value(femo_ctx* c, int tag, double* val);                       // outputs.cu
int pnorm_grad(femo_ctx* c, int tag, int arg, double* g);                 // outputs.cu
int stress_field(femo_ctx* c, int surface, double* out);                  // outputs.cu
// dynamic path: element kernels (outputs.cu) and dense factor (mg.cu)
inline const double* stiffness_vals(const femo_ctx* c) { return c->dyn_swapped ? c->vals_K : c->vals; }
int mass_elements(femo_ctx* c, double* me);                               // (nc,97): 81 Q2xQ2 + 16 Q1xQ1 scalars
int mass_param_T(femo_ctx* c, const double* acc, const double* lam, double g_acc, bool accumulate);   // -> cellbuf[0..3]
int stiff_param_T(femo_ctx* c, const double* w, const double* lam, bool accumulate);                  // -> cellbuf[0..3]
int mass_param_fwd(femo_ctx* c, const double* acc, const double* dh, double g_acc, double* out);     // (dM/dh dh) acc - dW/dh dh
int stiff_param_fwd(femo_ctx* c, const double* w, const double* dh, double* out);
int cellbuf_to_material(femo_ctx* c, double* out);
int self_weight_load(femo_ctx* c, double g_acc, double* out);
struct femo_mg;
int mg_num_levels(femo_mg* m);                                            // mg.cu
femo_ctx* mg_level_ctx(femo_mg* m, int l);
void mg_invalidate(femo_mg* m);
// coarse.cu: direct solver (block-tridiagonal factorisation on level sets) of the BC-applied c->vals
int coarse_analyse(femo_ctx* c, CoarseSolver** out);
int coarse_factor(femo_ctx* c, CoarseSolver* cs);
int coarse_apply(femo_ctx* c, CoarseSolver* cs, const double* b, double* x);     // device vectors; kernels only (capturable)
void coarse_free(CoarseSolver* cs);
int coarse_num_levels(const CoarseSolver* cs);
int64_t coarse_max_front(const CoarseSolver* cs);
double coarse_bytes(const CoarseSolver* cs);
bool coarse_is_dense(const CoarseSolver* cs);
int vec_mul(femo_ctx* c, const double* a, const double* m, int64_t n, double* out);                         // api.cu: out = a .* m
int vec_axpby(femo_ctx* c, double a, const double* x, double b, const double* y, int64_t n, double* out);   // api.cu

// ---- device maths --------------------------------------------------------------
#ifdef __CUDACC__
// host + device: the element maths can be exercised on the CPU by nvcc-compiled test harnesses (tests/shape_host.cu)
#define FEMO_HD __host__ __device__ __forceinline__
struct V3 { double x, y, z; };
FEMO_HD V3 v3(double x, double y, double z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
FEMO_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
FEMO_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
FEMO_HD V3 operator*(double s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
FEMO_HD double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
FEMO_HD V3 cross(V3 a, V3 b) {
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
FEMO_HD V3 ldv3(const double* p) { return v3(p[0], p[1], p[2]); }
// K[i][j] += a_i * b_j
FEMO_HD void outer_acc(double* K, V3 a, V3 b) {
  K[0] += a.x * b.x; K[1] += a.x * b.y; K[2] += a.x * b.z;
  K[3] += a.y * b.x; K[4] += a.y * b.y; K[5] += a.y * b.z;
  K[6] += a.z * b.x; K[7] += a.z * b.y; K[8] += a.z * b.z;
}

// position of block node I: Q2 vertices [0,nn), edge mid-points, cell centres, then the Q1 rotation nodes
FEMO_HD V3 block_coord(const double* xyz, const int32_t* edges, const int32_t* cv, int64_t nn, int64_t ne, int64_t nc, int64_t I) {
  const int64_t nq2 = nn + ne + nc;
  if (I < nn) return ldv3(xyz + 3 * I);
  if (I < nn + ne) {
    const int64_t e = I - nn;
    return 0.5 * (ldv3(xyz + 3 * (int64_t)edges[2 * e]) + ldv3(xyz + 3 * (int64_t)edges[2 * e + 1]));
  }
  if (I < nq2) {
    const int64_t c = I - nn - ne;
    return 0.25 * (ldv3(xyz + 3 * (int64_t)cv[4 * c]) + ldv3(xyz + 3 * (int64_t)cv[4 * c + 1]) +
                   ldv3(xyz + 3 * (int64_t)cv[4 * c + 2]) + ldv3(xyz + 3 * (int64_t)cv[4 * c + 3]));
  }
  return ldv3(xyz + 3 * (I - nq2));
}

// geometry of one cell at one reference point
struct QpGeom {
  V3 E0, E1, E2;        // local orthonormal frame           kinematics.py:54-70
  V3 Kx0, Kx1;          // rows of K F^-1: gradx(f) = df/dX0 Kx0 + df/dX1 Kx1   kinematics.py:12-21
  V3 Ki0, Ki1;          // rows of the plain pseudo-inverse (grad without uhat)
  V3 dE2_0, dE2_1;      // dE2/dX_alpha (zero on planar cells)
  double detJ;          // sqrt(det(J^T J))
  double Ju;            // det F(uhat)                        kinematics.py:24-32
  double Finv[9];
};

FEMO_HD void inv3x3(const double* A, double* Ai, double* det) {
  double c00 = A[4] * A[8] - A[5] * A[7], c01 = A[5] * A[6] - A[3] * A[8], c02 = A[3] * A[7] - A[4] * A[6];
  double d = A[0] * c00 + A[1] * c01 + A[2] * c02;
  double id = 1.0 / d;
  Ai[0] = c00 * id; Ai[1] = (A[2] * A[7] - A[1] * A[8]) * id; Ai[2] = (A[1] * A[5] - A[2] * A[4]) * id;
  Ai[3] = c01 * id; Ai[4] = (A[0] * A[8] - A[2] * A[6]) * id; Ai[5] = (A[2] * A[3] - A[0] * A[5]) * id;
  Ai[6] = c02 * id; Ai[7] = (A[1] * A[6] - A[0] * A[7]) * id; Ai[8] = (A[0] * A[4] - A[1] * A[3]) * id;
  *det = d;
}

// xv: 4 tensor-ordered vertex coordinates, uh: nodal mesh displacement (may be null -> zero)
FEMO_HD void qp_geometry(const V3* xv, const V3* uh, double X0, double X1, QpGeom& g) {
  // bilinear derivatives
  double dN0[4] = {-(1 - X1), (1 - X1), -X1, X1};
  double dN1[4] = {-(1 - X0), -X0, (1 - X0), X0};
  V3 a0 = v3(0, 0, 0), a1 = v3(0, 0, 0);
#pragma unroll
  for (int v = 0; v < 4; ++v) { a0 = a0 + dN0[v] * xv[v]; a1 = a1 + dN1[v] * xv[v]; }
  double g00 = dot(a0, a0), g01 = dot(a0, a1), g11 = dot(a1, a1);
  double dg = g00 * g11 - g01 * g01;
  double idg = 1.0 / dg;
  g.detJ = sqrt(dg);
  // pseudo-inverse rows: Kinv = G^-1 J^T
  g.Ki0 = (g11 * idg) * a0 - (g01 * idg) * a1;
  g.Ki1 = (g00 * idg) * a1 - (g01 * idg) * a0;
  V3 n = cross(a0, a1);
  double nl = sqrt(dot(n, n));
  g.E2 = (1.0 / nl) * n;
  g.E0 = (1.0 / sqrt(g00)) * a0;
  g.E1 = cross(g.E2, g.E0);
  V3 c = xv[0] - xv[1] - xv[2] + xv[3];
  V3 dn0 = cross(a0, c), dn1 = cross(c, a1);
  g.dE2_0 = (1.0 / nl) * (dn0 - dot(dn0, g.E2) * g.E2);
  g.dE2_1 = (1.0 / nl) * (dn1 - dot(dn1, g.E2) * g.E2);
  if (uh) {
    V3 u0 = v3(0, 0, 0), u1 = v3(0, 0, 0);
#pragma unroll
    for (int v = 0; v < 4; ++v) { u0 = u0 + dN0[v] * uh[v]; u1 = u1 + dN1[v] * uh[v]; }
    // F = I + u0 (x) Ki0 + u1 (x) Ki1
    double F[9] = {1 + u0.x * g.Ki0.x + u1.x * g.Ki1.x, u0.x * g.Ki0.y + u1.x * g.Ki1.y, u0.x * g.Ki0.z + u1.x * g.Ki1.z,
                   u0.y * g.Ki0.x + u1.y * g.Ki1.x, 1 + u0.y * g.Ki0.y + u1.y * g.Ki1.y, u0.y * g.Ki0.z + u1.y * g.Ki1.z,
                   u0.z * g.Ki0.x + u1.z * g.Ki1.x, u0.z * g.Ki0.y + u1.z * g.Ki1.y, 1 + u0.z * g.Ki0.z + u1.z * g.Ki1.z};
    inv3x3(F, g.Finv, &g.Ju);
    const double* Fi = g.Finv;
    g.Kx0 = v3(g.Ki0.x * Fi[0] + g.Ki0.y * Fi[3] + g.Ki0.z * Fi[6], g.Ki0.x * Fi[1] + g.Ki0.y * Fi[4] + g.Ki0.z * Fi[7],
               g.Ki0.x * Fi[2] + g.Ki0.y * Fi[5] + g.Ki0.z * Fi[8]);
    g.Kx1 = v3(g.Ki1.x * Fi[0] + g.Ki1.y * Fi[3] + g.Ki1.z * Fi[6], g.Ki1.x * Fi[1] + g.Ki1.y * Fi[4] + g.Ki1.z * Fi[7],
               g.Ki1.x * Fi[2] + g.Ki1.y * Fi[5] + g.Ki1.z * Fi[8]);
  } else {
    g.Ju = 1.0;
    g.Kx0 = g.Ki0; g.Kx1 = g.Ki1;
#pragma unroll
    for (int i = 0; i < 9; ++i) g.Finv[i] = (i % 4 == 0) ? 1.0 : 0.0;
  }
}

// material coefficients at a point (no quadrature weight), linear_shell_model.py:136-157,284-296
//   mode 0: value, 1: d/dh, 2: d/dE, 3: d/dnu
struct Coefs { double cm11, cm12, cm33, cb11, cb12, cb33, cs, cd; };
FEMO_HD Coefs clt_coefs(double h, double E, double nu, double hE, double Ju, int mode) {
  Coefs c;
  double om = 1.0 / (1.0 - nu * nu);
  double c0 = E * om;                 // E/(1-nu^2)
  double G = E / (2.0 * (1.0 + nu));
  double h3_12 = h * h * h / 12.0;
  double ihe2 = 1.0 / (hE * hE);
  if (mode == 0) {
    c.cm11 = h * c0; c.cm12 = h * c0 * nu; c.cm33 = h * c0 * 0.5 * (1 - nu);
    c.cb11 = h3_12 * c0; c.cb12 = h3_12 * c0 * nu; c.cb33 = h3_12 * c0 * 0.5 * (1 - nu);
    c.cs = FEMO_SHEAR_K * G * h * Ju; c.cd = E * h * h * h * ihe2 * Ju;
  } else if (mode == 1) {
    double h2_4 = h * h / 4.0;
    c.cm11 = c0; c.cm12 = c0 * nu; c.cm33 = c0 * 0.5 * (1 - nu);
    c.cb11 = h2_4 * c0; c.cb12 = h2_4 * c0 * nu; c.cb33 = h2_4 * c0 * 0.5 * (1 - nu);
    c.cs = FEMO_SHEAR_K * G * Ju; c.cd = 3.0 * E * h * h * ihe2 * Ju;
  } else if (mode == 2) {
    c.cm11 = h * om; c.cm12 = h * om * nu; c.cm33 = h * om * 0.5 * (1 - nu);
    c.cb11 = h3_12 * om; c.cb12 = h3_12 * om * nu; c.cb33 = h3_12 * om * 0.5 * (1 - nu);
    c.cs = FEMO_SHEAR_K * h * Ju / (2.0 * (1.0 + nu)); c.cd = h * h * h * ihe2 * Ju;
  } else {
    double dc = 2.0 * E * nu * om * om;
    double d11 = dc, d12 = dc * nu + c0, d33 = dc * 0.5 * (1 - nu) - 0.5 * c0;
    c.cm11 = h * d11; c.cm12 = h * d12; c.cm33 = h * d33;
    c.cb11 = h3_12 * d11; c.cb12 = h3_12 * d12; c.cb33 = h3_12 * d33;
    c.cs = -FEMO_SHEAR_K * E / (2.0 * (1.0 + nu) * (1.0 + nu)) * h * Ju; c.cd = 0.0;
  }
  return c;
}

// strain rows (as global 3-vectors) of one displacement node / one rotation node at a point
struct URows { V3 e0, e1, e2, g0, g1, dr; };
struct TRows { V3 k0, k1, k2, g0, g1, dr; };
FEMO_HD URows u_rows(double d0, double d1, V3 E0, V3 E1, V3 E2) {
  URows r;
  r.e0 = d0 * E0; r.e1 = d1 * E1; r.e2 = d1 * E0 + d0 * E1;     // membrane   model:232-239
  r.g0 = d0 * E2; r.g1 = d1 * E2;                                // shear      model:245-258
  r.dr = 0.5 * (d1 * E0 - d0 * E1);                              // drilling   model:288-289
  return r;
}
// gq0,gq1: g_q = sum_alpha k_alpha,q dE2/dX_alpha (zero on planar cells)
FEMO_HD TRows t_rows(double N, double d0, double d1, V3 E0, V3 E1, V3 E2, V3 gq0, V3 gq1) {
  TRows r;
  V3 r00 = N * cross(E0, gq0) - d0 * E1;
  V3 r11 = d1 * E0 + N * cross(E1, gq1);
  V3 r01 = N * cross(E0, gq1) - d1 * E1;
  V3 r10 = d0 * E0 + N * cross(E1, gq0);
  r.k0 = r00; r.k1 = r11; r.k2 = r01 + r10;                      // bending    model:241-243
  r.g0 = N * E1; r.g1 = (-N) * E0;                               // shear (-cross(E2,theta) part)
  r.dr = N * E2;                                                 // drilling (theta . E2)
  return r;
}
#endif
