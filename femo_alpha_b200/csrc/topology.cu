// Mesh topology, Q2xQ1 block dofmap, BSR pattern of dR/du and the deterministic
// cell->nnz contribution permutation (built on the GPU with CUB radix sorts).
//
// Replaces what dolfinx does inside FunctionSpace(mesh, MixedElement[P2^3, P1^3]) and
// create_matrix(form) for femo_alpha (rm_shell/linear_shell_fenicsx/linear_shell_model.py:60-65,
// fea/utils_dolfinx.py:200-206).
#include <cub/cub.cuh>
#include <algorithm>
#include <map>
#include <cstdlib>
#include "femo_internal.cuh"

namespace {

__constant__ int c_local_edges[4][2] = {{0, 1}, {0, 2}, {1, 3}, {2, 3}};

__global__ void k_edge_keys(const int32_t* cv, int64_t nc, int64_t nn, uint64_t* keys, int32_t* vals) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= nc * 4) return;
  int64_t cell = i >> 2; int le = (int)(i & 3);
  int32_t a = cv[cell * 4 + c_local_edges[le][0]], b = cv[cell * 4 + c_local_edges[le][1]];
  int32_t lo = min(a, b), hi = max(a, b);
  keys[i] = (uint64_t)lo * (uint64_t)nn + (uint64_t)hi;
  vals[i] = (int32_t)i;
}

__global__ void k_head_flags(const uint64_t* keys, int64_t n, int32_t* flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  flag[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

__global__ void k_edge_scatter(const uint64_t* keys, const int32_t* vals, const int32_t* rank_incl, int64_t n,
                               int64_t nn, int32_t* ce, int32_t* edges) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t id = rank_incl[i] - 1;
  ce[vals[i]] = id;
  if (i == 0 || keys[i] != keys[i - 1]) {
    edges[2 * (int64_t)id] = (int32_t)(keys[i] / (uint64_t)nn);
    edges[2 * (int64_t)id + 1] = (int32_t)(keys[i] % (uint64_t)nn);
  }
}

__global__ void k_cell_blocks(const int32_t* cv, const int32_t* ce, int64_t nc, int64_t nn, int64_t ne,
                              int64_t nq2, int32_t* cb) {
  int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= nc) return;
  int32_t* o = cb + c * 13;
  for (int v = 0; v < 4; ++v) {
    o[v] = cv[c * 4 + v];
    o[4 + v] = (int32_t)(nn + ce[c * 4 + v]);
    o[9 + v] = (int32_t)(nq2 + cv[c * 4 + v]);
  }
  o[8] = (int32_t)(nn + ne + c);
}

// CellDiameter (max vertex distance) and planarity flag
__global__ void k_cell_size(const double* xyz, const int32_t* cv, int64_t nc, double* hE, int* not_planar) {
  int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= nc) return;
  V3 x[4];
  for (int v = 0; v < 4; ++v) x[v] = ldv3(xyz + 3 * (int64_t)cv[c * 4 + v]);
  double d2 = 0;
  for (int i = 0; i < 4; ++i)
    for (int j = i + 1; j < 4; ++j) { V3 d = x[i] - x[j]; d2 = fmax(d2, dot(d, d)); }
  double h = sqrt(d2);
  hE[c] = h;
  V3 a0 = 0.5 * ((x[1] - x[0]) + (x[3] - x[2]));
  V3 a1 = 0.5 * ((x[2] - x[0]) + (x[3] - x[1]));
  V3 n = cross(a0, a1);
  n = (1.0 / sqrt(dot(n, n))) * n;
  V3 cc = x[0] - x[1] - x[2] + x[3];
  if (fabs(dot(cc, n)) > 1e-12 * h) atomicOr(not_planar, 1);
}

__global__ void k_pair_keys(const int32_t* cb, int64_t ncontrib, uint64_t* keys, int32_t* vals) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= ncontrib) return;
  int64_t cell = i / 169; int pair = (int)(i - cell * 169);
  int a = pair / 13, b = pair - a * 13;
  keys[i] = ((uint64_t)(uint32_t)cb[cell * 13 + a] << 32) | (uint64_t)(uint32_t)cb[cell * 13 + b];
  vals[i] = (int32_t)i;
}

__global__ void k_pattern_fill(const uint64_t* keys, const int32_t* rank_incl, int64_t n, int64_t nn, int64_t nq2,
                               int32_t* cdest, int32_t* colidx, int32_t* cptr, int32_t* rowptr, int32_t* diag_pos,
                               int32_t* ut_pos, int32_t* tu_pos) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t d = rank_incl[i] - 1;
  cdest[i] = d;
  uint64_t k = keys[i];
  bool head = (i == 0) || (k != keys[i - 1]);
  if (!head) return;
  int32_t row = (int32_t)(k >> 32), col = (int32_t)(k & 0xffffffffu);
  colidx[d] = col;
  cptr[d] = (int32_t)i;
  if (i == 0 || (int32_t)(keys[i - 1] >> 32) != row) rowptr[row] = d;
  if (row == col) diag_pos[row] = d;
  if (row < nn && col == nq2 + row) ut_pos[row] = d;
  if (row >= nq2 && col == row - nq2) tu_pos[col] = d;
}

__global__ void k_adj_keys(const int32_t* cb, int64_t n, uint32_t* keys, int32_t* vals) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  keys[i] = (uint32_t)cb[i];
  vals[i] = (int32_t)i;
}

__global__ void k_adj_ptr(const uint32_t* keys, int64_t n, int64_t nb, int32_t* ptr) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (i == 0 || keys[i] != keys[i - 1]) ptr[keys[i]] = (int32_t)i;
  if (i == n - 1) ptr[nb] = (int32_t)n;
}

__global__ void k_find_pos(const int32_t* rowptr, const int32_t* colidx, const int32_t* rows, const int32_t* cols,
                           int64_t n, int32_t* pos) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t r = rows[i], cidx = cols[i];
  int32_t lo = rowptr[r], hi = rowptr[r + 1] - 1, res = -1;
  while (lo <= hi) {
    int32_t mid = (lo + hi) >> 1;
    int32_t v = colidx[mid];
    if (v == cidx) { res = mid; break; }
    if (v < cidx) lo = mid + 1; else hi = mid - 1;
  }
  pos[i] = res;
}

// ---- chunked symmetric assembly plan -------------------------------------------------------
__device__ __forceinline__ int pair_type(int a, int b) { return (a < 9 && b < 9) ? 0 : ((a >= 9 && b >= 9) ? 1 : 2); }

// flag contributions whose destination lies in the upper block triangle (row <= col)
__global__ void k_plan_flag(const int32_t* centry, const int32_t* cb, int64_t n, uint8_t* flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t e = centry[i];
  int32_t cell = e / 169; int pair = e - cell * 169;
  int a = pair / 13, b = pair - a * 13;
  flag[i] = cb[(int64_t)cell * 13 + a] <= cb[(int64_t)cell * 13 + b] ? 1 : 0;
}

// sort key = (owner chunk << 2) | type; owner chunk = chunk of the first (lowest) contributing cell
__global__ void k_plan_keys(const int32_t* sel, const int32_t* centry, const int32_t* cdest, const int32_t* cptr,
                            int64_t n, int C, uint32_t* key) {
  int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= n) return;
  int32_t i = sel[j];
  int32_t e = centry[i];
  int pair = e % 169;
  int32_t ocell = centry[cptr[cdest[i]]] / 169;
  key[j] = ((uint32_t)(ocell / C) << 2) | (uint32_t)pair_type(pair / 13, pair % 13);
}

__global__ void k_plan_halo_keys(const int32_t* src, const uint32_t* key, const int32_t* centry, int64_t n, int C,
                                 uint64_t* hkey, uint8_t* hflag) {
  int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= n) return;
  int32_t cell = centry[src[j]] / 169;
  uint32_t k = key[j] >> 2;
  bool halo = (uint32_t)(cell / C) != k;
  hkey[j] = ((uint64_t)k << 32) | (uint32_t)cell;
  hflag[j] = halo ? 1 : 0;
}

__global__ void k_plan_unique64(const uint64_t* in, int64_t n, uint8_t* flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  flag[i] = (i == 0 || in[i] != in[i - 1]) ? 1 : 0;
}

// ptr[k] = first index whose (key >> shift) >= k, k = 0..nk
template <typename K>
__global__ void k_plan_lower_bound(const K* keys, int64_t n, int shift, int64_t nk, int32_t* ptr) {
  int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k > nk) return;
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if ((int64_t)(keys[mid] >> shift) < k) lo = mid + 1; else hi = mid;
  }
  ptr[k] = (int32_t)lo;
}

__global__ void k_plan_items(const int32_t* src, const uint32_t* key, const int32_t* centry, const int32_t* cdest,
                             const int32_t* cptr, const uint64_t* hk, const int32_t* halo_ptr, int64_t n, int C,
                             uint32_t* it_pack, int32_t* it_dest, int* bad) {
  int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= n) return;
  int32_t i = src[j];
  int32_t e = centry[i];
  int32_t cell = e / 169; int pair = e - cell * 169;
  int32_t d = cdest[i];
  uint32_t k = key[j] >> 2;
  int slot;
  if ((uint32_t)(cell / C) == k) slot = cell - (int32_t)k * C;
  else {
    uint64_t want = ((uint64_t)k << 32) | (uint32_t)cell;
    int32_t lo = halo_ptr[k], hi = halo_ptr[k + 1];
    const int32_t base = lo;
    while (lo < hi) { int32_t mid = (lo + hi) >> 1; if (hk[mid] < want) lo = mid + 1; else hi = mid; }
    slot = C + (lo - base);
  }
  int32_t first = cptr[d], len = cptr[d + 1] - first;
  if (slot > 255 || len > 127) atomicOr(bad, 1);
  it_pack[j] = (uint32_t)slot | ((uint32_t)pair << 8) | ((first == i) ? (1u << 16) : 0u) | ((uint32_t)len << 17);
  it_dest[j] = d;
}

__global__ void k_plan_halo_cells(const uint64_t* hk, int64_t n, int32_t* cells) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) cells[i] = (int32_t)(hk[i] & 0xffffffffu);
}

// tpos[k] = position of block (col,row) for k = (row,col), row != col
__global__ void k_plan_tpos(const int32_t* rowptr, const int32_t* colidx, int64_t nb, int32_t* tpos) {
  int64_t row = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (row >= nb) return;
  for (int32_t k = rowptr[row]; k < rowptr[row + 1]; ++k) {
    int32_t col = colidx[k];
    int32_t res = -1;
    if (col != row) {
      int32_t lo = rowptr[col], hi = rowptr[col + 1] - 1;
      while (lo <= hi) {
        int32_t mid = (lo + hi) >> 1; int32_t v = colidx[mid];
        if (v == (int32_t)row) { res = mid; break; }
        if (v < (int32_t)row) lo = mid + 1; else hi = mid - 1;
      }
    }
    tpos[k] = res;
  }
}

int bits_for(uint64_t v) { int b = 0; while (v) { ++b; v >>= 1; } return b < 1 ? 1 : b; }

}  // namespace

int topo_build(femo_ctx* c, const int32_t* conn) {
  const int64_t nc = c->nc, nn = c->nn;
  // VTK cyclic -> dolfinx tensor vertex order (io.cells.perm_vtk: [0,1,3,2])
  std::vector<int32_t> cvh((size_t)nc * 4);
  for (int64_t i = 0; i < nc; ++i) {
    const int32_t* q = conn + 4 * i;
    for (int k = 0; k < 4; ++k)
      if (q[k] < 0 || q[k] >= nn) return femo_fail("femo_mesh_set: connectivity index out of range");
    cvh[4 * i + 0] = q[0]; cvh[4 * i + 1] = q[1]; cvh[4 * i + 2] = q[3]; cvh[4 * i + 3] = q[2];
  }
  TRY(dev_alloc(&c->cv, (size_t)nc * 4));
  CUDA_TRY(cudaMemcpyAsync(c->cv, cvh.data(), sizeof(int32_t) * nc * 4, cudaMemcpyHostToDevice, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));

  // ---- edges: unique sorted vertex pairs (non-manifold safe) ----
  const int64_t n4 = nc * 4;
  uint64_t *k_in = nullptr, *k_out = nullptr; int32_t *v_in = nullptr, *v_out = nullptr, *flag = nullptr, *rank = nullptr;
  TRY(dev_alloc(&k_in, n4)); TRY(dev_alloc(&k_out, n4)); TRY(dev_alloc(&v_in, n4)); TRY(dev_alloc(&v_out, n4));
  TRY(dev_alloc(&flag, n4)); TRY(dev_alloc(&rank, n4));
  k_edge_keys<<<(unsigned)cdiv(n4, 256), 256, 0, c->stream>>>(c->cv, nc, nn, k_in, v_in);
  TRY(check_launch(c, "k_edge_keys"));
  size_t tb = 0; void* tmp = nullptr;
  int ebits = bits_for((uint64_t)nn * (uint64_t)nn);
  cub::DeviceRadixSort::SortPairs(nullptr, tb, k_in, k_out, v_in, v_out, (int)n4, 0, ebits, c->stream);
  size_t tb2 = 0;
  cub::DeviceScan::InclusiveSum(nullptr, tb2, flag, rank, (int)n4, c->stream);
  tb = std::max(tb, tb2);
  CUDA_TRY(cudaMalloc(&tmp, tb));
  cub::DeviceRadixSort::SortPairs(tmp, tb, k_in, k_out, v_in, v_out, (int)n4, 0, ebits, c->stream);
  k_head_flags<<<(unsigned)cdiv(n4, 256), 256, 0, c->stream>>>(k_out, n4, flag);
  cub::DeviceScan::InclusiveSum(tmp, tb, flag, rank, (int)n4, c->stream);
  int32_t ne32 = 0;
  CUDA_TRY(cudaMemcpyAsync(&ne32, rank + n4 - 1, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  c->ne = ne32;
  TRY(dev_alloc(&c->ce, (size_t)n4)); TRY(dev_alloc(&c->edges, (size_t)c->ne * 2));
  k_edge_scatter<<<(unsigned)cdiv(n4, 256), 256, 0, c->stream>>>(k_out, v_out, rank, n4, nn, c->ce, c->edges);
  TRY(check_launch(c, "k_edge_scatter"));
  c->launches += 4;
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  cudaFree(tmp); dev_free(&k_in); dev_free(&k_out); dev_free(&v_in); dev_free(&v_out); dev_free(&flag); dev_free(&rank);

  c->nq2 = c->nn + c->ne + c->nc;
  c->nb = c->nq2 + c->nn;
  c->ndof = 3 * c->nb;
  if (c->nb >= (int64_t)1 << 31 || c->nc * 169 >= (int64_t)1 << 31)
    return femo_fail("mesh too large for 32-bit block / contribution indices");

  TRY(dev_alloc(&c->cell_blocks, (size_t)nc * 13));
  k_cell_blocks<<<(unsigned)cdiv(nc, 256), 256, 0, c->stream>>>(c->cv, c->ce, nc, c->nn, c->ne, c->nq2, c->cell_blocks);
  TRY(check_launch(c, "k_cell_blocks"));
  TRY(dev_alloc(&c->hE, (size_t)nc));
  int* np_flag = nullptr; TRY(dev_alloc(&np_flag, 1));
  CUDA_TRY(cudaMemsetAsync(np_flag, 0, sizeof(int), c->stream));
  k_cell_size<<<(unsigned)cdiv(nc, 256), 256, 0, c->stream>>>(c->xyz, c->cv, nc, c->hE, np_flag);
  TRY(check_launch(c, "k_cell_size"));
  int npf = 0;
  CUDA_TRY(cudaMemcpyAsync(&npf, np_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  dev_free(&np_flag);
  c->all_planar = npf ? 0 : 1;
  c->launches += 2;
  return 0;
}

int pattern_build(femo_ctx* c) {
  const int64_t nc = c->nc, n = nc * 169;
  c->ncontrib = n;
  uint64_t *k_in = nullptr, *k_out = nullptr; int32_t *v_in = nullptr, *flag = nullptr, *rank = nullptr;
  TRY(dev_alloc(&k_in, n)); TRY(dev_alloc(&k_out, n)); TRY(dev_alloc(&v_in, n));
  TRY(dev_alloc(&c->centry, n));
  k_pair_keys<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(c->cell_blocks, n, k_in, v_in);
  TRY(check_launch(c, "k_pair_keys"));
  size_t tb = 0, tb2 = 0; void* tmp = nullptr;
  int kb = 32 + bits_for((uint64_t)c->nb);
  cub::DeviceRadixSort::SortPairs(nullptr, tb, k_in, k_out, v_in, c->centry, (int)n, 0, kb, c->stream);
  cub::DeviceScan::InclusiveSum(nullptr, tb2, (int32_t*)nullptr, (int32_t*)nullptr, (int)n, c->stream);
  tb = std::max(tb, tb2);
  CUDA_TRY(cudaMalloc(&tmp, tb));
  // stable LSD radix sort: contributions to one block stay in ascending (cell, pair) order
  cub::DeviceRadixSort::SortPairs(tmp, tb, k_in, k_out, v_in, c->centry, (int)n, 0, kb, c->stream);
  dev_free(&k_in); dev_free(&v_in);
  TRY(dev_alloc(&flag, n)); TRY(dev_alloc(&rank, n));
  k_head_flags<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(k_out, n, flag);
  cub::DeviceScan::InclusiveSum(tmp, tb, flag, rank, (int)n, c->stream);
  int32_t nnzb32 = 0;
  CUDA_TRY(cudaMemcpyAsync(&nnzb32, rank + n - 1, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  c->nnzb = nnzb32;
  dev_free(&flag);
  TRY(dev_alloc(&c->cdest, n)); TRY(dev_alloc(&c->colidx, (size_t)c->nnzb)); TRY(dev_alloc(&c->cptr, (size_t)c->nnzb + 1));
  TRY(dev_alloc(&c->rowptr, (size_t)c->nb + 1)); TRY(dev_alloc(&c->diag_pos, (size_t)c->nb));
  TRY(dev_alloc(&c->ut_pos, (size_t)c->nn)); TRY(dev_alloc(&c->tu_pos, (size_t)c->nn));
  k_pattern_fill<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(k_out, rank, n, c->nn, c->nq2, c->cdest, c->colidx, c->cptr,
                                                              c->rowptr, c->diag_pos, c->ut_pos, c->tu_pos);
  TRY(check_launch(c, "k_pattern_fill"));
  int32_t last = (int32_t)n, lastb = nnzb32;
  CUDA_TRY(cudaMemcpyAsync(c->cptr + c->nnzb, &last, sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
  CUDA_TRY(cudaMemcpyAsync(c->rowptr + c->nb, &lastb, sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  dev_free(&k_out); dev_free(&rank); cudaFree(tmp);
  TRY(dev_alloc(&c->vals, (size_t)c->nnzb * 9));
  c->launches += 5;

  // ---- block node -> (cell, local index) adjacency ----
  const int64_t n13 = nc * 13;
  uint32_t *a_in = nullptr, *a_out = nullptr; int32_t* av_in = nullptr;
  TRY(dev_alloc(&a_in, n13)); TRY(dev_alloc(&a_out, n13)); TRY(dev_alloc(&av_in, n13));
  TRY(dev_alloc(&c->nc_adj, n13)); TRY(dev_alloc(&c->nc_ptr, (size_t)c->nb + 1));
  k_adj_keys<<<(unsigned)cdiv(n13, 256), 256, 0, c->stream>>>(c->cell_blocks, n13, a_in, av_in);
  tb = 0; tmp = nullptr;
  cub::DeviceRadixSort::SortPairs(nullptr, tb, a_in, a_out, av_in, c->nc_adj, (int)n13, 0, bits_for((uint64_t)c->nb), c->stream);
  CUDA_TRY(cudaMalloc(&tmp, tb));
  cub::DeviceRadixSort::SortPairs(tmp, tb, a_in, a_out, av_in, c->nc_adj, (int)n13, 0, bits_for((uint64_t)c->nb), c->stream);
  k_adj_ptr<<<(unsigned)cdiv(n13, 256), 256, 0, c->stream>>>(a_out, n13, c->nb, c->nc_ptr);
  TRY(check_launch(c, "k_adj_ptr"));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  dev_free(&a_in); dev_free(&a_out); dev_free(&av_in); cudaFree(tmp);
  c->launches += 3;
  return 0;
}

// Penalty Dirichlet BC bookkeeping (linear_shell_model.py:323-333; utils_dolfinx.py:555-565).
// Facets = mesh edges whose two vertices satisfy the predicate; each adjacent cell adds
// beta/h_E(cell) * int (w.dw) ds on the edge trace (exterior: ds, interior: '+' and '-' of dS).
int penalty_build(femo_ctx* c, const uint8_t* von) {
  const int64_t nc = c->nc;
  std::vector<int32_t> cv((size_t)nc * 4), cb((size_t)nc * 13);
  CUDA_TRY(cudaMemcpy(cv.data(), c->cv, sizeof(int32_t) * nc * 4, cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemcpy(cb.data(), c->cell_blocks, sizeof(int32_t) * nc * 13, cudaMemcpyDeviceToHost));
  static const int LE[4][2] = {{0, 1}, {0, 2}, {1, 3}, {2, 3}};
  static const int EQ2[4][3] = {{0, 4, 1}, {0, 5, 2}, {1, 6, 3}, {2, 7, 3}};
  std::vector<int32_t> pen;
  for (int64_t cell = 0; cell < nc; ++cell)
    for (int le = 0; le < 4; ++le)
      if (von[cv[4 * cell + LE[le][0]]] && von[cv[4 * cell + LE[le][1]]]) pen.push_back((int32_t)(cell * 4 + le));
  c->n_pen = (int64_t)pen.size();
  dev_free(&c->pen_cell); dev_free(&c->pen_dest_pos); dev_free(&c->pen_dest_ptr); dev_free(&c->pen_src); dev_free(&c->pen_me);
  dev_free(&c->pen_val); dev_free(&c->pen_dest_col); dev_free(&c->pen_row); dev_free(&c->pen_row_ptr);
  dev_free(&c->pen_vkey); dev_free(&c->pen_vptr); dev_free(&c->pen_vadj); dev_free(&c->pen_bkey); dev_free(&c->pen_bptr);
  dev_free(&c->pen_badj); dev_free(&c->pen_buf);
  c->n_pen_dest = 0; c->n_pen_rows = 0; c->n_pen_vkey = 0; c->n_pen_bkey = 0;
  c->jac_valid = false; c->prec_valid = false;
  if (c->n_pen == 0) return 0;
  // destinations: (row block, col block) -> list of sources (pair*25 + i*5 + j); u-u (3x3) and th-th (2x2) traces
  std::map<std::pair<int32_t, int32_t>, std::vector<int32_t>> dst;
  for (size_t p = 0; p < pen.size(); ++p) {
    int64_t cell = pen[p] >> 2; int le = pen[p] & 3;
    int32_t blk[5] = {cb[13 * cell + EQ2[le][0]], cb[13 * cell + EQ2[le][1]], cb[13 * cell + EQ2[le][2]],
                      cb[13 * cell + 9 + LE[le][0]], cb[13 * cell + 9 + LE[le][1]]};
    for (int i = 0; i < 5; ++i)
      for (int j = 0; j < 5; ++j) {
        if ((i < 3) != (j < 3)) continue;
        dst[{blk[i], blk[j]}].push_back((int32_t)(p * 25 + i * 5 + j));
      }
  }
  std::vector<int32_t> rows, cols, ptr(1, 0), src;
  for (auto& kv : dst) {
    rows.push_back(kv.first.first); cols.push_back(kv.first.second);
    for (int32_t s : kv.second) src.push_back(s);
    ptr.push_back((int32_t)src.size());
  }
  c->n_pen_dest = (int64_t)rows.size();
  std::vector<int32_t> urow, urow_ptr;
  for (size_t i = 0; i < rows.size(); ++i)
    if (i == 0 || rows[i] != rows[i - 1]) { urow.push_back(rows[i]); urow_ptr.push_back((int32_t)i); }
  urow_ptr.push_back((int32_t)rows.size());
  c->n_pen_rows = (int64_t)urow.size();
  TRY(dev_alloc(&c->pen_val, rows.size())); TRY(dev_alloc(&c->pen_dest_col, cols.size()));
  TRY(dev_alloc(&c->pen_row, urow.size())); TRY(dev_alloc(&c->pen_row_ptr, urow_ptr.size()));
  CUDA_TRY(cudaMemcpy(c->pen_dest_col, cols.data(), sizeof(int32_t) * cols.size(), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(c->pen_row, urow.data(), sizeof(int32_t) * urow.size(), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(c->pen_row_ptr, urow_ptr.data(), sizeof(int32_t) * urow_ptr.size(), cudaMemcpyHostToDevice));
  int32_t *drows = nullptr, *dcols = nullptr;
  TRY(dev_alloc(&drows, rows.size())); TRY(dev_alloc(&dcols, cols.size()));
  TRY(dev_alloc(&c->pen_cell, pen.size())); TRY(dev_alloc(&c->pen_dest_pos, rows.size()));
  TRY(dev_alloc(&c->pen_dest_ptr, ptr.size())); TRY(dev_alloc(&c->pen_src, src.size()));
  TRY(dev_alloc(&c->pen_me, pen.size() * 25));
  CUDA_TRY(cudaMemcpy(drows, rows.data(), sizeof(int32_t) * rows.size(), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(dcols, cols.data(), sizeof(int32_t) * cols.size(), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(c->pen_cell, pen.data(), sizeof(int32_t) * pen.size(), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(c->pen_dest_ptr, ptr.data(), sizeof(int32_t) * ptr.size(), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(c->pen_src, src.data(), sizeof(int32_t) * src.size(), cudaMemcpyHostToDevice));
  k_find_pos<<<(unsigned)cdiv(c->n_pen_dest, 256), 256, 0, c->stream>>>(c->rowptr, c->colidx, drows, dcols, c->n_pen_dest,
                                                                       c->pen_dest_pos);
  TRY(check_launch(c, "k_find_pos"));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  c->launches += 1;
  dev_free(&drows); dev_free(&dcols);
  return 0;
}

// Chunked symmetric assembly plan.  The contribution list of pattern_build
// (sorted by destination, ascending cell inside a destination) is filtered to destinations with
// row <= col and stable-sorted by (owner chunk, type): inside one chunk the items are grouped by
// type (uu | theta-theta | u-theta), then by destination, then by cell -> fixed summation order.
int asm_plan_build(femo_ctx* c) {
  c->plan_ok = false;
  dev_free(&c->it_pack); dev_free(&c->it_dest); dev_free(&c->chunk_ptr); dev_free(&c->halo_ptr);
  dev_free(&c->halo_cells); dev_free(&c->tpos);
  const char* envc = getenv("FEMO_ASM_CHUNK");
  int C = envc ? atoi(envc) : 8;
  if (C < 1) return 0;                       // FEMO_ASM_CHUNK=0 disables the plan (gather kernel is used)
  if (C > 64) C = 64;
  c->plan_C = C;
  const int64_t n = c->ncontrib;
  const int64_t nchunks = cdiv(c->nc, C);
  c->nchunks = nchunks;
  uint8_t* flag = nullptr; int32_t *sel = nullptr, *src = nullptr, *dnum = nullptr;
  uint32_t *key_in = nullptr, *key = nullptr;
  TRY(dev_alloc(&flag, n)); TRY(dev_alloc(&sel, n)); TRY(dev_alloc(&dnum, 1));
  k_plan_flag<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(c->centry, c->cell_blocks, n, flag);
  TRY(check_launch(c, "k_plan_flag"));
  size_t tb = 0; void* tmp = nullptr;
  cub::CountingInputIterator<int32_t> count_it(0);
  cub::DeviceSelect::Flagged(nullptr, tb, count_it, flag, sel, dnum, (int)n, c->stream);
  CUDA_TRY(cudaMalloc(&tmp, tb));
  cub::DeviceSelect::Flagged(tmp, tb, count_it, flag, sel, dnum, (int)n, c->stream);
  int32_t ni32 = 0;
  CUDA_TRY(cudaMemcpyAsync(&ni32, dnum, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  cudaFree(tmp); tmp = nullptr; dev_free(&flag);
  const int64_t ni = ni32;
  c->n_items = ni;
  TRY(dev_alloc(&key_in, ni)); TRY(dev_alloc(&key, ni)); TRY(dev_alloc(&src, ni));
  k_plan_keys<<<(unsigned)cdiv(ni, 256), 256, 0, c->stream>>>(sel, c->centry, c->cdest, c->cptr, ni, C, key_in);
  TRY(check_launch(c, "k_plan_keys"));
  const int kbits = bits_for((uint64_t)nchunks) + 2;
  tb = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tb, key_in, key, sel, src, (int)ni, 0, kbits, c->stream);
  CUDA_TRY(cudaMalloc(&tmp, tb));
  cub::DeviceRadixSort::SortPairs(tmp, tb, key_in, key, sel, src, (int)ni, 0, kbits, c->stream);   // stable
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  cudaFree(tmp); tmp = nullptr; dev_free(&key_in); dev_free(&sel);
  TRY(dev_alloc(&c->chunk_ptr, (size_t)nchunks + 1));
  k_plan_lower_bound<uint32_t><<<(unsigned)cdiv(nchunks + 1, 256), 256, 0, c->stream>>>(key, ni, 2, nchunks, c->chunk_ptr);
  TRY(check_launch(c, "k_plan_lower_bound"));
  // halo cells: unique (chunk, cell) pairs with cell outside the chunk
  uint64_t *hkey = nullptr, *hsel = nullptr, *hsorted = nullptr, *hk = nullptr; uint8_t* hflag = nullptr;
  TRY(dev_alloc(&hkey, ni)); TRY(dev_alloc(&hflag, ni)); TRY(dev_alloc(&hsel, ni));
  k_plan_halo_keys<<<(unsigned)cdiv(ni, 256), 256, 0, c->stream>>>(src, key, c->centry, ni, C, hkey, hflag);
  TRY(check_launch(c, "k_plan_halo_keys"));
  tb = 0;
  cub::DeviceSelect::Flagged(nullptr, tb, hkey, hflag, hsel, dnum, (int)ni, c->stream);
  CUDA_TRY(cudaMalloc(&tmp, tb));
  cub::DeviceSelect::Flagged(tmp, tb, hkey, hflag, hsel, dnum, (int)ni, c->stream);
  int32_t nh32 = 0;
  CUDA_TRY(cudaMemcpyAsync(&nh32, dnum, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  cudaFree(tmp); tmp = nullptr; dev_free(&hkey);
  int64_t nh = nh32, nu = 0;
  TRY(dev_alloc(&hsorted, nh)); TRY(dev_alloc(&hk, nh));
  if (nh > 0) {
    tb = 0;
    const int hbits = 32 + bits_for((uint64_t)nchunks);
    cub::DeviceRadixSort::SortKeys(nullptr, tb, hsel, hsorted, (int)nh, 0, hbits, c->stream);
    CUDA_TRY(cudaMalloc(&tmp, tb));
    cub::DeviceRadixSort::SortKeys(tmp, tb, hsel, hsorted, (int)nh, 0, hbits, c->stream);
    cudaFree(tmp); tmp = nullptr;
    k_plan_unique64<<<(unsigned)cdiv(nh, 256), 256, 0, c->stream>>>(hsorted, nh, hflag);
    tb = 0;
    cub::DeviceSelect::Flagged(nullptr, tb, hsorted, hflag, hk, dnum, (int)nh, c->stream);
    CUDA_TRY(cudaMalloc(&tmp, tb));
    cub::DeviceSelect::Flagged(tmp, tb, hsorted, hflag, hk, dnum, (int)nh, c->stream);
    int32_t nu32 = 0;
    CUDA_TRY(cudaMemcpyAsync(&nu32, dnum, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    cudaFree(tmp); tmp = nullptr;
    nu = nu32;
  }
  dev_free(&hsel); dev_free(&hsorted); dev_free(&hflag);
  c->n_halo = nu;
  TRY(dev_alloc(&c->halo_ptr, (size_t)nchunks + 1)); TRY(dev_alloc(&c->halo_cells, (size_t)nu));
  k_plan_lower_bound<uint64_t><<<(unsigned)cdiv(nchunks + 1, 256), 256, 0, c->stream>>>(hk, nu, 32, nchunks, c->halo_ptr);
  if (nu > 0) k_plan_halo_cells<<<(unsigned)cdiv(nu, 256), 256, 0, c->stream>>>(hk, nu, c->halo_cells);
  TRY(check_launch(c, "k_plan_halo"));
  TRY(dev_alloc(&c->it_pack, (size_t)ni)); TRY(dev_alloc(&c->it_dest, (size_t)ni));
  int* bad = nullptr; TRY(dev_alloc(&bad, 1));
  CUDA_TRY(cudaMemsetAsync(bad, 0, sizeof(int), c->stream));
  k_plan_items<<<(unsigned)cdiv(ni, 256), 256, 0, c->stream>>>(src, key, c->centry, c->cdest, c->cptr, hk, c->halo_ptr, ni, C,
                                                            c->it_pack, c->it_dest, bad);
  TRY(check_launch(c, "k_plan_items"));
  TRY(dev_alloc(&c->tpos, (size_t)c->nnzb));
  k_plan_tpos<<<(unsigned)cdiv(c->nb, 128), 128, 0, c->stream>>>(c->rowptr, c->colidx, c->nb, c->tpos);
  TRY(check_launch(c, "k_plan_tpos"));
  int badh = 0;
  std::vector<int32_t> hp((size_t)nchunks + 1);
  CUDA_TRY(cudaMemcpyAsync(&badh, bad, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaMemcpyAsync(hp.data(), c->halo_ptr, sizeof(int32_t) * (nchunks + 1), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  dev_free(&bad); dev_free(&src); dev_free(&key); dev_free(&hk);
  int maxh = 0;
  for (int64_t k = 0; k < nchunks; ++k) maxh = std::max(maxh, hp[k + 1] - hp[k]);
  c->plan_max_slots = C + maxh;
  c->launches += 10;
  c->plan_ok = !badh;
  return 0;
}
