// Shape derivative of the penalty-BC facet terms on the device (maths in penalty_shape.cuh): one thread per
// (edge, cell) pair writes its 12 (vertex) or 15 (trace node) values, then one thread per vertex / block node adds the
// values of its pairs in ascending pair order (lists built by penalty_build) -- deterministic, no atomics.
// The element maths is covered on the CPU by tests/test_shape_host.py against the complex-step oracle, the launch by
// tests/test_gpu_shape.py::test_penalty_shape.
#include <cstdlib>
#include <map>
#include <vector>

#include "penalty_shape.cuh"

namespace {

__device__ __forceinline__ void load_pair(const double* __restrict__ xyz, const int32_t* __restrict__ cv,
                                          const double* __restrict__ uhat, const int32_t* __restrict__ cell_blocks, int32_t pc,
                                          V3* xv, V3* uh, int32_t* blk) {
  const int LE[4][2] = {{0, 1}, {0, 2}, {1, 3}, {2, 3}};
  const int EQ2[4][3] = {{0, 4, 1}, {0, 5, 2}, {1, 6, 3}, {2, 7, 3}};
  const int64_t cell = pc >> 2;
  const int le = pc & 3;
  for (int v = 0; v < 4; ++v) {
    const int32_t id = cv[cell * 4 + v];
    xv[v] = ldv3(xyz + 3 * (int64_t)id);
    uh[v] = uhat ? ldv3(uhat + 3 * (int64_t)id) : v3(0, 0, 0);
  }
  for (int i = 0; i < 3; ++i) blk[i] = cell_blocks[13 * cell + EQ2[le][i]];
  blk[3] = cell_blocks[13 * cell + 9 + LE[le][0]];
  blk[4] = cell_blocks[13 * cell + 9 + LE[le][1]];
}

// mode 0: buf[p][0..12) = (dR_pen/duhat)^T lam of the pair;  mode 1: buf[p][0..15) = (dR_pen/duhat) duhat
__global__ void k_penalty_shape_pairs(int mode, const double* __restrict__ xyz, const int32_t* __restrict__ cv,
                                      const double* __restrict__ uhat, const double* __restrict__ hE,
                                      const int32_t* __restrict__ cell_blocks, const int32_t* __restrict__ pen_cell, int64_t n_pen,
                                      const double* __restrict__ w, const double* __restrict__ x, double* __restrict__ buf) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= n_pen) return;
  const int32_t pc = pen_cell[p];
  V3 xv[4], uh[4];
  int32_t blk[5];
  load_pair(xyz, cv, uhat, cell_blocks, pc, xv, uh, blk);
  double we[15];
  for (int i = 0; i < 5; ++i)
    for (int d = 0; d < 3; ++d) we[3 * i + d] = w[3 * (int64_t)blk[i] + d];
  double out[15];
  if (mode == 0) {
    double le_[15];
    for (int i = 0; i < 5; ++i)
      for (int d = 0; d < 3; ++d) le_[3 * i + d] = x[3 * (int64_t)blk[i] + d];
    penalty_shape_T_pair(xv, uh, pc & 3, hE[pc >> 2], we, le_, out);
    for (int i = 0; i < 12; ++i) buf[p * 15 + i] = out[i];
  } else {
    V3 du[4];
    for (int v = 0; v < 4; ++v) du[v] = ldv3(x + 3 * (int64_t)cv[(int64_t)(pc >> 2) * 4 + v]);
    penalty_shape_fwd_pair(xv, uh, pc & 3, hE[pc >> 2], we, du, out);
    for (int i = 0; i < 15; ++i) buf[p * 15 + i] = out[i];
  }
}

// out[3 key + d] += sum over the entries e = pair * per + i of the group:  buf[pair * 15 + 3 i + d]
__global__ void k_penalty_shape_gather(const int32_t* __restrict__ key, const int32_t* __restrict__ ptr,
                                       const int32_t* __restrict__ adj, int64_t nkey, int per, const double* __restrict__ buf,
                                       double* __restrict__ out) {
  const int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (g >= nkey) return;
  double s0 = 0, s1 = 0, s2 = 0;
  for (int32_t e = ptr[g]; e < ptr[g + 1]; ++e) {
    const int32_t pair = adj[e] / per, i = adj[e] - pair * per;
    const double* b = buf + (int64_t)pair * 15 + 3 * i;
    s0 += b[0]; s1 += b[1]; s2 += b[2];
  }
  double* o = out + 3 * (int64_t)key[g];
  o[0] += s0; o[1] += s1; o[2] += s2;
}

// pairs grouped by vertex and by trace block node (ascending pair order inside a group), built on first use from the
// lists of penalty_build (which frees them whenever the penalised edges change)
int ensure_lists(femo_ctx* c) {
  if (c->pen_buf) return 0;
  const int LE[4][2] = {{0, 1}, {0, 2}, {1, 3}, {2, 3}};
  const int EQ2[4][3] = {{0, 4, 1}, {0, 5, 2}, {1, 6, 3}, {2, 7, 3}};
  std::vector<int32_t> pen((size_t)c->n_pen), cv((size_t)c->nc * 4), cb((size_t)c->nc * 13);
  CUDA_TRY(cudaMemcpy(pen.data(), c->pen_cell, sizeof(int32_t) * pen.size(), cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemcpy(cv.data(), c->cv, sizeof(int32_t) * cv.size(), cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemcpy(cb.data(), c->cell_blocks, sizeof(int32_t) * cb.size(), cudaMemcpyDeviceToHost));
  auto group = [&](int per, auto key_of, int32_t** dkey, int32_t** dptr, int32_t** dadj, int64_t* nkey) -> int {
    std::map<int32_t, std::vector<int32_t>> g;
    for (size_t p = 0; p < pen.size(); ++p)
      for (int i = 0; i < per; ++i) g[key_of(p, i)].push_back((int32_t)(p * per + i));
    std::vector<int32_t> key, ptr(1, 0), adj;
    for (auto& kv : g) {
      key.push_back(kv.first);
      adj.insert(adj.end(), kv.second.begin(), kv.second.end());
      ptr.push_back((int32_t)adj.size());
    }
    *nkey = (int64_t)key.size();
    TRY(dev_alloc(dkey, key.size())); TRY(dev_alloc(dptr, ptr.size())); TRY(dev_alloc(dadj, adj.size()));
    CUDA_TRY(cudaMemcpy(*dkey, key.data(), sizeof(int32_t) * key.size(), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(*dptr, ptr.data(), sizeof(int32_t) * ptr.size(), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(*dadj, adj.data(), sizeof(int32_t) * adj.size(), cudaMemcpyHostToDevice));
    return 0;
  };
  TRY(group(4, [&](size_t p, int v) { return cv[4 * (int64_t)(pen[p] >> 2) + v]; }, &c->pen_vkey, &c->pen_vptr, &c->pen_vadj,
            &c->n_pen_vkey));
  TRY(group(5, [&](size_t p, int i) {
        const int64_t cell = pen[p] >> 2; const int le = pen[p] & 3;
        return i < 3 ? cb[13 * cell + EQ2[le][i]] : cb[13 * cell + 9 + LE[le][i - 3]];
      }, &c->pen_bkey, &c->pen_bptr, &c->pen_badj, &c->n_pen_bkey));
  TRY(dev_alloc(&c->pen_buf, pen.size() * 15));
  return 0;
}

int run(femo_ctx* c, int mode, const double* x, double* out) {
  if (c->n_pen == 0) return 0;
  // On by default since its first green hardware run (round 1 driver box: tests/test_gpu_shape.py::test_penalty_shape);
  // FEMO_PENALTY_SHAPE=0 drops the term again (it is O(reaction^2 / beta)).
  const char* env = getenv("FEMO_PENALTY_SHAPE");
  if (env && env[0] == '0') return 0;
  TRY(ensure_lists(c));
  const double* uh = c->uhat_zero ? nullptr : c->uhat;
  k_penalty_shape_pairs<<<(unsigned)cdiv(c->n_pen, 128), 128, 0, c->stream>>>(mode, c->xyz, c->cv, uh, c->hE, c->cell_blocks,
                                                                            c->pen_cell, c->n_pen, c->w, x, c->pen_buf);
  TRY(check_launch(c, "k_penalty_shape_pairs"));
  if (mode == 0)
    k_penalty_shape_gather<<<(unsigned)cdiv(c->n_pen_vkey, 128), 128, 0, c->stream>>>(c->pen_vkey, c->pen_vptr, c->pen_vadj,
                                                                                    c->n_pen_vkey, 4, c->pen_buf, out);
  else
    k_penalty_shape_gather<<<(unsigned)cdiv(c->n_pen_bkey, 128), 128, 0, c->stream>>>(c->pen_bkey, c->pen_bptr, c->pen_badj,
                                                                                    c->n_pen_bkey, 5, c->pen_buf, out);
  TRY(check_launch(c, "k_penalty_shape_gather"));
  c->launches += 2;
  return 0;
}

}  // namespace

int penalty_shape_T(femo_ctx* c, const double* lam, double* out_nn3) { return run(c, 0, lam, out_nn3); }
int penalty_shape_fwd(femo_ctx* c, const double* duhat, double* out_ndof) { return run(c, 1, duhat, out_ndof); }
