// Host-side plan of the tiled planar assembly kernel (assemble_tiles.cu).
// Pure C++ (no CUDA) so that the CPU test-suite can build and check it without a GPU
// (tests/test_tile_plan.py compiles this header with g++ through tests/tile_plan_host.cpp).
//
// The matrix written is the one dolfinx' assemble_matrix(dR_du) produces
// (femo_alpha/fea/utils_dolfinx.py:200-221); the plan only fixes WHO writes each 3x3 block and in
// which order the element contributions are added, so that the result is bitwise reproducible.
//
//   processing order   cells are visited in `order` (a permutation; chunk k = slots [16k, 16k+16))
//   destination        a BSR block (row <= col); it is owned by the chunk of its first contributing
//                      cell in processing order; the owner also writes the transposed block
//   ref                one element contribution (cell, a, b) to a destination: where its local-frame
//                      moments lie in the CTA's shared memory, which frame rotates them, transposed?
//   halo item          a contribution whose cell belongs to another chunk: its moments are computed
//                      one pair at a time by the owner (the cell's own chunk never needs to talk)
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <algorithm>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

// threads of the host-side plan construction: FEMO_HOST_THREADS if set (launchers such as torchrun export
// OMP_NUM_THREADS=1 for every rank), else OpenMP's default
inline int femo_host_threads() {
  const char* e = getenv("FEMO_HOST_THREADS");
  int n = e ? atoi(e) : 0;
#ifdef _OPENMP
  if (n <= 0) n = omp_get_max_threads();
#else
  n = 1;
#endif
  return n < 1 ? 1 : n;
}

namespace femo_tiles {

constexpr int C = 16;                 // cells per chunk = lanes of a half-warp (a warp runs two tiles)
constexpr int MOM_UU = 0;             // 45 symmetric Q2 pairs x 5 moments
constexpr int MOM_TT = 225;           // 10 symmetric Q1 pairs x 5 moments
constexpr int MOM_UT = 275;           // 9 x 4 (Q2, Q1) pairs x 4 moments
constexpr int MOM_CELL = 419;         // doubles per cell (odd -> conflict-free lane stride)
constexpr int MOM_HALO_BASE = C * MOM_CELL;
constexpr int MOM_HALO = 5;           // doubles per halo item
// ref word: [15:0] moment offset (doubles)  [17:16] type 0 uu 1 tt 2 ut  [18] transposed  [31:19] frame slot
// dest info: [19:0] first ref relative to the chunk's ref base  [31:20] number of refs
// halo item: [31:8] processing index of the cell  [7:0] lo*13 + hi (local node indices, lo <= hi)

struct Plan {
  int64_t nchunks = 0;
  std::vector<int32_t> cell_order;    // nchunks*C : slot -> cell, -1 = padding
  std::vector<int32_t> dest_ptr;      // nchunks+1
  std::vector<int32_t> dest_pos, dest_tpos;
  std::vector<uint32_t> dest_info;
  std::vector<int32_t> ref_ptr;       // nchunks+1
  std::vector<uint32_t> refs;
  std::vector<int32_t> hitem_ptr;     // nchunks+1
  std::vector<uint32_t> hitems;
  std::vector<int32_t> hcell_ptr;     // nchunks+1
  std::vector<int32_t> hcells;        // processing indices of the distinct halo cells of a chunk
  int max_dests = 0, max_hitems = 0, max_hcells = 0, max_refs = 0;
};

inline int sym_index(int lo, int hi, int n) { return lo * n - lo * (lo - 1) / 2 + (hi - lo); }

// a, b: local node indices (0..8 Q2, 9..12 Q1) of the ROW and COLUMN block of the destination
inline bool pair_code(int a, int b, int* type, int* off, int* tr, int* lohi) {
  if (a < 9 && b < 9) {
    int lo = std::min(a, b), hi = std::max(a, b);
    *type = 0; *tr = a > b; *off = MOM_UU + 5 * sym_index(lo, hi, 9); *lohi = lo * 13 + hi;
  } else if (a >= 9 && b >= 9) {
    int lo = std::min(a, b) - 9, hi = std::max(a, b) - 9;
    *type = 1; *tr = a > b; *off = MOM_TT + 5 * sym_index(lo, hi, 4); *lohi = (lo + 9) * 13 + hi + 9;
  } else if (a < 9) {
    *type = 2; *tr = 0; *off = MOM_UT + 4 * (a * 4 + (b - 9)); *lohi = a * 13 + b;
  } else {
    return false;                     // (theta row, u column) with row <= col cannot happen: Q1 blocks are numbered last
  }
  return true;
}

struct Match { int32_t p, a, b; };

// contributions to block (I, J): cells holding both nodes.  adjacency lists are sorted by cell.
inline int matches(const int32_t* nc_ptr, const int32_t* nc_adj, const int32_t* inv, int32_t I, int32_t J, Match* out,
                   int cap) {
  int n = 0;
  int32_t i = nc_ptr[I], ie = nc_ptr[I + 1], j = nc_ptr[J], je = nc_ptr[J + 1];
  while (i < ie && j < je) {
    int32_t ci = nc_adj[i] / 13, cj = nc_adj[j] / 13;
    if (ci < cj) ++i;
    else if (cj < ci) ++j;
    else {
      if (n < cap) { out[n].p = inv[ci]; out[n].a = nc_adj[i] - ci * 13; out[n].b = nc_adj[j] - cj * 13; }
      ++n; ++i; ++j;
    }
  }
  return n;
}

// order: nc cell ids in processing order (null = identity).  Returns "" or an error text.
inline std::string build(int64_t nc, int64_t nb, const int32_t* rowptr, const int32_t* colidx, const int32_t* nc_ptr,
                         const int32_t* nc_adj, const int32_t* order, Plan& P, int sort_mode = 2) {
  const int CAP = 64;
  const int64_t nchunks = (nc + C - 1) / C;
  P = Plan();
  P.nchunks = nchunks;
  P.cell_order.assign((size_t)nchunks * C, -1);
  std::vector<int32_t> inv((size_t)nc, -1);
  for (int64_t p = 0; p < nc; ++p) {
    int32_t cell = order ? order[p] : (int32_t)p;
    if (cell < 0 || cell >= nc || inv[cell] >= 0) return "tile plan: processing order is not a permutation";
    P.cell_order[p] = cell; inv[cell] = (int32_t)p;
  }
  const int64_t nnzb = rowptr[nb];
  // pass 1: owner chunk of every destination in the upper block triangle
  std::vector<int32_t> owner((size_t)nnzb, -1);
  std::vector<int32_t> row_of((size_t)nnzb);
  int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad) num_threads(femo_host_threads())
  for (int64_t I = 0; I < nb; ++I) {
    Match m[CAP];
    for (int32_t k = rowptr[I]; k < rowptr[I + 1]; ++k) {
      row_of[k] = (int32_t)I;
      const int32_t J = colidx[k];
      if (J < I) continue;
      int n = matches(nc_ptr, nc_adj, inv.data(), (int32_t)I, J, m, CAP);
      if (n < 1 || n > CAP) { bad |= 1; continue; }
      int32_t pmin = m[0].p;
      for (int r = 1; r < n; ++r) pmin = std::min(pmin, m[r].p);
      owner[k] = pmin / C;
    }
  }
  if (bad) return "tile plan: a block of the pattern has no (or more than 64) contributing cells";
  // stable counting sort of the destinations by owner chunk (ascending BSR position inside a chunk)
  P.dest_ptr.assign((size_t)nchunks + 1, 0);
  for (int64_t k = 0; k < nnzb; ++k)
    if (owner[k] >= 0) P.dest_ptr[owner[k] + 1]++;
  for (int64_t c = 0; c < nchunks; ++c) P.dest_ptr[c + 1] += P.dest_ptr[c];
  const int64_t ndest = P.dest_ptr[nchunks];
  P.dest_pos.resize((size_t)ndest);
  {
    std::vector<int32_t> fill(P.dest_ptr.begin(), P.dest_ptr.end() - 1);
    for (int64_t k = 0; k < nnzb; ++k)
      if (owner[k] >= 0) P.dest_pos[fill[owner[k]]++] = (int32_t)k;
  }
  std::vector<int32_t>().swap(owner);
  P.dest_tpos.resize((size_t)ndest);
  P.dest_info.resize((size_t)ndest);
  // pass 2: per chunk, count (fill = 0) then fill (fill = 1)
  P.ref_ptr.assign((size_t)nchunks + 1, 0);
  P.hitem_ptr.assign((size_t)nchunks + 1, 0);
  P.hcell_ptr.assign((size_t)nchunks + 1, 0);
  for (int fill = 0; fill < 2; ++fill) {
    int err = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(| : err) num_threads(femo_host_threads())
    for (int64_t c = 0; c < nchunks; ++c) {
      Match m[CAP];
      int32_t hc[512]; int nhc = 0;
      std::vector<uint32_t> dkey;
      int64_t nref = 0, nhit = 0;
      const int64_t rbase = fill ? P.ref_ptr[c] : 0, hbase = fill ? P.hitem_ptr[c] : 0;
      for (int32_t d = P.dest_ptr[c]; d < P.dest_ptr[c + 1]; ++d) {
        const int32_t k = P.dest_pos[d];
        const int32_t I = row_of[k], J = colidx[k];
        int n = matches(nc_ptr, nc_adj, inv.data(), I, J, m, CAP);
        std::sort(m, m + n, [](const Match& x, const Match& y) { return x.p < y.p; });
        if (!fill) {
          int type = 0, off, tr, lohi;
          pair_code(m[0].a, m[0].b, &type, &off, &tr, &lohi);
          dkey.push_back(sort_mode == 2 ? (((uint32_t)type << 12) | (uint32_t)std::min(n, 4095)) : (sort_mode == 1 ? ((uint32_t)type << 12) : 0u));
        }
        if (fill) {
          int32_t tp = -1;
          if (I != J) {
            int32_t lo = rowptr[J], hi = rowptr[J + 1] - 1;
            while (lo <= hi) {
              int32_t mid = (lo + hi) >> 1, v = colidx[mid];
              if (v == I) { tp = mid; break; }
              if (v < I) lo = mid + 1; else hi = mid - 1;
            }
            if (tp < 0) err |= 2;
          }
          P.dest_tpos[d] = tp;
          if (nref >= (1 << 20) || n >= (1 << 12)) err |= 4;
          P.dest_info[d] = (uint32_t)nref | ((uint32_t)n << 20);
        }
        for (int r = 0; r < n; ++r) {
          int type, off, tr, lohi;
          if (!pair_code(m[r].a, m[r].b, &type, &off, &tr, &lohi)) { err |= 8; continue; }
          int fslot;
          if (m[r].p / C == c) {
            fslot = m[r].p % C;
            off += fslot * MOM_CELL;
          } else {
            int s = 0;
            while (s < nhc && hc[s] != m[r].p) ++s;
            if (s == nhc) {
              if (nhc >= 512) { err |= 16; continue; }
              hc[nhc++] = m[r].p;
            }
            fslot = C + s;
            off = MOM_HALO_BASE + MOM_HALO * (int)nhit;
            if (fill) P.hitems[hbase + nhit] = ((uint32_t)m[r].p << 8) | (uint32_t)lohi;
            ++nhit;
          }
          if (off >= (1 << 16) || fslot >= (1 << 13)) { err |= 32; continue; }
          if (fill) P.refs[rbase + nref] = (uint32_t)off | ((uint32_t)type << 16) | ((uint32_t)tr << 18) | ((uint32_t)fslot << 19);
          ++nref;
        }
      }
      if (!fill) {
        P.ref_ptr[c + 1] = (int32_t)nref; P.hitem_ptr[c + 1] = (int32_t)nhit; P.hcell_ptr[c + 1] = nhc;
        // warp-uniform phase 2: order the chunk's destinations by (pair type, number of contributions), then position
        const int32_t db = P.dest_ptr[c], de = P.dest_ptr[c + 1];
        std::vector<std::pair<uint32_t, int32_t>> kd((size_t)(de - db));
        for (int32_t d = db; d < de; ++d) kd[d - db] = {dkey[d - db], P.dest_pos[d]};
        std::sort(kd.begin(), kd.end());
        for (int32_t d = db; d < de; ++d) P.dest_pos[d] = kd[d - db].second;
      }
      else for (int s = 0; s < nhc; ++s) P.hcells[P.hcell_ptr[c] + s] = hc[s];
    }
    if (err) return "tile plan: inconsistent pattern (code " + std::to_string(err) + ")";
    if (!fill) {
      for (int64_t c = 0; c < nchunks; ++c) {
        P.max_refs = std::max(P.max_refs, P.ref_ptr[c + 1]);
        P.max_hitems = std::max(P.max_hitems, P.hitem_ptr[c + 1]);
        P.max_hcells = std::max(P.max_hcells, P.hcell_ptr[c + 1]);
        P.max_dests = std::max(P.max_dests, P.dest_ptr[c + 1] - P.dest_ptr[c]);
        int64_t a = (int64_t)P.ref_ptr[c] + P.ref_ptr[c + 1];
        if (a >= ((int64_t)1 << 31)) return "tile plan: more than 2^31 contributions";
        P.ref_ptr[c + 1] = (int32_t)a;
        P.hitem_ptr[c + 1] += P.hitem_ptr[c];
        P.hcell_ptr[c + 1] += P.hcell_ptr[c];
      }
      P.refs.resize((size_t)P.ref_ptr[nchunks]);
      P.hitems.resize((size_t)P.hitem_ptr[nchunks]);
      P.hcells.resize((size_t)P.hcell_ptr[nchunks]);
    }
  }
  return "";
}

// Morton (Z-curve) processing order from cell centroids: consecutive cells form compact patches, so a
// chunk shares few blocks with other chunks (fewer halo items).  Deterministic (ties broken by cell id).
inline uint64_t spread3(uint64_t v) {
  v &= 0x1fffff;
  v = (v | (v << 32)) & 0x1f00000000ffffULL;
  v = (v | (v << 16)) & 0x1f0000ff0000ffULL;
  v = (v | (v << 8)) & 0x100f00f00f00f00fULL;
  v = (v | (v << 4)) & 0x10c30c30c30c30c3ULL;
  v = (v | (v << 2)) & 0x1249249249249249ULL;
  return v;
}

inline void morton_order(int64_t nc, const double* xyz, const int32_t* cv /* nc x 4 vertex ids */, double cell_size,
                         std::vector<int32_t>& order) {
  double lo[3] = {1e300, 1e300, 1e300};
  std::vector<double> cen((size_t)nc * 3);
  for (int64_t c = 0; c < nc; ++c)
    for (int d = 0; d < 3; ++d) {
      double s = 0;
      for (int v = 0; v < 4; ++v) s += xyz[3 * (int64_t)cv[4 * c + v] + d];
      cen[3 * c + d] = 0.25 * s;
      lo[d] = std::min(lo[d], cen[3 * c + d]);
    }
  const double inv = cell_size > 0 ? 1.0 / cell_size : 1.0;
  std::vector<std::pair<uint64_t, int32_t>> key((size_t)nc);
#pragma omp parallel for schedule(static) num_threads(femo_host_threads())
  for (int64_t c = 0; c < nc; ++c) {
    uint64_t q[3];
    for (int d = 0; d < 3; ++d) {
      double t = (cen[3 * c + d] - lo[d]) * inv + 0.25;
      q[d] = t < 0 ? 0 : (t > 2097151.0 ? 2097151 : (uint64_t)t);
    }
    key[c] = {spread3(q[0]) | (spread3(q[1]) << 1) | (spread3(q[2]) << 2), (int32_t)c};
  }
  std::sort(key.begin(), key.end());
  order.resize((size_t)nc);
  for (int64_t c = 0; c < nc; ++c) order[c] = key[c].second;
}

}  // namespace femo_tiles
