// Host side of femo_hex_setup: node -> (cell, local index) adjacency and the BSR pattern of the node graph of a
// hexahedral mesh.  Plain C++ (shared by hex.cu and the CPU harness tests/hex_host.cu).
#pragma once
#include <algorithm>
#include <cstdint>
#include <vector>

namespace femo_hex {

struct Pattern {
  std::vector<int32_t> nc_ptr, nc_adj;          // nb + 1; nc * 8 entries = cell * 8 + a, ascending per node
  std::vector<int32_t> rowptr, colidx, diag;    // nb + 1; nnzb (sorted per row); nb
};

// returns 0, or 1 + the index of a node that belongs to no cell, or -1 when the pattern exceeds 2e9 blocks
inline int64_t build_pattern(int64_t nb, int64_t nc, const int32_t* cells, Pattern& P) {
  P.nc_ptr.assign(nb + 1, 0);
  P.nc_adj.resize((size_t)nc * 8);
  for (int64_t e = 0; e < nc * 8; ++e) P.nc_ptr[cells[e] + 1]++;
  for (int64_t i = 0; i < nb; ++i) P.nc_ptr[i + 1] += P.nc_ptr[i];
  {
    std::vector<int32_t> cur(P.nc_ptr.begin(), P.nc_ptr.end() - 1);
    for (int64_t e = 0; e < nc * 8; ++e) P.nc_adj[cur[cells[e]]++] = (int32_t)e;
  }
  P.rowptr.assign(nb + 1, 0);
  P.colidx.clear();
  P.colidx.reserve((size_t)nb * 27);
  P.diag.resize(nb);
  std::vector<int32_t> row;
  for (int64_t i = 0; i < nb; ++i) {
    row.clear();
    for (int32_t k = P.nc_ptr[i]; k < P.nc_ptr[i + 1]; ++k) {
      const int32_t* cn = cells + 8 * (int64_t)(P.nc_adj[k] >> 3);
      row.insert(row.end(), cn, cn + 8);
    }
    if (row.empty()) return 1 + i;
    std::sort(row.begin(), row.end());
    row.erase(std::unique(row.begin(), row.end()), row.end());
    if ((int64_t)P.colidx.size() + (int64_t)row.size() > 2000000000LL) return -1;
    P.diag[i] = (int32_t)(P.colidx.size() + (std::lower_bound(row.begin(), row.end(), (int32_t)i) - row.begin()));
    P.colidx.insert(P.colidx.end(), row.begin(), row.end());
    P.rowptr[i + 1] = (int32_t)P.colidx.size();
  }
  return 0;
}

}  // namespace femo_hex
