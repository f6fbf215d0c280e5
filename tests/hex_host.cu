// CPU harness around femo_alpha_b200/csrc/hex_element.cuh for tests/test_hex_host.py: the per-thread bodies of the
// kernels of hex.cu (k_hex_geom, k_hex_assemble, k_hex_cellvec, k_hex_gather, k_hex_energy) run in host loops with the
// same index arithmetic -- compiled by nvcc into a shared object, no GPU needed.
#include <cstdint>
#include <algorithm>
#include <cmath>
#include <vector>

#include "../femo_alpha_b200/csrc/hex_element.cuh"

using namespace femo_hex;

extern "C" {

// geo (nc*8*10), vol (nc); returns the number of non-positive Jacobians
int hex_host_geom(int64_t nc, const double* xyz, const int32_t* cells, double* geo, double* vol) {
  int bad = 0;
  for (int64_t cell = 0; cell < nc; ++cell) {
    double v = 0;
    for (int q = 0; q < 8; ++q) {
      const double det = geom_qp(xyz, cells + 8 * cell, q, geo + (cell * 8 + q) * GEO);
      if (!(det > 0.0)) ++bad;
      v += det;
    }
    vol[cell] = v;
  }
  return bad;
}

// the warp-per-row algorithm of k_hex_assemble: passes of HEX_PASS staged items, then the destination lanes collect
int hex_host_assemble(int form, double c1, double c2, double p, int64_t nb, const double* geo, const int32_t* cells,
                      const double* rho, int64_t nc, const int32_t* nc_ptr, const int32_t* nc_adj, const int32_t* rowptr,
                      const int32_t* colidx, double* vals) {
  std::vector<double> scale(nc);
  for (int64_t e = 0; e < nc; ++e) scale[e] = pow(rho[e], p);
  double s_val[HEX_PASS * 9];
  int32_t s_col[HEX_PASS] = {0};
  for (int64_t i = 0; i < nb; ++i) {
    const int n_items = 8 * (nc_ptr[i + 1] - nc_ptr[i]);
    for (int t0 = 0; t0 < n_items; t0 += HEX_PASS) {
      const int n_staged = std::min(HEX_PASS, n_items - t0);
      for (int lane = 0; lane < 32; ++lane) {
        const int t = 2 * lane;
        double B0[9], B1[9];
        int32_t col[2];
        if (t < n_staged && row_item_pair(form, c1, c2, geo, cells, scale.data(), nc_ptr, nc_adj, i, t0 + t, B0, B1, col)) {
          s_col[t] = col[0]; s_col[t + 1] = col[1];
          for (int k = 0; k < 9; ++k) { s_val[k * HEX_PASS + t] = B0[k]; s_val[k * HEX_PASS + t + 1] = B1[k]; }
        }
      }
      for (int lane = 0; lane < 32; ++lane)
        for (int32_t k = rowptr[i] + lane; k < rowptr[i + 1]; k += 32) {
          double acc[9];
          row_collect(s_val, s_col, n_staged, colidx[k], acc);
          for (int e = 0; e < 9; ++e) {
            if (t0 == 0) vals[(int64_t)k * 9 + e] = acc[e];
            else vals[(int64_t)k * 9 + e] += acc[e];
          }
        }
    }
  }
  return 0;
}

// k_hex_cellvec + k_hex_gather
int hex_host_cellvec_gather(int mode, int form, double c1, double c2, int64_t nb, int64_t nc, const double* geo,
                            const int32_t* cells, const double* x, const double* cellcoef, double coef, const int32_t* nc_ptr,
                            const int32_t* nc_adj, double* out) {
  std::vector<double> cellvec((size_t)nc * 24);
  for (int64_t cell = 0; cell < nc; ++cell) {
    const double cf = coef * (cellcoef ? cellcoef[cell] : 1.0);
    if (mode == 0) cell_mass_apply(geo, cells, cell, x, cf, cellvec.data() + cell * 24);
    else cell_apply(form, c1, c2, geo, cells, cell, x, cf, cellvec.data() + cell * 24);
  }
  for (int64_t i = 0; i < nb; ++i) node_gather(nc_ptr, nc_adj, cellvec.data(), i, out + 3 * i);
  return 0;
}

// k_hex_energy
int hex_host_energy(int form, double c1, double c2, double p, int64_t nc, const double* geo, const int32_t* cells, const double* u,
                    const double* lm, const double* rho, double* out) {
  for (int64_t cell = 0; cell < nc; ++cell)
    out[cell] = p * pow(rho[cell], p - 1.0) * cell_energy(form, c1, c2, geo, cells, cell, u, lm);
  return 0;
}

}  // extern "C"

#include "../femo_alpha_b200/csrc/hex_pattern.hpp"

extern "C" {

// pattern of femo_hex_setup; call once with null outputs for the sizes (returns nnzb), then with buffers
int64_t hex_host_pattern(int64_t nb, int64_t nc, const int32_t* cells, int32_t* nc_ptr, int32_t* nc_adj, int32_t* rowptr,
                         int32_t* colidx, int32_t* diag) {
  Pattern P;
  const int64_t r = build_pattern(nb, nc, cells, P);
  if (r != 0) return -r;
  if (nc_ptr) {
    std::copy(P.nc_ptr.begin(), P.nc_ptr.end(), nc_ptr);
    std::copy(P.nc_adj.begin(), P.nc_adj.end(), nc_adj);
    std::copy(P.rowptr.begin(), P.rowptr.end(), rowptr);
    std::copy(P.colidx.begin(), P.colidx.end(), colidx);
    std::copy(P.diag.begin(), P.diag.end(), diag);
  }
  return (int64_t)P.colidx.size();
}

}  // extern "C"

extern "C" {

// k_hex_geom2 + k_hex_assemble_g (FEMO_HEX_GRAD=1): sub-passes of 4 cells, one item per lane, two sub-passes staged per collect
int hex_host_assemble_g(int form, double c1, double c2, double p, int64_t nb, const double* geo, const int32_t* cells,
                        const double* rho, int64_t nc, const int32_t* nc_ptr, const int32_t* nc_adj, const int32_t* rowptr,
                        const int32_t* colidx, double* vals) {
  std::vector<double> scale(nc), geo2((size_t)nc * 8 * GEO2);
  for (int64_t e = 0; e < nc; ++e) scale[e] = pow(rho[e], p);
  for (int64_t r = 0; r < nc * 8; ++r) geom2_qp(geo + r * GEO, (int)(r & 7), geo2.data() + r * GEO2);
  double stg[HEX_PASS * 9];
  int32_t s_col[HEX_PASS] = {0};
  for (int64_t i = 0; i < nb; ++i) {
    const int32_t e0 = nc_ptr[i], n_cells_i = nc_ptr[i + 1] - e0;
    for (int c0 = 0; c0 < n_cells_i; c0 += 8) {
      const int n_chunk = std::min(8, n_cells_i - c0);
      for (int h = 0; h < 2 && 4 * h < n_chunk; ++h) {
        const int n_pass = std::min(4, n_chunk - 4 * h);
        for (int lane = 0; lane < 32; ++lane) {
          const int cl = lane >> 3, b = lane & 7;
          if (cl >= n_pass) continue;
          const int32_t ca = nc_adj[e0 + c0 + 4 * h + cl];
          const int64_t cell = ca >> 3;
          double B[9];
          item_body_g(form, c1, c2, geo2.data() + cell * 8 * GEO2, scale[cell], ca & 7, b, B);
          const int t = 32 * h + lane;
          s_col[t] = cells[8 * cell + b];
          for (int k = 0; k < 9; ++k) stg[k * HEX_PASS + t] = B[k];
        }
      }
      const int n_staged = 8 * n_chunk;
      for (int32_t k = rowptr[i]; k < rowptr[i + 1]; ++k) {
        double acc[9];
        row_collect(stg, s_col, n_staged, colidx[k], acc);
        for (int e = 0; e < 9; ++e) {
          if (c0 == 0) vals[(int64_t)k * 9 + e] = acc[e];
          else vals[(int64_t)k * 9 + e] += acc[e];
        }
      }
    }
  }
  return 0;
}

}  // extern "C"
