// CPU harness around femo_alpha_b200/csrc/tile_plan.hpp for tests/test_tile_plan.py (no CUDA needed).
#include <cstring>
#include "../femo_alpha_b200/csrc/tile_plan.hpp"

extern "C" {
void* tp_build(int64_t nc, int64_t nb, const int32_t* rowptr, const int32_t* colidx, const int32_t* nc_ptr,
               const int32_t* nc_adj, const int32_t* order, char* err, int errlen) {
  auto* P = new femo_tiles::Plan();
  std::string e = femo_tiles::build(nc, nb, rowptr, colidx, nc_ptr, nc_adj, order, *P);
  if (!e.empty()) { strncpy(err, e.c_str(), errlen - 1); err[errlen - 1] = 0; delete P; return nullptr; }
  err[0] = 0;
  return P;
}
void tp_sizes(void* h, int64_t* o) {
  auto* P = (femo_tiles::Plan*)h;
  o[0] = P->nchunks; o[1] = (int64_t)P->dest_pos.size(); o[2] = (int64_t)P->refs.size(); o[3] = (int64_t)P->hitems.size();
  o[4] = (int64_t)P->hcells.size(); o[5] = P->max_dests; o[6] = P->max_hitems; o[7] = P->max_hcells;
}
#define CP(dst, v) memcpy(dst, P->v.data(), sizeof(P->v[0]) * P->v.size())
void tp_copy(void* h, int32_t* cell_order, int32_t* dest_ptr, int32_t* dest_pos, int32_t* dest_tpos, uint32_t* dest_info,
             int32_t* ref_ptr, uint32_t* refs, int32_t* hitem_ptr, uint32_t* hitems, int32_t* hcell_ptr, int32_t* hcells) {
  auto* P = (femo_tiles::Plan*)h;
  CP(cell_order, cell_order); CP(dest_ptr, dest_ptr); CP(dest_pos, dest_pos); CP(dest_tpos, dest_tpos);
  CP(dest_info, dest_info); CP(ref_ptr, ref_ptr); CP(refs, refs); CP(hitem_ptr, hitem_ptr); CP(hitems, hitems);
  CP(hcell_ptr, hcell_ptr); CP(hcells, hcells);
}
void tp_free(void* h) { delete (femo_tiles::Plan*)h; }
void tp_morton(int64_t nc, const double* xyz, const int32_t* cv, double cell_size, int32_t* order) {
  std::vector<int32_t> o;
  femo_tiles::morton_order(nc, xyz, cv, cell_size, o);
  memcpy(order, o.data(), sizeof(int32_t) * nc);
}
}
