// Host-side check of warped_uu / warped_ut / warped_tt (assemble.cu, the cell-centric kernel for warped cells) against
// block_general (the round-1 formulation the GPU parity tests pinned to the oracle): random records of one group of 32
// cells in the three layouts (per-cell records, grouped frame records, grouped "U" records with the M matrices), all 91
// pairs a <= b, NB = 1 .. 4, plus the task list covering every pair exactly once.
// Compiled and run by tests/test_warped_host.py (CPU suite; the same __host__ __device__ functions the kernel calls):
//   nvcc -std=c++17 -O1 -gencode arch=compute_100a,code=sm_100a -I femo_alpha_b200/csrc tests/warped_host.cu \
//        -L femo_alpha_b200 -lfemo_b200 -o tests/_build/warped_host
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "assemble.cu"

template <int NB>
static double check(const std::vector<double>& aos, const std::vector<double>& grp, const std::vector<double>& grpu, const SmemTab& T,
                    int nq, int stride) {
  WpTasks tasks = wp_make_tasks(NB, NB);
  std::vector<int> seen(91, 0);
  double worst = 0;
  for (int lane = 0; lane < 32; ++lane)
    for (int t = 0; t < tasks.n; ++t) {
      double K[NB < 4 ? 4 : NB][9];
      const int a = tasks.a[t], b0 = tasks.b0[t], nb = tasks.nb[t];
      if (t < tasks.n_uu) warped_uu<NB>(grpu.data() + lane, T, nq, a, b0, nb, (double (*)[9])K);
      else if (a < 9) warped_ut(grp.data() + lane, T, nq, a, K);
      else warped_tt(grp.data() + lane, T, nq, a, b0, nb, K);
      for (int j = 0; j < nb; ++j) {
        double R[9];
        block_general(aos.data() + (size_t)lane * stride, T, nq, a, b0 + j, R);
        double sc = 0;
        for (int k = 0; k < 9; ++k) sc = fmax(sc, fabs(R[k]));
        for (int k = 0; k < 9; ++k) worst = fmax(worst, fabs(K[j][k] - R[k]) / (sc + 1e-300));
        if (lane == 0) seen[wp_pair(a, b0 + j)]++;
      }
    }
  for (int p = 0; p < 91; ++p)
    if (seen[p] != 1) { printf("NB=%d: pair %d covered %d times\n", NB, p, seen[p]); return 1.0; }
  printf("NB=%d: %d tasks (%d displacement-displacement), worst relative difference %.2e\n", NB, tasks.n, tasks.n_uu, worst);
  return worst;
}

int main() {
  const int nq = 16, stride = REC_HEADER + nq * REC_NQF_WARPED;
  std::vector<double> aos(32 * (size_t)stride), grp((size_t)nq * REC_NQF_WARPED * 32), grpu(grp.size());
  srand(7);
  for (int cell = 0; cell < 32; ++cell)
    for (int q = 0; q < nq; ++q) {
      double* rec = &aos[(size_t)cell * stride + REC_HEADER + q * REC_NQF_WARPED];
      for (int f = 0; f < REC_NQF_WARPED; ++f) {
        rec[f] = rand() / (double)RAND_MAX - 0.3;
        grp[((size_t)q * REC_NQF_WARPED + f) * 32 + cell] = rec[f];
      }
      double m[21];
      wp_m_tensors(ldv3(rec + REC_E0), ldv3(rec + REC_E0 + 3), ldv3(rec + REC_E0 + 6), rec[REC_CM11], rec[REC_CM12], rec[REC_CM33],
                   rec[REC_CS], rec[REC_CD], m);
      for (int f = 0; f < 4; ++f) grpu[((size_t)q * REC_NQF_WARPED + f) * 32 + cell] = rec[f];
      for (int f = 0; f < 21; ++f) grpu[((size_t)q * REC_NQF_WARPED + REC_U_M00 + f) * 32 + cell] = m[f];
    }
  SmemTab T;
  for (size_t i = 0; i < sizeof(T) / sizeof(double); ++i) ((double*)&T)[i] = rand() / (double)RAND_MAX - 0.5;
  double w = fmax(fmax(check<1>(aos, grp, grpu, T, nq, stride), check<2>(aos, grp, grpu, T, nq, stride)),
                  fmax(check<3>(aos, grp, grpu, T, nq, stride), check<4>(aos, grp, grpu, T, nq, stride)));
  std::vector<int> hit(91, 0);                     // pair index: bijection onto [0, 91)
  for (int a = 0; a < 13; ++a) for (int b = a; b < 13; ++b) hit[wp_pair(a, b)]++;
  for (int p = 0; p < 91; ++p) if (hit[p] != 1) { printf("wp_pair is not a bijection\n"); return 1; }
  printf(w <= 1e-13 ? "OK\n" : "FAILED\n");
  return w <= 1e-13 ? 0 : 1;
}
