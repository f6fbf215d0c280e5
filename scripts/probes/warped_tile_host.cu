// Host-side check of warped_tile<NB> (assemble.cu, the cell-centric kernel for warped cells) against block_general (the
// round-1 formulation the GPU parity tests pinned to the oracle): random records of one group of 32 cells in both layouts,
// all 91 pairs a <= b, NB = 1, 2, 3, plus the task list covering every pair exactly once.
//   nvcc -std=c++17 -O1 -I femo_alpha_b200/csrc scripts/probes/warped_tile_host.cu -L femo_alpha_b200 -lfemo_b200 -o /tmp/wt && LD_LIBRARY_PATH=femo_alpha_b200 /tmp/wt
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "assemble.cu"

template <int NB>
static double check(const std::vector<double>& aos, const std::vector<double>& grp, const SmemTab& T, int nq, int stride) {
  WpTasks tasks = wp_make_tasks(NB);
  std::vector<int> seen(91, 0);
  double worst = 0;
  for (int lane = 0; lane < 32; ++lane)
    for (int t = 0; t < tasks.n; ++t) {
      double K[NB][9];
      warped_tile<NB>(grp.data() + lane, T, nq, tasks.a[t], tasks.b0[t], tasks.nb[t], K);
      for (int j = 0; j < tasks.nb[t]; ++j) {
        double R[9];
        block_general(aos.data() + (size_t)lane * stride, T, nq, tasks.a[t], tasks.b0[t] + j, R);
        double sc = 0;
        for (int k = 0; k < 9; ++k) sc = fmax(sc, fabs(R[k]));
        for (int k = 0; k < 9; ++k) worst = fmax(worst, fabs(K[j][k] - R[k]) / (sc + 1e-300));
        if (lane == 0) seen[wp_pair(tasks.a[t], tasks.b0[t] + j)]++;
      }
    }
  for (int p = 0; p < 91; ++p)
    if (seen[p] != 1) { printf("NB=%d: pair %d covered %d times\n", NB, p, seen[p]); return 1.0; }
  printf("NB=%d: %d tasks, worst relative difference %.2e\n", NB, tasks.n, worst);
  return worst;
}

int main() {
  const int nq = 16, stride = REC_HEADER + nq * REC_NQF_WARPED;
  std::vector<double> aos(32 * (size_t)stride), grp((size_t)nq * REC_NQF_WARPED * 32);
  srand(7);
  for (int cell = 0; cell < 32; ++cell)
    for (int q = 0; q < nq; ++q)
      for (int f = 0; f < REC_NQF_WARPED; ++f) {
        double v = rand() / (double)RAND_MAX - 0.3;
        aos[(size_t)cell * stride + REC_HEADER + q * REC_NQF_WARPED + f] = v;
        grp[((size_t)q * REC_NQF_WARPED + f) * 32 + cell] = v;
      }
  SmemTab T;
  for (size_t i = 0; i < sizeof(T) / sizeof(double); ++i) ((double*)&T)[i] = rand() / (double)RAND_MAX - 0.5;
  double w = fmax(check<1>(aos, grp, T, nq, stride), fmax(check<2>(aos, grp, T, nq, stride), check<3>(aos, grp, T, nq, stride)));
  // pair index: bijection onto [0, 91)
  std::vector<int> hit(91, 0);
  for (int a = 0; a < 13; ++a) for (int b = a; b < 13; ++b) hit[wp_pair(a, b)]++;
  for (int p = 0; p < 91; ++p) if (hit[p] != 1) { printf("wp_pair is not a bijection\n"); return 1; }
  printf(w <= 1e-13 ? "OK\n" : "FAILED\n");
  return w <= 1e-13 ? 0 : 1;
}
