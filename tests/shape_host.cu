// CPU harness around femo_alpha_b200/csrc/shape_partials.cuh for tests/test_shape_host.py: the same
// __host__ __device__ functions the CUDA kernels call (k_cell_op<OP_SHAPE_*> in outputs.cu), run on the host by an
// nvcc-compiled shared object (no GPU needed), one cell at a time with the quadrature sum of the kernel.
#include "../femo_alpha_b200/csrc/shape_partials.cuh"
#include "../femo_alpha_b200/csrc/tables.hpp"

// stubs: femo_internal.cuh declares these for the library
int femo_fail(const std::string&) { return 1; }

namespace {
struct CellIn {
  V3 xv[4], uh[4];
  double h[4], E[4], nu[4], rho[4];
  V3 f[4];
};
void load_cell(int64_t c, const double* xv, const double* uh, const double* hv, const double* Ev, const double* nuv,
               const double* rhov, const double* fv, CellIn& I) {
  for (int v = 0; v < 4; ++v) {
    I.xv[v] = ldv3(xv + (c * 4 + v) * 3);
    I.uh[v] = ldv3(uh + (c * 4 + v) * 3);
    I.h[v] = hv[c * 4 + v]; I.E[v] = Ev[c * 4 + v]; I.nu[v] = nuv[c * 4 + v]; I.rho[v] = rhov[c * 4 + v];
    I.f[v] = ldv3(fv + (c * 4 + v) * 3);
  }
}
}  // namespace

extern "C" {

// per-cell inputs: xv, uh, fv (nc,4,3); hv, Ev, nuv, rhov (nc,4) (vertex values; repeat the cell value for DG0);
// we, le (nc,39); hE (nc).  op 0: shape_T (out nc x 12, with_load = flag), 1: shape_fwd (duh (nc,4,3), out nc x 39),
// 2: measure (flag 0: u.u, 1: rho h; out nc x 12), 3: pnorm integrand (flag: element-wise material; pm, prho; degree-4 rule
// is the caller's choice of m)
int shape_host(int op, int m, int64_t nc, const double* xv, const double* uh, const double* we, const double* le,
               const double* hv, const double* Ev, const double* nuv, const double* rhov, const double* hE,
               const double* fv, const double* duh, int flag, double pm, double prho, double* out) {
  Tables T;
  build_tables(T, m);
  const int nout = op == 1 ? 39 : 12;
  for (int64_t c = 0; c < nc; ++c) {
    CellIn I;
    load_cell(c, xv, uh, hv, Ev, nuv, rhov, fv, I);
    double acc[39];
    for (int i = 0; i < nout; ++i) acc[i] = 0.0;
    for (int q = 0; q < T.nq; ++q) {
      QpGeom g;
      qp_geometry(I.xv, I.uh, T.X0[q], T.X1[q], g);
      const double wd = T.W[q] * g.detJ;
      double hq = 0, Eq = 0, nuq = 0, rhoq = 0;
      V3 fq = v3(0, 0, 0);
      for (int v = 0; v < 4; ++v) {
        const double N = T.n1[q][v];
        hq += N * I.h[v]; Eq += N * I.E[v]; nuq += N * I.nu[v]; rhoq += N * I.rho[v];
        fq = fq + N * I.f[v];
      }
      const Coefs cf = clt_coefs(hq, Eq, nuq, hE[c], g.Ju, 0);
      double o[39];
      if (op == 0) shape_T_qp(g, &T, q, we + c * 39, le + c * 39, cf, fq, wd, flag, o);
      else if (op == 1) {
        V3 d[4];
        for (int v = 0; v < 4; ++v) d[v] = ldv3(duh + (c * 4 + v) * 3);
        shape_fwd_qp(g, &T, q, we + c * 39, cf, fq, wd, d, o);
      } else if (op == 3) {
        VmPoint P;
        vm_at(g, &T, q, we + c * 39, hq, Eq, nuq, I.h, flag, 1.0, P);
        shape_pnorm_qp(g, &T, q, we + c * 39, I.h, flag, 1.0, P, wd, pm, prho, o);
      } else {
        double cq;
        if (flag == 0) {
          V3 u = v3(0, 0, 0);
          for (int a = 0; a < 9; ++a) u = u + T.phi2[q][a] * ldv3(we + c * 39 + 3 * a);
          cq = dot(u, u);
        } else cq = rhoq * hq;
        shape_measure_qp(g, &T, q, wd * cq, o);
      }
      for (int i = 0; i < nout; ++i) acc[i] += o[i];
    }
    for (int i = 0; i < nout; ++i) out[c * nout + i] = acc[i];
  }
  return 0;
}

}  // extern "C"

#include "../femo_alpha_b200/csrc/penalty_shape.cuh"

extern "C" {

// penalty-BC facet terms, one (edge, cell) pair per row: xv, uh (np,4,3); le (np) local edge; hE (np); we, lam (np,5,3)
// trace-node values; duh (np,4,3).  op 0: (dR_pen/duhat)^T lam -> out (np,12); 1: (dR_pen/duhat) duhat -> out (np,15);
// 2: the 5x5 trace matrix of the forward penalty (what k_penalty_me computes) -> out (np,25)
int penalty_shape_host(int op, int64_t np_, const double* xv, const double* uh, const int32_t* le, const double* hE,
                       const double* we, const double* lam, const double* duh, double* out) {
  for (int64_t p = 0; p < np_; ++p) {
    V3 x[4], u[4], d[4];
    for (int v = 0; v < 4; ++v) {
      x[v] = ldv3(xv + (p * 4 + v) * 3); u[v] = ldv3(uh + (p * 4 + v) * 3); d[v] = ldv3(duh + (p * 4 + v) * 3);
    }
    if (op == 0) penalty_shape_T_pair(x, u, le[p], hE[p], we + p * 15, lam + p * 15, out + p * 12);
    else if (op == 1) penalty_shape_fwd_pair(x, u, le[p], hE[p], we + p * 15, d, out + p * 15);
    else {
      for (int i = 0; i < 25; ++i) out[p * 25 + i] = 0.0;
      for (int k = 0; k < 3; ++k) {
        PenQp q;
        penalty_qp(x, u, le[p], k, hE[p], q);
        for (int i = 0; i < 5; ++i)
          for (int j = 0; j < 5; ++j)
            if ((i < 3) == (j < 3)) out[p * 25 + i * 5 + j] += q.coef * q.sc * q.tr[i] * q.tr[j];
      }
    }
  }
  return 0;
}

}  // extern "C"
