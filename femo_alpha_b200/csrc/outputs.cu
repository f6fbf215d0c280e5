// Scalar outputs, their partials, and the matrix-free partials of the residual.
//
// Replaces OutputOperation.compute / compute_derivatives (femo_alpha/csdl_alpha_opt/
// output_operation.py:51-69) for the forms compliance / mass / elastic_energy
// (rm_shell/rm_shell_pde.py:64-110), and assembleMatrix(dR_df) + MatMult[Transpose]
// (csdl_alpha_opt/state_operation.py:163-184,282-286; fea/utils_dolfinx.py:275-306).
//
// Pattern: one warp per cell, one lane per quadrature point -> per-cell element vector in
// `cellbuf` (lane reduction by shuffles, fixed order), then a per-node gather that sums the
// adjacent cells in ascending cell order.  No atomics: results are bitwise reproducible.
#include "femo_internal.cuh"
#include "stress_point.cuh"
#include "shape_partials.cuh"

namespace {

enum CellOp {
  OP_PARAM_T = 0,    // nodal/cell param gradient: sum_q N_v g(q), g = s(lam)^T D' s(w) w detJ   -> 4 (or 1)
  OP_PARAM_FWD = 1,  // sum_q dp(q) B^T D' s(w) w detJ                                           -> 39
  OP_F_T = 2,        // -(sum_q wJ N_v(q) lam_u(q))                                              -> 12 (or 3)
  OP_F_FWD = 3,      // -(sum_q wJ phi_a(q) df(q))                                               -> 27
  OP_MASS_U = 4,     // sum_q wJ phi_a(q) u(q)                                                   -> 27
  OP_UU = 5,         // sum_q wJ u.u                                                             -> 1
  OP_MASS = 6,       // sum_q wJ rho h                                                           -> 1
  OP_REG = 7,        // regulariser value                                                        -> 1
  OP_REG_GRAD = 8,   // regulariser gradient                                                     -> 4 (or 1)
  OP_DMASS_DH = 9,   // sum_q wJ rho N_v                                                         -> 4 (or 1)
  OP_DMASS_DRHO = 10, // sum_q wJ h N_v                                                          -> 4 (or 1)
  OP_PNORM = 11,      // sum_q wJ (m vm)^rho   (degree-4 rule)                                       -> 1
  OP_PNORM_DW = 12,   // d/dw of the above                                                           -> 39
  OP_PNORM_DP = 13,   // d/d(h|E|nu) nodal (mode 1|2|3)                                              -> 4 (or 1)
  OP_VM_PROJ = 14,    // cell-wise L2 projection of von Mises onto DG1                               -> 4
  OP_AREA = 15,       // sum_q wJ                                                                    -> 1
  // dynamic path (dynamic.cu; femo_alpha/dynamic_rm_shell): mass form of linear_shell_model.py:335-348, 4x4 rule
  OP_MASS_ELEM = 16,     // coef = rho h: int coef phi_a phi_b J (81), int coef hE^2 N_a N_b J (16)                -> 97
  OP_MASS_PARAM_T = 17,  // sum_q N_v [wJ rho (a_u.l_u + hE^2 a_t.l_t) - wd rho g l_uz]                           -> 4 (or 1)
  OP_MASS_PARAM_FWD = 18, // sum_q dh(q) [wJ rho (phi_a a_u ; hE^2 N_b a_t) - wd rho g phi_a e_z]                  -> 39
  OP_WEIGHT = 19,        // self weight  sum_q wd rho h g phi_a e_z  (plate_sim.py:203-213)                        -> 27
  OP_VOLUME = 20,        // sum_q wd h  (volume_operation.py:68)                                                   -> 1
  OP_DVOLUME_DH = 21,    // sum_q wd N_v                                                                           -> 4 (or 1)
  // shape partials (shape_partials.cuh)
  OP_SHAPE_T = 22,       // d(lam^T R)/d uhat_v, mode 1: with the load term                                        -> 12
  OP_SHAPE_FWD = 23,     // (dR_e/duhat) duhat                                                                      -> 39
  OP_SHAPE_MEASURE = 24, // d/duhat_v sum_q wd J c(q), mode 0: c = u.u (compliance), 1: c = rho h (mass)            -> 12
  OP_SHAPE_PNORM = 25    // d/duhat_v sum_q wd J (m vm)^rho  (degree-4 rule, optional sub-domain)                   -> 12
};

struct CellArgs {
  const double* xyz; const int32_t* cv; const int32_t* cb; const double* uhat;
  const double *h, *E, *nu, *rho, *f, *hE;
  const Tables* tab;
  int64_t nc; int nq, ewm, ewp;
  const double* vec_w;    // state-like vector (ndof)
  const double* vec_l;    // adjoint-like vector (ndof) or parameter perturbation
  int mode;               // coefficient derivative mode 1:h 2:E 3:nu
  double* out; int out_stride;
  double sm, srho, xi_sign;  // p-norm parameters (m, rho) and through-thickness position xi2 = xi_sign * h/2
  const int32_t* tags; int tag;   // optional sub-domain restriction
  int accumulate;         // OP_PARAM_T / OP_MASS_PARAM_T: add to `out` instead of overwriting (sum over time levels)
  double g_acc;           // self-weight acceleration g_factor * 9.81 (0: none)
};

__device__ __forceinline__ double lane_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}

// sig = D s
__device__ __forceinline__ void apply_D(const Coefs& c, const double* s, double* sig) {
  sig[0] = c.cm11 * s[0] + c.cm12 * s[1]; sig[1] = c.cm12 * s[0] + c.cm11 * s[1]; sig[2] = c.cm33 * s[2];
  sig[3] = c.cb11 * s[3] + c.cb12 * s[4]; sig[4] = c.cb12 * s[3] + c.cb11 * s[4]; sig[5] = c.cb33 * s[5];
  sig[6] = c.cs * s[6]; sig[7] = c.cs * s[7]; sig[8] = c.cd * s[8];
}

template <int OP>
__global__ void __launch_bounds__(128) k_cell_op(CellArgs A) {
  __shared__ double s_w[4][40], s_l[4][40];
  const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t cell = blockIdx.x * 4LL + wid;
  if (cell >= A.nc) return;
  const Tables* tab = A.tab;
  int32_t vid[4];
  V3 xv[4], uh[4];
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    vid[v] = A.cv[cell * 4 + v];
    xv[v] = ldv3(A.xyz + 3 * (int64_t)vid[v]);
    if (A.uhat) uh[v] = ldv3(A.uhat + 3 * (int64_t)vid[v]);
  }
  constexpr bool need_w = (OP == OP_PARAM_T || OP == OP_PARAM_FWD || OP == OP_MASS_U || OP == OP_UU || OP == OP_PNORM ||
                            OP == OP_PNORM_DW || OP == OP_PNORM_DP || OP == OP_VM_PROJ || OP == OP_MASS_PARAM_T ||
                            OP == OP_MASS_PARAM_FWD || OP == OP_SHAPE_T || OP == OP_SHAPE_FWD || OP == OP_SHAPE_MEASURE ||
                            OP == OP_SHAPE_PNORM);
  constexpr bool need_l = (OP == OP_PARAM_T || OP == OP_F_T || OP == OP_MASS_PARAM_T || OP == OP_SHAPE_T);
  if (need_w || need_l) {
    for (int i = lane; i < 39; i += 32) {
      int64_t d = 3 * (int64_t)A.cb[cell * 13 + i / 3] + (i % 3);
      if (need_w) s_w[wid][i] = A.vec_w[d];
      if (need_l) s_l[wid][i] = A.vec_l[d];
    }
    __syncwarp();
  }
  const bool active = lane < A.nq;
  const int q = active ? lane : 0;
  QpGeom g;
  qp_geometry(xv, (A.uhat && OP != OP_REG && OP != OP_REG_GRAD) ? uh : nullptr, tab->X0[q], tab->X1[q], g);
  const double wd = active ? tab->W[q] * g.detJ : 0.0;
  const double wJ = wd * g.Ju;
  double hq = 0, Eq = 0, nuq = 0, rhoq = 0;
  if (A.ewm) { hq = A.h[cell]; Eq = A.E[cell]; nuq = A.nu[cell]; rhoq = A.rho[cell]; }
  else {
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      double N = tab->n1[q][v];
      hq += N * A.h[vid[v]]; Eq += N * A.E[vid[v]]; nuq += N * A.nu[vid[v]]; rhoq += N * A.rho[vid[v]];
    }
  }
  double* out = A.out + cell * A.out_stride;

  if (OP == OP_PARAM_T || OP == OP_PARAM_FWD) {
    Coefs cf = clt_coefs(hq, Eq, nuq, A.hE[cell], g.Ju, A.mode);
    double s[9], sig[9];
    strains_at(g, tab, q, s_w[wid], s);
    apply_D(cf, s, sig);
    if (OP == OP_PARAM_T) {
      double sl[9];
      strains_at(g, tab, q, s_l[wid], sl);
      double gq = 0;
#pragma unroll
      for (int i = 0; i < 9; ++i) gq += sl[i] * sig[i];
      gq *= wd;
      if (A.ewm) { double t = lane_sum(gq); if (lane == 0) out[0] = A.accumulate ? out[0] + t : t; }
      else
        for (int v = 0; v < 4; ++v) {
          double t = lane_sum(gq * tab->n1[q][v]);
          if (lane == 0) out[v] = A.accumulate ? out[v] + t : t;
        }
    } else {
      // perturbation of the parameter at this point
      double dp;
      if (A.ewm) dp = A.vec_l[cell];
      else { dp = 0; for (int v = 0; v < 4; ++v) dp += tab->n1[q][v] * A.vec_l[vid[v]]; }
      double sc = wd * dp;
#pragma unroll
      for (int i = 0; i < 9; ++i) sig[i] *= sc;
      const double k00 = dot(g.Kx0, g.E0), k01 = dot(g.Kx0, g.E1), k10 = dot(g.Kx1, g.E0), k11 = dot(g.Kx1, g.E1);
      const V3 gq0 = k00 * g.dE2_0 + k10 * g.dE2_1, gq1 = k01 * g.dE2_0 + k11 * g.dE2_1;
      for (int a = 0; a < 9; ++a) {
        double d0 = tab->dphi2[q][a][0] * k00 + tab->dphi2[q][a][1] * k10, d1 = tab->dphi2[q][a][0] * k01 + tab->dphi2[q][a][1] * k11;
        URows R = u_rows(d0, d1, g.E0, g.E1, g.E2);
        V3 r = sig[0] * R.e0 + sig[1] * R.e1 + sig[2] * R.e2 + sig[6] * R.g0 + sig[7] * R.g1 + sig[8] * R.dr;
        double rx = lane_sum(r.x), ry = lane_sum(r.y), rz = lane_sum(r.z);
        if (lane == 0) { out[3 * a] = rx; out[3 * a + 1] = ry; out[3 * a + 2] = rz; }
      }
      for (int b = 0; b < 4; ++b) {
        double d0 = tab->dn1[q][b][0] * k00 + tab->dn1[q][b][1] * k10, d1 = tab->dn1[q][b][0] * k01 + tab->dn1[q][b][1] * k11;
        TRows R = t_rows(tab->n1[q][b], d0, d1, g.E0, g.E1, g.E2, gq0, gq1);
        V3 r = sig[3] * R.k0 + sig[4] * R.k1 + sig[5] * R.k2 + sig[6] * R.g0 + sig[7] * R.g1 + sig[8] * R.dr;
        double rx = lane_sum(r.x), ry = lane_sum(r.y), rz = lane_sum(r.z);
        if (lane == 0) { out[27 + 3 * b] = rx; out[27 + 3 * b + 1] = ry; out[27 + 3 * b + 2] = rz; }
      }
    }
  } else if (OP == OP_F_T) {
    V3 lu = v3(0, 0, 0);
    for (int a = 0; a < 9; ++a) lu = lu + tab->phi2[q][a] * ldv3(s_l[wid] + 3 * a);
    if (A.ewp) {
      double x = lane_sum(-wJ * lu.x), y = lane_sum(-wJ * lu.y), z = lane_sum(-wJ * lu.z);
      if (lane == 0) { out[0] = x; out[1] = y; out[2] = z; }
    } else
      for (int v = 0; v < 4; ++v) {
        double sN = -wJ * tab->n1[q][v];
        double x = lane_sum(sN * lu.x), y = lane_sum(sN * lu.y), z = lane_sum(sN * lu.z);
        if (lane == 0) { out[3 * v] = x; out[3 * v + 1] = y; out[3 * v + 2] = z; }
      }
  } else if (OP == OP_F_FWD || OP == OP_MASS_U) {
    V3 fq = v3(0, 0, 0);
    if (OP == OP_F_FWD) {
      if (A.ewp) fq = ldv3(A.vec_l + 3 * cell);
      else for (int v = 0; v < 4; ++v) fq = fq + tab->n1[q][v] * ldv3(A.vec_l + 3 * (int64_t)vid[v]);
      fq = (-1.0) * fq;
    } else {
      for (int a = 0; a < 9; ++a) fq = fq + tab->phi2[q][a] * ldv3(s_w[wid] + 3 * a);
    }
    for (int a = 0; a < 9; ++a) {
      double sN = wJ * tab->phi2[q][a];
      double x = lane_sum(sN * fq.x), y = lane_sum(sN * fq.y), z = lane_sum(sN * fq.z);
      if (lane == 0) { out[3 * a] = x; out[3 * a + 1] = y; out[3 * a + 2] = z; }
    }
  } else if (OP == OP_UU) {
    V3 u = v3(0, 0, 0);
    for (int a = 0; a < 9; ++a) u = u + tab->phi2[q][a] * ldv3(s_w[wid] + 3 * a);
    double t = lane_sum(wJ * dot(u, u));
    if (lane == 0) out[0] = t;
  } else if (OP == OP_MASS) {
    double t = lane_sum(wJ * rhoq * hq);
    if (lane == 0) out[0] = t;
  } else if (OP == OP_DMASS_DH || OP == OP_DMASS_DRHO) {
    double cq = wJ * (OP == OP_DMASS_DH ? rhoq : hq);
    if (A.ewm) { double t = lane_sum(cq); if (lane == 0) out[0] = t; }
    else
      for (int v = 0; v < 4; ++v) { double t = lane_sum(cq * tab->n1[q][v]); if (lane == 0) out[v] = t; }
  } else if (OP == OP_REG || OP == OP_REG_GRAD) {
    // rm_shell_pde.py:64-83: 'L2' (DG0 thickness) 0.5 a1 h^2 dx ; 'H1' (CG1) 0.5 a1 |grad h|^2 dx
    if (A.ewm) {
      double t = lane_sum(OP == OP_REG ? 0.5 * FEMO_REG_ALPHA1 * wd * hq * hq : FEMO_REG_ALPHA1 * wd * hq);
      if (lane == 0) out[0] = t;
    } else {
      V3 gN[4], gh = v3(0, 0, 0);
      for (int v = 0; v < 4; ++v) {
        gN[v] = tab->dn1[q][v][0] * g.Ki0 + tab->dn1[q][v][1] * g.Ki1;
        gh = gh + A.h[vid[v]] * gN[v];
      }
      if (OP == OP_REG) {
        double t = lane_sum(0.5 * FEMO_REG_ALPHA1 * wd * dot(gh, gh));
        if (lane == 0) out[0] = t;
      } else
        for (int v = 0; v < 4; ++v) { double t = lane_sum(FEMO_REG_ALPHA1 * wd * dot(gN[v], gh)); if (lane == 0) out[v] = t; }
    }
  } else if (OP == OP_AREA) {
    double t = lane_sum(wJ);
    if (lane == 0) out[0] = t;
  } else if (OP == OP_MASS_ELEM) {
    const double cu = wJ * rhoq * hq, ct = cu * A.hE[cell] * A.hE[cell];
    for (int a = 0; a < 9; ++a)
      for (int b = 0; b < 9; ++b) {
        double t = lane_sum(cu * tab->phi2[q][a] * tab->phi2[q][b]);
        if (lane == 0) out[9 * a + b] = t;
      }
    for (int a = 0; a < 4; ++a)
      for (int b = 0; b < 4; ++b) {
        double t = lane_sum(ct * tab->n1[q][a] * tab->n1[q][b]);
        if (lane == 0) out[81 + 4 * a + b] = t;
      }
  } else if (OP == OP_MASS_PARAM_T || OP == OP_MASS_PARAM_FWD) {
    // acceleration (vec_w) at the point
    V3 au = v3(0, 0, 0), at = v3(0, 0, 0);
    for (int a = 0; a < 9; ++a) au = au + tab->phi2[q][a] * ldv3(s_w[wid] + 3 * a);
    for (int b = 0; b < 4; ++b) at = at + tab->n1[q][b] * ldv3(s_w[wid] + 27 + 3 * b);
    const double cu = wJ * rhoq, ct = cu * A.hE[cell] * A.hE[cell], cg = wd * rhoq * A.g_acc;
    if (OP == OP_MASS_PARAM_T) {
      V3 lu = v3(0, 0, 0), lt = v3(0, 0, 0);
      for (int a = 0; a < 9; ++a) lu = lu + tab->phi2[q][a] * ldv3(s_l[wid] + 3 * a);
      for (int b = 0; b < 4; ++b) lt = lt + tab->n1[q][b] * ldv3(s_l[wid] + 27 + 3 * b);
      const double gq = cu * dot(au, lu) + ct * dot(at, lt) - cg * lu.z;
      if (A.ewm) { double t = lane_sum(gq); if (lane == 0) out[0] = A.accumulate ? out[0] + t : t; }
      else
        for (int v = 0; v < 4; ++v) {
          double t = lane_sum(gq * tab->n1[q][v]);
          if (lane == 0) out[v] = A.accumulate ? out[v] + t : t;
        }
    } else {
      double dp;
      if (A.ewm) dp = A.vec_l[cell];
      else { dp = 0; for (int v = 0; v < 4; ++v) dp += tab->n1[q][v] * A.vec_l[vid[v]]; }
      const V3 ru = (dp * cu) * au - v3(0, 0, dp * cg), rt = (dp * ct) * at;
      for (int a = 0; a < 9; ++a) {
        double sN = tab->phi2[q][a];
        double x = lane_sum(sN * ru.x), y = lane_sum(sN * ru.y), z = lane_sum(sN * ru.z);
        if (lane == 0) { out[3 * a] = x; out[3 * a + 1] = y; out[3 * a + 2] = z; }
      }
      for (int b = 0; b < 4; ++b) {
        double sN = tab->n1[q][b];
        double x = lane_sum(sN * rt.x), y = lane_sum(sN * rt.y), z = lane_sum(sN * rt.z);
        if (lane == 0) { out[27 + 3 * b] = x; out[27 + 3 * b + 1] = y; out[27 + 3 * b + 2] = z; }
      }
    }
  } else if (OP == OP_WEIGHT) {
    const double cg = wd * rhoq * hq * A.g_acc;
    for (int a = 0; a < 9; ++a) {
      double z = lane_sum(cg * tab->phi2[q][a]);
      if (lane == 0) { out[3 * a] = 0.0; out[3 * a + 1] = 0.0; out[3 * a + 2] = z; }
    }
  } else if (OP == OP_SHAPE_T || OP == OP_SHAPE_FWD) {
    const Coefs cf = clt_coefs(hq, Eq, nuq, A.hE[cell], g.Ju, 0);
    V3 fq = v3(0, 0, 0);
    if (A.ewp) fq = ldv3(A.f + 3 * cell);
    else for (int v = 0; v < 4; ++v) fq = fq + tab->n1[q][v] * ldv3(A.f + 3 * (int64_t)vid[v]);
    if (OP == OP_SHAPE_T) {
      double o12[12];
      shape_T_qp(g, tab, q, s_w[wid], s_l[wid], cf, fq, wd, A.mode, o12);
#pragma unroll
      for (int i = 0; i < 12; ++i) { double t = lane_sum(o12[i]); if (lane == 0) out[i] = t; }
    } else {
      V3 duh[4];
#pragma unroll
      for (int v = 0; v < 4; ++v) duh[v] = ldv3(A.vec_l + 3 * (int64_t)vid[v]);
      double o39[39];
      shape_fwd_qp(g, tab, q, s_w[wid], cf, fq, wd, duh, o39);
#pragma unroll
      for (int i = 0; i < 39; ++i) { double t = lane_sum(o39[i]); if (lane == 0) out[i] = t; }
    }
  } else if (OP == OP_SHAPE_MEASURE) {
    double cq;
    if (A.mode == 0) {
      V3 u = v3(0, 0, 0);
      for (int a = 0; a < 9; ++a) u = u + tab->phi2[q][a] * ldv3(s_w[wid] + 3 * a);
      cq = dot(u, u);
    } else cq = rhoq * hq;
    double o12[12];
    shape_measure_qp(g, tab, q, wd * cq, o12);
#pragma unroll
    for (int i = 0; i < 12; ++i) { double t = lane_sum(o12[i]); if (lane == 0) out[i] = t; }
  } else if (OP == OP_VOLUME) {
    double t = lane_sum(wd * hq);
    if (lane == 0) out[0] = t;
  } else if (OP == OP_DVOLUME_DH) {
    if (A.ewm) { double t = lane_sum(wd); if (lane == 0) out[0] = t; }
    else
      for (int v = 0; v < 4; ++v) { double t = lane_sum(wd * tab->n1[q][v]); if (lane == 0) out[v] = t; }
  } else if (OP == OP_SHAPE_PNORM) {
    const bool in_dom = (A.tags == nullptr || A.tag < 0 || A.tags[cell] == A.tag);
    double hv[4] = {0, 0, 0, 0};
    if (!A.ewm) for (int v = 0; v < 4; ++v) hv[v] = A.h[vid[v]];
    VmPoint P;
    vm_at(g, tab, q, s_w[wid], hq, Eq, nuq, hv, A.ewm, A.xi_sign, P);
    double o12[12];
    shape_pnorm_qp(g, tab, q, s_w[wid], hv, A.ewm, A.xi_sign, P, in_dom ? wd : 0.0, A.sm, A.srho, o12);
#pragma unroll
    for (int i = 0; i < 12; ++i) { double t = lane_sum(o12[i]); if (lane == 0) out[i] = t; }
  } else if (OP == OP_PNORM || OP == OP_PNORM_DW || OP == OP_PNORM_DP || OP == OP_VM_PROJ) {
    const bool in_dom = (A.tags == nullptr || A.tag < 0 || A.tags[cell] == A.tag);
    double hv[4] = {0, 0, 0, 0};
    if (!A.ewm) for (int v = 0; v < 4; ++v) hv[v] = A.h[vid[v]];
    VmPoint P;
    vm_at(g, tab, q, s_w[wid], hq, Eq, nuq, hv, A.ewm, A.xi_sign, P);
    const double wdom = in_dom ? wJ : 0.0;
    if (OP == OP_PNORM) {
      double t = lane_sum(wdom * pow(A.sm * P.vm, A.srho));
      if (lane == 0) out[0] = t;
    } else if (OP == OP_VM_PROJ) {
      // M s = b, M_ij = int N_i N_j dx, b_i = int vm N_i dx  (utils_dolfinx.py:568-602, plain dx)
      double M[16], bb[4];
      for (int i = 0; i < 4; ++i) {
        bb[i] = lane_sum(wd * tab->n1[q][i] * P.vm);
        for (int j = 0; j < 4; ++j) M[4 * i + j] = lane_sum(wd * tab->n1[q][i] * tab->n1[q][j]);
      }
      if (lane == 0) {
        for (int k = 0; k < 4; ++k) {
          double ip = 1.0 / M[4 * k + k];
          for (int i = k + 1; i < 4; ++i) {
            double f = M[4 * i + k] * ip;
            for (int j = k; j < 4; ++j) M[4 * i + j] -= f * M[4 * k + j];
            bb[i] -= f * bb[k];
          }
        }
        for (int k = 3; k >= 0; --k) {
          double v = bb[k];
          for (int j = k + 1; j < 4; ++j) v -= M[4 * k + j] * bb[j];
          bb[k] = v / M[4 * k + k];
        }
        for (int i = 0; i < 4; ++i) out[i] = bb[i];
      }
    } else {
      // G = wJ rho m^rho vm^(rho-1)
      const double G = (P.vm > 0) ? wdom * A.srho * A.sm * pow(A.sm * P.vm, A.srho - 1.0) : 0.0;
      const double t0 = G * P.t[0], t1 = G * P.t[1], t2 = G * P.t[2];
      if (OP == OP_PNORM_DW) {
        const double k00 = dot(g.Kx0, g.E0), k01 = dot(g.Kx0, g.E1), k10 = dot(g.Kx1, g.E0), k11 = dot(g.Kx1, g.E1);
        const V3 gq0 = k00 * g.dE2_0 + k10 * g.dE2_1, gq1 = k01 * g.dE2_0 + k11 * g.dE2_1;
        for (int a = 0; a < 9; ++a) {
          double d0 = tab->dphi2[q][a][0] * k00 + tab->dphi2[q][a][1] * k10, d1 = tab->dphi2[q][a][0] * k01 + tab->dphi2[q][a][1] * k11;
          URows R = u_rows(d0, d1, g.E0, g.E1, g.E2);
          V3 r = t0 * R.e0 + t1 * R.e1 + t2 * R.e2;
          double rx = lane_sum(r.x), ry = lane_sum(r.y), rz = lane_sum(r.z);
          if (lane == 0) { out[3 * a] = rx; out[3 * a + 1] = ry; out[3 * a + 2] = rz; }
        }
        for (int b = 0; b < 4; ++b) {
          double d0 = tab->dn1[q][b][0] * k00 + tab->dn1[q][b][1] * k10, d1 = tab->dn1[q][b][0] * k01 + tab->dn1[q][b][1] * k11;
          double N = tab->n1[q][b];
          TRows R = t_rows(N, d0, d1, g.E0, g.E1, g.E2, gq0, gq1);
          V3 r = (-P.xi) * (t0 * R.k0 + t1 * R.k1 + t2 * R.k2) +
                 N * ((t0 * P.gh0 + t2 * P.gh1) * g.E1 - (t1 * P.gh1 + t2 * P.gh0) * g.E0);
          double rx = lane_sum(r.x), ry = lane_sum(r.y), rz = lane_sum(r.z);
          if (lane == 0) { out[27 + 3 * b] = rx; out[27 + 3 * b + 1] = ry; out[27 + 3 * b + 2] = rz; }
        }
      } else {
        const double k00 = dot(g.Kx0, g.E0), k01 = dot(g.Kx0, g.E1), k10 = dot(g.Kx1, g.E0), k11 = dot(g.Kx1, g.E1);
        for (int v = 0; v < (A.ewm ? 1 : 4); ++v) {
          double N = A.ewm ? 1.0 : tab->n1[q][v];
          double val;
          if (A.mode == 1) {
            double dv0 = A.ewm ? 0.0 : (tab->dn1[q][v][0] * k00 + tab->dn1[q][v][1] * k10);
            double dv1 = A.ewm ? 0.0 : (tab->dn1[q][v][0] * k01 + tab->dn1[q][v][1] * k11);
            double hs = 0.5 * A.xi_sign;
            double de0 = -hs * N * P.kap[0] - P.b0 * hs * dv0;
            double de1 = -hs * N * P.kap[1] - P.b1 * hs * dv1;
            double de2 = -hs * N * P.kap[2] - hs * (P.b0 * dv1 + P.b1 * dv0);
            val = t0 * de0 + t1 * de1 + t2 * de2;
          } else if (A.mode == 2) {
            val = G * P.vm / Eq * N;
          } else {
            double nu_ = P.nu, c0 = P.c0, dc = 2.0 * Eq * nu_ / ((1 - nu_ * nu_) * (1 - nu_ * nu_));
            double s0 = c0 * (P.e[0] + nu_ * P.e[1]), s1 = c0 * (nu_ * P.e[0] + P.e[1]), s2 = c0 * 0.5 * (1 - nu_) * P.e[2];
            double iv = P.vm > 0 ? 0.5 / P.vm : 0.0;
            double a0 = (2 * s0 - s1) * iv, a1 = (2 * s1 - s0) * iv, a2 = 6 * s2 * iv;
            double ds0 = dc * (P.e[0] + nu_ * P.e[1]) + c0 * P.e[1];
            double ds1 = dc * (nu_ * P.e[0] + P.e[1]) + c0 * P.e[0];
            double ds2 = (dc * 0.5 * (1 - nu_) - 0.5 * c0) * P.e[2];
            val = G * (a0 * ds0 + a1 * ds1 + a2 * ds2) * N;
          }
          double tt = lane_sum(val);
          if (lane == 0) out[v] = tt;
        }
      }
    }
  }
}

// out[block I, 0..2] = sum over adjacent cells (ascending) of cellbuf[cell*stride + off(a)]
__global__ void k_node_gather(const int32_t* __restrict__ nc_ptr, const int32_t* __restrict__ nc_adj,
                              const double* __restrict__ cellbuf, int stride, int64_t first_block, int64_t nblocks,
                              int local_shift, int ncomp, double* __restrict__ out) {
  int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= nblocks) return;
  int64_t I = first_block + t;
  double a0 = 0, a1 = 0, a2 = 0;
  for (int32_t k = nc_ptr[I]; k < nc_ptr[I + 1]; ++k) {
    int32_t e = nc_adj[k];
    int64_t cell = e / 13; int a = e - (int32_t)cell * 13 - local_shift;
    const double* s = cellbuf + cell * stride + a * ncomp;
    a0 += s[0];
    if (ncomp == 3) { a1 += s[1]; a2 += s[2]; }
  }
  if (ncomp == 3) { out[3 * t] = a0; out[3 * t + 1] = a1; out[3 * t + 2] = a2; }
  else out[t] = a0;
}

__global__ void k_copy_strided(const double* __restrict__ src, int stride, int64_t n, double* __restrict__ dst) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[i * stride];
}
__global__ void k_copy_cell3(const double* __restrict__ src, int stride, int64_t n, double* __restrict__ dst) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) { dst[3 * i] = src[i * stride]; dst[3 * i + 1] = src[i * stride + 1]; dst[3 * i + 2] = src[i * stride + 2]; }
}
__global__ void k_scale(double* x, int64_t n, double s) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) x[i] *= s;
}
// deterministic sum of cellbuf[cell*stride] over all cells: fixed grid, fixed order
__global__ void __launch_bounds__(256) k_sum_cells(const double* __restrict__ buf, int stride, int64_t n, double* __restrict__ part) {
  __shared__ double s[8];
  double v = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += gridDim.x * (int64_t)blockDim.x) v += buf[i * stride];
  v = lane_sum(v);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) { double t = 0; for (int i = 0; i < 8; ++i) t += s[i]; part[blockIdx.x] = t; }
}
__global__ void k_sum_final(const double* __restrict__ part, int n, double* out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) { double t = 0; for (int i = 0; i < n; ++i) t += part[i]; *out = t; }
}

__global__ void k_mask_cells(double* buf, int stride, const int32_t* tags, int tag, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n && tags[i] != tag) buf[i * stride] = 0.0;
}
__global__ void k_copy_cell4(const double* __restrict__ src, int stride, int64_t n, double* __restrict__ dst) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) for (int k = 0; k < 4; ++k) dst[4 * i + k] = src[i * stride + k];
}

CellArgs base_args(femo_ctx* c) {
  CellArgs A;
  A.xyz = c->xyz; A.cv = c->cv; A.cb = c->cell_blocks; A.uhat = c->uhat_zero ? nullptr : c->uhat;
  A.h = c->h; A.E = c->E; A.nu = c->nu; A.rho = c->rho; A.f = c->f; A.hE = c->hE; A.tab = c->tab;
  A.nc = c->nc; A.nq = c->nq; A.ewm = c->ewm; A.ewp = c->ewp;
  A.vec_w = c->w; A.vec_l = nullptr; A.mode = 0; A.out = c->cellbuf; A.out_stride = 39;
  A.sm = c->stress_m; A.srho = c->stress_rho; A.xi_sign = 1.0; A.tags = c->cell_tags; A.tag = -1;
  A.accumulate = 0; A.g_acc = 0.0;
  return A;
}

template <int OP>
int run_cell_op(femo_ctx* c, const CellArgs& A) {
  k_cell_op<OP><<<(unsigned)cdiv(c->nc, 4), 128, 0, c->stream>>>(A);
  TRY(check_launch(c, "k_cell_op"));
  c->launches += 1;
  return 0;
}

int sum_cells(femo_ctx* c, double* val) {
  const int NB = 296;
  double* part = c->red + 6 * 1184;
  // multi-GPU: every cell is integrated by exactly one rank (owned cells come first), then summed over ranks
  const int64_t ncs = c->world > 1 ? c->nc_owned : c->nc;
  k_sum_cells<<<NB, 256, 0, c->stream>>>(c->cellbuf, 39, ncs, part);
  k_sum_final<<<1, 32, 0, c->stream>>>(part, NB, c->scal + 9);
  TRY(check_launch(c, "k_sum_cells"));
  c->launches += 2;
  if (c->world > 1) TRY(dist_allreduce_sum(c, c->scal + 9, 1));
  CUDA_TRY(cudaMemcpyAsync(val, c->scal + 9, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return 0;
}

// material-parameter shaped result: vertex gather (CG1) or per-cell copy (DG0)
int gather_material(femo_ctx* c, double* out) {
  if (c->ewm) k_copy_strided<<<(unsigned)cdiv(c->nc, 256), 256, 0, c->stream>>>(c->cellbuf, 39, c->nc, out);
  else k_node_gather<<<(unsigned)cdiv(c->nn, 256), 256, 0, c->stream>>>(c->nc_ptr, c->nc_adj, c->cellbuf, 39, c->nq2, c->nn, 9, 1, out);
  TRY(check_launch(c, "gather_material"));
  c->launches += 1;
  return 0;
}

int gather_dofs(femo_ctx* c, double* out, bool with_theta) {
  k_node_gather<<<(unsigned)cdiv(c->nq2, 256), 256, 0, c->stream>>>(c->nc_ptr, c->nc_adj, c->cellbuf, 39, 0, c->nq2, 0, 3, out);
  if (with_theta)
    k_node_gather<<<(unsigned)cdiv(c->nn, 256), 256, 0, c->stream>>>(c->nc_ptr, c->nc_adj, c->cellbuf, 39, c->nq2, c->nn, 0, 3,
                                                                    out + 3 * c->nq2);
  else CUDA_TRY(cudaMemsetAsync(out + 3 * c->nq2, 0, sizeof(double) * 3 * c->nn, c->stream));
  TRY(check_launch(c, "gather_dofs"));
  c->launches += with_theta ? 2 : 1;
  return 0;
}

int mode_of(int arg) { return arg == FEMO_THICKNESS ? 1 : arg == FEMO_E ? 2 : arg == FEMO_NU ? 3 : -1; }

}  // namespace

// ---- dynamic path helpers (called from dynamic.cu) ---------------------------------------------------
static CellArgs mass_args(femo_ctx* c) {
  CellArgs A = base_args(c);
  A.tab = c->tabm; A.nq = 16;       // 4x4 rule of the mass form
  return A;
}
int mass_elements(femo_ctx* c, double* me) {
  CellArgs A = mass_args(c);
  A.out = me; A.out_stride = 97;
  return run_cell_op<OP_MASS_ELEM>(c, A);
}
int mass_param_T(femo_ctx* c, const double* acc, const double* lam, double g_acc, bool accumulate) {
  CellArgs A = mass_args(c);
  A.vec_w = acc; A.vec_l = lam; A.g_acc = g_acc; A.accumulate = accumulate ? 1 : 0;
  return run_cell_op<OP_MASS_PARAM_T>(c, A);
}
int stiff_param_T(femo_ctx* c, const double* w, const double* lam, bool accumulate) {
  CellArgs A = base_args(c);
  A.vec_w = w; A.vec_l = lam; A.mode = 1; A.accumulate = accumulate ? 1 : 0;
  return run_cell_op<OP_PARAM_T>(c, A);
}
int mass_param_fwd(femo_ctx* c, const double* acc, const double* dh, double g_acc, double* out) {
  CellArgs A = mass_args(c);
  A.vec_w = acc; A.vec_l = dh; A.g_acc = g_acc;
  TRY(run_cell_op<OP_MASS_PARAM_FWD>(c, A));
  return gather_dofs(c, out, true);
}
int stiff_param_fwd(femo_ctx* c, const double* w, const double* dh, double* out) {
  CellArgs A = base_args(c);
  A.vec_w = w; A.vec_l = dh; A.mode = 1;
  TRY(run_cell_op<OP_PARAM_FWD>(c, A));
  return gather_dofs(c, out, true);
}
int cellbuf_to_material(femo_ctx* c, double* out) { return gather_material(c, out); }
int self_weight_load(femo_ctx* c, double g_acc, double* out) {
  CellArgs A = mass_args(c);
  A.g_acc = g_acc;
  TRY(run_cell_op<OP_WEIGHT>(c, A));
  return gather_dofs(c, out, false);
}

// cellbuf[cell][v][3] -> vertex vector (nn,3), adjacent cells summed in ascending order
static int gather_vertices3(femo_ctx* c, double* out) {
  k_node_gather<<<(unsigned)cdiv(c->nn, 256), 256, 0, c->stream>>>(c->nc_ptr, c->nc_adj, c->cellbuf, 39, c->nq2, c->nn, 9, 3, out);
  TRY(check_launch(c, "gather_vertices3"));
  c->launches += 1;
  return 0;
}

int dRdf_T_apply(femo_ctx* c, int arg, const double* lam, double* out) {
  CellArgs A = base_args(c);
  A.vec_l = lam;
  if (arg == FEMO_UHAT) {                       // shape partial: cell terms, then the penalty-BC facet terms
    A.mode = 1;
    TRY(run_cell_op<OP_SHAPE_T>(c, A));
    TRY(gather_vertices3(c, out));
    return penalty_shape_T(c, lam, out);
  }
  if (arg == FEMO_F_SOLID) {
    TRY(run_cell_op<OP_F_T>(c, A));
    if (c->ewp) k_copy_cell3<<<(unsigned)cdiv(c->nc, 256), 256, 0, c->stream>>>(c->cellbuf, 39, c->nc, out);
    else k_node_gather<<<(unsigned)cdiv(c->nn, 256), 256, 0, c->stream>>>(c->nc_ptr, c->nc_adj, c->cellbuf, 39, c->nq2, c->nn, 9, 3, out);
    TRY(check_launch(c, "dRdF^T"));
    c->launches += 1;
    return 0;
  }
  A.mode = mode_of(arg);
  if (A.mode < 0) return femo_fail("femo_dRdf_T_apply: argument not supported (thickness, F_solid, E, nu)");
  TRY(run_cell_op<OP_PARAM_T>(c, A));
  return gather_material(c, out);
}

int dRdf_apply(femo_ctx* c, int arg, const double* df, double* out) {
  CellArgs A = base_args(c);
  A.vec_l = df;
  if (arg == FEMO_UHAT) {
    TRY(run_cell_op<OP_SHAPE_FWD>(c, A));
    TRY(gather_dofs(c, out, true));
    return penalty_shape_fwd(c, df, out);
  }
  if (arg == FEMO_F_SOLID) {
    TRY(run_cell_op<OP_F_FWD>(c, A));
    return gather_dofs(c, out, false);
  }
  A.mode = mode_of(arg);
  if (A.mode < 0) return femo_fail("femo_dRdf_apply: argument not supported (thickness, F_solid, E, nu)");
  TRY(run_cell_op<OP_PARAM_FWD>(c, A));
  return gather_dofs(c, out, true);
}

int out_scalar(femo_ctx* c, int which, double* val) {
  CellArgs A = base_args(c);
  if (which == FEMO_COMPLIANCE) {
    double uu = 0, reg = 0;
    TRY(run_cell_op<OP_UU>(c, A)); TRY(sum_cells(c, &uu));
    TRY(run_cell_op<OP_REG>(c, A)); TRY(sum_cells(c, &reg));
    *val = uu + reg;
    return 0;
  }
  if (which == FEMO_MASS) { TRY(run_cell_op<OP_MASS>(c, A)); return sum_cells(c, val); }
  if (which == FEMO_ELASTIC_ENERGY) {
    if (!c->jac_valid) TRY(assemble_jacobian(c));
    TRY(bsr_spmv_vals(c, stiffness_vals(c), c->w, c->tmp, 0));
    TRY(penalty_apply(c, c->w, c->tmp, -1.0));
    double d = 0;
    TRY(dev_dot(c, c->w, c->tmp, c->ndof, &d));
    *val = 0.5 * d;
    return 0;
  }
  if (which == FEMO_PNORM_STRESS) return pnorm_value(c, -1, val);
  if (which == FEMO_VOLUME) { TRY(run_cell_op<OP_VOLUME>(c, A)); return sum_cells(c, val); }
  return femo_fail("femo_output: output not implemented");
}

int out_grad(femo_ctx* c, int which, int arg, double* g) {
  CellArgs A = base_args(c);
  const int64_t n_arg = (arg == FEMO_ARG_STATE) ? c->ndof : (arg == FEMO_F_SOLID) ? c->n_f : (arg == FEMO_UHAT) ? 3 * c->nn : c->n_mat;
  if (arg == FEMO_UHAT) {                       // shape partials of the outputs
    if (which == FEMO_COMPLIANCE || which == FEMO_MASS) {
      A.mode = (which == FEMO_COMPLIANCE) ? 0 : 1;
      TRY(run_cell_op<OP_SHAPE_MEASURE>(c, A));
      return gather_vertices3(c, g);
    }
    if (which == FEMO_ELASTIC_ENERGY) {         // 1/2 w^T dK/duhat w
      A.vec_l = c->w; A.mode = 0;
      TRY(run_cell_op<OP_SHAPE_T>(c, A));
      TRY(gather_vertices3(c, g));
      k_scale<<<(unsigned)cdiv(n_arg, 256), 256, 0, c->stream>>>(g, n_arg, 0.5);
      c->launches += 1;
      return check_launch(c, "k_scale");
    }
    if (which == FEMO_VOLUME) { CUDA_TRY(cudaMemsetAsync(g, 0, sizeof(double) * n_arg, c->stream)); return 0; }
    if (which == FEMO_PNORM_STRESS) return pnorm_grad(c, -1, arg, g);
    return femo_fail("femo_output_grad: output not implemented");
  }
  if (which == FEMO_COMPLIANCE) {
    if (arg == FEMO_ARG_STATE) {
      TRY(run_cell_op<OP_MASS_U>(c, A));
      TRY(gather_dofs(c, g, false));
      k_scale<<<(unsigned)cdiv(c->ndof, 256), 256, 0, c->stream>>>(g, c->ndof, 2.0);
      c->launches += 1;
      return check_launch(c, "k_scale");
    }
    if (arg == FEMO_THICKNESS) { TRY(run_cell_op<OP_REG_GRAD>(c, A)); return gather_material(c, g); }
    CUDA_TRY(cudaMemsetAsync(g, 0, sizeof(double) * n_arg, c->stream));
    return 0;
  }
  if (which == FEMO_MASS) {
    if (arg == FEMO_THICKNESS) { TRY(run_cell_op<OP_DMASS_DH>(c, A)); return gather_material(c, g); }
    if (arg == FEMO_DENSITY) { TRY(run_cell_op<OP_DMASS_DRHO>(c, A)); return gather_material(c, g); }
    CUDA_TRY(cudaMemsetAsync(g, 0, sizeof(double) * n_arg, c->stream));
    return 0;
  }
  if (which == FEMO_ELASTIC_ENERGY) {
    if (arg == FEMO_ARG_STATE) {
      if (!c->jac_valid) TRY(assemble_jacobian(c));
      TRY(bsr_spmv_vals(c, stiffness_vals(c), c->w, g, 0));
      return penalty_apply(c, c->w, g, -1.0);
    }
    if (arg == FEMO_THICKNESS || arg == FEMO_E || arg == FEMO_NU) {
      TRY(dRdf_T_apply(c, arg, c->w, g));
      k_scale<<<(unsigned)cdiv(c->n_mat, 256), 256, 0, c->stream>>>(g, c->n_mat, 0.5);
      c->launches += 1;
      return check_launch(c, "k_scale");
    }
    CUDA_TRY(cudaMemsetAsync(g, 0, sizeof(double) * n_arg, c->stream));
    return 0;
  }
  if (which == FEMO_PNORM_STRESS) return pnorm_grad(c, -1, arg, g);
  if (which == FEMO_VOLUME) {
    if (arg == FEMO_THICKNESS) { TRY(run_cell_op<OP_DVOLUME_DH>(c, A)); return gather_material(c, g); }
    CUDA_TRY(cudaMemsetAsync(g, 0, sizeof(double) * n_arg, c->stream));
    return 0;
  }
  return femo_fail("femo_output_grad: output not implemented");
}

// ---- stress functionals (rm_shell_pde.py:112-166; rm_shell_model.py:200-253) -----------------
static CellArgs stress_args(femo_ctx* c, int tag) {
  CellArgs A = base_args(c);
  A.tab = c->tab4; A.nq = c->nq4; A.tag = tag;     // degree-4 measure (rm_shell_model.py:200-201)
  return A;
}

int stress_area(femo_ctx* c, int tag, double* area) {
  // frozen at first use with uhat as registered then (pde:124-127: evaluated once at construction)
  auto it = c->area_cache.find(tag);
  if (it != c->area_cache.end()) { *area = it->second; return 0; }
  CellArgs A = base_args(c);
  A.uhat = nullptr;
  TRY(run_cell_op<OP_AREA>(c, A));
  if (tag >= 0) {
    if (!c->cell_tags) return femo_fail("pnorm_stress: no cell tags set");
    k_mask_cells<<<(unsigned)cdiv(c->nc, 256), 256, 0, c->stream>>>(c->cellbuf, 39, c->cell_tags, tag, c->nc);
    c->launches += 1;
  }
  TRY(sum_cells(c, area));
  c->area_cache[tag] = *area;
  return 0;
}

int pnorm_value(femo_ctx* c, int tag, double* val) {
  double area = 0;
  TRY(stress_area(c, tag, &area));
  CellArgs A = stress_args(c, tag);
  TRY(run_cell_op<OP_PNORM>(c, A));
  double v = 0;
  TRY(sum_cells(c, &v));
  *val = v / area;
  return 0;
}

int pnorm_grad(femo_ctx* c, int tag, int arg, double* g) {
  double area = 0;
  TRY(stress_area(c, tag, &area));
  CellArgs A = stress_args(c, tag);
  int64_t n;
  if (arg == FEMO_ARG_STATE) {
    TRY(run_cell_op<OP_PNORM_DW>(c, A));
    TRY(gather_dofs(c, g, true));
    n = c->ndof;
  } else if (arg == FEMO_THICKNESS || arg == FEMO_E || arg == FEMO_NU) {
    A.mode = mode_of(arg);
    TRY(run_cell_op<OP_PNORM_DP>(c, A));
    TRY(gather_material(c, g));
    n = c->n_mat;
  } else if (arg == FEMO_UHAT) {             // shape partial; the area normalisation is a constant of the form
    TRY(run_cell_op<OP_SHAPE_PNORM>(c, A));
    TRY(gather_vertices3(c, g));
    n = 3 * c->nn;
  } else {
    n = (arg == FEMO_F_SOLID) ? c->n_f : c->n_mat;
    CUDA_TRY(cudaMemsetAsync(g, 0, sizeof(double) * n, c->stream));
    return 0;
  }
  k_scale<<<(unsigned)cdiv(n, 256), 256, 0, c->stream>>>(g, n, 1.0 / area);
  c->launches += 1;
  return check_launch(c, "k_scale");
}

int stress_field(femo_ctx* c, int surface, double* out) {
  CellArgs A = base_args(c);
  A.xi_sign = (surface == 0) ? 1.0 : (surface == 1 ? 0.0 : -1.0);
  TRY(run_cell_op<OP_VM_PROJ>(c, A));
  k_copy_cell4<<<(unsigned)cdiv(c->nc, 256), 256, 0, c->stream>>>(c->cellbuf, 39, c->nc, out);
  c->launches += 1;
  return check_launch(c, "k_copy_cell4");
}

// ---- allocation-free variant of the matrix-free (dR/dparam)^T lam product ------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA-graph capture, tight serving loops, harnesses that forbid allocation).
//
// The regular path (dRdf_T_apply above) lets k_cell_op<OP_PARAM_T> write one element vector per cell into the context's
// cudaMalloc'ed `cellbuf` and gathers it per node / per cell.  Here that per-cell scratch is a file-scope device array of
// FEMO_NOALLOC_MAX_CELLS x 4 doubles (2 MiB; every GPU that loads the library holds it until the process exits).
// In HOST code the name `g_noalloc_cellbuf` is only a host-side stand-in for the symbol, not a device address: the
// address the kernels receive comes from cudaGetSymbolAddress, resolved once per device with that device current.
// All calls share the array and it keeps the previous call's contents: the part a call uses is cleared first, and two
// calls that overlap on the GPU (different streams, different contexts on one device) race.
static __device__ __align__(256) double g_noalloc_cellbuf[(size_t)FEMO_NOALLOC_MAX_CELLS * 4];

namespace {
constexpr int NOALLOC_MAX_DEV = 64;
double* g_noalloc_ptr[NOALLOC_MAX_DEV] = {};        // device address of the scratch array, per device
std::string g_noalloc_err[NOALLOC_MAX_DEV];

// the caller has made `device` current
int noalloc_resolve(int device, double** out) {
  if (device < 0 || device >= NOALLOC_MAX_DEV) return femo_fail("allocation-free scratch: device index out of range");
  if (!g_noalloc_ptr[device]) {
    double* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_cellbuf);
    if (e != cudaSuccess || !p) {
      g_noalloc_err[device] = std::string("cudaGetSymbolAddress(g_noalloc_cellbuf) failed: ") + cudaGetErrorString(e);
      return femo_fail(g_noalloc_err[device]);
    }
    g_noalloc_ptr[device] = p;
  }
  *out = g_noalloc_ptr[device];
  return 0;
}
}  // namespace

// called by femo_shell_setup (device current): resolves the scratch address at set-up so that the entry point below does
// nothing but launch.  A failure is kept for the entry point to report; it does not fail the set-up.
void noalloc_prepare(int device) {
  double* p = nullptr;
  (void)noalloc_resolve(device, &p);
}

extern "C" int femo_dRdf_T_apply_noalloc(femo_ctx* c, int arg, const double* lam_dev, double* out_dev) {
  if (!c || !c->setup_done) return femo_fail("femo_dRdf_T_apply_noalloc: context not set up (femo_mesh_set + femo_shell_setup)");
  if (c->kind != 0) return femo_fail("femo_dRdf_T_apply_noalloc: RM-shell contexts only");
  if (!lam_dev || !out_dev) return femo_fail("femo_dRdf_T_apply_noalloc: null device pointer");
  const int mode = mode_of(arg);
  if (mode < 0) return femo_fail("femo_dRdf_T_apply_noalloc: argument not supported (thickness, E, nu)");
  if (c->nc > FEMO_NOALLOC_MAX_CELLS)
    return femo_fail("femo_dRdf_T_apply_noalloc: " + std::to_string(c->nc) + " cells exceed the static scratch of " +
                     std::to_string(FEMO_NOALLOC_MAX_CELLS) + " cells (FEMO_NOALLOC_MAX_CELLS); use femo_dRdf_T_apply");
  int prev = 0;
  cudaGetDevice(&prev);
  if (prev != c->device) cudaSetDevice(c->device);
  struct Restore { int p, d; ~Restore() { if (p != d) cudaSetDevice(p); } } restore{prev, c->device};
  double* buf = nullptr;
  TRY(noalloc_resolve(c->device, &buf));
  CUDA_TRY(cudaMemsetAsync(buf, 0, sizeof(double) * 4 * (size_t)c->nc, c->stream));     // shared array: start from zero
  CellArgs A = base_args(c);
  A.vec_l = lam_dev; A.mode = mode; A.out = buf; A.out_stride = 4;
  TRY(run_cell_op<OP_PARAM_T>(c, A));
  if (c->ewm) k_copy_strided<<<(unsigned)cdiv(c->nc, 256), 256, 0, c->stream>>>(buf, 4, c->nc, out_dev);
  else k_node_gather<<<(unsigned)cdiv(c->nn, 256), 256, 0, c->stream>>>(c->nc_ptr, c->nc_adj, buf, 4, c->nq2, c->nn, 9, 1, out_dev);
  c->launches += 1;
  return check_launch(c, "femo_dRdf_T_apply_noalloc");
}
