// rldl_device_fn.cuh — device functions shared by the generic kernels (rldl_kernels.cu)
// and the streaming kernels (rldl_stream.cu).  Every routine is written for a "group" of
// G consecutive lanes (8, 16 or 32) owning one QP instance.
//
// Reference semantics mirrored here (file:line into the reference tree):
//   stage recursion ................ src/recursive_ldl.c:1139-1318 (LDL_factorize_recursive)
//   KKT solve + z_tilde step ....... lin_sys/direct/qdldl/qdldl_interface.c:559-585
//   ADMM step ...................... src/auxil.c:164-228, src/proj.c:4-14
//   residuals / tolerances ......... src/auxil.c:243-362
//   infeasibility certificates ..... src/auxil.c:364-515
//   termination .................... src/auxil.c:684-789
//   rho estimate / adaptation ...... src/auxil.c:13-77, src/osqp.c:1268-1319
//   constraint typing .............. src/auxil.c:79-145
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include "rldl_device.h"

namespace rldl {

static constexpr double kInfty = 1e30;
static constexpr double kMinScaling = 1e-4, kMaxScaling = 1e4;
static constexpr double kInfBound = kInfty * kMinScaling;  // 1e26, auxil.c:85
static constexpr double kRhoMin = 1e-6, kRhoMax = 1e6, kRhoEq = 1e3, kRhoTol = 1e-4;
static constexpr double kOsqpNan = (double)0x7fc00000UL;  // constants.h:95 (sic)

#define FULL 0xffffffffu

// A "group" is G consecutive lanes of a warp (G = 8, 16 or 32) that own one QP instance.
// gl = lane index inside the group, gm = the group's lane mask; all reductions and
// barriers are group-scoped so the groups of a warp may diverge from each other.
template <int G>
__device__ __forceinline__ double grp_max(double v, unsigned gm) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(gm, v, o));
  return v;
}
template <int G>
__device__ __forceinline__ double grp_sum(double v, unsigned gm) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(gm, v, o);
  return v;
}
template <int G>
__device__ __forceinline__ int grp_or(int v, unsigned gm) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v |= __shfl_xor_sync(gm, v, o);
  return v;
}

// Row `row` of a shared ELL matrix times a vector held in shared memory.
__device__ __forceinline__ double ell_dot(const ModelDev &M, const Ell &E, int row, const double *v) {
  const int *ix = M.idx + E.off + row * E.width;
  const double *vv = M.val + E.off + row * E.width;
  double s = 0.0;
  for (int e = 0; e < E.width; e++) s = fma(__ldg(vv + e), v[__ldg(ix + e)], s);
  return s;
}

// out_i = sum_j Mx[j*dim+i] * v[j]  (Mx symmetric: lanes read consecutive addresses)
__device__ __forceinline__ double sym_mv(const double *__restrict__ Mx, int dim, int i, const double *v) {
  double s0 = 0.0, s1 = 0.0;
  int j = 0;
  for (; j + 1 < dim; j += 2) {
    s0 = fma(__ldcs(Mx + j * dim + i), v[j], s0);
    s1 = fma(__ldcs(Mx + (j + 1) * dim + i), v[j + 1], s1);
  }
  if (j < dim) s0 = fma(__ldcs(Mx + j * dim + i), v[j], s0);
  return s0 + s1;
}
__device__ __forceinline__ double sym_mv_model(const double *Mx, int dim, int i, const double *v) {
  double s = 0.0;
  for (int j = 0; j < dim; j++) s = fma(__ldg(Mx + j * dim + i), v[j], s);
  return s;
}

// Constraint typing of one row (auxil.c:79-101): returns rho for that row.
__device__ __forceinline__ double row_rho(double l, double u, double rho) {
  if (l < -kInfBound && u > kInfBound) return kRhoMin;
  if (u - l < kRhoTol) return kRhoEq * rho;
  return rho;
}

// The same with the Ruiz scaling of the instance (B.sE != null): the reference types the SCALED bounds
// (src/osqp.c:815-817 then auxil.c:103-145) and iterates on scaled data; in unscaled variables that is the step size
// rho_class * E_r^2 / c on row r.  ri = flat index of the row (inst * mmax + r).
__device__ __forceinline__ double row_rho_eff(const BatchDev &B, int inst, size_t ri, double rho) {
  if (!B.sE) return row_rho(B.l[ri], B.u[ri], rho);
  const double e = B.sE[ri];
  return row_rho(e * B.l[ri], e * B.u[ri], rho) * e * e / B.sC[inst];
}
// per-variable sigma of the unscaled form of the scaled iteration: sigma / (c D_i^2); vi = inst * nmax + i
__device__ __forceinline__ double var_sigma(const BatchDev &B, int inst, size_t vi, double sigma) {
  if (!B.sD) return sigma;
  const double d = B.sD[vi];
  return sigma / (B.sC[inst] * d * d);
}

// ---------------------------------------------------------------------------
// In-place inversion of a dim x dim SPD matrix held in shared memory
// (Gauss-Jordan without pivoting, lane i owns column i => conflict-free).
// Returns non-zero in every lane if a pivot was not strictly positive.
// ---------------------------------------------------------------------------
template <int G>
__device__ int spd_invert(double *Gm, int dim, int lane, unsigned gm) {
  int bad = 0;
  for (int p = 0; p < dim; p++) {
    const double piv = Gm[p * dim + p];
    if (!(piv > 0.0)) bad = 1;
    const double d = 1.0 / piv;
    __syncwarp(gm);
    // scale pivot "row" (memory column p)
    for (int j = lane; j < dim; j += G) Gm[j * dim + p] = (j == p) ? d : Gm[j * dim + p] * d;
    __syncwarp(gm);
    for (int i = lane; i < dim; i += G) {
      if (i == p) continue;
      const double f = Gm[p * dim + i];
      for (int j = 0; j < dim; j++) {
        if (j == p) continue;
        Gm[j * dim + i] = fma(-f, Gm[j * dim + p], Gm[j * dim + i]);
      }
      Gm[p * dim + i] = -f * d;
    }
    __syncwarp(gm);
  }
  return bad;
}

// ---------------------------------------------------------------------------
// Stage recursion for one instance, stages k0..N (k0 > 0 restarts from the stored
// Sinv_{k0-1}: the horizon-change path, reference src/recursive_ldl.c:946-1110).
// smem: G[a*a], W[a*nx], Sm[nx*nx], rp[max(ny,nt,nx)]
// ---------------------------------------------------------------------------
template <int G>
__device__ int factor_instance(const ModelDev &M, const BatchDev &B, double sigma, int inst, int lane, unsigned gm,
                               int k0, double *Gm, double *W, double *Sm, double *rp) {
  const int nx = M.nx, a = M.a, b = M.b, ny = M.ny, nt = M.nt, N = B.N;
  const double *rv = B.rho_vec + (size_t)inst * B.mmax;
  double *Xi = B.Xinv + (size_t)inst * B.xstride;
  double *Si = B.Sinv + (size_t)inst * B.sstride;
  int flag = 0;
  if (k0 > 0) {
    for (int e = lane; e < nx * nx; e += G) Sm[e] = Si[(size_t)(k0 - 1) * nx * nx + e];
    __syncwarp(gm);
  }
  for (int k = k0; k <= N; k++) {
    const bool term = (k == N);
    const StageClass &sc = M.sc[term ? 2 : (k == 0 ? 0 : 1)];
    const int dim = term ? nx : a;
    const int np = term ? nt : ny;
    const int roff = k * b;
    for (int r = lane; r < np; r += G) rp[r] = rv[roff + r];
    // G = Q + sigma I  (sigma per variable under Ruiz scaling; the phantom variables of stage 0 keep sigma)
    const int xoff_f = (N == 0) ? 0 : k * a - nx, ilo_f = (k == 0 && !term) ? nx : 0;
    for (int e = lane; e < dim * dim; e += G) {
      const int i = e / dim, j = e - i * dim;
      double v = __ldg(M.val + sc.Q + e);
      if (i == j) v += (i >= ilo_f) ? var_sigma(B, inst, (size_t)inst * B.nmax + xoff_f + i, sigma) : sigma;
      Gm[e] = v;
    }
    __syncwarp(gm);
    // G += C' diag(rho_p) C   (lane i owns memory column i; symmetric update)
    for (int i = lane; i < dim; i += G) {
      const int *cti = M.idx + sc.Ct.off + i * sc.Ct.width;
      const double *ctv = M.val + sc.Ct.off + i * sc.Ct.width;
      for (int e = 0; e < sc.Ct.width; e++) {
        const double cv = __ldg(ctv + e);
        if (cv == 0.0) continue;
        const int r = __ldg(cti + e);
        const double wgt = cv * rp[r];
        const int *ci = M.idx + sc.C.off + r * sc.C.width;
        const double *cvv = M.val + sc.C.off + r * sc.C.width;
        for (int e2 = 0; e2 < sc.C.width; e2++) {
          const double c2 = __ldg(cvv + e2);
          if (c2 != 0.0) Gm[__ldg(ci + e2) * dim + i] += wgt * c2;
        }
      }
      if (k > 0 && i < nx)
        for (int j = 0; j < nx; j++) Gm[j * dim + i] += Sm[j * nx + i];
    }
    __syncwarp(gm);
    if (spd_invert<G>(Gm, dim, lane, gm) && !flag) flag = 1 + k;
    double *Xk = Xi + (size_t)k * a * a;
    for (int e = lane; e < dim * dim; e += G) Xk[e] = Gm[e];
    if (term) break;
    // W = Xinv F'  (a x nx, W[i*nx + r]);  S = diag(1/rho_d) + F W
    for (int r = lane; r < nx; r += G) rp[r] = rv[roff + ny + r];
    __syncwarp(gm);
    for (int i = lane; i < a; i += G) {
      for (int r = 0; r < nx; r++) {
        const int *fi = M.idx + sc.F.off + r * sc.F.width;
        const double *fv = M.val + sc.F.off + r * sc.F.width;
        double s = 0.0;
        for (int e = 0; e < sc.F.width; e++) s = fma(__ldg(fv + e), Gm[__ldg(fi + e) * a + i], s);
        W[r * a + i] = s;  // stored transposed: W[r][i], lanes contiguous
      }
    }
    __syncwarp(gm);
    for (int c = lane; c < nx; c += G) {
      for (int r = 0; r < nx; r++) {
        const int *fi = M.idx + sc.F.off + r * sc.F.width;
        const double *fv = M.val + sc.F.off + r * sc.F.width;
        double s = (r == c) ? 1.0 / rp[r] : 0.0;
        for (int e = 0; e < sc.F.width; e++) s = fma(__ldg(fv + e), W[c * a + __ldg(fi + e)], s);
        Sm[r * nx + c] = s;
      }
    }
    __syncwarp(gm);
    if (spd_invert<G>(Sm, nx, lane, gm) && !flag) flag = 1 + k;
    double *Sk = Si + (size_t)k * nx * nx;
    for (int e = lane; e < nx * nx; e += G) Sk[e] = Sm[e];
    __syncwarp(gm);
  }
  __syncwarp(gm);
  return flag;
}

// ---------------------------------------------------------------------------
// Per-warp scratch vectors for the sweeps.
// ---------------------------------------------------------------------------
struct Scratch {
  double *g, *w, *yv, *td, *sp, *nud, *vn, *v;
};
__device__ __forceinline__ Scratch make_scratch(double *base, int vmax) {
  Scratch s;
  s.g = base; s.w = base + vmax; s.yv = base + 2 * vmax; s.td = base + 3 * vmax;
  s.sp = base + 4 * vmax; s.nud = base + 5 * vmax; s.vn = base + 6 * vmax; s.v = base + 7 * vmax;
  return s;
}

// One KKT solve (KKT=true: explicit rhs in rb, overwritten with [x~; z~]) or one
// full ADMM iteration (KKT=false: rhs built from the iterates, iterates updated
// in place stage by stage on the way back).
template <int G, bool KKT>
__device__ void kkt_sweep(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int inst,
                          int lane, unsigned gm, const Scratch &sm, double *rb, bool store_delta) {
  const int nx = M.nx, a = M.a, b = M.b, ny = M.ny, nt = M.nt, N = B.N;
  const int n = B.n;
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  const double *q = B.q + vo, *l = B.l + ro, *u = B.u + ro, *rv = B.rho_vec + ro;
  double *x = B.x + vo, *z = B.z + ro, *y = B.y + ro, *dx = B.dx + vo, *dy = B.dy + ro;
  const double *Xi = B.Xinv + (size_t)inst * B.xstride;
  const double *Si = B.Sinv + (size_t)inst * B.sstride;
  double *ykg = B.yk + (size_t)inst * B.ykstride, *skg = B.sk + (size_t)inst * B.skstride;
  const double sigma = S.sigma, alpha = S.alpha;

  // ------------------------------ forward ------------------------------
  for (int k = 0; k <= N; k++) {
    const bool term = (k == N);
    const StageClass &sc = M.sc[term ? 2 : (k == 0 ? 0 : 1)];
    const int dim = term ? nx : a, np = term ? nt : ny;
    const int xoff = (N == 0) ? 0 : k * a - nx, roff = k * b, ilo = (k == 0 && !term) ? nx : 0;
    for (int i = lane; i < dim; i += G) {
      double gv = 0.0;
      if (i >= ilo) gv = KKT ? rb[xoff + i] : var_sigma(B, inst, vo + xoff + i, sigma) * x[xoff + i] - q[xoff + i];
      sm.g[i] = gv;
    }
    for (int r = lane; r < np; r += G) {
      const double rho = rv[roff + r];
      sm.w[r] = KKT ? rho * rb[n + roff + r] : rho * z[roff + r] - y[roff + r];
    }
    __syncwarp(gm);
    for (int i = lane; i < dim; i += G) {
      double t = sm.g[i] + ell_dot(M, sc.Ct, i, sm.w);
      if (k > 0 && i < nx) t -= sm.sp[i];
      sm.v[i] = t;
    }
    __syncwarp(gm);
    const double *Xk = Xi + (size_t)k * a * a;
    for (int i = lane; i < dim; i += G) {
      const double yv = sym_mv(Xk, dim, i, sm.v);
      sm.yv[i] = yv;
      ykg[k * a + i] = yv;
    }
    __syncwarp(gm);
    if (!term) {
      for (int r = lane; r < nx; r += G) {
        const int ri = roff + ny + r;
        const double rhs = KKT ? rb[n + ri] : z[ri] - y[ri] / rv[ri];
        sm.td[r] = rhs - ell_dot(M, sc.F, r, sm.yv);
      }
      __syncwarp(gm);
      const double *Sk = Si + (size_t)k * nx * nx;
      for (int r = lane; r < nx; r += G) {
        const double s = sym_mv(Sk, nx, r, sm.td);
        sm.sp[r] = s;
        skg[k * nx + r] = s;
      }
      __syncwarp(gm);
    }
  }
  // ------------------------------ backward -----------------------------
  for (int k = N; k >= 0; k--) {
    const bool term = (k == N);
    const StageClass &sc = M.sc[term ? 2 : (k == 0 ? 0 : 1)];
    const int dim = term ? nx : a, np = term ? nt : ny;
    const int xoff = (N == 0) ? 0 : k * a - nx, roff = k * b, ilo = (k == 0 && !term) ? nx : 0;
    if (term) {
      for (int i = lane; i < dim; i += G) sm.v[i] = sm.yv[i];
      __syncwarp(gm);
    } else {
      const double *Sk = Si + (size_t)k * nx * nx;
      for (int r = lane; r < nx; r += G) sm.nud[r] = -(skg[k * nx + r] + sym_mv(Sk, nx, r, sm.vn));
      __syncwarp(gm);
      for (int i = lane; i < dim; i += G) sm.g[i] = ell_dot(M, sc.Ft, i, sm.nud);
      __syncwarp(gm);
      const double *Xk = Xi + (size_t)k * a * a;
      for (int i = lane; i < dim; i += G) sm.v[i] = ykg[k * a + i] - sym_mv(Xk, dim, i, sm.g);
      __syncwarp(gm);
    }
    // path rows: z~ = C v
    for (int r = lane; r < np; r += G) {
      const int ri = roff + r;
      const double zt = ell_dot(M, sc.C, r, sm.v);
      if (KKT) {
        rb[n + ri] = zt;
      } else {
        const double rho = rv[ri], zo = z[ri], yo = y[ri];
        const double zh = alpha * zt + (1.0 - alpha) * zo;
        const double zn = fmin(fmax(zh + yo / rho, l[ri]), u[ri]);
        const double d = rho * (zh - zn);
        z[ri] = zn; y[ri] = yo + d;
        if (store_delta) dy[ri] = d;
      }
    }
    // dynamics rows: z~ = rhs + nu/rho
    if (!term) {
      for (int r = lane; r < nx; r += G) {
        const int ri = roff + ny + r;
        const double rho = rv[ri];
        if (KKT) {
          rb[n + ri] = rb[n + ri] + sm.nud[r] / rho;
        } else {
          const double zo = z[ri], yo = y[ri];
          const double zt = (zo - yo / rho) + sm.nud[r] / rho;
          const double zh = alpha * zt + (1.0 - alpha) * zo;
          const double zn = fmin(fmax(zh + yo / rho, l[ri]), u[ri]);
          const double d = rho * (zh - zn);
          z[ri] = zn; y[ri] = yo + d;
          if (store_delta) dy[ri] = d;
        }
      }
    }
    // variables
    for (int i = lane; i < dim; i += G) {
      if (i < ilo) continue;
      const double xt = sm.v[i];
      if (KKT) {
        rb[xoff + i] = xt;
      } else {
        const double xo = x[xoff + i];
        const double xn = alpha * xt + (1.0 - alpha) * xo;
        x[xoff + i] = xn;
        if (store_delta) dx[xoff + i] = xn - xo;
      }
    }
    __syncwarp(gm);
    for (int i = lane; i < nx; i += G) sm.vn[i] = sm.v[i];
    __syncwarp(gm);
  }
}

// Stage sweep computing A*xv (per row), P*xv and A'*yv (per variable) and handing
// them to the callbacks.  rowf(ri, ax), varf(xi, xval, px, aty).
template <int G, class RowF, class VarF>
__device__ void product_sweep(const ModelDev &M, const BatchDev &B, int lane, unsigned gm, const Scratch &sm,
                              const double *xv, const double *yv, RowF rowf, VarF varf) {
  const int nx = M.nx, a = M.a, b = M.b, ny = M.ny, nt = M.nt, N = B.N;
  for (int k = 0; k <= N; k++) {
    const bool term = (k == N);
    const StageClass &sc = M.sc[term ? 2 : (k == 0 ? 0 : 1)];
    const int dim = term ? nx : a, np = term ? nt : ny;
    const int xoff = (N == 0) ? 0 : k * a - nx, roff = k * b, ilo = (k == 0 && !term) ? nx : 0;
    for (int i = lane; i < dim; i += G) sm.v[i] = (i >= ilo) ? xv[xoff + i] : 0.0;
    for (int r = lane; r < np; r += G) sm.w[r] = yv[roff + r];
    if (!term) {
      for (int r = lane; r < nx; r += G) {
        sm.vn[r] = xv[xoff + a + r];
        sm.td[r] = yv[roff + ny + r];
      }
    }
    __syncwarp(gm);
    for (int r = lane; r < np; r += G) rowf(roff + r, ell_dot(M, sc.C, r, sm.v));
    if (!term)
      for (int r = lane; r < nx; r += G) rowf(roff + ny + r, ell_dot(M, sc.F, r, sm.v) - sm.vn[r]);
    for (int i = lane; i < dim; i += G) {
      if (i < ilo) continue;
      const double px = sym_mv_model(M.val + sc.Q, dim, i, sm.v);
      double aty = ell_dot(M, sc.Ct, i, sm.w);
      if (!term) aty += ell_dot(M, sc.Ft, i, sm.td);
      if (k > 0 && i < nx) aty -= sm.sp[i];
      varf(xoff + i, sm.v[i], px, aty);
    }
    __syncwarp(gm);
    if (!term)
      for (int r = lane; r < nx; r += G) sm.sp[r] = sm.td[r];
    __syncwarp(gm);
  }
}

struct Resid {
  double pri, dua, nz, nAx, nq, nAty, nPx, obj;
};

// update_info (auxil.c:567-632): residuals, their normalisers and the objective — always of the UNSCALED problem
// (scaled_termination = 0: auxil.c:251, 263, 314, 327 unscale them).  Under Ruiz scaling `scaled`, when given,
// receives the same seven norms of the scaled problem (||E (Ax - z)||, ||c D (Px + q + A'y)||, ...): the rho
// estimate reads those (compute_rho_estimate works on the scaled vectors, auxil.c:13-55).
template <int G, bool WITH_SCALED = true>
__device__ Resid compute_residuals(const ModelDev &M, const BatchDev &B, int inst, int lane, unsigned gm,
                                   const Scratch &sm, Resid *scaled = nullptr) {
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  const double *q = B.q + vo, *x = B.x + vo, *z = B.z + ro, *y = B.y + ro;
  double pri = 0, nz = 0, nAx = 0, dua = 0, nq = 0, nAty = 0, nPx = 0, oq = 0, oP = 0;
  const bool ws = WITH_SCALED && scaled != nullptr && B.sD != nullptr;  // (WITH_SCALED = false: no code for it)
  const double *sD = ws ? B.sD + vo : nullptr, *sE = ws ? B.sE + ro : nullptr;
  const double sc = ws ? B.sC[inst] : 1.0;
  double s_pri = 0, s_nz = 0, s_nAx = 0, s_dua = 0, s_nq = 0, s_nAty = 0, s_nPx = 0;
  product_sweep<G>(
      M, B, lane, gm, sm, x, y,
      [&](int ri, double ax) {
        const double zz = z[ri];
        pri = fmax(pri, fabs(ax - zz));
        nz = fmax(nz, fabs(zz));
        nAx = fmax(nAx, fabs(ax));
        if (ws) {
          const double e = sE[ri];
          s_pri = fmax(s_pri, e * fabs(ax - zz)); s_nz = fmax(s_nz, e * fabs(zz)); s_nAx = fmax(s_nAx, e * fabs(ax));
        }
      },
      [&](int xi, double xval, double px, double aty) {
        const double qv = q[xi];
        dua = fmax(dua, fabs((qv + px) + aty));
        nq = fmax(nq, fabs(qv));
        nAty = fmax(nAty, fabs(aty));
        nPx = fmax(nPx, fabs(px));
        oq = fma(qv, xval, oq);
        oP = fma(xval, px, oP);
        if (ws) {
          const double w = sc * sD[xi];
          s_dua = fmax(s_dua, w * fabs((qv + px) + aty)); s_nq = fmax(s_nq, w * fabs(qv));
          s_nAty = fmax(s_nAty, w * fabs(aty)); s_nPx = fmax(s_nPx, w * fabs(px));
        }
      });
  Resid r;
  r.pri = grp_max<G>(pri, gm); r.nz = grp_max<G>(nz, gm); r.nAx = grp_max<G>(nAx, gm);
  r.dua = grp_max<G>(dua, gm); r.nq = grp_max<G>(nq, gm); r.nAty = grp_max<G>(nAty, gm); r.nPx = grp_max<G>(nPx, gm);
  r.obj = 0.5 * grp_sum<G>(oP, gm) + grp_sum<G>(oq, gm);
  if (scaled) {
    *scaled = r;
    if (ws) {
      scaled->pri = grp_max<G>(s_pri, gm); scaled->nz = grp_max<G>(s_nz, gm); scaled->nAx = grp_max<G>(s_nAx, gm);
      scaled->dua = grp_max<G>(s_dua, gm); scaled->nq = grp_max<G>(s_nq, gm); scaled->nAty = grp_max<G>(s_nAty, gm);
      scaled->nPx = grp_max<G>(s_nPx, gm);
    }
  }
  return r;
}

// auxil.c:364-427
template <int G>
__device__ int primal_infeasible(const ModelDev &M, const BatchDev &B, int inst, int lane, unsigned gm,
                                 const Scratch &sm, double eps, int m) {
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  const double *l = B.l + ro, *u = B.u + ro;
  double *dy = B.dy + ro;
  // Under Ruiz scaling the reference tests the finiteness of the SCALED bounds and ||E dy_scaled|| = c ||dy||
  // against eps; the other two conditions are invariant (the factors cancel).
  const double *sE = B.sE ? B.sE + ro : nullptr;
  const double cs = B.sE ? B.sC[inst] : 1.0;
  double nrm = 0.0;
  for (int r = lane; r < m; r += G) {
    double d = dy[r];
    const double e = sE ? sE[r] : 1.0;
    if (e * u[r] > kInfBound) {
      d = (e * l[r] < -kInfBound) ? 0.0 : fmin(d, 0.0);
    } else if (e * l[r] < -kInfBound) {
      d = fmax(d, 0.0);
    }
    dy[r] = d;
    nrm = fmax(nrm, fabs(d));
  }
  nrm = grp_max<G>(nrm, gm);
  __syncwarp(gm);
  if (!(cs * nrm > eps)) return 0;
  double lhs = 0.0;
  for (int r = lane; r < m; r += G) {
    const double d = dy[r];
    lhs += u[r] * fmax(d, 0.0) + l[r] * fmin(d, 0.0);
  }
  lhs = grp_sum<G>(lhs, gm);
  if (!(lhs < -eps * nrm)) return 0;
  double nAtdy = 0.0;
  product_sweep<G>(
      M, B, lane, gm, sm, B.dx + vo /*unused values*/, dy, [&](int, double) {},
      [&](int, double, double, double aty) { nAtdy = fmax(nAtdy, fabs(aty)); });
  nAtdy = grp_max<G>(nAtdy, gm);
  return nAtdy < eps * nrm;
}

// auxil.c:429-515
template <int G>
__device__ int dual_infeasible(const ModelDev &M, const BatchDev &B, int inst, int lane, unsigned gm,
                               const Scratch &sm, double eps, int n) {
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  const double *l = B.l + ro, *u = B.u + ro, *q = B.q + vo, *dx = B.dx + vo;
  double nrm = 0.0, qdx = 0.0;
  for (int i = lane; i < n; i += G) {
    nrm = fmax(nrm, fabs(dx[i]));
    qdx = fma(q[i], dx[i], qdx);
  }
  nrm = grp_max<G>(nrm, gm);
  qdx = grp_sum<G>(qdx, gm);
  if (!(nrm > eps)) return 0;
  if (!(qdx < -eps * nrm)) return 0;
  double nPdx = 0.0;
  int viol = 0;
  const double thr = eps * nrm;
  const double *sE = B.sE ? B.sE + ro : nullptr;  // finiteness of the scaled bounds (auxil.c:496-499)
  product_sweep<G>(
      M, B, lane, gm, sm, dx, B.dy + ro /*unused values*/,
      [&](int ri, double adx) {
        const double e = sE ? sE[ri] : 1.0;
        if ((e * u[ri] < kInfBound && adx > thr) || (e * l[ri] > -kInfBound && adx < -thr)) viol = 1;
      },
      [&](int, double, double px, double) { nPdx = fmax(nPdx, fabs(px)); });
  nPdx = grp_max<G>(nPdx, gm);
  viol = grp_or<G>(viol, gm);
  if (!(nPdx < thr)) return 0;
  return viol ? 0 : 1;
}

// check_termination (auxil.c:684-789).  Returns 1 if the instance stops; sets status/obj.
template <int G>
__device__ int check_termination(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int inst,
                                 int lane, unsigned gm, const Scratch &sm, const Resid &R, int approximate,
                                 int *status, double *obj) {
  const int n = B.n, m = B.m;
  if (R.pri > kInfty || R.dua > kInfty) {
    *status = -7;  // OSQP_NON_CVX
    *obj = kOsqpNan;
    return 1;
  }
  double ea = S.eps_abs, er = S.eps_rel, epi = S.eps_prim_inf, edi = S.eps_dual_inf;
  if (approximate) { ea *= 10; er *= 10; epi *= 10; edi *= 10; }
  int prim_ok = 0, dual_ok = 0, prim_inf = 0, dual_inf = 0;
  if (m == 0) {
    prim_ok = 1;
  } else {
    const double eps_prim = ea + er * fmax(R.nz, R.nAx);
    if (R.pri < eps_prim) prim_ok = 1;
    else prim_inf = primal_infeasible<G>(M, B, inst, lane, gm, sm, epi, m);
  }
  const double eps_dual = ea + er * fmax(fmax(R.nq, R.nAty), R.nPx);
  if (R.dua < eps_dual) dual_ok = 1;
  else dual_inf = dual_infeasible<G>(M, B, inst, lane, gm, sm, edi, n);
  if (prim_ok && dual_ok) {
    *status = approximate ? 2 : 1;
    return 1;
  }
  if (prim_inf) {
    *status = approximate ? 3 : -3;
    *obj = kInfty;
    return 1;
  }
  if (dual_inf) {
    *status = approximate ? 4 : -4;
    *obj = -kInfty;
    return 1;
  }
  return 0;
}

// time_limit (src/osqp.c:389-405).  The reference tests its run time at the top of every iteration; the batch
// kernels test at the iterations that run a termination check (every 25th when the checks are off), where the
// residuals are at hand, so a time-limited instance stops at a multiple of check_termination.
__device__ __forceinline__ bool time_is_up(const BatchDev &B, const SettingsDev &S) {
  if (S.time_limit_ns <= 0) return false;
  unsigned long long now;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
  const unsigned long long t0 = *reinterpret_cast<const volatile unsigned long long *>(B.sched + kSchedT0);
  return (long long)(now - t0) >= S.time_limit_ns;
}
__device__ __forceinline__ bool time_check_due(const SettingsDev &S, int iter, bool can_check) {
  return S.time_limit_ns > 0 && (can_check || (!S.check_termination && iter % 25 == 0));
}

__device__ __forceinline__ void trace_finish(const BatchDev &B, int inst, int engine) {
  if (!B.trace) return;
  unsigned long long now;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
  const unsigned long long t0 = *reinterpret_cast<const volatile unsigned long long *>(B.sched + kSchedT0);
  B.trace[inst] = 1e-6 * (double)(long long)(now - t0) + 1e6 * engine;
}

// store_solution (auxil.c:527-563) + the info scalars of a finished instance, written straight to the caller's pinned
// host arrays (BatchDev::hx ...) by `nthr` cooperating threads (tid = 0 .. nthr-1) that can all see the final
// iterates in global memory.  No-op unless the solve was started through rldl_batch_solve_to_host.
__device__ __forceinline__ void deliver_to_host(const BatchDev &B, int inst, int tid, int nthr, int n, int m, int status,
                                                long long iter, double obj, double pri, double dua, double rho_est) {
  if (!B.h_status) return;
  const bool has_sol = !(status == -3 || status == 3 || status == -4 || status == 4 || status == -7);
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  if (B.hx)
    for (int i = tid; i < n; i += nthr) B.hx[vo + i] = has_sol ? __ldcg(B.x + vo + i) : kOsqpNan;
  if (B.hy)
    for (int r = tid; r < m; r += nthr) B.hy[ro + r] = has_sol ? __ldcg(B.y + ro + r) : kOsqpNan;
  if (tid == 0) {
    B.h_status[inst] = status;
    if (B.h_iter) B.h_iter[inst] = iter;
    if (B.h_obj) B.h_obj[inst] = obj;
    if (B.h_pri) B.h_pri[inst] = pri;
    if (B.h_dua) B.h_dua[inst] = dua;
    if (B.h_rho_est) B.h_rho_est[inst] = rho_est;
    if (B.h_rho_updates) B.h_rho_updates[inst] = B.rho_updates[inst];
  }
}

// compute_rho_estimate (auxil.c:13-55)
__device__ __forceinline__ double rho_estimate(const Resid &R, double rho) {
  const double pr = R.pri / (fmax(R.nz, R.nAx) + 1e-10);
  const double dr = R.dua / (fmax(fmax(R.nq, R.nAty), R.nPx) + 1e-10);
  const double e = rho * sqrt(pr / (dr + 1e-10));
  return fmin(fmax(e, kRhoMin), kRhoMax);
}


}  // namespace rldl
