// rldl_kernels.cu — FP64 sm_100a kernels: batched stage-wise KKT factorisation,
// KKT solve and the fused ADMM loop, one QP instance per warp.
//
// Generic (runtime-shape) kernels.  They define the arithmetic every shape uses;
// the shape-specialised streaming kernels in rldl_stream.cu compute the same
// recurrences with the factor staged through shared memory by bulk async copies.
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
//   solve loop control flow ........ src/osqp.c:354-627
//   warm start ..................... src/osqp.c:929-994
#include <cuda_runtime.h>
#include <math.h>
#include "rldl_device.h"
#include "rldl_kernels.h"

namespace rldl {

#define FULL 0xffffffffu
static constexpr double kInfty = 1e30;
static constexpr double kMinScaling = 1e-4;
static constexpr double kInfBound = kInfty * kMinScaling;  // 1e26, auxil.c:85
static constexpr double kRhoMin = 1e-6, kRhoMax = 1e6, kRhoEq = 1e3, kRhoTol = 1e-4;
static constexpr double kOsqpNan = (double)0x7fc00000UL;  // constants.h:95 (sic)

__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(FULL, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ int warp_or(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v |= __shfl_xor_sync(FULL, v, o);
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

// ---------------------------------------------------------------------------
// In-place inversion of a dim x dim SPD matrix held in shared memory
// (Gauss-Jordan without pivoting, lane i owns column i => conflict-free).
// Returns non-zero in every lane if a pivot was not strictly positive.
// ---------------------------------------------------------------------------
__device__ int spd_invert(double *G, int dim, int lane) {
  int bad = 0;
  for (int p = 0; p < dim; p++) {
    const double piv = G[p * dim + p];
    if (!(piv > 0.0)) bad = 1;
    const double d = 1.0 / piv;
    __syncwarp();
    // scale pivot "row" (memory column p)
    for (int j = lane; j < dim; j += 32) G[j * dim + p] = (j == p) ? d : G[j * dim + p] * d;
    __syncwarp();
    for (int i = lane; i < dim; i += 32) {
      if (i == p) continue;
      const double f = G[p * dim + i];
      for (int j = 0; j < dim; j++) {
        if (j == p) continue;
        G[j * dim + i] = fma(-f, G[j * dim + p], G[j * dim + i]);
      }
      G[p * dim + i] = -f * d;
    }
    __syncwarp();
  }
  return bad;
}

// ---------------------------------------------------------------------------
// Stage recursion for one instance, stages k0..N (k0 > 0 restarts from the stored
// Sinv_{k0-1}: the horizon-change path, reference src/recursive_ldl.c:946-1110).
// smem: G[a*a], W[a*nx], Sm[nx*nx], rp[max(ny,nt,nx)]
// ---------------------------------------------------------------------------
__device__ int factor_instance(const ModelDev &M, const BatchDev &B, double sigma, int inst, int lane,
                               int k0, double *G, double *W, double *Sm, double *rp) {
  const int nx = M.nx, a = M.a, b = M.b, ny = M.ny, nt = M.nt, N = B.N;
  const double *rv = B.rho_vec + (size_t)inst * B.mmax;
  double *Xi = B.Xinv + (size_t)inst * B.xstride;
  double *Si = B.Sinv + (size_t)inst * B.sstride;
  int flag = 0;
  if (k0 > 0) {
    for (int e = lane; e < nx * nx; e += 32) Sm[e] = Si[(size_t)(k0 - 1) * nx * nx + e];
    __syncwarp();
  }
  for (int k = k0; k <= N; k++) {
    const bool term = (k == N);
    const StageClass &sc = M.sc[term ? 2 : (k == 0 ? 0 : 1)];
    const int dim = term ? nx : a;
    const int np = term ? nt : ny;
    const int roff = k * b;
    for (int r = lane; r < np; r += 32) rp[r] = rv[roff + r];
    // G = Q + sigma I
    for (int e = lane; e < dim * dim; e += 32) {
      const int i = e / dim, j = e - i * dim;
      G[e] = __ldg(M.val + sc.Q + e) + (i == j ? sigma : 0.0);
    }
    __syncwarp();
    // G += C' diag(rho_p) C   (lane i owns memory column i; symmetric update)
    for (int i = lane; i < dim; i += 32) {
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
          if (c2 != 0.0) G[__ldg(ci + e2) * dim + i] += wgt * c2;
        }
      }
      if (k > 0 && i < nx)
        for (int j = 0; j < nx; j++) G[j * dim + i] += Sm[j * nx + i];
    }
    __syncwarp();
    if (spd_invert(G, dim, lane) && !flag) flag = 1 + k;
    double *Xk = Xi + (size_t)k * a * a;
    for (int e = lane; e < dim * dim; e += 32) Xk[e] = G[e];
    if (term) break;
    // W = Xinv F'  (a x nx, W[i*nx + r]);  S = diag(1/rho_d) + F W
    for (int r = lane; r < nx; r += 32) rp[r] = rv[roff + ny + r];
    __syncwarp();
    for (int i = lane; i < a; i += 32) {
      for (int r = 0; r < nx; r++) {
        const int *fi = M.idx + sc.F.off + r * sc.F.width;
        const double *fv = M.val + sc.F.off + r * sc.F.width;
        double s = 0.0;
        for (int e = 0; e < sc.F.width; e++) s = fma(__ldg(fv + e), G[__ldg(fi + e) * a + i], s);
        W[r * a + i] = s;  // stored transposed: W[r][i], lanes contiguous
      }
    }
    __syncwarp();
    for (int c = lane; c < nx; c += 32) {
      for (int r = 0; r < nx; r++) {
        const int *fi = M.idx + sc.F.off + r * sc.F.width;
        const double *fv = M.val + sc.F.off + r * sc.F.width;
        double s = (r == c) ? 1.0 / rp[r] : 0.0;
        for (int e = 0; e < sc.F.width; e++) s = fma(__ldg(fv + e), W[c * a + __ldg(fi + e)], s);
        Sm[r * nx + c] = s;
      }
    }
    __syncwarp();
    if (spd_invert(Sm, nx, lane) && !flag) flag = 1 + k;
    double *Sk = Si + (size_t)k * nx * nx;
    for (int e = lane; e < nx * nx; e += 32) Sk[e] = Sm[e];
    __syncwarp();
  }
  __syncwarp();
  return flag;
}

__global__ void k_set_rho_vec(BatchDev B, int m, int *changed) {
  // one warp per instance; rewrites rho_vec rows [0,mmax) from (l,u,rho) and flags
  // instances whose active rows [0,m) changed (auxil.c:103-145 semantics).
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B.batch) return;
  const size_t off = (size_t)warp * B.mmax;
  double rho = fmin(fmax(B.rho[warp], kRhoMin), kRhoMax);
  int ch = 0;
  for (int r = lane; r < B.mmax; r += 32) {
    const double nv = row_rho(B.l[off + r], B.u[off + r], rho);
    if (r < m && nv != B.rho_vec[off + r]) ch = 1;
    B.rho_vec[off + r] = nv;
  }
  ch = warp_or(ch);
  if (lane == 0) {
    B.rho[warp] = rho;
    if (changed) changed[warp] = ch;
  }
}

size_t factor_smem_doubles(int a, int nx, int ny, int nt) {
  int vm = ny > nt ? ny : nt;
  if (nx > vm) vm = nx;
  return (size_t)a * a + (size_t)a * nx + (size_t)nx * nx + vm;
}

__global__ void k_factor(ModelDev M, BatchDev B, double sigma, int k0, const int *only_flagged) {
  extern __shared__ double smem[];
  const int wpb = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int inst = blockIdx.x * wpb + wid;
  if (inst >= B.batch) return;
  if (only_flagged && !only_flagged[inst]) return;
  int vm = M.ny > M.nt ? M.ny : M.nt;
  if (M.nx > vm) vm = M.nx;
  const size_t per = (size_t)M.a * M.a + (size_t)M.a * M.nx + (size_t)M.nx * M.nx + vm;
  double *G = smem + per * wid, *W = G + M.a * M.a, *Sm = W + M.a * M.nx, *rp = Sm + M.nx * M.nx;
  const int flag = factor_instance(M, B, sigma, inst, lane, k0, G, W, Sm, rp);
  if (lane == 0) B.fact_flag[inst] = flag;
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
template <bool KKT>
__device__ void kkt_sweep(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int inst,
                          int lane, const Scratch &sm, double *rb, bool store_delta) {
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
    for (int i = lane; i < dim; i += 32) {
      double gv = 0.0;
      if (i >= ilo) gv = KKT ? rb[xoff + i] : sigma * x[xoff + i] - q[xoff + i];
      sm.g[i] = gv;
    }
    for (int r = lane; r < np; r += 32) {
      const double rho = rv[roff + r];
      sm.w[r] = KKT ? rho * rb[n + roff + r] : rho * z[roff + r] - y[roff + r];
    }
    __syncwarp();
    for (int i = lane; i < dim; i += 32) {
      double t = sm.g[i] + ell_dot(M, sc.Ct, i, sm.w);
      if (k > 0 && i < nx) t -= sm.sp[i];
      sm.v[i] = t;
    }
    __syncwarp();
    const double *Xk = Xi + (size_t)k * a * a;
    for (int i = lane; i < dim; i += 32) {
      const double yv = sym_mv(Xk, dim, i, sm.v);
      sm.yv[i] = yv;
      ykg[k * a + i] = yv;
    }
    __syncwarp();
    if (!term) {
      for (int r = lane; r < nx; r += 32) {
        const int ri = roff + ny + r;
        const double rhs = KKT ? rb[n + ri] : z[ri] - y[ri] / rv[ri];
        sm.td[r] = rhs - ell_dot(M, sc.F, r, sm.yv);
      }
      __syncwarp();
      const double *Sk = Si + (size_t)k * nx * nx;
      for (int r = lane; r < nx; r += 32) {
        const double s = sym_mv(Sk, nx, r, sm.td);
        sm.sp[r] = s;
        skg[k * nx + r] = s;
      }
      __syncwarp();
    }
  }
  // ------------------------------ backward -----------------------------
  for (int k = N; k >= 0; k--) {
    const bool term = (k == N);
    const StageClass &sc = M.sc[term ? 2 : (k == 0 ? 0 : 1)];
    const int dim = term ? nx : a, np = term ? nt : ny;
    const int xoff = (N == 0) ? 0 : k * a - nx, roff = k * b, ilo = (k == 0 && !term) ? nx : 0;
    if (term) {
      for (int i = lane; i < dim; i += 32) sm.v[i] = sm.yv[i];
      __syncwarp();
    } else {
      const double *Sk = Si + (size_t)k * nx * nx;
      for (int r = lane; r < nx; r += 32) sm.nud[r] = -(skg[k * nx + r] + sym_mv(Sk, nx, r, sm.vn));
      __syncwarp();
      for (int i = lane; i < dim; i += 32) sm.g[i] = ell_dot(M, sc.Ft, i, sm.nud);
      __syncwarp();
      const double *Xk = Xi + (size_t)k * a * a;
      for (int i = lane; i < dim; i += 32) sm.v[i] = ykg[k * a + i] - sym_mv(Xk, dim, i, sm.g);
      __syncwarp();
    }
    // path rows: z~ = C v
    for (int r = lane; r < np; r += 32) {
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
      for (int r = lane; r < nx; r += 32) {
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
    for (int i = lane; i < dim; i += 32) {
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
    __syncwarp();
    for (int i = lane; i < nx; i += 32) sm.vn[i] = sm.v[i];
    __syncwarp();
  }
}

// Stage sweep computing A*xv (per row), P*xv and A'*yv (per variable) and handing
// them to the callbacks.  rowf(ri, ax), varf(xi, xval, px, aty).
template <class RowF, class VarF>
__device__ void product_sweep(const ModelDev &M, const BatchDev &B, int lane, const Scratch &sm,
                              const double *xv, const double *yv, RowF rowf, VarF varf) {
  const int nx = M.nx, a = M.a, b = M.b, ny = M.ny, nt = M.nt, N = B.N;
  for (int k = 0; k <= N; k++) {
    const bool term = (k == N);
    const StageClass &sc = M.sc[term ? 2 : (k == 0 ? 0 : 1)];
    const int dim = term ? nx : a, np = term ? nt : ny;
    const int xoff = (N == 0) ? 0 : k * a - nx, roff = k * b, ilo = (k == 0 && !term) ? nx : 0;
    for (int i = lane; i < dim; i += 32) sm.v[i] = (i >= ilo) ? xv[xoff + i] : 0.0;
    for (int r = lane; r < np; r += 32) sm.w[r] = yv[roff + r];
    if (!term) {
      for (int r = lane; r < nx; r += 32) {
        sm.vn[r] = xv[xoff + a + r];
        sm.td[r] = yv[roff + ny + r];
      }
    }
    __syncwarp();
    for (int r = lane; r < np; r += 32) rowf(roff + r, ell_dot(M, sc.C, r, sm.v));
    if (!term)
      for (int r = lane; r < nx; r += 32) rowf(roff + ny + r, ell_dot(M, sc.F, r, sm.v) - sm.vn[r]);
    for (int i = lane; i < dim; i += 32) {
      if (i < ilo) continue;
      const double px = sym_mv_model(M.val + sc.Q, dim, i, sm.v);
      double aty = ell_dot(M, sc.Ct, i, sm.w);
      if (!term) aty += ell_dot(M, sc.Ft, i, sm.td);
      if (k > 0 && i < nx) aty -= sm.sp[i];
      varf(xoff + i, sm.v[i], px, aty);
    }
    __syncwarp();
    if (!term)
      for (int r = lane; r < nx; r += 32) sm.sp[r] = sm.td[r];
    __syncwarp();
  }
}

struct Resid {
  double pri, dua, nz, nAx, nq, nAty, nPx, obj;
};

// update_info (auxil.c:567-632): residuals, their normalisers and the objective.
__device__ Resid compute_residuals(const ModelDev &M, const BatchDev &B, int inst, int lane,
                                   const Scratch &sm) {
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  const double *q = B.q + vo, *x = B.x + vo, *z = B.z + ro, *y = B.y + ro;
  double pri = 0, nz = 0, nAx = 0, dua = 0, nq = 0, nAty = 0, nPx = 0, oq = 0, oP = 0;
  product_sweep(
      M, B, lane, sm, x, y,
      [&](int ri, double ax) {
        const double zz = z[ri];
        pri = fmax(pri, fabs(ax - zz));
        nz = fmax(nz, fabs(zz));
        nAx = fmax(nAx, fabs(ax));
      },
      [&](int xi, double xval, double px, double aty) {
        const double qv = q[xi];
        dua = fmax(dua, fabs((qv + px) + aty));
        nq = fmax(nq, fabs(qv));
        nAty = fmax(nAty, fabs(aty));
        nPx = fmax(nPx, fabs(px));
        oq = fma(qv, xval, oq);
        oP = fma(xval, px, oP);
      });
  Resid r;
  r.pri = warp_max(pri); r.nz = warp_max(nz); r.nAx = warp_max(nAx);
  r.dua = warp_max(dua); r.nq = warp_max(nq); r.nAty = warp_max(nAty); r.nPx = warp_max(nPx);
  r.obj = 0.5 * warp_sum(oP) + warp_sum(oq);
  return r;
}

// auxil.c:364-427
__device__ int primal_infeasible(const ModelDev &M, const BatchDev &B, int inst, int lane,
                                 const Scratch &sm, double eps, int m) {
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  const double *l = B.l + ro, *u = B.u + ro;
  double *dy = B.dy + ro;
  double nrm = 0.0;
  for (int r = lane; r < m; r += 32) {
    double d = dy[r];
    if (u[r] > kInfBound) {
      d = (l[r] < -kInfBound) ? 0.0 : fmin(d, 0.0);
    } else if (l[r] < -kInfBound) {
      d = fmax(d, 0.0);
    }
    dy[r] = d;
    nrm = fmax(nrm, fabs(d));
  }
  nrm = warp_max(nrm);
  __syncwarp();
  if (!(nrm > eps)) return 0;
  double lhs = 0.0;
  for (int r = lane; r < m; r += 32) {
    const double d = dy[r];
    lhs += u[r] * fmax(d, 0.0) + l[r] * fmin(d, 0.0);
  }
  lhs = warp_sum(lhs);
  if (!(lhs < -eps * nrm)) return 0;
  double nAtdy = 0.0;
  product_sweep(
      M, B, lane, sm, B.dx + vo /*unused values*/, dy, [&](int, double) {},
      [&](int, double, double, double aty) { nAtdy = fmax(nAtdy, fabs(aty)); });
  nAtdy = warp_max(nAtdy);
  return nAtdy < eps * nrm;
}

// auxil.c:429-515
__device__ int dual_infeasible(const ModelDev &M, const BatchDev &B, int inst, int lane,
                               const Scratch &sm, double eps, int n) {
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  const double *l = B.l + ro, *u = B.u + ro, *q = B.q + vo, *dx = B.dx + vo;
  double nrm = 0.0, qdx = 0.0;
  for (int i = lane; i < n; i += 32) {
    nrm = fmax(nrm, fabs(dx[i]));
    qdx = fma(q[i], dx[i], qdx);
  }
  nrm = warp_max(nrm);
  qdx = warp_sum(qdx);
  if (!(nrm > eps)) return 0;
  if (!(qdx < -eps * nrm)) return 0;
  double nPdx = 0.0;
  int viol = 0;
  const double thr = eps * nrm;
  product_sweep(
      M, B, lane, sm, dx, B.dy + ro /*unused values*/,
      [&](int ri, double adx) {
        if ((u[ri] < kInfBound && adx > thr) || (l[ri] > -kInfBound && adx < -thr)) viol = 1;
      },
      [&](int, double, double px, double) { nPdx = fmax(nPdx, fabs(px)); });
  nPdx = warp_max(nPdx);
  viol = warp_or(viol);
  if (!(nPdx < thr)) return 0;
  return viol ? 0 : 1;
}

// check_termination (auxil.c:684-789).  Returns 1 if the instance stops; sets status/obj.
__device__ int check_termination(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int inst,
                                 int lane, const Scratch &sm, const Resid &R, int approximate,
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
    else prim_inf = primal_infeasible(M, B, inst, lane, sm, epi, m);
  }
  const double eps_dual = ea + er * fmax(fmax(R.nq, R.nAty), R.nPx);
  if (R.dua < eps_dual) dual_ok = 1;
  else dual_inf = dual_infeasible(M, B, inst, lane, sm, edi, n);
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

// compute_rho_estimate (auxil.c:13-55)
__device__ __forceinline__ double rho_estimate(const Resid &R, double rho) {
  const double pr = R.pri / (fmax(R.nz, R.nAx) + 1e-10);
  const double dr = R.dua / (fmax(fmax(R.nq, R.nAty), R.nPx) + 1e-10);
  const double e = rho * sqrt(pr / (dr + 1e-10));
  return fmin(fmax(e, kRhoMin), kRhoMax);
}

size_t admm_smem_doubles(int a, int nx, int ny, int nt) {
  int vm = a;
  if (ny > vm) vm = ny;
  if (nt > vm) vm = nt;
  if (nx > vm) vm = nx;
  return 8 * (size_t)vm + factor_smem_doubles(a, nx, ny, nt);
}

// The whole osqp_solve loop (osqp.c:354-627) for one instance per warp.
__global__ void k_admm(ModelDev M, BatchDev B, SettingsDev S, int warm_start) {
  extern __shared__ double smem[];
  const int wpb = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int inst = blockIdx.x * wpb + wid;
  if (inst >= B.batch) return;
  int vm = M.a;
  if (M.ny > vm) vm = M.ny;
  if (M.nt > vm) vm = M.nt;
  if (M.nx > vm) vm = M.nx;
  int fvm = M.ny > M.nt ? M.ny : M.nt;
  if (M.nx > fvm) fvm = M.nx;
  const size_t per = 8 * (size_t)vm + (size_t)M.a * M.a + (size_t)M.a * M.nx + (size_t)M.nx * M.nx + fvm;
  double *base = smem + per * wid;
  const Scratch sm = make_scratch(base, vm);
  double *G = base + 8 * vm, *W = G + M.a * M.a, *Sm = W + M.a * M.nx, *rp = Sm + M.nx * M.nx;

  const int n = B.n, m = B.m;
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  if (B.fact_flag[inst]) {  // failed factorisation: report non-convex, like a diverged solve
    if (lane == 0) { B.status[inst] = -7; B.iter[inst] = 0; B.obj[inst] = kOsqpNan; }
    return;
  }
  if (!warm_start) {
    for (int i = lane; i < n; i += 32) B.x[vo + i] = 0.0;
    for (int r = lane; r < m; r += 32) { B.z[ro + r] = 0.0; B.y[ro + r] = 0.0; }
    __syncwarp();
  }
  double rho = B.rho[inst];
  int status = -10, rho_updates = 0, can_check = 0, iter;
  double obj = 0.0, rho_est = rho;
  Resid R = {0, 0, 0, 0, 0, 0, 0, 0};
  for (iter = 1; iter <= S.max_iter; iter++) {
    can_check = S.check_termination && (iter % S.check_termination == 0);
    const bool store_delta = can_check || iter == S.max_iter;
    kkt_sweep<false>(M, B, S, inst, lane, sm, nullptr, store_delta);
    __syncwarp();
    if (can_check) {
      R = compute_residuals(M, B, inst, lane, sm);
      obj = R.obj;
      if (check_termination(M, B, S, inst, lane, sm, R, 0, &status, &obj)) break;
    }
    if (S.adaptive_rho && S.adaptive_rho_interval && (iter % S.adaptive_rho_interval == 0)) {
      if (!can_check) { R = compute_residuals(M, B, inst, lane, sm); obj = R.obj; }
      const double rn = rho_estimate(R, rho);
      rho_est = rn;
      if (rn > rho * S.adaptive_rho_tolerance || rn < rho / S.adaptive_rho_tolerance) {
        // osqp_update_rho (osqp.c:1268-1319): retype-free rho_vec rewrite + full refactor
        rho = fmin(fmax(rn, kRhoMin), kRhoMax);
        for (int r = lane; r < B.mmax; r += 32)
          B.rho_vec[ro + r] = row_rho(B.l[ro + r], B.u[ro + r], rho);
        __syncwarp();
        const int flag = factor_instance(M, B, S.sigma, inst, lane, 0, G, W, Sm, rp);
        rho_updates++;
        if (flag) { status = -7; obj = kOsqpNan; if (lane == 0) B.fact_flag[inst] = flag; break; }
      }
    }
  }
  int it_done = (iter > S.max_iter) ? S.max_iter : iter;
  if (status == -10) {
    if (!can_check) {  // osqp.c:524-546
      R = compute_residuals(M, B, inst, lane, sm);
      obj = R.obj;
      check_termination(M, B, S, inst, lane, sm, R, 0, &status, &obj);
    }
    if (status == -10) {  // osqp.c:562-567
      if (!check_termination(M, B, S, inst, lane, sm, R, 1, &status, &obj)) status = -2;
    }
  }
  const bool has_sol = !(status == -3 || status == 3 || status == -4 || status == 4 || status == -7);
  if (has_sol && status != -10) obj = R.obj;
  rho_est = rho_estimate(R, rho);
  if (!has_sol) {
    // store_solution (auxil.c:527-563): normalise the certificate, cold-start the iterates
    if (status == -3 || status == 3) {
      double nrm = 0.0;
      for (int r = lane; r < m; r += 32) nrm = fmax(nrm, fabs(B.dy[ro + r]));
      nrm = warp_max(nrm);
      for (int r = lane; r < m; r += 32) B.dy[ro + r] *= 1.0 / nrm;
    }
    if (status == -4 || status == 4) {
      double nrm = 0.0;
      for (int i = lane; i < n; i += 32) nrm = fmax(nrm, fabs(B.dx[vo + i]));
      nrm = warp_max(nrm);
      for (int i = lane; i < n; i += 32) B.dx[vo + i] *= 1.0 / nrm;
    }
    for (int i = lane; i < n; i += 32) B.x[vo + i] = 0.0;
    for (int r = lane; r < m; r += 32) { B.z[ro + r] = 0.0; B.y[ro + r] = 0.0; }
  }
  if (lane == 0) {
    B.iter[inst] = it_done;
    B.status[inst] = status;
    B.obj[inst] = obj;
    B.pri[inst] = R.pri;
    B.dua[inst] = R.dua;
    B.rho_est[inst] = rho_est;
    B.rho_updates[inst] += rho_updates;
    B.rho[inst] = rho;
  }
}

__global__ void k_kkt_solve(ModelDev M, BatchDev B, SettingsDev S, double *rhs, int ld) {
  extern __shared__ double smem[];
  const int wpb = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int inst = blockIdx.x * wpb + wid;
  if (inst >= B.batch) return;
  int vm = M.a;
  if (M.ny > vm) vm = M.ny;
  if (M.nt > vm) vm = M.nt;
  if (M.nx > vm) vm = M.nx;
  const Scratch sm = make_scratch(smem + 8 * (size_t)vm * wid, vm);
  kkt_sweep<true>(M, B, S, inst, lane, sm, rhs + (size_t)inst * ld, false);
}

// osqp_warm_start (osqp.c:929-994): x,y already copied to the device; z = A x.
__global__ void k_warm_start_z(ModelDev M, BatchDev B) {
  extern __shared__ double smem[];
  const int wpb = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int inst = blockIdx.x * wpb + wid;
  if (inst >= B.batch) return;
  int vm = M.a;
  if (M.ny > vm) vm = M.ny;
  if (M.nt > vm) vm = M.nt;
  if (M.nx > vm) vm = M.nx;
  const Scratch sm = make_scratch(smem + 8 * (size_t)vm * wid, vm);
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  double *z = B.z + ro;
  product_sweep(
      M, B, lane, sm, B.x + vo, B.y + ro, [&](int ri, double ax) { z[ri] = ax; },
      [&](int, double, double, double) {});
}

__global__ void k_fill(double *p, double v, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t st = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) p[i] = v;
}

// strided scatter of a [batch][len] host-staged block into rows [start,start+len)
__global__ void k_scatter_rows(double *dst, int ld, const double *src, int len, int start, int batch) {
  const size_t tot = (size_t)batch * len;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t st = (size_t)gridDim.x * blockDim.x;
  for (; i < tot; i += st) {
    const size_t b = i / len, c = i - b * len;
    dst[b * ld + start + c] = src[i];
  }
}

// ------------------------------- launchers --------------------------------
static inline int pick_wpb(size_t per_warp_bytes) {
  int wpb = 4;
  while (wpb > 1 && per_warp_bytes * wpb > 200 * 1024) wpb >>= 1;
  return wpb;
}

cudaError_t launch_set_rho_vec(const BatchDev &B, int m, int *changed, cudaStream_t st) {
  const int threads = 128, wpb = threads / 32;
  k_set_rho_vec<<<(B.batch + wpb - 1) / wpb, threads, 0, st>>>(B, m, changed);
  return cudaGetLastError();
}

cudaError_t launch_factor(const ModelDev &M, const BatchDev &B, double sigma, int k0,
                          const int *only_flagged, cudaStream_t st) {
  const size_t per = factor_smem_doubles(M.a, M.nx, M.ny, M.nt) * sizeof(double);
  if (per > 220 * 1024) return cudaErrorInvalidValue;
  const int wpb = pick_wpb(per);
  const size_t shm = per * wpb;
  cudaError_t e = cudaFuncSetAttribute(k_factor, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm);
  if (e != cudaSuccess) return e;
  k_factor<<<(B.batch + wpb - 1) / wpb, wpb * 32, shm, st>>>(M, B, sigma, k0, only_flagged);
  return cudaGetLastError();
}

cudaError_t launch_admm(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                        cudaStream_t st) {
  const size_t per = admm_smem_doubles(M.a, M.nx, M.ny, M.nt) * sizeof(double);
  if (per > 220 * 1024) return cudaErrorInvalidValue;
  int wpb = pick_wpb(per);
  if (wpb > 2) wpb = 2;
  const size_t shm = per * wpb;
  cudaError_t e = cudaFuncSetAttribute(k_admm, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm);
  if (e != cudaSuccess) return e;
  k_admm<<<(B.batch + wpb - 1) / wpb, wpb * 32, shm, st>>>(M, B, S, warm_start);
  return cudaGetLastError();
}

cudaError_t launch_kkt_solve(const ModelDev &M, const BatchDev &B, const SettingsDev &S, double *rhs,
                             int ld, cudaStream_t st) {
  int vm = M.a;
  if (M.ny > vm) vm = M.ny;
  if (M.nt > vm) vm = M.nt;
  if (M.nx > vm) vm = M.nx;
  const size_t per = 8 * (size_t)vm * sizeof(double);
  const int wpb = pick_wpb(per);
  const size_t shm = per * wpb;
  cudaError_t e = cudaFuncSetAttribute(k_kkt_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm);
  if (e != cudaSuccess) return e;
  k_kkt_solve<<<(B.batch + wpb - 1) / wpb, wpb * 32, shm, st>>>(M, B, S, rhs, ld);
  return cudaGetLastError();
}

cudaError_t launch_warm_start_z(const ModelDev &M, const BatchDev &B, cudaStream_t st) {
  int vm = M.a;
  if (M.ny > vm) vm = M.ny;
  if (M.nt > vm) vm = M.nt;
  if (M.nx > vm) vm = M.nx;
  const size_t per = 8 * (size_t)vm * sizeof(double);
  const int wpb = pick_wpb(per);
  const size_t shm = per * wpb;
  cudaError_t e = cudaFuncSetAttribute(k_warm_start_z, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm);
  if (e != cudaSuccess) return e;
  k_warm_start_z<<<(B.batch + wpb - 1) / wpb, wpb * 32, shm, st>>>(M, B);
  return cudaGetLastError();
}

cudaError_t launch_fill(double *p, double v, size_t n, cudaStream_t st) {
  if (!n) return cudaSuccess;
  size_t blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_fill<<<(unsigned)blocks, 256, 0, st>>>(p, v, n);
  return cudaGetLastError();
}

cudaError_t launch_scatter_rows(double *dst, int ld, const double *src, int len, int start, int batch,
                                cudaStream_t st) {
  const size_t tot = (size_t)batch * len;
  if (!tot) return cudaSuccess;
  size_t blocks = (tot + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_scatter_rows<<<(unsigned)blocks, 256, 0, st>>>(dst, ld, src, len, start, batch);
  return cudaGetLastError();
}

}  // namespace rldl
