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
#include "rldl_device_fn.cuh"
#include "rldl_kernels.h"

namespace rldl {

__global__ void k_set_rho_vec(BatchDev B, int m, int *changed) {
  // one warp per instance; rewrites rho_vec rows [0,mmax) from (l,u,rho) and flags
  // instances whose active rows [0,m) changed (auxil.c:103-145 semantics).
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B.batch) return;
  const size_t off = (size_t)warp * B.mmax;
  double rho = fmin(fmax(B.rho[warp], kRhoMin), kRhoMax);
  int ch = 0;
  for (int r = lane; r < B.mmax; r += 32) {
    const double nv = row_rho_eff(B, warp, off + r, rho);
    if (r < m && nv != B.rho_vec[off + r]) ch = 1;
    B.rho_vec[off + r] = nv;
  }
  ch = grp_or<32>(ch, FULL);
  if (lane == 0) {
    B.rho[warp] = rho;
    if (changed) changed[warp] = ch;
    if (B.status[warp] == kStatusParkedRho) B.status[warp] = -10;  // parked for this refresh: resumable again
  }
}

// ---------------------------------------------------------------------------
// Ruiz equilibration (reference src/scaling.c:44-156, scale_data), one warp per instance, on the stage
// structure: `iters` rounds of  D <- D / sqrt(||KKT column||_inf), E <- E / sqrt(||KKT row||_inf)  followed by the
// cost normalisation c <- c / max(mean_j ||P column j||_inf, ||q||_inf).  The scaled matrices are never formed:
// entry (i,j) of the current scaled P is c D_i P_ij D_j, of the scaled A it is E_r A_rj D_j.  Norms below 1e-4
// count as 1, above 1e4 as 1e4 (limit_scaling, scaling.c:7-14).  dx / dy serve as the D_temp / E_temp of the
// reference (free at setup and update time).  Rows and variables beyond the current horizon keep scaling 1.
// ---------------------------------------------------------------------------
__device__ __forceinline__ double limit_scaling(double v) {
  v = v < kMinScaling ? 1.0 : v;
  return v > kMaxScaling ? kMaxScaling : v;
}

__global__ void k_ruiz(ModelDev M, BatchDev B, int iters) {
  const int inst = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (inst >= B.batch) return;
  const int nx = M.nx, a = M.a, b = M.b, ny = M.ny, nt = M.nt, N = B.N;
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  double *D = B.sD + vo, *E = B.sE + ro, *dD = B.dx + vo, *dE = B.dy + ro;
  const double *q = B.q + vo;
  for (int i = lane; i < B.nmax; i += 32) D[i] = 1.0;
  for (int r = lane; r < B.mmax; r += 32) E[r] = 1.0;
  double c = 1.0;
  __syncwarp();
  // ||column i of the stage cost block||_inf weighted by D (without the c D_i factor)
  auto pcol = [&](const StageClass &sc, int dim, int ilo, int xoff, int i) {
    double pn = 0.0;
    for (int j = ilo; j < dim; j++) pn = fmax(pn, fabs(__ldg(M.val + sc.Q + j * dim + i)) * D[xoff + j]);
    return pn;
  };
  for (int t = 0; t < iters; t++) {
    for (int k = 0; k <= N; k++) {
      const bool term = (k == N);
      const StageClass &sc = M.sc[term ? 2 : (k == 0 ? 0 : 1)];
      const int dim = term ? nx : a, np = term ? nt : ny;
      const int xoff = (N == 0) ? 0 : k * a - nx, roff = k * b, ilo = (k == 0 && !term) ? nx : 0;
      for (int i = lane; i < dim; i += 32) {
        if (i < ilo) continue;
        const double pn = c * pcol(sc, dim, ilo, xoff, i);
        double an = 0.0;
        for (int e = 0; e < sc.Ct.width; e++)
          an = fmax(an, fabs(__ldg(M.val + sc.Ct.off + i * sc.Ct.width + e)) * E[roff + __ldg(M.idx + sc.Ct.off + i * sc.Ct.width + e)]);
        if (!term)
          for (int e = 0; e < sc.Ft.width; e++)
            an = fmax(an, fabs(__ldg(M.val + sc.Ft.off + i * sc.Ft.width + e)) *
                              E[roff + ny + __ldg(M.idx + sc.Ft.off + i * sc.Ft.width + e)]);
        if (k > 0 && i < nx) an = fmax(an, E[roff - b + ny + i]);  // the -I of stage k-1's dynamics rows
        dD[xoff + i] = 1.0 / sqrt(limit_scaling(D[xoff + i] * fmax(pn, an)));
      }
      for (int r = lane; r < np; r += 32) {
        double rn = 0.0;
        for (int e = 0; e < sc.C.width; e++) {
          const int j = __ldg(M.idx + sc.C.off + r * sc.C.width + e);
          if (j >= ilo) rn = fmax(rn, fabs(__ldg(M.val + sc.C.off + r * sc.C.width + e)) * D[xoff + j]);
        }
        dE[roff + r] = 1.0 / sqrt(limit_scaling(E[roff + r] * rn));
      }
      if (!term)
        for (int r = lane; r < nx; r += 32) {
          double rn = D[xoff + a + r];  // the -I on x_{k+1}
          for (int e = 0; e < sc.F.width; e++) {
            const int j = __ldg(M.idx + sc.F.off + r * sc.F.width + e);
            if (j >= ilo) rn = fmax(rn, fabs(__ldg(M.val + sc.F.off + r * sc.F.width + e)) * D[xoff + j]);
          }
          dE[roff + ny + r] = 1.0 / sqrt(limit_scaling(E[roff + ny + r] * rn));
        }
    }
    __syncwarp();
    for (int i = lane; i < B.n; i += 32) D[i] *= dD[i];
    for (int r = lane; r < B.m; r += 32) E[r] *= dE[r];
    __syncwarp();
    // cost normalisation (scaling.c:111-138) with the new D and the c of the earlier rounds
    double sum = 0.0, qn = 0.0;
    for (int k = 0; k <= N; k++) {
      const bool term = (k == N);
      const StageClass &sc = M.sc[term ? 2 : (k == 0 ? 0 : 1)];
      const int dim = term ? nx : a;
      const int xoff = (N == 0) ? 0 : k * a - nx, ilo = (k == 0 && !term) ? nx : 0;
      for (int i = lane; i < dim; i += 32) {
        if (i < ilo) continue;
        sum += c * D[xoff + i] * pcol(sc, dim, ilo, xoff, i);
        qn = fmax(qn, c * fabs(D[xoff + i] * q[xoff + i]));
      }
    }
    sum = grp_sum<32>(sum, FULL);
    qn = grp_max<32>(qn, FULL);
    const double ct = limit_scaling(fmax(sum / (double)B.n, limit_scaling(qn)));
    c *= 1.0 / ct;
    __syncwarp();
  }
  if (lane == 0) B.sC[inst] = c;
}

// The iterates between the unscaled coordinates the kernels work in and the scaled coordinates of the current
// equilibration (x_s = D^-1 x, z_s = E z, y_s = c E^-1 y; scaling.c:177-192, osqp.c:942-945).  The reference keeps
// its iterates in scaled coordinates and leaves them numerically untouched when osqp_update_P / _A re-runs
// scale_data (osqp.c:1033-1051), i.e. a warm start after a matrix update continues from D_new (D_old^-1 x): the
// same is done here by going to scaled coordinates before the new equilibration and back after it.
__global__ void k_rescale_iterates(BatchDev B, int to_scaled) {
  const int inst = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (inst >= B.batch) return;
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  const double c = B.sC[inst];
  for (int i = lane; i < B.n; i += 32) {
    const double d = B.sD[vo + i];
    B.x[vo + i] = to_scaled ? B.x[vo + i] / d : B.x[vo + i] * d;
  }
  for (int r = lane; r < B.m; r += 32) {
    const double e = B.sE[ro + r];
    B.z[ro + r] = to_scaled ? B.z[ro + r] * e : B.z[ro + r] / e;
    B.y[ro + r] = to_scaled ? c * B.y[ro + r] / e : e * B.y[ro + r] / c;
  }
}
cudaError_t launch_rescale_iterates(const BatchDev &B, int to_scaled, cudaStream_t st) {
  const int threads = 128, wpb = threads / 32;
  k_rescale_iterates<<<(B.batch + wpb - 1) / wpb, threads, 0, st>>>(B, to_scaled);
  return cudaGetLastError();
}

cudaError_t launch_ruiz(const ModelDev &M, const BatchDev &B, int iters, cudaStream_t st) {
  const int threads = 128, wpb = threads / 32;
  k_ruiz<<<(B.batch + wpb - 1) / wpb, threads, 0, st>>>(M, B, iters);
  return cudaGetLastError();
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
  const int flag = factor_instance<32>(M, B, sigma, inst, lane, FULL, k0, G, W, Sm, rp);
  if (lane == 0) B.fact_flag[inst] = flag;
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
    deliver_to_host(B, inst, lane, 32, n, m, -7, 0, kOsqpNan, 0.0, 0.0, 0.0);
    return;
  }
  if (!warm_start) {
    for (int i = lane; i < n; i += 32) B.x[vo + i] = 0.0;
    for (int r = lane; r < m; r += 32) { B.z[ro + r] = 0.0; B.y[ro + r] = 0.0; }
    __syncwarp();
  }
  double rho = B.rho[inst];
  int status = -10, rho_updates = 0, can_check = 0, iter;
  bool timed_out = false;
  double obj = 0.0, rho_est = rho;
  Resid R = {0, 0, 0, 0, 0, 0, 0, 0}, Rs = R;  // Rs: the norms of the scaled problem (rho estimate under Ruiz scaling)
  for (iter = 1; iter <= S.max_iter; iter++) {
    can_check = S.check_termination && (iter % S.check_termination == 0);
    const bool time_due = time_check_due(S, iter, can_check);
    const bool store_delta = can_check || iter == S.max_iter || time_due;
    kkt_sweep<32, false>(M, B, S, inst, lane, FULL, sm, nullptr, store_delta);
    __syncwarp();
    if (can_check) {
      R = compute_residuals<32>(M, B, inst, lane, FULL, sm, &Rs);
      obj = R.obj;
      if (check_termination<32>(M, B, S, inst, lane, FULL, sm, R, 0, &status, &obj)) break;
    }
    if (S.adaptive_rho && S.adaptive_rho_interval && (iter % S.adaptive_rho_interval == 0)) {
      if (!can_check) { R = compute_residuals<32>(M, B, inst, lane, FULL, sm, &Rs); obj = R.obj; }
      const double rn = rho_estimate(Rs, rho);
      rho_est = rn;
      if (rn > rho * S.adaptive_rho_tolerance || rn < rho / S.adaptive_rho_tolerance) {
        // osqp_update_rho (osqp.c:1268-1319): retype-free rho_vec rewrite + full refactor
        rho = fmin(fmax(rn, kRhoMin), kRhoMax);
        for (int r = lane; r < B.mmax; r += 32) B.rho_vec[ro + r] = row_rho_eff(B, inst, ro + r, rho);
        __syncwarp();
        const int flag = factor_instance<32>(M, B, S.sigma, inst, lane, FULL, 0, G, W, Sm, rp);
        rho_updates++;
        if (flag) { status = -7; obj = kOsqpNan; if (lane == 0) B.fact_flag[inst] = flag; break; }
      }
    }
    if (time_due) {  // osqp.c:389-405
      int up = lane == 0 ? (int)time_is_up(B, S) : 0;
      up = __shfl_sync(FULL, up, 0);
      if (up) { timed_out = true; break; }
    }
  }
  int it_done = (iter > S.max_iter) ? S.max_iter : iter;
  if (status == -10) {
    if (!can_check) {  // osqp.c:524-546
      R = compute_residuals<32>(M, B, inst, lane, FULL, sm, &Rs);
      obj = R.obj;
      check_termination<32>(M, B, S, inst, lane, FULL, sm, R, 0, &status, &obj);
    }
    if (status == -10) {  // osqp.c:562-567 (max_iter) and :570-576 (time limit)
      if (!check_termination<32>(M, B, S, inst, lane, FULL, sm, R, 1, &status, &obj)) status = timed_out ? -6 : -2;
    }
  }
  const bool has_sol = !(status == -3 || status == 3 || status == -4 || status == 4 || status == -7);
  if (has_sol && status != -10) obj = R.obj;
  rho_est = rho_estimate(Rs, rho);
  if (!has_sol) {
    // store_solution (auxil.c:527-563): normalise the certificate, cold-start the iterates
    if (status == -3 || status == 3) {
      double nrm = 0.0;
      for (int r = lane; r < m; r += 32) nrm = fmax(nrm, fabs(B.dy[ro + r]));
      nrm = grp_max<32>(nrm, FULL);
      for (int r = lane; r < m; r += 32) B.dy[ro + r] *= 1.0 / nrm;
    }
    if (status == -4 || status == 4) {
      double nrm = 0.0;
      for (int i = lane; i < n; i += 32) nrm = fmax(nrm, fabs(B.dx[vo + i]));
      nrm = grp_max<32>(nrm, FULL);
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
  __syncwarp();
  deliver_to_host(B, inst, lane, 32, n, m, status, it_done, obj, R.pri, R.dua, rho_est);
}

// ---------------------------------------------------------------------------
// Polish (reference src/polish.c:212-350), one warp per instance, for the instances that ended OSQP_SOLVED.
// The reduced KKT system of the guessed active set
//     [ P + delta I   Ared' ] [x]   [ -q            ]
//     [ Ared        -delta I ] [nu] = [ l_low ; u_upp ]
// is the stage-ordered KKT with sigma = delta, rho = 1/delta on the active rows and the inactive rows removed
// (path rows: rho = 0, they drop out of C' rho C; dynamics rows: 1/rho = 1e200, their row and column of
// S_k^-1 vanish), so it is factored by the same stage recursion into a second factor (P.Xinv / P.Sinv) and solved
// by the same block substitution; polish_refine_iter rounds of iterative refinement against the unregularised
// products (polish.c:120-168) remove the delta error.  Then z = A x, the duals of the active rows, the normal-cone
// projection (proj.c:16-29), the residuals of the polished point and the acceptance test (polish.c:300-330).
// Under Ruiz scaling the reference does all of this on scaled data; in unscaled variables that is
// sigma_i = delta / (c D_i^2), rho_j = E_j^2 / (c delta), the activity tests w_j (z_j - l_j) < -y_j and
// w_j (u_j - z_j) < y_j with w_j = E_j^2 / c, and the projection z' = clip(z + y / w, l, u), y' = w (z + y / w - z').
// P = the batch with Xinv / Sinv / rho_vec pointing at the polish buffers; pb, ps, pr: [batch][nmax + mmax]
// work vectors (rhs b; solution [x ; nu]; refinement vector, then the polished z).
// ---------------------------------------------------------------------------
__global__ void k_polish(ModelDev M, BatchDev B, BatchDev P, SettingsDev S, double delta, int refine_iter,
                         double *pb, double *ps, double *pr, int *status_polish) {
  extern __shared__ double smem[];
  const int wpb = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int inst = blockIdx.x * wpb + wid;
  if (inst >= B.batch) return;
  if (B.status[inst] != 1) {  // polish only follows OSQP_SOLVED (osqp.c:590-591)
    if (lane == 0) status_polish[inst] = 0;
    return;
  }
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
  const int n = B.n, m = B.m, ld = B.nmax + B.mmax, ny = M.ny, b = M.b, N = B.N;
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  const double *q = B.q + vo, *l = B.l + ro, *u = B.u + ro;
  double *x = B.x + vo, *z = B.z + ro, *y = B.y + ro;
  double *bb = pb + (size_t)inst * ld, *sol = ps + (size_t)inst * ld, *res = pr + (size_t)inst * ld;
  double *prho = P.rho_vec + ro;
  const double cs = B.sE ? B.sC[inst] : 1.0;
  auto wgt = [&](int r) -> double { return B.sE ? B.sE[ro + r] * B.sE[ro + r] / cs : 1.0; };
  auto is_dyn = [&](int r) -> bool { const int k = r / b; return k < N && (r - k * b) >= ny; };
  // ---- active set (form_Ared, polish.c:20-49), reduced rhs (form_rhs_red :103-118), step sizes ----
  for (int i = lane; i < n; i += 32) bb[i] = -q[i];
  for (int r = lane; r < m; r += 32) {
    const double w = wgt(r);
    double rhs = 0.0, rho = is_dyn(r) ? 1e-200 : 0.0;
    if (w * (z[r] - l[r]) < -y[r]) { rhs = l[r]; rho = w / delta; }
    else if (w * (u[r] - z[r]) < y[r]) { rhs = u[r]; rho = w / delta; }
    bb[n + r] = rhs;
    prho[r] = rho;
  }
  __syncwarp();
  // ---- factor the reduced KKT by the stage recursion ----
  if (factor_instance<32>(M, P, delta, inst, lane, FULL, 0, G, W, Sm, rp)) {
    if (lane == 0) status_polish[inst] = -1;
    return;
  }
  // v <- K_reg^-1 v = [x ; nu], nu = 0 on the inactive rows.  The block substitution returns z~ = rhs + nu / rho
  // on the rows (qdldl_interface.c:572-579); nu = rho (z~ - rhs).
  double *rowc = B.dy + ro;  // scratch (the deltas are not needed after OSQP_SOLVED)
  auto solve = [&](double *v) {
    for (int r = lane; r < m; r += 32) rowc[r] = v[n + r];
    __syncwarp();
    kkt_sweep<32, true>(M, P, S, inst, lane, FULL, sm, v, false);
    __syncwarp();
    for (int r = lane; r < m; r += 32) {
      const double rho = prho[r];
      v[n + r] = rho > 1e-100 ? rho * (v[n + r] - rowc[r]) : 0.0;
    }
    __syncwarp();
  };
  for (int i = lane; i < n + m; i += 32) sol[i] = bb[i];
  __syncwarp();
  solve(sol);
  // ---- iterative refinement (polish.c:120-168): z += K_reg^-1 (b - K z), K = [P Ared'; Ared 0] ----
  for (int it = 0; it < refine_iter; it++) {
    product_sweep<32>(
        M, B, lane, FULL, sm, sol, sol + n,
        [&](int ri, double ax) { res[n + ri] = prho[ri] > 1e-100 ? bb[n + ri] - ax : 0.0; },
        [&](int xi, double, double px, double aty) { res[xi] = (bb[xi] - px) - aty; });
    __syncwarp();
    solve(res);
    for (int i = lane; i < n + m; i += 32) sol[i] += res[i];
    __syncwarp();
  }
  // ---- polished point: x, z = A x, y (get_ypol_from_yred), normal-cone projection, residuals ----
  double *pz = res + n;   // polished z (the refinement vector is free now)
  double *py = sol + n;   // polished y, in place over nu
  double pri = 0.0, dua = 0.0, oq = 0.0, oP = 0.0;
  product_sweep<32>(
      M, B, lane, FULL, sm, sol, py, [&](int ri, double ax) { pz[ri] = ax; }, [&](int, double, double, double) {});
  __syncwarp();
  for (int r = lane; r < m; r += 32) {  // project_normalcone (proj.c:16-29) on the scaled pair
    const double w = wgt(r);
    const double t = pz[r] + py[r] / w;
    const double zn = fmin(fmax(t, l[r]), u[r]);
    py[r] = w * (t - zn);
    pz[r] = zn;
  }
  __syncwarp();
  product_sweep<32>(
      M, B, lane, FULL, sm, sol, py,
      [&](int ri, double ax) { pri = fmax(pri, fabs(ax - pz[ri])); },
      [&](int xi, double xval, double px, double aty) {
        const double qv = q[xi];
        dua = fmax(dua, fabs((qv + px) + aty));
        oq = fma(qv, xval, oq);
        oP = fma(xval, px, oP);
      });
  pri = grp_max<32>(pri, FULL); dua = grp_max<32>(dua, FULL);
  const double obj = 0.5 * grp_sum<32>(oP, FULL) + grp_sum<32>(oq, FULL);
  if (m == 0) pri = 0.0;
  const double ipri = B.pri[inst], idua = B.dua[inst];
  const bool ok = (pri < ipri && dua < idua) || (pri < ipri && idua < 1e-10) || (dua < idua && ipri < 1e-10);
  if (ok) {
    for (int i = lane; i < n; i += 32) x[i] = sol[i];
    for (int r = lane; r < m; r += 32) { z[r] = pz[r]; y[r] = py[r]; }
    if (lane == 0) { B.obj[inst] = obj; B.pri[inst] = pri; B.dua[inst] = dua; }
  }
  if (lane == 0) status_polish[inst] = ok ? 1 : -1;
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
  kkt_sweep<32, true>(M, B, S, inst, lane, FULL, sm, rhs + (size_t)inst * ld, false);
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
  product_sweep<32>(
      M, B, lane, FULL, sm, B.x + vo, B.y + ro, [&](int ri, double ax) { z[ri] = ax; },
      [&](int, double, double, double) {});
}

// l <= u over the active rows [0,m) of every instance (osqp_update_bounds precondition, osqp.c:793-801)
__global__ void k_check_bounds(const double *l, const double *u, int ld, int m, int batch, int *bad) {
  const size_t tot = (size_t)batch * m;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t st = (size_t)gridDim.x * blockDim.x;
  int b = 0;
  for (; i < tot; i += st) {
    const size_t inst = i / m, r = i - inst * m;
    if (l[inst * ld + r] > u[inst * ld + r]) b = 1;
  }
  if (__any_sync(FULL, b) && (threadIdx.x & 31) == 0) atomicOr(bad, 1);
}

// In-place shift of every instance's row by `sh` entries towards index 0 over [0,len): one warp per
// instance, 32-entry chunks in ascending order (a chunk is read completely before it is overwritten,
// and later chunks only read entries that have not been written yet).  Entries [len,ld) are zeroed,
// as they are in the vectors rldl_batch_get_solution hands to a host-side shift.
__global__ void k_shift_rows(double *v, int ld, int len, int sh, int batch) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= batch) return;
  double *row = v + (size_t)warp * ld;
  for (int base = 0; base < len - sh; base += 32) {
    const int i = base + lane;
    double t = 0.0;
    if (i < len - sh) t = row[i + sh];
    __syncwarp();
    if (i < len - sh) row[i] = t;
    __syncwarp();
  }
  for (int i = len + lane; i < ld; i += 32) row[i] = 0.0;  // entries beyond the current horizon start cold
}

__global__ void k_fill(double *p, double v, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t st = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) p[i] = v;
}

// %globaltimer at the start of a solve: the reference point of the device-side time_limit test
__global__ void k_mark_time(int *sched) {
  unsigned long long now;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
  *reinterpret_cast<unsigned long long *>(sched + kSchedT0) = now;
}
cudaError_t launch_mark_time(const BatchDev &B, cudaStream_t st) {
  k_mark_time<<<1, 1, 0, st>>>(B.sched);
  return cudaGetLastError();
}

__global__ void k_fill_int(int *p, int v, size_t n) {
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

cudaError_t launch_polish(const ModelDev &M, const BatchDev &B, const BatchDev &P, const SettingsDev &S, double delta,
                          int refine_iter, double *pb, double *ps, double *pr, int *status_polish, cudaStream_t st) {
  const size_t per = admm_smem_doubles(M.a, M.nx, M.ny, M.nt) * sizeof(double);
  if (per > 220 * 1024) return cudaErrorInvalidValue;
  int wpb = pick_wpb(per);
  if (wpb > 2) wpb = 2;
  const size_t shm = per * wpb;
  cudaError_t e = cudaFuncSetAttribute(k_polish, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm);
  if (e != cudaSuccess) return e;
  k_polish<<<(B.batch + wpb - 1) / wpb, wpb * 32, shm, st>>>(M, B, P, S, delta, refine_iter, pb, ps, pr, status_polish);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Allocation-free KKT solve (rldl_batch_kkt_solve_static).  The existing entry point stages the right-hand sides in
// a cudaMalloc'ed buffer that grows on demand; this variant stages them in a fixed file-scope device array, so the
// call path contains no cudaMalloc / cudaFree.  Every GPU that loads the library holds the array (4 MiB) until exit.
// All calls share it: it keeps the last call's contents, calls that overlap on the GPU race.  Nothing in it has to
// start at zero — the kernel reads exactly the count doubles the call has just copied in.
// In host code `g_kkt_scratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, looked up once per device with that device current.
// ---------------------------------------------------------------------------
static __device__ __align__(256) double g_kkt_scratch[kKktStaticDoubles];

static cudaError_t kkt_scratch_address(double **out) {
  static double *addr[64] = {nullptr};  // per device; racing first calls write the same value
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  if (!addr[dev]) {
    double *p = nullptr;
    e = cudaGetSymbolAddress((void **)&p, g_kkt_scratch);
    if (e != cudaSuccess) return e;
    addr[dev] = p;
  }
  *out = addr[dev];
  return cudaSuccess;
}

// rhs_in / sol_out: host arrays of count = batch * ld doubles (ld = n + m); the caller has checked count against
// kKktStaticDoubles.  Copies in, runs k_kkt_solve on the static scratch, copies out; stream-ordered, not synchronised.
cudaError_t launch_kkt_solve_static(const ModelDev &M, const BatchDev &B, const SettingsDev &S, const double *rhs_in,
                                    double *sol_out, size_t count, int ld, cudaStream_t st) {
  if (count > (size_t)kKktStaticDoubles) return cudaErrorInvalidValue;
  double *scratch = nullptr;
  cudaError_t e = kkt_scratch_address(&scratch);
  if (e != cudaSuccess) return e;
  e = cudaMemcpyAsync(scratch, rhs_in, count * sizeof(double), cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return e;
  e = launch_kkt_solve(M, B, S, scratch, ld, st);
  if (e != cudaSuccess) return e;
  return cudaMemcpyAsync(sol_out, scratch, count * sizeof(double), cudaMemcpyDeviceToHost, st);
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

cudaError_t launch_check_bounds(const double *l, const double *u, int ld, int m, int batch, int *bad,
                                cudaStream_t st, bool reset) {
  cudaError_t e = reset ? cudaMemsetAsync(bad, 0, sizeof(int), st) : cudaSuccess;
  if (e != cudaSuccess) return e;
  k_check_bounds<<<148 * 8, 256, 0, st>>>(l, u, ld, m, batch, bad);
  return cudaGetLastError();
}

cudaError_t launch_shift_rows(double *v, int ld, int len, int sh, int batch, cudaStream_t st) {
  if (len <= sh || sh <= 0) return cudaSuccess;
  const int threads = 128, wpb = threads / 32;
  k_shift_rows<<<(batch + wpb - 1) / wpb, threads, 0, st>>>(v, ld, len, sh, batch);
  return cudaGetLastError();
}

cudaError_t launch_fill(double *p, double v, size_t n, cudaStream_t st) {
  if (!n) return cudaSuccess;
  size_t blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_fill<<<(unsigned)blocks, 256, 0, st>>>(p, v, n);
  return cudaGetLastError();
}

cudaError_t launch_fill_int(int *p, int v, size_t n, cudaStream_t st) {
  if (!n) return cudaSuccess;
  size_t blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_fill_int<<<(unsigned)blocks, 256, 0, st>>>(p, v, n);
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
