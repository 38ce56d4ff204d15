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
    const double nv = row_rho(B.l[off + r], B.u[off + r], rho);
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
  Resid R = {0, 0, 0, 0, 0, 0, 0, 0};
  for (iter = 1; iter <= S.max_iter; iter++) {
    can_check = S.check_termination && (iter % S.check_termination == 0);
    const bool time_due = time_check_due(S, iter, can_check);
    const bool store_delta = can_check || iter == S.max_iter || time_due;
    kkt_sweep<32, false>(M, B, S, inst, lane, FULL, sm, nullptr, store_delta);
    __syncwarp();
    if (can_check) {
      R = compute_residuals<32>(M, B, inst, lane, FULL, sm);
      obj = R.obj;
      if (check_termination<32>(M, B, S, inst, lane, FULL, sm, R, 0, &status, &obj)) break;
    }
    if (S.adaptive_rho && S.adaptive_rho_interval && (iter % S.adaptive_rho_interval == 0)) {
      if (!can_check) { R = compute_residuals<32>(M, B, inst, lane, FULL, sm); obj = R.obj; }
      const double rn = rho_estimate(R, rho);
      rho_est = rn;
      if (rn > rho * S.adaptive_rho_tolerance || rn < rho / S.adaptive_rho_tolerance) {
        // osqp_update_rho (osqp.c:1268-1319): retype-free rho_vec rewrite + full refactor
        rho = fmin(fmax(rn, kRhoMin), kRhoMax);
        for (int r = lane; r < B.mmax; r += 32)
          B.rho_vec[ro + r] = row_rho(B.l[ro + r], B.u[ro + r], rho);
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
      R = compute_residuals<32>(M, B, inst, lane, FULL, sm);
      obj = R.obj;
      check_termination<32>(M, B, S, inst, lane, FULL, sm, R, 0, &status, &obj);
    }
    if (status == -10) {  // osqp.c:562-567 (max_iter) and :570-576 (time limit)
      if (!check_termination<32>(M, B, S, inst, lane, FULL, sm, R, 1, &status, &obj)) status = timed_out ? -6 : -2;
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
                                cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(bad, 0, sizeof(int), st);
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
