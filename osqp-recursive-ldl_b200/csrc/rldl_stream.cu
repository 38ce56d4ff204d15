// rldl_stream.cu — the HBM-streaming ADMM kernel (sm_100a).
//
// One QP instance per group of G lanes (G = 8/16/32, the smallest power of two >= nx+nu),
// persistent CTAs pulling instances from a device-side queue, the whole osqp_solve loop
// (reference src/osqp.c:354-627) inside one launch.  Per ADMM iteration an instance walks
// its N+1 stages forward and backward (the block substitution of SURVEY §8a "block form");
// each stage step needs that stage's record
//     [ q_k | Xinv_k packed | Sinv_k packed | rho_k | l_k | u_k ]
// which is immutable during a solve and is streamed HBM -> shared memory by the TMA engine
// (cp.async.bulk + mbarrier complete_tx) into a per-group ring, D stage steps ahead of the
// arithmetic, so the dependent chain of small FP64 mat-vecs only ever touches shared memory.
// The mutable iterates (x, z, y and the forward intermediates) stay in HBM/L2 in the flat
// per-instance arrays and are register-prefetched one stage step ahead by the lane that
// owns (wrote) them.
//
// The arithmetic of a stage step is the same as kkt_sweep<> in rldl_device_fn.cuh (same
// formulas, same summation order); residuals, termination and infeasibility detection are
// the shared group routines of that header.
#include <cuda_runtime.h>
#include <stdint.h>
#include "rldl_device_fn.cuh"
#include "rldl_kernels.h"

namespace rldl {

// ------------------------------ PTX helpers ------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// 1-D bulk async copy global -> shared (TMA engine, SASS UBLKCP); completes on `bar`.
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// Row `row` of an ELL matrix whose idx/val live in shared memory.
__device__ __forceinline__ double ell_dot_s(const int *sidx, const double *sval, const Ell &E, int row,
                                            const double *v) {
  const int *ix = sidx + E.off + row * E.width;
  const double *vv = sval + E.off + row * E.width;
  double s = 0.0;
  for (int e = 0; e < E.width; e++) s = fma(vv[e], v[ix[e]], s);
  return s;
}

// y_i = sum_j M(i,j) v_j for a symmetric n x n matrix stored as its lower triangle,
// column-major packed: (i,j), i >= j, at j*n - j*(j+1)/2 + i.  Two accumulators over even
// and odd j, like sym_mv().
__device__ __forceinline__ double sym_mv_packed(const double *__restrict__ P, int n, int i, const double *v) {
  const int ci = i * n - (i * (i + 1)) / 2;
  double s0 = 0.0, s1 = 0.0;
  int cj = 0;  // j*n - j*(j+1)/2
  int j = 0;
  for (; j + 1 < n; j += 2) {
    const int c1 = cj + n - (j + 1);
    const int i0 = (j <= i) ? cj + i : ci + j;
    const int i1 = (j + 1 <= i) ? c1 + i : ci + j + 1;
    s0 = fma(P[i0], v[j], s0);
    s1 = fma(P[i1], v[j + 1], s1);
    cj = c1 + n - (j + 2);
  }
  if (j < n) s0 = fma(P[(j <= i) ? cj + i : ci + j], v[j], s0);
  return s0 + s1;
}

// ---------------------------------------------------------------------------
// Record packing: (Xinv, Sinv, q, l, u, rho_vec) -> rec, stages k0..N of every instance.
// One thread per record element.
// ---------------------------------------------------------------------------
__global__ void k_pack_records(ModelDev M, BatchDev B, int k0, int parts /*bit0 factor, bit1 vectors*/) {
  const int a = M.a, nx = M.nx, b = M.b, nt = M.nt, N = B.N;
  const int ns = N + 1 - k0;
  const long long total = (long long)B.batch * ns * B.RS;
  long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long step = (long long)gridDim.x * blockDim.x;
  for (; t < total; t += step) {
    const int e = (int)(t % B.RS);
    const long long rest = t / B.RS;
    const int k = k0 + (int)(rest % ns);
    const int inst = (int)(rest / ns);
    const bool term = (k == N);
    double *dst = B.rec + (long long)inst * B.rec_stride + (long long)k * B.RS + e;
    const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
    int o = e;
    if (o < B.qa) {  // q_k
      if (!(parts & 2)) continue;
      const int dim = term ? nx : a, ilo = (k == 0 && !term) ? nx : 0, xoff = k * a - nx;
      *dst = (o >= ilo && o < dim) ? B.q[vo + xoff + o] : 0.0;
      continue;
    }
    o -= B.qa;
    if (o < B.PX) {  // Xinv_k, a x a packed (terminal: [Xinv_N 0; 0 I])
      if (!(parts & 1)) continue;
      double v = 0.0;
      if (o < (a * (a + 1)) / 2) {
        int j = 0, cs = 0;  // column start
        while (o >= cs + (a - j)) { cs += a - j; j++; }
        const int i = j + (o - cs);
        const double *Xk = B.Xinv + (size_t)inst * B.xstride + (size_t)k * a * a;
        if (!term) v = Xk[j * a + i];
        else v = (i < nx && j < nx) ? Xk[j * nx + i] : (i == j ? 1.0 : 0.0);
      }
      *dst = v;
      continue;
    }
    o -= B.PX;
    if (o < B.PS) {  // Sinv_k, nx x nx packed
      if (!(parts & 1)) continue;
      double v = 0.0;
      if (!term && o < (nx * (nx + 1)) / 2) {
        int j = 0, cs = 0;
        while (o >= cs + (nx - j)) { cs += nx - j; j++; }
        const int i = j + (o - cs);
        v = B.Sinv[(size_t)inst * B.sstride + (size_t)k * nx * nx + j * nx + i];
      }
      *dst = v;
      continue;
    }
    o -= B.PS;
    if (!(parts & 2)) continue;
    const int which = o / B.bb, r = o - which * B.bb;
    const int rows = term ? nt : b;
    double v = (which == 0) ? 1.0 : 0.0;
    if (r < rows) {
      const size_t ri = ro + (size_t)k * b + r;
      v = which == 0 ? B.rho_vec[ri] : (which == 1 ? B.l[ri] : B.u[ri]);
    }
    *dst = v;
  }
}

cudaError_t launch_pack_records(const ModelDev &M, const BatchDev &B, int k0, int parts, cudaStream_t st) {
  const long long total = (long long)B.batch * (B.N + 1 - k0) * B.RS;
  if (total <= 0) return cudaSuccess;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  k_pack_records<<<(unsigned)blocks, 256, 0, st>>>(M, B, k0, parts);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// The persistent streaming ADMM kernel.
// ---------------------------------------------------------------------------
template <int NYT>
struct MutVec {  // the mutable per-stage values a lane owns
  double x, zd, yd, yk, sk;
  double zp[NYT], yp[NYT];
};

struct StreamSmem {  // per-CTA shared memory carve-up (in doubles unless noted)
  int model_val;     // ELL values copy
  int model_idx;     // ELL indices copy (ints, stored after the doubles)
  int group0;        // first per-group block
  int group_stride;  // doubles per group
  int ring;          // offset of ring inside the group block
  int bars;          // offset of the D mbarriers inside the group block
  int scratch;       // offset of scratch vectors inside the group block
};

__host__ __device__ inline int even_up(int v) { return (v + 1) & ~1; }

__host__ __device__ inline StreamSmem stream_smem_layout(int nval, int RS, int D, int G, int NYT, int vm,
                                                         int groups) {
  StreamSmem L;
  L.model_val = 0;
  L.model_idx = even_up(nval);
  L.group0 = L.model_idx + even_up((nval + 1) / 2);
  L.ring = 0;
  L.bars = D * RS;
  L.scratch = L.bars + D;
  int sc = (NYT + 9) * G;
  if (8 * vm > sc) sc = 8 * vm;
  L.group_stride = even_up(L.scratch + sc);
  (void)groups;
  return L;
}

template <int G, int NYT>
__global__ void __launch_bounds__(128) k_admm_stream(ModelDev M, BatchDev B, SettingsDev S, int warm_start, int D) {
  extern __shared__ __align__(16) double smem[];
  constexpr int GPW = 32 / G;  // groups per warp
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int gl = lane & (G - 1), gw = lane / G;
  const unsigned gm = (G == 32) ? FULL : (((1u << G) - 1u) << (gw * G));
  const int grp = wid * GPW + gw;
  const int nx = M.nx, a = M.a, b = M.b, ny = M.ny, nt = M.nt, N = B.N, n = B.n, m = B.m;
  int vm = a;
  if (ny > vm) vm = ny;
  if (nt > vm) vm = nt;
  if (nx > vm) vm = nx;
  const StreamSmem L = stream_smem_layout(M.nval, B.RS, D, G, NYT, vm, (blockDim.x >> 5) * GPW);

  // CTA-shared copy of the model's ELL part
  double *sval = smem + L.model_val;
  int *sidx = reinterpret_cast<int *>(smem + L.model_idx);
  for (int e = threadIdx.x; e < M.nval; e += blockDim.x) {
    sval[e] = M.val[e];
    sidx[e] = M.idx[e];
  }
  double *gb = smem + L.group0 + (size_t)grp * L.group_stride;
  double *ring = gb + L.ring;
  uint64_t *bars = reinterpret_cast<uint64_t *>(gb + L.bars);
  double *sc = gb + L.scratch;
  double *w = sc, *tv = sc + NYT * G, *yvs = tv + G, *td = yvs + G, *sp = td + G, *nudv = sp + G,
         *gg = nudv + G, *vb0 = gg + G, *vb1 = vb0 + G;
  const Scratch rs = make_scratch(sc, vm);  // residual routines alias the sweep scratch
  if (gl == 0)
    for (int d = 0; d < D; d++) mbar_init(&bars[d], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();

  const int qa = B.qa, PX = B.PX, PS = B.PS, bb = B.bb, RS = B.RS;
  const int npos = 2 * (N + 1);
  const uint32_t fwd_bytes = 8u * (qa + PX + PS + bb), bwd_bytes = 8u * (PX + PS + 3 * bb);
  const double sigma = S.sigma, alpha = S.alpha;
  uint32_t pk = 0;  // packets consumed so far by this group (ring position)

  // per-instance state
  int inst = -1, iter = 0;
  bool have = false;
  const double *recb = nullptr;
  double *x = nullptr, *z = nullptr, *y = nullptr, *dx = nullptr, *dy = nullptr, *ykg = nullptr, *skg = nullptr;
  MutVec<NYT> cur;

  auto issue = [&](int pos, uint32_t pidx) {  // group leader only
    const int slot = pidx % D;
    const bool fwd = pos <= N;
    const int k = fwd ? pos : npos - 1 - pos;
    const double *src = recb + (size_t)k * RS + (fwd ? 0 : qa);
    const uint32_t bytes = fwd ? fwd_bytes : bwd_bytes;
    mbar_expect_tx(&bars[slot], bytes);
    bulk_g2s(ring + slot * RS + (fwd ? 0 : qa), src, bytes, &bars[slot]);
  };
  auto load_mv = [&](MutVec<NYT> &r, int pos) {
    const bool fwd = pos <= N;
    const int k = fwd ? pos : npos - 1 - pos;
    const bool term = (k == N);
    const int dim = term ? nx : a, np = term ? nt : ny;
    const int xoff = k * a - nx, roff = k * b, ilo = (k == 0) ? nx : 0;
    r.x = (gl >= ilo && gl < dim) ? x[xoff + gl] : 0.0;
#pragma unroll
    for (int t = 0; t < NYT; t++) {
      const int rr = gl + G * t;
      r.zp[t] = 0.0; r.yp[t] = 0.0;
      if (rr < np) { r.zp[t] = z[roff + rr]; r.yp[t] = y[roff + rr]; }
    }
    r.zd = r.yd = r.yk = r.sk = 0.0;
    if (!term && gl < nx) { r.zd = z[roff + ny + gl]; r.yd = y[roff + ny + gl]; }
    if (!fwd && !term) {
      if (gl < a) r.yk = ykg[k * a + gl];
      if (gl < nx) r.sk = skg[k * nx + gl];
    }
  };
  // claim the next instance of the queue; sets up pointers, cold start, ring and registers
  auto fetch = [&]() -> bool {
    for (;;) {
      int i = 0;
      if (gl == 0) i = atomicAdd(B.work_counter, 1);
      i = __shfl_sync(gm, i, 0, G);
      if (i >= B.batch) return false;
      inst = i;
      if (B.fact_flag[inst]) {  // failed factorisation: report non-convex, like a diverged solve
        if (gl == 0) { B.status[inst] = -7; B.iter[inst] = 0; B.obj[inst] = kOsqpNan; }
        continue;
      }
      break;
    }
    const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
    x = B.x + vo; z = B.z + ro; y = B.y + ro; dx = B.dx + vo; dy = B.dy + ro;
    ykg = B.yk + (size_t)inst * B.ykstride; skg = B.sk + (size_t)inst * B.skstride;
    recb = B.rec + (size_t)inst * B.rec_stride;
    if (gl == 0)
      for (int d = 0; d < D; d++) issue(d % npos, pk + d);
    if (!warm_start) {
      for (int i = gl; i < n; i += G) x[i] = 0.0;
      for (int r = gl; r < m; r += G) { z[r] = 0.0; y[r] = 0.0; }
    }
    __syncwarp(gm);
    iter = 0;
    load_mv(cur, 0);
    return true;
  };
  auto drain = [&]() {  // consume the D packets still in flight for the finished instance
    for (int d = 0; d < D; d++) {
      mbar_wait(&bars[pk % D], (pk / D) & 1);
      pk++;
    }
    __syncwarp(gm);
  };

  have = fetch();
  while (__any_sync(FULL, have)) {
    if (have) {
      iter++;
      const bool can_check = S.check_termination && (iter % S.check_termination == 0);
      const bool store_delta = can_check || iter == S.max_iter;
      double yv_reg = 0.0;  // this lane's y-vector entry of the last forward stage
      int par = 0;
      for (int pos = 0; pos < npos; pos++) {
        const bool fwd = pos <= N;
        const int k = fwd ? pos : npos - 1 - pos;
        const bool term = (k == N);
        const StageClass &scl = M.sc[term ? 2 : (k == 0 ? 0 : 1)];
        const int dim = term ? nx : a, np = term ? nt : ny;
        const int xoff = k * a - nx, roff = k * b, ilo = (k == 0) ? nx : 0;
        MutVec<NYT> nxt;
        if (pos + 1 < npos) load_mv(nxt, pos + 1);
        const double *slot = ring + (pk % D) * RS;
        const double *sq = slot, *sX = slot + qa, *sS = sX + PX, *srho = sS + PS, *sl = srho + bb, *su = sl + bb;
        mbar_wait(&bars[pk % D], (pk / D) & 1);
        if (fwd) {
          // ---- forward stage step (kkt_sweep<> forward, rldl_device_fn.cuh) ----
          double g = 0.0;
          if (gl >= ilo && gl < dim) g = sigma * cur.x - sq[gl];
#pragma unroll
          for (int t = 0; t < NYT; t++) {
            const int rr = gl + G * t;
            if (rr < np) w[rr] = srho[rr] * cur.zp[t] - cur.yp[t];
          }
          __syncwarp(gm);
          double tt = 0.0;
          if (gl < dim) {
            tt = g + ell_dot_s(sidx, sval, scl.Ct, gl, w);
            if (k > 0 && gl < nx) tt -= sp[gl];
          }
          tv[gl] = tt;
          __syncwarp(gm);
          double yvv = 0.0;
          if (gl < a) yvv = sym_mv_packed(sX, a, gl, tv);
          yvs[gl] = yvv;
          yv_reg = yvv;
          if (gl < dim) ykg[k * a + gl] = yvv;
          __syncwarp(gm);
          if (!term) {
            if (gl < nx) {
              const double rd = srho[ny + gl];
              td[gl] = (cur.zd - cur.yd / rd) - ell_dot_s(sidx, sval, scl.F, gl, yvs);
            }
            __syncwarp(gm);
            if (gl < nx) {
              const double s = sym_mv_packed(sS, nx, gl, td);
              sp[gl] = s;
              skg[k * nx + gl] = s;
            }
            __syncwarp(gm);
          }
        } else {
          // ---- backward stage step (kkt_sweep<> backward) ----
          double *vcur = par ? vb1 : vb0;
          const double *vprev = par ? vb0 : vb1;
          double vv = 0.0, nud = 0.0;
          if (term) {
            vv = yv_reg;
            vcur[gl] = vv;
            __syncwarp(gm);
          } else {
            if (gl < nx) {
              nud = -(cur.sk + sym_mv_packed(sS, nx, gl, vprev));
              nudv[gl] = nud;
            }
            __syncwarp(gm);
            if (gl < a) gg[gl] = ell_dot_s(sidx, sval, scl.Ft, gl, nudv);
            __syncwarp(gm);
            if (gl < a) vv = cur.yk - sym_mv_packed(sX, a, gl, gg);
            vcur[gl] = vv;
            __syncwarp(gm);
          }
          MutVec<NYT> nw = cur;  // new iterate values of this stage (carried into forward stage 0)
#pragma unroll
          for (int t = 0; t < NYT; t++) {
            const int rr = gl + G * t;
            if (rr < np) {
              const int ri = roff + rr;
              const double zt = ell_dot_s(sidx, sval, scl.C, rr, vcur);
              const double rho = srho[rr], zo = cur.zp[t], yo = cur.yp[t];
              const double zh = alpha * zt + (1.0 - alpha) * zo;
              const double zn = fmin(fmax(zh + yo / rho, sl[rr]), su[rr]);
              const double d = rho * (zh - zn);
              z[ri] = zn; y[ri] = yo + d;
              nw.zp[t] = zn; nw.yp[t] = yo + d;
              if (store_delta) dy[ri] = d;
            }
          }
          if (!term && gl < nx) {
            const int ri = roff + ny + gl;
            const double rho = srho[ny + gl];
            const double zo = cur.zd, yo = cur.yd;
            const double zt = (zo - yo / rho) + nud / rho;
            const double zh = alpha * zt + (1.0 - alpha) * zo;
            const double zn = fmin(fmax(zh + yo / rho, sl[ny + gl]), su[ny + gl]);
            const double d = rho * (zh - zn);
            z[ri] = zn; y[ri] = yo + d;
            nw.zd = zn; nw.yd = yo + d;
            if (store_delta) dy[ri] = d;
          }
          if (gl >= ilo && gl < dim) {
            const double xo = cur.x;
            const double xn = alpha * vv + (1.0 - alpha) * xo;
            x[xoff + gl] = xn;
            nw.x = xn;
            if (store_delta) dx[xoff + gl] = xn - xo;
          }
          par ^= 1;
          if (pos + 1 == npos) nxt = nw;  // stage 0 feeds the next iteration's forward stage 0
          __syncwarp(gm);
        }
        // slot consumed by the whole group: refill it D packets ahead
        if (gl == 0) {
          int ip = pos + D;
          while (ip >= npos) ip -= npos;
          issue(ip, pk + D);
        }
        pk++;
        cur = nxt;
      }
      // ---- iteration epilogue: residuals, termination (osqp.c:426-457, 524-567) ----
      bool done = false;
      int status = -10;
      double obj = 0.0;
      Resid R = {0, 0, 0, 0, 0, 0, 0, 0};
      if (can_check) {
        R = compute_residuals<G>(M, B, inst, gl, gm, rs);
        obj = R.obj;
        if (check_termination<G>(M, B, S, inst, gl, gm, rs, R, 0, &status, &obj)) done = true;
      }
      if (!done && iter == S.max_iter) {
        if (!can_check) {
          R = compute_residuals<G>(M, B, inst, gl, gm, rs);
          obj = R.obj;
          check_termination<G>(M, B, S, inst, gl, gm, rs, R, 0, &status, &obj);
        }
        if (status == -10) {
          if (!check_termination<G>(M, B, S, inst, gl, gm, rs, R, 1, &status, &obj)) status = -2;
        }
        done = true;
      }
      if (done) {
        const double rho = B.rho[inst];
        const bool has_sol = !(status == -3 || status == 3 || status == -4 || status == 4 || status == -7);
        if (has_sol) obj = R.obj;
        const double rho_est = rho_estimate(R, rho);
        if (!has_sol) {  // store_solution (auxil.c:527-563)
          if (status == -3 || status == 3) {
            double nrm = 0.0;
            for (int r = gl; r < m; r += G) nrm = fmax(nrm, fabs(dy[r]));
            nrm = grp_max<G>(nrm, gm);
            for (int r = gl; r < m; r += G) dy[r] *= 1.0 / nrm;
          }
          if (status == -4 || status == 4) {
            double nrm = 0.0;
            for (int i = gl; i < n; i += G) nrm = fmax(nrm, fabs(dx[i]));
            nrm = grp_max<G>(nrm, gm);
            for (int i = gl; i < n; i += G) dx[i] *= 1.0 / nrm;
          }
          for (int i = gl; i < n; i += G) x[i] = 0.0;
          for (int r = gl; r < m; r += G) { z[r] = 0.0; y[r] = 0.0; }
        }
        if (gl == 0) {
          B.iter[inst] = iter;
          B.status[inst] = status;
          B.obj[inst] = obj;
          B.pri[inst] = R.pri;
          B.dua[inst] = R.dua;
          B.rho_est[inst] = rho_est;
        }
        drain();
        have = fetch();
      } else {
        __syncwarp(gm);  // residual scratch aliases the sweep scratch
      }
    }
    __syncwarp();
  }
}

// ------------------------------- launcher --------------------------------
int stream_group_lanes(int a) { return a <= 8 ? 8 : (a <= 16 ? 16 : (a <= 32 ? 32 : 0)); }

bool stream_supported(const ModelDev &M, int N) {
  const int G = stream_group_lanes(M.a);
  if (!G || N < 1) return false;
  const int rows = M.ny > M.nt ? M.ny : M.nt;
  return (rows + G - 1) / G <= 4 && M.nx <= G && M.nu >= 0;
}

template <int G, int NYT>
static cudaError_t launch_stream_t(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                                   cudaStream_t st) {
  const int threads = 128, groups = (threads / 32) * (32 / G);
  int vm = M.a;
  if (M.ny > vm) vm = M.ny;
  if (M.nt > vm) vm = M.nt;
  if (M.nx > vm) vm = M.nx;
  int D = 4;
  if (D > 2 * (B.N + 1)) D = 2 * (B.N + 1);
  StreamSmem L = stream_smem_layout(M.nval, B.RS, D, G, NYT, vm, groups);
  size_t shm = ((size_t)L.group0 + (size_t)groups * L.group_stride) * sizeof(double);
  while (shm > 220 * 1024 && D > 2) {
    D--;
    L = stream_smem_layout(M.nval, B.RS, D, G, NYT, vm, groups);
    shm = ((size_t)L.group0 + (size_t)groups * L.group_stride) * sizeof(double);
  }
  if (shm > 220 * 1024) return cudaErrorInvalidValue;
  auto kern = k_admm_stream<G, NYT>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm);
  if (e != cudaSuccess) return e;
  int dev = 0, sms = 148, per_sm = 1;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, shm);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  long long grid = (long long)sms * per_sm;
  const long long need = ((long long)B.batch + groups - 1) / groups;
  if (grid > need) grid = need;
  e = cudaMemsetAsync(B.work_counter, 0, sizeof(int), st);
  if (e != cudaSuccess) return e;
  kern<<<(unsigned)grid, threads, shm, st>>>(M, B, S, warm_start, D);
  return cudaGetLastError();
}

cudaError_t launch_admm_stream(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                               cudaStream_t st) {
  const int G = stream_group_lanes(M.a);
  const int rows = M.ny > M.nt ? M.ny : M.nt;
  const int nyt = (rows + G - 1) / G;
#define RLDL_CASE(g, t) \
  if (G == g && nyt == t) return launch_stream_t<g, t>(M, B, S, warm_start, st);
  RLDL_CASE(8, 1) RLDL_CASE(8, 2) RLDL_CASE(8, 3) RLDL_CASE(8, 4)
  RLDL_CASE(16, 1) RLDL_CASE(16, 2) RLDL_CASE(16, 3) RLDL_CASE(16, 4)
  RLDL_CASE(32, 1) RLDL_CASE(32, 2) RLDL_CASE(32, 3) RLDL_CASE(32, 4)
#undef RLDL_CASE
  return cudaErrorInvalidValue;
}

}  // namespace rldl
