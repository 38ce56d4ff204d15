// rldl_stream.cu — the HBM-streaming ADMM kernel (sm_100a), shape-specialised.
//
// One QP instance per group of G lanes (G = 8/16/32, the smallest power of two >= nx+nu),
// persistent CTAs pulling instances from a device-side queue, the whole osqp_solve loop
// (reference src/osqp.c:354-627) inside one launch.  Per ADMM iteration an instance walks
// its N+1 stages forward and backward (the block substitution of SURVEY §8a "block form");
// each stage step needs that stage's record
//     [ q_k | Xinv_k | Sinv_k | row classes | l_k | u_k ]
// which is immutable during a solve and is streamed HBM -> shared memory by the TMA engine
// (cp.async.bulk + mbarrier complete_tx, SASS UBLKCP) into a per-group ring, D stage steps
// ahead of the arithmetic, so the dependent chain of small FP64 mat-vecs only ever touches
// shared memory.  The mutable iterates (x, z, y and the forward intermediates) stay in
// HBM/L2 in the flat per-instance arrays and are register-prefetched one stage step ahead
// by the lane that owns (wrote) them.  Because all instance state lives in HBM between
// iterations, instances are preemptible: a device-side priority scheduler (below) decides
// at every termination check which instance a group serves next, and an adaptive-rho update
// parks the instance for a batched re-factorisation launch and a resume.
//
// Symmetric pivot inverses are stored "diagonal-packed": DP[d][i] = M(i, (i+d) mod n) for
// d = 0..n/2 — half the bytes of the full matrix, and lane i (owner of row i) reads
// DP[d][i] for the upper wrap and DP[d][(i-d) mod n] for the lower wrap, both conflict-free;
// the input vector is kept doubled in shared memory so v[(i+d) mod n] is v2[i+d].
//
// The kernel is a template over the stage shape (nx+nu, nx, ny, nt) and the ELL width of the
// path-constraint matrices so that every inner loop unrolls; shapes outside the table at the
// bottom of this file use the generic kernel in rldl_kernels.cu.  The arithmetic follows
// kkt_sweep<> in rldl_device_fn.cuh (reference src/auxil.c:164-228, qdldl_interface.c:559-585)
// except that rho^-1 is a per-class precomputed reciprocal, as in the reference
// (rho_inv_vec, src/auxil.c:96), instead of a division.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include "rldl_device_fn.cuh"
#include "rldl_kernels.h"
#include "rldl_stream_common.cuh"

namespace rldl {

// ---------------------------------------------------------------------------
// Vector part of the stage records, stages k0..N of every instance, straight from the flat per-instance
// vectors (one CTA per instance, one warp per stage, the rows of a stage across the lanes: coalesced reads of
// q, l, u and coalesced writes of the record parts):
//   parts bit 0: q_k
//   parts bit 1: constraint typing of every row from (l, u) (reference src/auxil.c:79-145: loose / equality /
//                inequality), written as the 2-bit class words of the record and as rho_vec (the per-row values
//                the runtime-shape kernels and the scalar API read), l_k, u_k; the instance's rho is clipped to
//                [RHO_MIN, RHO_MAX] and an instance parked by an adaptive-rho update becomes resumable again.
//                changed[inst] = 1 iff a class differs from the one the record held (update_rho_vec's
//                "constr_type_changed": the caller re-runs the stage recursion for those instances only) or the
//                instance was parked by an adaptive-rho update.
//   parts bit 2: touch only the instances parked by an adaptive-rho update (the others get changed = 0).
// Replaces the k_set_rho_vec + k_pack_records pair of round 1 (two passes over rho_vec with per-element
// index decoding) by one pass at streaming speed.
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long spread_bits(unsigned v) {  // bit r -> bit 2r
  unsigned long long x = v;
  x = (x | (x << 16)) & 0x0000ffff0000ffffull;
  x = (x | (x << 8)) & 0x00ff00ff00ff00ffull;
  x = (x | (x << 4)) & 0x0f0f0f0f0f0f0f0full;
  x = (x | (x << 2)) & 0x3333333333333333ull;
  x = (x | (x << 1)) & 0x5555555555555555ull;
  return x;
}

template <class SH>
__global__ void __launch_bounds__(256) k_vec_records(BatchDev B, int k0, int parts, const int *only_flagged,
                                                     int *changed, double sigma) {
  // one CTA per instance, one warp per stage (stages k0 + w, k0 + w + 8, ...): at every step the CTA reads
  // 8 consecutive stages of q / l / u (2 KB runs) — a warp walking its own row alone leaves HBM at a tenth of
  // its bandwidth (one 224-byte access per DRAM page and 9 000 concurrent streams)
  constexpr int A = SH::A, NX = SH::NX, NT = SH::NT, BQ = SH::B, RS = SH::RS;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int inst = blockIdx.x;
  if (only_flagged && !only_flagged[inst]) return;
  const bool parked = B.status[inst] == kStatusParkedRho;
  if ((parts & 4) && !parked) {  // adaptive-rho refresh: only the parked instances
    if (changed && threadIdx.x == 0) changed[inst] = 0;
    return;
  }
  __shared__ int s_ch;
  if (threadIdx.x == 0) s_ch = parked ? 1 : 0;  // a parked instance has a new rho: its recursion is re-run anyway
  __syncthreads();
  const int N = B.N;
  const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
  const double rho = fmin(fmax(B.rho[inst], kRhoMin), kRhoMax);
  double *rec = B.rec + (size_t)inst * B.rec_stride;
  int ch = 0;
  // Each warp takes its stages U at a time: all loads of the U stages are issued before the first dependent
  // instruction (the stores of one stage may alias the loads of the next as far as the compiler knows, so a plain
  // loop runs one DRAM round trip per stage and leaves the memory system idle)
  constexpr int U = 4, W = (SH::BB + 31) / 32;
  static_assert(W <= kClsWords, "class words");
  RLDL_ASSERT(k0 >= 0 && k0 <= N && N <= B.Nmax && (long long)(N + 1) * RS <= B.rec_stride);
  for (int kb = k0 + wid; kb <= N; kb += U * nw) {
    double lv[U][W], uv[U][W], qv[U];
    double sg[U], ev[U][W], orho[U][W];  // Ruiz-scaled variants: sigma_i, E_r, the step size the record held
    unsigned long long old[U][W];
    const double cs = SH::RZ ? B.sC[inst] : 1.0;
#pragma unroll
    for (int j = 0; j < U; j++) {
      const int k = kb + j * nw;
      const bool on = k <= N, term = (k == N);
      const int rows = term ? NT : BQ;
      const int dim = term ? NX : A, ilo = (k == 0 && !term) ? NX : 0, xoff = k * A - NX;
      qv[j] = 0.0; sg[j] = sigma;
      if (on && (parts & 1) && lane >= ilo && lane < dim) {
        qv[j] = __ldcs(B.q + vo + xoff + lane);
        if constexpr (SH::RZ) {
          const double d = B.sD[vo + xoff + lane];
          sg[j] = sigma / (cs * d * d);
        }
      }
#pragma unroll
      for (int w = 0; w < W; w++) {
        const int r = 32 * w + lane;
        lv[j][w] = uv[j][w] = 0.0; old[j][w] = 0ull; ev[j][w] = 1.0; orho[j][w] = 0.0;
        if (on && (parts & 2)) {
          if (r < rows) {
            const size_t ri = ro + (size_t)k * BQ + r;
            lv[j][w] = __ldcs(B.l + ri); uv[j][w] = __ldcs(B.u + ri);
            if constexpr (SH::RZ) { ev[j][w] = B.sE[ri]; orho[j][w] = rec[(size_t)k * RS + SH::CLS + r]; }
          }
          if (!SH::RZ && lane == 0)
            old[j][w] = (unsigned long long)__double_as_longlong(rec[(size_t)k * RS + SH::CLS + w]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < U; j++) {
      const int k = kb + j * nw;
      if (k > N) break;
      const bool term = (k == N);
      double *rk = rec + (size_t)k * RS;
      if (parts & 1) {
        static_assert(SH::QA <= 32, "q part of a record fits one warp pass");
        if (lane < SH::QA) {
          rk[lane] = qv[j];
          if constexpr (SH::RZ) rk[SH::SIG + lane] = sg[j];
        }
      }
      if (parts & 2) {
        const int rows = term ? NT : BQ;
#pragma unroll
        for (int w = 0; w < W; w++) {
          const int r = 32 * w + lane;
          if constexpr (SH::RZ) {
            // rows are typed on the SCALED bounds (osqp.c:815-817, auxil.c:79-145); in unscaled variables the step
            // size of the row is rho_class * E_r^2 / c
            double rv = rho;
            if (r < rows) {
              const double e = ev[j][w];
              rv = row_rho(e * lv[j][w], e * uv[j][w], rho) * e * e / cs;
              B.rho_vec[ro + (size_t)k * BQ + r] = rv;
              if (rv != orho[j][w]) ch = 1;
            }
            if (r < SH::BB) { rk[SH::CLS + r] = rv; rk[SH::LO + r] = lv[j][w]; rk[SH::UP + r] = uv[j][w]; }
          } else {
            int c = 0;
            if (r < rows) {
              c = (lv[j][w] < -kInfBound && uv[j][w] > kInfBound) ? 2 : ((uv[j][w] - lv[j][w] < kRhoTol) ? 1 : 0);
              B.rho_vec[ro + (size_t)k * BQ + r] = c == 2 ? kRhoMin : (c == 1 ? kRhoEq * rho : rho);
            }
            const unsigned long long wv = spread_bits(__ballot_sync(FULL, c == 1)) |
                                          (spread_bits(__ballot_sync(FULL, c == 2)) << 1);
            if (lane == 0) {
              if (old[j][w] != wv) ch = 1;
              rk[SH::CLS + w] = __longlong_as_double((long long)wv);
            }
            if (r < SH::BB) { rk[SH::LO + r] = lv[j][w]; rk[SH::UP + r] = uv[j][w]; }
          }
        }
        if constexpr (!SH::RZ) {
#pragma unroll
          for (int w = W; w < kClsWords; w++)
            if (lane == 0) rk[SH::CLS + w] = 0.0;
        }
      }
    }
  }
  if (parts & 2) {
    if (ch) atomicOr(&s_ch, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
      B.rho[inst] = rho;
      if (changed) changed[inst] = s_ch;
      if (parked) B.status[inst] = -10;  // parked for this refresh: resumable again
    }
  }
}

template <class SH>
static cudaError_t launch_vec_records_t(const BatchDev &B, int k0, int parts, const int *only_flagged, int *changed,
                                        double sigma, cudaStream_t st) {
  const int stages = B.N + 1 - k0;
  const int threads = stages >= 8 ? 256 : (stages >= 4 ? 128 : (stages >= 2 ? 64 : 32));
  k_vec_records<SH><<<B.batch, threads, 0, st>>>(B, k0, parts, only_flagged, changed, sigma);
  return cudaGetLastError();
}

// Record layout of the specialisation that serves M (the first table entry that does; the full a x a layout
// when none does and the records are unused).
void stream_record_dims(const ModelDev &M, int *qa, int *PX, int *PS, int *bb, int *bd, int *PXX, int *sigw,
                        int *clsw) {
  const int cw = max_c_width(M);
  int use_bd = 0, use_sc = 0;
#define RLDL_X(a_, nx_, ny_, nt_, cw_, bd_, sc_) \
  if (shape_serves(M, cw, a_, nx_, ny_, nt_, cw_, bd_, sc_)) { use_bd = bd_; use_sc = sc_; goto found; }
  RLDL_SHAPES(RLDL_X)
#undef RLDL_X
found:
  *bd = use_bd;
  *qa = even_up(M.a);
  *PXX = even_up(dp_size(M.nx));
  *PX = use_bd ? *PXX + even_up(dp_size(M.a - M.nx)) : even_up(dp_size(M.a));
  *PS = even_up(dp_size(M.nx));
  *bb = even_up(M.b);
  *sigw = use_sc ? *qa : 0;
  *clsw = use_sc ? *bb : kClsWords;
}

// ---------------------------------------------------------------------------
// Device work scheduler.  Iteration counts inside a batch are heavy-tailed (C5: mean 330,
// max 4000), and an instance is a strictly sequential chain, so the order in which instances
// are started decides how long the kernel runs half-empty at the end.  Instances are therefore
// preemptible at every termination check (all their state already lives in HBM): a group that
// finds a more urgent instance waiting suspends its own into a priority bucket and resumes the
// other.  Priority = how far the instance is from its tolerances, floor(2*log2(max(r_p/eps_p,
// r_d/eps_d))) — the instances that will run longest are served first (measured: within 5 % of
// the longest-first schedule computed with hindsight, 20 % shorter than first-come order).
// Fresh instances outrank everything until their first check so that every instance gets a
// priority early.  Buckets are multi-producer/multi-consumer rings of capacity `batch` (an
// instance sits in at most one ring), heads claimed by CAS, slots handed over with atomicExch.
// ---------------------------------------------------------------------------
// highest non-empty bucket (-1 if none); every lane of the group scans a strided share
template <int G>
__device__ __forceinline__ int sched_best(const BatchDev &B, int gl, unsigned gm) {
  int best = -1;
  for (int b = gl; b < kSchedBuckets; b += G)
    if (sched_ring_nonempty(B, b)) best = b;
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(gm, best, o));
  return best;
}

// ---------------------------------------------------------------------------
// The persistent streaming ADMM kernel
//   resume = 0: every instance starts at iteration 0 (cold or warm start)
//   resume = 1: only instances still OSQP_UNSOLVED continue from B.iter (after the host
//               re-factorised the instances parked by an adaptive-rho update)
//   s0     > 0: preemptive scheduling on; fresh instances run s0 iterations before they yield
// ---------------------------------------------------------------------------
template <class SH>
__global__ void __launch_bounds__(128, 4) k_admm_stream(ModelDev M, BatchDev B, SettingsDev S, int warm_start, int D,
                                                        int resume, int s0, int park_below, int l2_hints, int k_l,
                                                        int retire_below) {
  constexpr int A = SH::A, NX = SH::NX, NY = SH::NY, NT = SH::NT, CW = SH::CW, G = SH::G, NYT = SH::NYT,
                ROWS = SH::ROWS, RS = SH::RS, BQ = SH::B;
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int gl = lane & (G - 1), gw = lane / G;
  const unsigned gm = (G == 32) ? FULL : (((1u << G) - 1u) << (gw * G));
  const int grp = wid * SH::GPW + gw;
  const int N = B.N, n = B.n, m = B.m;

  // ---- CTA-shared model block: dense dynamics rows F (class 0 / 1) and padded ELL of C, C' ----
  double *Fd = smem + SH::M_FD, *FdT = smem + SH::M_FDT, *Cv = smem + SH::M_CV, *Ctv = smem + SH::M_CTV;
  int *Ci = reinterpret_cast<int *>(smem + SH::M_CI), *Cti = reinterpret_cast<int *>(smem + SH::M_CTI);
  for (int e = threadIdx.x; e < SH::M_END; e += blockDim.x) smem[e] = 0.0;
  __syncthreads();
  for (int t = threadIdx.x; t < 2 * NX; t += blockDim.x) {  // F rows of class 0 and 1
    const int c = t / NX, r = t - c * NX;
    const Ell &E = M.sc[c].F;
    for (int e = 0; e < E.width; e++) {
      const double v = M.val[E.off + r * E.width + e];
      const int j = M.idx[E.off + r * E.width + e];
      if (v != 0.0) {
        Fd[c * A * NX + j * NX + r] = v;
        FdT[c * A * NX + r * A + j] = v;
      }
    }
  }
  for (int t = threadIdx.x; t < 3 * ROWS; t += blockDim.x) {  // C rows
    const int c = t / ROWS, r = t - c * ROWS;
    const Ell &E = M.sc[c].C;
    if (r < E.rows)
      for (int e = 0; e < E.width && e < CW; e++) {
        Cv[t * CW + e] = M.val[E.off + r * E.width + e];
        Ci[t * CW + e] = M.idx[E.off + r * E.width + e];
      }
  }
  for (int t = threadIdx.x; t < 3 * SH::AE; t += blockDim.x) {  // C' rows
    const int c = t / SH::AE, r = t - c * SH::AE;
    const Ell &E = M.sc[c].Ct;
    if (r < E.rows)
      for (int e = 0; e < E.width && e < CW; e++) {
        Ctv[t * CW + e] = M.val[E.off + r * E.width + e];
        Cti[t * CW + e] = M.idx[E.off + r * E.width + e];
      }
  }

  // ---- per-group block: ring, barriers, scratch ----
  double *gb = smem + SH::M_END + (size_t)grp * SH::group_doubles(D);
  double *ring = gb;
  uint64_t *bars = reinterpret_cast<uint64_t *>(gb + D * RS);
  double *sc = gb + D * RS + D;
  double *w = sc + SH::O_W, *tv2 = sc + SH::O_TV2, *yvs = sc + SH::O_YV, *td2 = sc + SH::O_TD2, *sp = sc + SH::O_SP,
         *vn2 = sc + SH::O_VN2, *vcur = sc + SH::O_VC, *nudv = sc + SH::O_NUD, *gg2 = sc + SH::O_GG2;
  Scratch rs;  // residual routines alias the sweep scratch; product_sweep touches v, w, vn, td, sp only
  rs.v = sc; rs.w = sc + SH::VM; rs.vn = sc + 2 * SH::VM; rs.td = sc + 3 * SH::VM; rs.sp = sc + 4 * SH::VM;
  rs.g = rs.yv = rs.nud = nullptr;
  if (gl == 0)
    for (int d = 0; d < D; d++) mbar_init(&bars[d], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const uint32_t ring_s = smem_u32(ring), bars_s = smem_u32(bars);
  // Concurrent latency-mode lane (k_l > 0): as the launch runs out of instances the CTAs on SMs 0 .. k_l-1 retire one
  // SM after the other and latency-mode CTAs (k_admm_cta, launched beside this kernel and pending until an SM is
  // free) take those SMs; the streaming groups hand them their most-unconverged instances, whose chains would
  // otherwise bound the tail, and they also resume suspended instances, most-unconverged first.
  unsigned smid = 0, nsmid = 0;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  asm volatile("mov.u32 %0, %%nsmid;" : "=r"(nsmid));
  if (threadIdx.x == 0 && k_l > 0) atomicAdd(B.sched + kSchedStarted, 1);

  // lane constants of the diagonal-packed mat-vecs
  Wraps<SH> wr;
  wr.init(gl);
  const int (&wrapS)[SH::WNS] = wr.S;

  const int npos = 2 * (N + 1);
  const double sigma = S.sigma, alpha = S.alpha, oma = 1.0 - S.alpha;
  const bool sched_on = s0 > 0 && S.check_termination > 0;
  // ring state: consumer slot/phase, and the position (0..npos-1) of the next packet to issue
  int slot = 0, ipos = 0;
  uint32_t phase = 0;

  // per-instance state
  int inst = -1, iter = 0;
  bool have = false;
  const double *recb = nullptr;
  double *x = nullptr, *z = nullptr, *y = nullptr, *dx = nullptr, *dy = nullptr, *ykg = nullptr, *skg = nullptr;
  double rho_in = 0.0;
  RhoClasses rcl;
  MutVec<NYT> cur;

  const uint64_t pol_first = l2_policy_evict_first(), pol_last = l2_policy_evict_last();
  // leader only: issue the packet at position ipos into ring slot `s`, advance ipos
  auto issue = [&](int s) {
    const bool fwd = ipos <= N;
    const int k = fwd ? ipos : npos - 1 - ipos;
    const double *src = recb + (size_t)k * RS + (fwd ? 0 : SH::XOFF);
    const uint32_t bytes = fwd ? SH::FWD_BYTES : SH::BWD_BYTES;
    RLDL_ASSERT(s >= 0 && s < D);                                             // ring slot
    RLDL_ASSERT(k >= 0 && k <= N && ipos >= 0 && ipos < npos);                // stage of the packet
    RLDL_ASSERT((long long)k * RS + (fwd ? 0 : SH::XOFF) + bytes / 8 <= B.rec_stride);  // inside the instance's records
    RLDL_ASSERT(recb >= B.rec && recb + B.rec_stride <= B.rec + (size_t)B.batch * B.rec_stride);
    const uint32_t bar = bars_s + 8u * s;
    mbar_expect_tx(bar, bytes);
    if (l2_hints & 1) bulk_g2s_hint(ring_s + 8u * (s * RS + (fwd ? 0 : SH::XOFF)), src, bytes, bar, pol_first);
    else bulk_g2s(ring_s + 8u * (s * RS + (fwd ? 0 : SH::XOFF)), src, bytes, bar);
    ipos = (ipos + 1 == npos) ? 0 : ipos + 1;
  };
  // Mutable per-stage values are register-prefetched one stage step ahead by the lane that owns
  // (and wrote) them.  L2-only loads (ld.global.cg): an instance may have been advanced by a group
  // on another SM since this SM last touched it, and L1 is not coherent across SMs.
  auto load_mv = [&](MutVec<NYT> &r, int k, bool bwd) {
    const bool term = (k == N);
    const int dim = term ? NX : A, np = term ? NT : NY;
    const int xoff = k * A - NX, roff = k * BQ, ilo = (k == 0) ? NX : 0;
    r.x = (gl >= ilo && gl < dim) ? __ldcg(x + xoff + gl) : 0.0;
#pragma unroll
    for (int t = 0; t < NYT; t++) {
      const int rr = gl + G * t;
      r.zp[t] = 0.0; r.yp[t] = 0.0;
      if (rr < np) { r.zp[t] = __ldcg(z + roff + rr); r.yp[t] = __ldcg(y + roff + rr); }
    }
    r.zd = r.yd = r.yk = r.sk = 0.0;
    if (!term && gl < NX) { r.zd = __ldcg(z + roff + NY + gl); r.yd = __ldcg(y + roff + NY + gl); }
    if (bwd && !term) {
      if (l2_hints & 2) {  // last use of the round trip: the line may go
        if (gl < A) r.yk = ldcg_hint(ykg + k * A + gl, pol_first);
        if (gl < NX) r.sk = ldcg_hint(skg + k * NX + gl, pol_first);
      } else {
        if (gl < A) r.yk = __ldcg(ykg + k * A + gl);
        if (gl < NX) r.sk = __ldcg(skg + k * NX + gl);
      }
    }
  };
  // claim an instance: fresh ones first, then the most urgent suspended one, provided its
  // bucket is above `above`.  Returns the instance (group-uniform) or -1; *fresh tells which.
  auto tail_reached = [&]() -> bool {  // few enough instances left: they go to the latency-mode kernel
    int t = 0;
    if (park_below > 0) {
      // only once every instance has been started by this kernel: the latency-mode kernel resumes suspended
      // instances, it never starts one (an instance that was never claimed has no saved iteration count)
      if (gl == 0) t = ld_volatile(B.sched) >= B.batch && ld_volatile(B.sched + kSchedUnfinished) <= park_below;
      t = __shfl_sync(gm, t, 0, G);
    }
    return t != 0;
  };
  auto retiring = [&]() -> bool {  // this SM goes to a latency-mode CTA
    int t = 0;
    if (k_l > 0 && (int)smid < k_l) {
      // (a progressive variant — one SM given up per 16 unfinished instances, up to 132 lane CTAs — was measured and
      // rejected: a latency-mode CTA sustains 60 k instance-iterations/s against 127-155 k of a streaming SM, so SMs
      // given up while the streaming kernel still has work lengthen it by more than the shorter tail returns:
      // 8 192 instances 190 -> 209 ms, 16 384: 299 -> 334 ms; profiles/r02_i_lane_progressive_trace.txt)
      if (gl == 0) t = ld_volatile(B.sched) >= B.batch && ld_volatile(B.sched + kSchedUnfinished) <= retire_below;
      t = __shfl_sync(gm, t, 0, G);
    }
    return t != 0;
  };
  auto acquire = [&](int above, bool *fresh) -> int {
    for (;;) {
      if (tail_reached() || retiring()) return -1;
      int i = -1;
      if (gl == 0 && ld_volatile(B.sched) < B.batch) {
        i = atomicAdd(B.sched, 1);
        if (i >= B.batch) i = -1;
      }
      i = __shfl_sync(gm, i, 0, G);
      if (i >= 0) {
        bool skip = false;
        if (B.fact_flag[i]) {  // failed factorisation: report non-convex, like a diverged solve
          const bool report = !resume || __ldcg(B.status + i) == -10;
          if (gl == 0 && report) {
            B.status[i] = -7; B.iter[i] = 0; B.obj[i] = kOsqpNan;
            atomicSub(B.sched + kSchedUnfinished, 1);
          }
          if (report) deliver_to_host(B, i, gl, G, n, m, -7, 0, kOsqpNan, 0.0, 0.0, 0.0);
          skip = true;
        } else if (resume && __ldcg(B.status + i) != -10) {
          skip = true;  // finished in an earlier launch of this solve
        }
        if (skip) continue;
        *fresh = true;
        return i;
      }
      if (!sched_on) return -1;
      const int best = sched_best<G>(B, gl, gm);
      if (best <= above) return -1;
      int v = -1;
      if (gl == 0) v = sched_pop(B, best);
      v = __shfl_sync(gm, v, 0, G);
      if (v >= 0) {
        *fresh = false;
        return v;
      }
    }
  };
  // take the next instance and set up pointers, iterates, ring and registers
  auto start_next = [&]() -> bool {
    bool fresh = true;
    const int i = acquire(-1, &fresh);
    if (i < 0) return false;
    inst = i;
    RLDL_ASSERT(inst >= 0 && inst < B.batch);
    __threadfence();  // order the loads below after the hand-over of the instance
    const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
    x = B.x + vo; z = B.z + ro; y = B.y + ro; dx = B.dx + vo; dy = B.dy + ro;
    ykg = B.yk + (size_t)inst * B.ykstride; skg = B.sk + (size_t)inst * B.skstride;
    recb = B.rec + (size_t)inst * B.rec_stride;
    rho_in = __ldcg(B.rho + inst);
    rcl.set(rho_in);
    ipos = 0;
    if (gl == 0) {
      int s = slot;
      for (int d = 0; d < D; d++) {
        issue(s);
        s = (s + 1 == D) ? 0 : s + 1;
      }
    }
    ipos = __shfl_sync(gm, ipos, 0, G);
    if (fresh && !resume) {
      if (!warm_start) {
        for (int i2 = gl; i2 < n; i2 += G) x[i2] = 0.0;
        for (int r = gl; r < m; r += G) { z[r] = 0.0; y[r] = 0.0; }
      }
      if (gl == 0) { B.status[inst] = -10; B.ratio_prev[inst] = 0.0; }
      iter = 0;
    } else {
      iter = (int)__ldcg(B.iter + inst);
    }
    __syncwarp(gm);
    load_mv(cur, 0, false);
    return true;
  };
  auto consume_wait = [&]() { mbar_wait(bars_s + 8u * slot, phase); };
  auto consume_done = [&]() {  // whole group has finished reading the slot: refill it D packets ahead
    if (gl == 0) issue(slot);
    else ipos = (ipos + 1 == npos) ? 0 : ipos + 1;
    slot++;
    if (slot == D) { slot = 0; phase ^= 1u; }
  };
  auto drain = [&]() {  // consume the D packets still in flight for the instance being left
    for (int d = 0; d < D; d++) {
      consume_wait();
      slot++;
      if (slot == D) { slot = 0; phase ^= 1u; }
    }
    __syncwarp(gm);
  };

  unsigned long long my_iters = 0;  // ADMM iterations this group executed (summed into sched[kSchedStreamIters] at exit)
  have = start_next();
  while (__any_sync(FULL, have)) {
    if (have) {
      iter++;
      my_iters++;
      const bool can_check = S.check_termination && (iter % S.check_termination == 0);
      const bool can_adapt = S.adaptive_rho && S.adaptive_rho_interval && (iter % S.adaptive_rho_interval == 0);
      const bool time_due = time_check_due(S, iter, can_check);
      const bool store_delta = can_check || iter == S.max_iter || time_due;
      double yv_reg = 0.0;  // this lane's entry of Xinv*t of the last forward stage
      MutVec<NYT> nxt;
      // =============================== forward sweep ===============================
      for (int k = 0; k <= N; k++) {
        const bool term = (k == N);
        const int cls = term ? 2 : (k == 0 ? 0 : 1);
        const int dim = term ? NX : A, np = term ? NT : NY;
        const int ilo = (k == 0) ? NX : 0;
        if (!term) load_mv(nxt, k + 1, false);
        else load_mv(nxt, N, true);
        const double *sl_ = ring + slot * RS;
        consume_wait();
        double g = 0.0;
        if (gl >= ilo && gl < dim) g = rec_sigma<SH>(sl_, gl, sigma) * cur.x - sl_[gl];
#pragma unroll
        for (int t = 0; t < NYT; t++) {
          const int rr = gl + G * t;
          if (rr < np) w[rr] = rec_rho<SH>(sl_, rr, rcl) * cur.zp[t] - cur.yp[t];
        }
        __syncwarp(gm);
        double tt = 0.0;
        if (gl < dim) {
          tt = g + ell_row<CW>(Ctv + cls * SH::AE * CW, Cti + cls * SH::AE * CW, gl, w);
          if (k > 0 && gl < NX) tt -= sp[gl];
        }
        if (gl < A) v2_store<SH>(tv2, gl, tt);
        __syncwarp(gm);
        double yvv = 0.0;
        if (gl < A) {
          yvv = xinv_mv<SH>(sl_, gl, tv2, wr);
          yvs[gl] = yvv;
          if (gl < dim) {
            if (l2_hints & 2) st_hint(ykg + k * A + gl, yvv, pol_last);
            else ykg[k * A + gl] = yvv;
          }
        }
        yv_reg = yvv;
        __syncwarp(gm);
        if (!term) {
          if (gl < NX) {
            const double tdv = (cur.zd - cur.yd * rec_rinv<SH>(sl_, NY + gl, rcl)) -
                               dense_mv<A, NX>(Fd + (k == 0 ? 0 : A * NX) + gl, yvs);
            td2[gl] = tdv; td2[gl + NX] = tdv;
          }
          __syncwarp(gm);
          if (gl < NX) {
            const double sv = dmv<NX, SH::WNS>(sl_ + SH::SOFF, sl_ + SH::SOFF + gl, td2 + gl, wrapS);
            sp[gl] = sv;
            if (l2_hints & 2) st_hint(skg + k * NX + gl, sv, pol_last);
            else skg[k * NX + gl] = sv;
          }
          __syncwarp(gm);
        }
        consume_done();
        cur = nxt;
      }
      // =============================== backward sweep ==============================
      for (int k = N; k >= 0; k--) {
        const bool term = (k == N);
        const int cls = term ? 2 : (k == 0 ? 0 : 1);
        const int dim = term ? NX : A, np = term ? NT : NY;
        const int xoff = k * A - NX, roff = k * BQ, ilo = (k == 0) ? NX : 0;
        if (k > 0) load_mv(nxt, k - 1, true);
        const double *sl_ = ring + slot * RS;
        consume_wait();
        double vv = 0.0, nud = 0.0;
        if (term) {
          vv = yv_reg;
        } else {
          if (gl < NX) {
            nud = -(cur.sk + dmv<NX, SH::WNS>(sl_ + SH::SOFF, sl_ + SH::SOFF + gl, vn2 + gl, wrapS));
            nudv[gl] = nud;
          }
          __syncwarp(gm);
          if (gl < A) {
            const double ggv = dense_mv<NX, A>(FdT + (k == 0 ? 0 : A * NX) + gl, nudv);
            v2_store<SH>(gg2, gl, ggv);
          }
          __syncwarp(gm);
          if (gl < A) vv = cur.yk - xinv_mv<SH>(sl_, gl, gg2, wr);
        }
        if (gl < A) vcur[gl] = vv;
        if (gl < NX) { vn2[gl] = vv; vn2[gl + NX] = vv; }
        __syncwarp(gm);
        MutVec<NYT> nw = cur;  // new iterate values of this stage (stage 0: carried into the next forward sweep)
#pragma unroll
        for (int t = 0; t < NYT; t++) {
          const int rr = gl + G * t;
          if (rr < np) {
            const int ri = roff + rr;
            const double zt = ell_row<CW>(Cv + cls * ROWS * CW, Ci + cls * ROWS * CW, rr, vcur);
            double rho, rinv;
            rec_rho_pair<SH>(sl_, rr, rcl, rho, rinv);
            const double zo = cur.zp[t], yo = cur.yp[t];
            const double zh = alpha * zt + oma * zo;
            const double zn = fmin(fmax(zh + yo * rinv, sl_[SH::LO + rr]), sl_[SH::UP + rr]);
            const double d = rho * (zh - zn);
            z[ri] = zn; y[ri] = yo + d;
            nw.zp[t] = zn; nw.yp[t] = yo + d;
            if (store_delta) dy[ri] = d;
          }
        }
        if (!term && gl < NX) {
          const int ri = roff + NY + gl;
          double rho, rinv;
          rec_rho_pair<SH>(sl_, NY + gl, rcl, rho, rinv);
          const double zo = cur.zd, yo = cur.yd;
          const double zt = (zo - yo * rinv) + nud * rinv;
          const double zh = alpha * zt + oma * zo;
          const double zn = fmin(fmax(zh + yo * rinv, sl_[SH::LO + NY + gl]), sl_[SH::UP + NY + gl]);
          const double d = rho * (zh - zn);
          z[ri] = zn; y[ri] = yo + d;
          nw.zd = zn; nw.yd = yo + d;
          if (store_delta) dy[ri] = d;
        }
        if (gl >= ilo && gl < dim) {
          const double xo = cur.x;
          const double xn = alpha * vv + oma * xo;
          x[xoff + gl] = xn;
          nw.x = xn;
          if (store_delta) dx[xoff + gl] = xn - xo;
        }
        if (k == 0) nxt = nw;
        __syncwarp(gm);
        consume_done();
        cur = nxt;
      }
      // ---- iteration epilogue: residuals, termination, rho adaptation (osqp.c:426-519, 524-567) ----
      bool done = false, rho_changed = false, at_tail = false;
      int status = -10;
      double obj = 0.0, rho_new = rho_in;
      Resid R = {0, 0, 0, 0, 0, 0, 0, 0}, Rs = R;  // Rs: the norms of the scaled problem (rho estimate), = R unless SC
      if (can_check) {
        R = compute_residuals<G, SH::RZ != 0>(M, B, inst, gl, gm, rs, &Rs);
        obj = R.obj;
        if (check_termination<G>(M, B, S, inst, gl, gm, rs, R, 0, &status, &obj)) done = true;
      }
      if (!done && can_adapt) {  // adapt_rho (auxil.c:57-77); the refactorisation itself is a separate launch
        if (!can_check) { R = compute_residuals<G, SH::RZ != 0>(M, B, inst, gl, gm, rs, &Rs); obj = R.obj; }
        const double rn = rho_estimate(Rs, rho_in);
        if (rn > rho_in * S.adaptive_rho_tolerance || rn < rho_in / S.adaptive_rho_tolerance) {
          rho_changed = true;
          rho_new = fmin(fmax(rn, kRhoMin), kRhoMax);
        }
      }
      bool timed_out = false;
      if (!done && time_due && iter != S.max_iter) {  // osqp.c:389-405
        int up = gl == 0 ? (int)time_is_up(B, S) : 0;
        timed_out = __shfl_sync(gm, up, 0, G) != 0;
      }
      if (!done && (iter == S.max_iter || timed_out)) {
        if (!can_check) {
          R = compute_residuals<G, SH::RZ != 0>(M, B, inst, gl, gm, rs, &Rs);
          obj = R.obj;
          check_termination<G>(M, B, S, inst, gl, gm, rs, R, 0, &status, &obj);
        }
        if (status == -10) {  // osqp.c:562-576: approximate check, else max_iter / time limit reached
          if (!check_termination<G>(M, B, S, inst, gl, gm, rs, R, 1, &status, &obj)) status = timed_out ? -6 : -2;
        }
        done = true;
      }
      if (rho_changed && gl == 0) {
        B.rho[inst] = rho_new;
        B.rho_updates[inst] += 1;
        atomicAdd(B.sched + 1, 1);
      }
      if (done) {
        const bool has_sol = !(status == -3 || status == 3 || status == -4 || status == 4 || status == -7);
        if (has_sol) obj = R.obj;
        const double rho_est = rho_estimate(Rs, rho_new);
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
          atomicSub(B.sched + kSchedUnfinished, 1);
          trace_finish(B, inst, 1);
        }
        if (B.h_status) {  // fused solve-and-deliver: this instance's results go to the host now
          __syncwarp(gm);
          deliver_to_host(B, inst, gl, G, n, m, status, iter, obj, R.pri, R.dua, rho_est);
        }
        drain();
        have = start_next();
      } else if (rho_changed) {
        // parked: the host re-types rho_vec, re-factorises and re-packs this instance, then resumes it;
        // until then no kernel of this solve may pick it up (k_set_rho_vec turns the mark back into UNSOLVED)
        if (gl == 0) { B.iter[inst] = iter; B.status[inst] = kStatusParkedRho; }
        drain();
        have = start_next();
      } else if (can_check && ((at_tail = tail_reached()) || retiring())) {
        // suspended as it stands (status stays OSQP_UNSOLVED) into its priority bucket: the latency-mode kernel
        // resumes the suspended instances most-unconverged first (the longest remaining chains start first);
        // a retiring SM's instances are picked up by the other groups like any suspended instance
        if (gl == 0) {
          B.iter[inst] = iter;
          if (k_l > 0 && at_tail) atomicExch(B.sched + kSchedTailFlag, 1);
          if (k_l > 0 && !at_tail) atomicAdd(B.sched + kSchedStats + 4, 1);  // instances suspended by retiring SMs
        }
        __threadfence();
        __syncwarp(gm);
        if (gl == 0) {
          const double ratio = urgency_ratio(S, R);
          sched_push(B, urgency_bucket(S, ratio, B.ratio_prev[inst], B.sched_key), inst);
          B.ratio_prev[inst] = ratio;
        }
        drain();
        have = false;
      } else {
        bool yield = false, handed = false;
        int myb = 0;
        if (sched_on && can_check) {
          // urgency of this instance: distance from its tolerances (auxil.c:259-288, 323-362)
          {
            const double ratio = urgency_ratio(S, R);
            if (gl == 0) {
              myb = urgency_bucket(S, ratio, B.ratio_prev[inst], B.sched_key);
              B.ratio_prev[inst] = ratio;  // (read again only after a hand-over of the instance, behind the fence)
            }
            myb = __shfl_sync(gm, myb, 0, G);
          }
          if (k_l > 0) {
            // hand-over to a waiting latency-mode CTA: only the most-unconverged instances qualify (kSchedTop is a
            // decaying maximum of the buckets seen at the checks; the waiting CTAs lower it until someone qualifies)
            int give = 0;
            if (gl == 0) {
              // kSchedTop holds (kSchedBuckets - 1 - threshold), so that an all-zero reset means "top bucket only"
              const int inv = kSchedBuckets - 1 - myb;
              atomicMin(B.sched + kSchedTop, inv);
              if (myb >= 2 && ld_volatile(B.sched + kSchedLWant) > 0 && inv <= ld_volatile(B.sched + kSchedTop)) {
                if (atomicSub(B.sched + kSchedLWant, 1) > 0) give = 1;
                else atomicAdd(B.sched + kSchedLWant, 1);
              }
            }
            give = __shfl_sync(gm, give, 0, G);
            if (give) {
              if (gl == 0) B.iter[inst] = iter;
              __threadfence();  // this lane's iterate stores precede the hand-over
              __syncwarp(gm);
              if (gl == 0) { sched_push(B, kSchedLQ, inst); atomicAdd(B.sched + kSchedStats, 1); }
              drain();
              have = start_next();
              handed = true;
            }
          }
          if (!handed) {
            int fresh_left = 0;
            if (gl == 0) fresh_left = ld_volatile(B.sched) < B.batch;
            fresh_left = __shfl_sync(gm, fresh_left, 0, G);
            if (fresh_left) yield = iter >= s0;
            else yield = sched_best<G>(B, gl, gm) > myb;
          }
        }
        if (yield) {
          if (gl == 0) B.iter[inst] = iter;
          __threadfence();  // this lane's iterate stores precede the hand-over
          __syncwarp(gm);
          if (gl == 0) sched_push(B, myb, inst);
          drain();
          have = start_next();
        } else {
          __syncwarp(gm);  // residual scratch aliases the sweep scratch
        }
      }
    }
    __syncwarp();
  }
  if (gl == 0 && my_iters)
    atomicAdd(reinterpret_cast<unsigned long long *>(B.sched + kSchedStreamIters), my_iters);
}

// ------------------------------- launcher --------------------------------
template <class SH>
static cudaError_t launch_stream_t(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                                   int resume, int park_below, const StreamTuning &tune, int k_l, int retire_below,
                                   cudaEvent_t reset_done, int *grid_out, cudaStream_t st) {
  const int threads = 128, groups = (threads / 32) * SH::GPW;
  int D = tune.ring_depth > 1 ? tune.ring_depth : 2;  // two stage steps of look-ahead cover the HBM latency
  if (D > 2 * (B.N + 1)) D = 2 * (B.N + 1);
  // fresh instances run up to their first check before they may yield (sched_s0 = 0 turns the scheduler off)
  const int s0 = tune.sched_s0 >= 0 ? tune.sched_s0 : S.check_termination;
  const size_t shm = ((size_t)SH::M_END + (size_t)groups * SH::group_doubles(D)) * sizeof(double);
  if (shm > 220 * 1024) return cudaErrorInvalidValue;
  auto kern = k_admm_stream<SH>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm);
  if (e != cudaSuccess) return e;
  int dev = 0, sms = 148, per_sm = 1;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, shm);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  // 3 CTAs/SM nearly saturate the memory system; a 4th adds 2.5 % of bulk throughput but lengthens every
  // instance's iteration (the machine is shared equally) and with it the chain of the longest instance — a net
  // loss (1 316 vs 1 273 ms at C5) unless the latency-mode kernel takes the tail over (1 147 vs 1 174 ms)
  const int cap = tune.ctas_per_sm > 0 ? tune.ctas_per_sm : (park_below > 0 ? 4 : 3);
  if (per_sm > cap) per_sm = cap;
  long long grid = (long long)sms * per_sm;
  const long long need = ((long long)B.batch + groups - 1) / groups;
  if (grid > need) grid = need;
  // scheduler state: counters to 0, every ring slot to -1 (empty)
  e = cudaMemsetAsync(B.sched, 0, (2 + 2 * kSchedRings) * sizeof(int), st);
  if (e != cudaSuccess) return e;
  // (instances suspended for the latency-mode kernel may be left in the rings of an earlier launch)
  e = cudaMemsetAsync(B.sched_slots, 0xff, (size_t)kSchedRings * B.batch * sizeof(int), st);
  if (e != cudaSuccess) return e;
  // concurrent latency-mode lane: waiting CTAs 0, hand-over threshold at the top bucket (stored inverted), CTAs
  // started 0, tail flag 0
  static_assert(kSchedTop == kSchedLWant + 1 && kSchedStarted == kSchedLWant + 2 && kSchedTailFlag == kSchedLWant + 3, "layout");
  e = cudaMemsetAsync(B.sched + kSchedLWant, 0, (4 + 6 + 2) * sizeof(int), st);  // lane state, statistics, iteration count
  if (e != cudaSuccess) return e;
  if (reset_done) {  // the concurrent latency-mode launch waits for this, not for the kernel
    e = cudaEventRecord(reset_done, st);
    if (e != cudaSuccess) return e;
  }
  if (grid_out) *grid_out = (int)grid;
  kern<<<(unsigned)grid, threads, shm, st>>>(M, B, S, warm_start, D, resume, s0, park_below, tune.l2_hints, k_l,
                                             retire_below);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Shape-specialised stage recursion (reference src/recursive_ldl.c:1139-1318; tail restart
// :946-1110): one instance per group of G lanes, lane i owns ROW i of the current pivot block in
// registers.  Each SPD pivot block is inverted in place by symmetric sweeps (sweep operator:
// after sweeping every pivot the block holds -inverse); the pivot row is broadcast through a
// double-buffered shared-memory line, so a sweep costs A/2 broadcast LDS.128 + A DFMA per lane
// and one __syncwarp.  The inverses are written straight into the stage records in the
// diagonal-packed form the streaming ADMM kernel reads (and, on request, unpacked for the
// generic kernels).  Requires ELL width 1 path rows (C' diag(rho) C diagonal), as the table.
//   M_k = Q_k + sigma I + C' rho C;  X_0 = M_0;  S_k = diag(1/rho_d) + F X_k^-1 F';
//   X_{k+1} = M_{k+1} + J' S_k^-1 J
// ---------------------------------------------------------------------------
template <int NP, int NE>
__device__ __forceinline__ void sweep_rows(double (&X)[NP], double *rowbuf, int row, unsigned gm, int &bad) {
  // X = this lane's row `row` of an NP x NP SPD block (lanes with row outside [0, NP) hold zero rows and stay
  // zero); NE = even_up(NP)
#pragma unroll
  for (int p = 0; p < NP; p++) {
    double *buf = rowbuf + (p & 1) * NE;
    if (row == p) {
#pragma unroll
      for (int j = 0; j < NP; j++) buf[j] = X[j];
    }
    __syncwarp(gm);
    double rp[NP];
#pragma unroll
    for (int j = 0; j < NP; j++) rp[j] = buf[j];
    if (!(rp[p] > 0.0)) bad = 1;
    const double d = 1.0 / rp[p];
    const double f = (row == p) ? (1.0 - d) : X[p] * d;
#pragma unroll
    for (int j = 0; j < NP; j++)
      if (j != p) X[j] = fma(-f, rp[j], X[j]);
    X[p] = (row == p) ? -d : f;
  }
#pragma unroll
  for (int j = 0; j < NP; j++) X[j] = -X[j];
}

// BD (block-diagonal pivots, every BASELINE model): X_k = blkdiag(X^xx_k, X^uu_k), so the state block (lanes < NX)
// and the input block (lanes NX..A-1) are swept separately — 12^3 + 4^3 instead of 16^3 multiply-adds at C5 — and
// W = Xinv F' only multiplies the non-zero block of each row.
template <class SH>
__global__ void __launch_bounds__(128, SH::A <= 16 ? 4 : 3) k_factor_stream(ModelDev M, BatchDev B, double sigma, int k0,
                                                          const int *only_flagged, int write_unpacked) {
  constexpr int A = SH::A, NX = SH::NX, NY = SH::NY, CW = SH::CW, G = SH::G, RS = SH::RS;
  constexpr int AE = SH::AE, NXE = SH::NXE, NU = SH::NU, BD = SH::BD;
  constexpr int XW = BD ? NX : A;                                 // width of the register row of the (state) block
  constexpr int TB = A * A > A * NXE ? A * A : A * NXE;          // transposition buffer (doubles)
  constexpr int M_F = 0, M_FT = M_F + 2 * A * NX, M_Q = M_FT + 2 * A * NX, M_C2 = M_Q + 3 * A * A,
                M_CI = M_C2 + 3 * AE, M_END = (M_CI + (3 * AE + 1) / 2 + 1) & ~1;
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int gl = lane & (G - 1), gw = lane / G;
  const unsigned gm = (G == 32) ? FULL : (((1u << G) - 1u) << (gw * G));
  const int grp = wid * SH::GPW + gw, groups = (blockDim.x >> 5) * SH::GPW;
  const int N = B.N;
  (void)CW;

  // ---- CTA-shared model: dense F / F' (class 0, 1), Q blocks padded to A x A, squared C' entries ----
  double *Fd = smem + M_F, *FdT = smem + M_FT, *Qs = smem + M_Q, *c2 = smem + M_C2;
  int *ci = reinterpret_cast<int *>(smem + M_CI);
  for (int e = threadIdx.x; e < M_END; e += blockDim.x) smem[e] = 0.0;
  __syncthreads();
  for (int t = threadIdx.x; t < 2 * NX; t += blockDim.x) {
    const int c = t / NX, r = t - c * NX;
    const Ell &E = M.sc[c].F;
    for (int e = 0; e < E.width; e++) {
      const double v = M.val[E.off + r * E.width + e];
      const int j = M.idx[E.off + r * E.width + e];
      if (v != 0.0) {
        Fd[c * A * NX + j * NX + r] = v;
        FdT[c * A * NX + r * A + j] = v;
      }
    }
  }
  for (int t = threadIdx.x; t < 3 * A * A; t += blockDim.x) {
    const int c = t / (A * A), e = t - c * A * A, j = e / A, i = e - j * A;
    const int dim = (c == 2) ? NX : A;
    Qs[t] = (i < dim && j < dim) ? M.val[M.sc[c].Q + j * dim + i] : (i == j ? 1.0 : 0.0);
  }
  for (int t = threadIdx.x; t < 3 * AE; t += blockDim.x) {
    const int c = t / AE, r = t - c * AE;
    const Ell &E = M.sc[c].Ct;
    if (r < E.rows && E.width > 0) {
      const double v = M.val[E.off + r * E.width];
      c2[t] = v * v;
      ci[t] = M.idx[E.off + r * E.width];
    }
  }
  __syncthreads();
  // per-group buffer stride = 2 (mod 16) doubles: the broadcast loads of the groups of one warp (pivot row, W, F')
  // then fall into different banks — one wavefront per LDS.128 instead of one per group
  constexpr int GSTRIDE = ((2 * AE + TB + 15) & ~15) + 2;
  double *gbuf = smem + M_END + (size_t)grp * GSTRIDE;
  double *rowbuf = gbuf, *tb = gbuf + 2 * AE;
  const int urow = gl - NX;  // this lane's row of the input block (BD), negative on the state lanes

  for (int inst = blockIdx.x * groups + grp; inst < B.batch; inst += gridDim.x * groups) {
    if (only_flagged && !only_flagged[inst]) continue;
    // per-row step sizes from the 2-bit row classes of the record (written by k_vec_records) and the
    // instance's rho: no rho_vec traffic
    RhoClasses rcl;
    rcl.set(B.rho[inst]);
    double *rec = B.rec + (size_t)inst * B.rec_stride;
    RLDL_ASSERT(k0 >= 0 && k0 <= N && (long long)(N + 1) * RS <= B.rec_stride);
    double Sx[NX];  // row gl of Sinv_{k-1} (lanes < NX)
#pragma unroll
    for (int j = 0; j < NX; j++) Sx[j] = 0.0;
    if (k0 > 0 && gl < NX) {  // tail restart: Sinv_{k0-1} from its diagonal-packed record
      const double *dp = rec + (size_t)(k0 - 1) * RS + SH::SOFF;
#pragma unroll
      for (int j = 0; j < NX; j++) {
        int d = j - gl;
        if (d < 0) d += NX;
        Sx[j] = (d <= NX / 2) ? dp[d * NX + gl] : dp[(NX - d) * NX + j];
      }
    }
    int flag = 0;
    for (int k = k0; k <= N; k++) {
      const bool term = (k == N);
      const int cls = term ? 2 : (k == 0 ? 0 : 1);
      const int dim = term ? NX : A;
      double *rk = rec + (size_t)k * RS;
      // ---- X_k row gl (BD: the row of the state block in Xr, the row of the input block in Xu) ----
      double Xr[XW], Xu[BD ? NU : 1];
      double dg = 0.0;
      if (gl < dim) {
        dg = rec_sigma<SH>(rk, gl, sigma);  // (Ruiz-scaled: sigma / (c D_i^2), written by k_vec_records)
        const double cc = c2[cls * AE + gl];
        if (cc != 0.0) dg = fma(cc, rec_rho<SH>(rk, ci[cls * AE + gl], rcl), dg);
      }
      if constexpr (BD) {
#pragma unroll
        for (int j = 0; j < NX; j++) {
          Xr[j] = (gl < NX) ? Qs[cls * A * A + j * A + gl] : 0.0;
          if (j == gl) Xr[j] += dg;
          if (k > 0 && gl < NX) Xr[j] += Sx[j];
        }
#pragma unroll
        for (int j = 0; j < NU; j++) {
          Xu[j] = (urow >= 0 && gl < A) ? Qs[cls * A * A + (NX + j) * A + gl] : 0.0;
          if (j == urow && gl < A) Xu[j] += dg;
        }
      } else {
#pragma unroll
        for (int j = 0; j < A; j++) {
          Xr[j] = (gl < A) ? Qs[cls * A * A + j * A + gl] : 0.0;
          if (j == gl) Xr[j] += dg;
          if (k > 0 && gl < NX && j < NX) Xr[j] += Sx[j < NX ? j : 0];
        }
        Xu[0] = 0.0;
      }
      int bad = 0;
      if constexpr (BD) {
        sweep_rows<NX, NXE>(Xr, rowbuf, gl, gm, bad);
        __syncwarp(gm);  // the second sweep reuses the pivot-row line
        sweep_rows<NU, ((NU + 1) & ~1)>(Xu, rowbuf, urow, gm, bad);
      } else {
        sweep_rows<A, AE>(Xr, rowbuf, gl, gm, bad);
      }
      if (bad && !flag) flag = 1 + k;
      // ---- store Xinv_k: diagonal-packed into the record (and unpacked on request) ----
      __syncwarp(gm);
      if constexpr (BD) {
        if (gl < NX) {
#pragma unroll
          for (int j = 0; j < NX; j++) tb[j * A + gl] = Xr[j];  // symmetric: column-major == row-major
        } else if (gl < A) {
#pragma unroll
          for (int j = 0; j < NU; j++) tb[(NX + j) * A + gl] = Xu[j];
        }
      } else if (gl < A) {
#pragma unroll
        for (int j = 0; j < A; j++) tb[j * A + gl] = Xr[j];
      }
      __syncwarp(gm);
      if constexpr (BD) {  // the nx x nx block, then the nu x nu block (the cross blocks are exactly zero)
        if (gl < NX) {
#pragma unroll
          for (int d = 0; d <= NX / 2; d++) {
            int j = gl + d;
            if (j >= NX) j -= NX;
            rk[SH::XOFF + d * NX + gl] = tb[j * A + gl];
          }
        } else if (gl < A) {
          const int i = gl - NX;
#pragma unroll
          for (int d = 0; d <= NU / 2; d++) {
            int j = i + d;
            if (j >= NU) j -= NU;
            rk[SH::UOFF + d * NU + i] = tb[(NX + j) * A + gl];
          }
        }
      } else if (gl < A) {
#pragma unroll
        for (int d = 0; d <= A / 2; d++) {
          int j = gl + d;
          if (j >= A) j -= A;
          rk[SH::XOFF + d * A + gl] = tb[j * A + gl];
        }
      }
      if (write_unpacked && gl < dim) {
        double *Xk = B.Xinv + (size_t)inst * B.xstride + (size_t)k * A * A + (size_t)gl * dim;
        if constexpr (BD) {
#pragma unroll
          for (int j = 0; j < A; j++)
            if (j < dim) Xk[j] = 0.0;
          if (gl < NX) {
#pragma unroll
            for (int j = 0; j < NX; j++) Xk[j] = Xr[j];
          } else {
#pragma unroll
            for (int j = 0; j < NU; j++) Xk[NX + j] = Xu[j];
          }
        } else {
#pragma unroll
          for (int j = 0; j < A; j++)
            if (j < dim) Xk[j] = Xr[j];
        }
      }
      if (term) break;
      // ---- W = Xinv F' (row gl), S = diag(1/rho_d) + F W (row gl, lanes < NX) ----
      // row gl of F is read from shared memory where it is used (lane-distinct, conflict-free): 32 registers less
      // than holding it across the stage loop, which buys a fourth CTA per SM
      const double *Fg = Fd + (k == 0 ? 0 : A * NX) + (gl < NX ? gl : 0);
      const double *Ft = FdT + (k == 0 ? 0 : A * NX);
      __syncwarp(gm);
      if (gl < A) {
        if constexpr (BD) {
          if (gl < NX) {
#pragma unroll
            for (int r = 0; r < NX; r++) {
              double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
              for (int j = 0; j < NX; j += 2) {
                acc0 = fma(Xr[j], Ft[r * A + j], acc0);
                if (j + 1 < NX) acc1 = fma(Xr[j + 1], Ft[r * A + j + 1], acc1);
              }
              tb[gl * NXE + r] = acc0 + acc1;
            }
          } else {
#pragma unroll
            for (int r = 0; r < NX; r++) {
              double acc = 0.0;
#pragma unroll
              for (int j = 0; j < NU; j++) acc = fma(Xu[j], Ft[r * A + NX + j], acc);
              tb[gl * NXE + r] = acc;
            }
          }
        } else {
#pragma unroll
          for (int r = 0; r < NX; r++) {
            double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
            for (int j = 0; j < A; j += 2) {
              acc0 = fma(Xr[j], Ft[r * A + j], acc0);
              if (j + 1 < A) acc1 = fma(Xr[j + 1], Ft[r * A + j + 1], acc1);
            }
            tb[gl * NXE + r] = acc0 + acc1;
          }
        }
      }
      __syncwarp(gm);
#pragma unroll
      for (int c = 0; c < NX; c++) Sx[c] = 0.0;
#pragma unroll
      for (int i = 0; i < A; i++) {
        const double fi = (gl < NX) ? Fg[i * NX] : 0.0;
#pragma unroll
        for (int c = 0; c < NX; c++) Sx[c] = fma(fi, tb[i * NXE + c], Sx[c]);
      }
      if (gl < NX) {
        const double rinv = rec_rinv<SH>(rk, NY + gl, rcl);
#pragma unroll
        for (int c = 0; c < NX; c++)
          if (c == gl) Sx[c] += rinv;
      }
      bad = 0;
      sweep_rows<NX, NXE>(Sx, rowbuf, gl, gm, bad);
      if (bad && !flag) flag = 1 + k;
      // ---- store Sinv_k ----
      __syncwarp(gm);
      if (gl < NX) {
#pragma unroll
        for (int j = 0; j < NX; j++) tb[j * NX + gl] = Sx[j];
      }
      __syncwarp(gm);
      if (gl < NX) {
#pragma unroll
        for (int d = 0; d <= NX / 2; d++) {
          int j = gl + d;
          if (j >= NX) j -= NX;
          rk[SH::SOFF + d * NX + gl] = tb[j * NX + gl];
        }
        if (write_unpacked) {
          double *Sk = B.Sinv + (size_t)inst * B.sstride + (size_t)k * NX * NX + (size_t)gl * NX;
#pragma unroll
          for (int j = 0; j < NX; j++) Sk[j] = Sx[j];
        }
      }
      __syncwarp(gm);
    }
    if (gl == 0) B.fact_flag[inst] = flag;
  }
}

template <class SH>
static cudaError_t launch_factor_stream_t(const ModelDev &M, const BatchDev &B, double sigma, int k0,
                                          const int *only_flagged, int write_unpacked, cudaStream_t st) {
  constexpr int A = SH::A, NX = SH::NX, AE = SH::AE, NXE = SH::NXE;
  constexpr int TB = A * A > A * NXE ? A * A : A * NXE;
  constexpr int M_END = ((2 * A * NX) * 2 + 3 * A * A + 3 * AE + (3 * AE + 1) / 2 + 1) & ~1;
  const int threads = 128, groups = (threads / 32) * SH::GPW;
  constexpr int GSTRIDE = ((2 * AE + TB + 15) & ~15) + 2;  // as in the kernel
  const size_t shm = ((size_t)M_END + (size_t)groups * GSTRIDE) * sizeof(double);
  auto kern = k_factor_stream<SH>;
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
  kern<<<(unsigned)grid, threads, shm, st>>>(M, B, sigma, k0, only_flagged, write_unpacked);
  return cudaGetLastError();
}

bool stream_supported(const ModelDev &M, int N) {
  if (N < 1) return false;
  const int cw = max_c_width(M);
#define RLDL_X(a_, nx_, ny_, nt_, cw_, bd_, sc_) \
  if (shape_serves(M, cw, a_, nx_, ny_, nt_, cw_, bd_, sc_)) return true;
  RLDL_SHAPES(RLDL_X)
#undef RLDL_X
  return false;
}

cudaError_t launch_admm_stream(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                               int resume, int park_below, const StreamTuning &tune, int k_l, int retire_below,
                               cudaEvent_t reset_done, int *grid_out, cudaStream_t st) {
  const int cw = max_c_width(M);
#define RLDL_X(a_, nx_, ny_, nt_, cw_, bd_, sc_)                                          \
  if (shape_serves(M, cw, a_, nx_, ny_, nt_, cw_, bd_, sc_))      \
    return launch_stream_t<Shape<a_, nx_, ny_, nt_, cw_, bd_, sc_>>(M, B, S, warm_start, resume, park_below, tune, k_l, \
                                                                retire_below, reset_done, grid_out, st);
  RLDL_SHAPES(RLDL_X)
#undef RLDL_X
  return cudaErrorInvalidValue;
}

__global__ void k_set_tail_flag(int *sched) { atomicExch(sched + kSchedTailFlag, 1); }
// stream-ordered "the streaming kernel has ended": latency-mode CTAs still waiting for a hand-over stop waiting
cudaError_t launch_set_tail_flag(const BatchDev &B, cudaStream_t st) {
  k_set_tail_flag<<<1, 1, 0, st>>>(B.sched);
  return cudaGetLastError();
}

cudaError_t preload_lane_kernels() {
  cudaFuncAttributes at;
  return cudaFuncGetAttributes(&at, k_set_tail_flag);
}

cudaError_t launch_vec_records(const ModelDev &M, const BatchDev &B, int k0, int parts, const int *only_flagged,
                               int *changed, double sigma, cudaStream_t st) {
  const int cw = max_c_width(M);
#define RLDL_X(a_, nx_, ny_, nt_, cw_, bd_, sc_)                                          \
  if (shape_serves(M, cw, a_, nx_, ny_, nt_, cw_, bd_, sc_))      \
    return launch_vec_records_t<Shape<a_, nx_, ny_, nt_, cw_, bd_, sc_>>(B, k0, parts, only_flagged, changed, sigma, st);
  RLDL_SHAPES(RLDL_X)
#undef RLDL_X
  return cudaErrorInvalidValue;
}

cudaError_t launch_factor_stream(const ModelDev &M, const BatchDev &B, double sigma, int k0,
                                 const int *only_flagged, int write_unpacked, cudaStream_t st) {
  const int cw = max_c_width(M);
#define RLDL_X(a_, nx_, ny_, nt_, cw_, bd_, sc_)                                          \
  if (shape_serves(M, cw, a_, nx_, ny_, nt_, cw_, bd_, sc_))      \
    return launch_factor_stream_t<Shape<a_, nx_, ny_, nt_, cw_, bd_, sc_>>(M, B, sigma, k0, only_flagged, write_unpacked, st);
  RLDL_SHAPES(RLDL_X)
#undef RLDL_X
  return cudaErrorInvalidValue;
}

}  // namespace rldl
