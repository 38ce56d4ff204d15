// rldl_stream_common.cuh — pieces shared by the shape-specialised kernels (rldl_stream.cu: the
// HBM-streaming ADMM kernel and the stage recursion; rldl_latency.cu: the CTA-per-instance
// latency-mode ADMM kernel): PTX helpers for mbarrier / bulk copies, the compile-time Shape with
// the stage-record layout, and the small mat-vec primitives on diagonal-packed symmetric blocks.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "rldl_device_fn.cuh"

namespace rldl {

// ------------------------------ PTX helpers ------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// 1-D bulk async copy global -> shared (TMA engine); completes on `bar`.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// L2 eviction-priority hints.  The stage records are read once per sweep and never reused before 6.8 GB of other
// records have passed (evict_first); the forward intermediates are written in the forward sweep and read back in
// the backward sweep of the same iteration (evict_last keeps that round trip inside the 126 MB L2).
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void bulk_g2s_hint(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar), "l"(pol)
      : "memory");
}
__device__ __forceinline__ void st_hint(double *p, double v, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.f64 [%0], %1, %2;" ::"l"(p), "d"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ double ldcg_hint(const double *p, uint64_t pol) {
  double v;
  asm volatile("ld.global.cg.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
  return v;
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "RLDL_WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra RLDL_DONE_%=;\n"
      "bra RLDL_WAIT_%=;\n"
      "RLDL_DONE_%=:\n"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}


// ------------------------------ device work scheduler (see rldl_stream.cu) ------------------------------
__device__ __forceinline__ int ld_volatile(const int *p) { return *reinterpret_cast<const volatile int *>(p); }

__device__ __forceinline__ void sched_push(const BatchDev &B, int b, int inst) {
  RLDL_ASSERT(b >= 0 && b < kSchedRings);
  RLDL_ASSERT(inst >= 0 && inst < B.batch);
  const int idx = atomicAdd(B.sched + 2 + kSchedRings + b, 1);
  RLDL_ASSERT(idx - ld_volatile(B.sched + 2 + b) < B.batch);  // an instance sits in at most one ring: never full
  int *sl = B.sched_slots + (size_t)b * B.batch + (idx % B.batch);
  while (atomicCAS(sl, -1, inst) != -1) __nanosleep(32);
}
__device__ __forceinline__ int sched_pop(const BatchDev &B, int b) {
  int *head = B.sched + 2 + b;
  const int *tail = B.sched + 2 + kSchedRings + b;
  for (;;) {
    const int h = ld_volatile(head), t = ld_volatile(tail);
    if (h >= t) return -1;
    if (atomicCAS(head, h, h + 1) == h) {
      int *sl = B.sched_slots + (size_t)b * B.batch + (h % B.batch);
      int v;
      while ((v = atomicExch(sl, -1)) < 0) __nanosleep(32);
      RLDL_ASSERT(v < B.batch);
      return v;
    }
  }
}
__device__ __forceinline__ bool sched_ring_nonempty(const BatchDev &B, int b) {
  return ld_volatile(B.sched + 2 + kSchedRings + b) > ld_volatile(B.sched + 2 + b);
}
// the most urgent suspended instance (thread-level; -1 if every bucket is empty)
__device__ __forceinline__ int sched_pop_best(const BatchDev &B) {
  for (int b = kSchedBuckets - 1; b >= 0; b--)
    while (sched_ring_nonempty(B, b)) {
      const int i = sched_pop(B, b);
      if (i >= 0) return i;
    }
  return -1;
}

// Urgency of an instance at a termination check.  `ratio` = how far it is from its tolerances (auxil.c:259-288,
// 323-362), max(r_p/eps_p, r_d/eps_d).
//  key 0: floor(2*log2(ratio)) — round 1.
//  key 1: the predicted number of remaining iterations.  ADMM converges linearly once the active set has settled, so
//         with the ratio of the previous check, `every` iterations ago, rate = ln(prev/ratio)/every and
//         remaining = ln(ratio)/rate.  On the recorded residual trajectories of 2 048 C5 instances its rank
//         correlation with the true remaining count is 0.94-0.99 from iteration 100 on (0.84-0.87 for key 0), and it
//         names the few 3 000-iteration instances by iteration 200-300, which key 0 does not.  Bucket =
//         floor(3*log2(remaining / every)), clamped; a first check (no previous ratio) assumes the ratio halves per
//         check, a ratio that did not fall counts as max_iter.
__device__ __forceinline__ double urgency_ratio(const SettingsDev &S, const Resid &R) {
  const double ep = S.eps_abs + S.eps_rel * fmax(R.nz, R.nAx);
  const double ed = S.eps_abs + S.eps_rel * fmax(fmax(R.nq, R.nAty), R.nPx);
  return fmax(R.pri / ep, R.dua / ed);
}
__device__ __forceinline__ int urgency_bucket(const SettingsDev &S, double ratio, double prev, int key) {
  if (!(ratio > 1.0)) return 0;
  if (key == 0) return min(kSchedBuckets - 1, (int)(2.0 * log2(ratio)));
  const double every = S.check_termination > 0 ? (double)S.check_termination : 25.0;
  double rem;
  if (!(prev > 0.0)) rem = every * log2(ratio);
  else if (prev > ratio * 1.0001) rem = every * log(ratio) / log(prev / ratio);
  else rem = (double)S.max_iter;
  rem = fmin(rem, (double)S.max_iter);
  return rem > every ? min(kSchedBuckets - 1, (int)(3.0 * log2(rem / every))) : 0;
}

__host__ __device__ inline int even_up(int v) { return (v + 1) & ~1; }
// The per-row step size of an instance takes one of three values (constraint typing, reference
// src/auxil.c:79-101): rho (inequality), 1e3*rho (equality), 1e-6 (loose).  The stage records carry the
// class of every row in 2 bits (kClsWords 64-bit words per stage) instead of the value: 26 fewer doubles
// per stage and sweep direction at the C5 shape.
__device__ __forceinline__ int row_class(const double *cls, int row) {
  const unsigned long long wv = reinterpret_cast<const unsigned long long *>(cls)[row >> 5];
  return (int)((wv >> (2 * (row & 31))) & 3ull);
}
// Step size of row `row` of a stage and its reciprocal, from the record: the row's class and the instance's class
// values (rho, 1e3 rho, 1e-6, reciprocals precomputed as in the reference's rho_inv_vec, auxil.c:96), or — Ruiz-scaled
// variants — the stored per-row value.
struct RhoClasses {
  double rho_in, rho_eq, inv_in, inv_eq, inv_min;
  __device__ __forceinline__ void set(double r) {
    rho_in = r; rho_eq = kRhoEq * r;
    inv_in = 1.0 / rho_in; inv_eq = 1.0 / rho_eq; inv_min = 1.0 / kRhoMin;
  }
};
template <class SH>
__device__ __forceinline__ void rec_rho_pair(const double *rec, int row, const RhoClasses &rc, double &rho, double &rinv) {
  if constexpr (SH::RZ) {
    rho = rec[SH::CLS + row];
    rinv = 1.0 / rho;
  } else {
    const int c = row_class(rec + SH::CLS, row);
    rho = c == 0 ? rc.rho_in : (c == 1 ? rc.rho_eq : kRhoMin);
    rinv = c == 0 ? rc.inv_in : (c == 1 ? rc.inv_eq : rc.inv_min);
  }
}
template <class SH>
__device__ __forceinline__ double rec_rho(const double *rec, int row, const RhoClasses &rc) {
  if constexpr (SH::RZ) return rec[SH::CLS + row];
  else {
    const int c = row_class(rec + SH::CLS, row);
    return c == 0 ? rc.rho_in : (c == 1 ? rc.rho_eq : kRhoMin);
  }
}
template <class SH>
__device__ __forceinline__ double rec_rinv(const double *rec, int row, const RhoClasses &rc) {
  if constexpr (SH::RZ) return 1.0 / rec[SH::CLS + row];
  else {
    const int c = row_class(rec + SH::CLS, row);
    return c == 0 ? rc.inv_in : (c == 1 ? rc.inv_eq : rc.inv_min);
  }
}
template <class SH>
__device__ __forceinline__ double rec_sigma(const double *rec, int i, double sigma) {
  if constexpr (SH::RZ) return rec[SH::SIG + i];
  else return sigma;
}

// doubles of the diagonal-packed form of a symmetric n x n matrix
__host__ __device__ inline int dp_size(int n) { return (n / 2 + 1) * n; }

// ---------------------------------------------------------------------------
// Compile-time shape of one specialisation
// ---------------------------------------------------------------------------
// BD = 1: the cost has no state-input cross terms (and the path rows touch one variable each, CW = 1), so every
// pivot block X_k = Q_k + sigma I + C' rho C + [Sinv_{k-1} 0; 0 0] is block diagonal, and so is its inverse: the
// record then carries the nx x nx and the nu x nu diagonal block instead of the a x a block (C5: 96 instead of
// 144 doubles per stage and sweep direction).
// RZ = 1: Ruiz-scaled handle (settings.scaling > 0).  The scaled iteration is the unscaled one with a sigma per
// variable and a step size per row, so the record carries sigma_i = sigma / (c D_i^2) behind q and the
// per-row step sizes rho_j in place of the 2-bit row classes (C5: 240 / 264 instead of 198 / 238 doubles per stage
// and sweep direction).
template <int A_, int NX_, int NY_, int NT_, int CW_, int BD_ = 0, int RZ_ = 0>
struct Shape {
  static constexpr int A = A_, NX = NX_, NY = NY_, NT = NT_, CW = CW_, BD = BD_, RZ = RZ_;
  static constexpr int NU = A - NX;
  static constexpr int B = NX + NY;
  static constexpr int G = A <= 8 ? 8 : (A <= 16 ? 16 : 32);
  static constexpr int GPW = 32 / G;
  static constexpr int RMAX = NY > NT ? NY : NT;
  static constexpr int NYT = (RMAX + G - 1) / G;
  static constexpr int ROWS = NYT * G;
  static constexpr int PXX = ((NX / 2 + 1) * NX + 1) & ~1, PUU = ((NU / 2 + 1) * NU + 1) & ~1;
  static constexpr int QA = (A + 1) & ~1, PX = BD ? PXX + PUU : (((A / 2 + 1) * A + 1) & ~1),
                       PS = ((NX / 2 + 1) * NX + 1) & ~1, BB = (B + 1) & ~1;
  // stage record: [ q | (RZ: sigma per variable) | Xinv (diagonal-packed; BD: its xx block, then its uu block) |
  //                 Sinv (diagonal-packed) | row classes (2 bits per row; RZ: step size per row) | l | u ]
  static_assert(B <= 32 * kClsWords, "row classes of a stage must fit the class words of the record");
  static constexpr int SIGW = RZ ? QA : 0, CLSW = RZ ? BB : kClsWords;
  static constexpr int RS = QA + SIGW + PX + PS + CLSW + 2 * BB;
  static constexpr int SIG = QA, XOFF = QA + SIGW, UOFF = XOFF + PXX, SOFF = XOFF + PX, CLS = XOFF + PX + PS,
                       LO = CLS + CLSW, UP = LO + BB;
  static constexpr uint32_t FWD_BYTES = 8u * (QA + SIGW + PX + PS + CLSW), BWD_BYTES = 8u * (PX + PS + CLSW + 2 * BB);
  // per-group scratch (doubles, every block even-sized)
  static constexpr int AE = (A + 1) & ~1, NXE = (NX + 1) & ~1;
  // the vectors of the forward sweep (w, tv2, yv, td2, sp) and of the backward sweep (vn2, vc, nud, gg2)
  // are never live together, so the two sets share one region (4 CTAs/SM need <= 56 KB per CTA)
  static constexpr int O_W = 0, O_TV2 = O_W + ROWS, O_YV = O_TV2 + 2 * AE, O_TD2 = O_YV + AE, O_SP = O_TD2 + 2 * NXE,
                       SC_FWD = O_SP + NXE;
  static constexpr int O_VN2 = 0, O_VC = O_VN2 + 2 * NXE, O_NUD = O_VC + AE, O_GG2 = O_NUD + NXE, SC_BWD = O_GG2 + 2 * AE;
  static constexpr int SC_SWEEP = SC_FWD > SC_BWD ? SC_FWD : SC_BWD;
  static constexpr int VM = (A > RMAX ? A : RMAX);
  // the residual routines (5 vectors of VM doubles) alias the sweep scratch between iterations
  static constexpr int SC = SC_SWEEP > 5 * VM ? SC_SWEEP : 5 * VM;
  // CTA-shared model block (doubles)
  static constexpr int M_FD = 0, M_FDT = M_FD + 2 * A * NX, M_CV = M_FDT + 2 * A * NX, M_CTV = M_CV + 3 * ROWS * CW,
                       M_CI = M_CTV + 3 * AE * CW,                       // ints from here: 3*ROWS*CW
                       M_CTI = M_CI + (3 * ROWS * CW + 1) / 2,            // ints: 3*AE*CW
                       M_END = (M_CTI + (3 * AE * CW + 1) / 2 + 1) & ~1;
  static constexpr int WNX = (A - 1 - A / 2) > 0 ? (A - 1 - A / 2) : 1;
  static constexpr int WNS = (NX - 1 - NX / 2) > 0 ? (NX - 1 - NX / 2) : 1;
  static constexpr int WNU = (NU - 1 - NU / 2) > 0 ? (NU - 1 - NU / 2) : 1;
  static_assert(NU >= 1, "a stage has at least one input");
  __host__ __device__ static int group_doubles(int D) { return ((D * RS + D + SC) + 1) & ~1; }
};

// y_i = sum_j M(i,j) v_j from the diagonal-packed form.  Dpi = DP + i, v2i = v2 + i (v doubled),
// wrap[d-1] = d*N + (i-d) mod N.
template <int N, int WN>
__device__ __forceinline__ double dmv(const double *__restrict__ Dp, const double *__restrict__ Dpi,
                                      const double *__restrict__ v2i, const int (&wrap)[WN]) {
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
  for (int d = 0; d <= N / 2; d++) acc[d & 3] = fma(Dpi[d * N], v2i[d], acc[d & 3]);
#pragma unroll
  for (int d = 1; d <= N - 1 - N / 2; d++)
    acc[(N / 2 + d) & 3] = fma(Dp[wrap[d - 1]], v2i[N - d], acc[(N / 2 + d) & 3]);
  return (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

// The doubled operand of an Xinv product: v2[i .. i+n) = v[(i+d) mod n] for the block lane gl belongs to.
template <class SH>
__device__ __forceinline__ void v2_store(double *v2, int gl, double v) {
  if constexpr (SH::BD) {
    if (gl < SH::NX) {
      v2[gl] = v; v2[gl + SH::NX] = v;
    } else {
      v2[2 * SH::NX + (gl - SH::NX)] = v; v2[2 * SH::NX + SH::NU + (gl - SH::NX)] = v;
    }
  } else {
    v2[gl] = v; v2[gl + SH::A] = v;
  }
}

// wrap-around offsets of lane gl for dmv: of the a x a block, of an nx x nx block, of the nu x nu block
template <class SH>
struct Wraps {
  int X[SH::WNX], S[SH::WNS], U[SH::WNU];
  __device__ __forceinline__ void init(int gl) {
    constexpr int A = SH::A, NX = SH::NX, NU = SH::NU;
#pragma unroll
    for (int d = 1; d <= A - 1 - A / 2; d++) X[d - 1] = d * A + ((gl < A ? gl : 0) - d + A) % A;
#pragma unroll
    for (int d = 1; d <= NX - 1 - NX / 2; d++) S[d - 1] = d * NX + ((gl < NX ? gl : 0) - d + NX) % NX;
    const int iu = (gl >= NX && gl < A) ? gl - NX : 0;
#pragma unroll
    for (int d = 1; d <= NU - 1 - NU / 2; d++) U[d - 1] = d * NU + (iu - d + NU) % NU;
  }
};

// (Xinv_k v)_gl for gl < A, v doubled in v2 by v2_store; rec = the stage record.
// BD: two passes (state lanes, then input lanes).  Running the input lanes predicated inside the state block's
// loop saves a quarter of the instructions of this product but measured 3 % slower (longer dependent chains per
// lane at the same shared-memory wavefront count).
template <class SH>
__device__ __forceinline__ double xinv_mv(const double *__restrict__ rec, int gl, const double *__restrict__ v2,
                                          const Wraps<SH> &wr) {
  if constexpr (SH::BD) {
    if (gl < SH::NX) return dmv<SH::NX, SH::WNS>(rec + SH::XOFF, rec + SH::XOFF + gl, v2 + gl, wr.S);
    const int i = gl - SH::NX;
    return dmv<SH::NU, SH::WNU>(rec + SH::UOFF, rec + SH::UOFF + i, v2 + 2 * SH::NX + i, wr.U);
  } else {
    return dmv<SH::A, SH::WNX>(rec + SH::XOFF, rec + SH::XOFF + gl, v2 + gl, wr.X);
  }
}

// entry (i, j) of the symmetric n x n block stored diagonal-packed at dp
template <int N>
__device__ __forceinline__ double dp_at(const double *__restrict__ dp, int i, int j) {
  int d = j - i;
  if (d < 0) d += N;
  return (d <= N / 2) ? dp[d * N + i] : dp[(N - d) * N + j];
}

// Xinv_k(i, j), i < A, j < NX (the state columns)
template <class SH>
__device__ __forceinline__ double xinv_state_col(const double *__restrict__ rec, int i, int j) {
  if constexpr (SH::BD) return i < SH::NX ? dp_at<SH::NX>(rec + SH::XOFF, i, j) : 0.0;
  else return dp_at<SH::A>(rec + SH::XOFF, i, j);
}

// sum_j Mcol[j*LD] * v[j], j < NJ  (Mcol already offset by the lane's row)
template <int NJ, int LD>
__device__ __forceinline__ double dense_mv(const double *__restrict__ Mcol, const double *__restrict__ v) {
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
  for (int j = 0; j < NJ; j++) acc[j & 3] = fma(Mcol[j * LD], v[j], acc[j & 3]);
  return (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

template <int CW>
__device__ __forceinline__ double ell_row(const double *__restrict__ val, const int *__restrict__ idx, int row,
                                          const double *__restrict__ v) {
  double s = 0.0;
#pragma unroll
  for (int e = 0; e < CW; e++) s = fma(val[row * CW + e], v[idx[row * CW + e]], s);
  return s;
}


template <int NYT>
struct MutVec {  // the mutable per-stage values a lane owns (iterates and forward intermediates)
  double x, zd, yd, yk, sk;
  double zp[NYT], yp[NYT];
};


// The specialisation table: (nx+nu, nx, ny, nt, ELL width of C).  First the BASELINE.json shapes, then
// other box-constrained MPC shapes (nx,nu) = (2,1) (4,1) (6,2) (6,3) (8,4) (16,4): odd and even block sizes
// in every lane-group width.  Any other shape runs on the runtime-shape kernels of rldl_kernels.cu.
// Last column: 1 = the block-diagonal-pivot variant (chosen first when the model qualifies, ModelDev::bd).
#define RLDL_SHAPES(X)       \
  X(6, 4, 6, 4, 1, 1, 0)     \
  X(16, 12, 16, 12, 1, 1, 0) \
  X(11, 8, 11, 8, 1, 1, 0)   \
  X(6, 4, 6, 4, 1, 1, 1)     \
  X(16, 12, 16, 12, 1, 1, 1) \
  X(11, 8, 11, 8, 1, 1, 1)   \
  X(6, 4, 6, 4, 1, 0, 0)     \
  X(16, 12, 16, 12, 1, 0, 0) \
  X(11, 8, 11, 8, 1, 0, 0)   \
  X(3, 2, 3, 2, 1, 0, 0)     \
  X(5, 4, 5, 4, 1, 0, 0)     \
  X(8, 6, 8, 6, 1, 0, 0)     \
  X(9, 6, 9, 6, 1, 0, 0)     \
  X(12, 8, 12, 8, 1, 0, 0)   \
  X(20, 16, 20, 16, 1, 0, 0)

// does table entry (a, nx, ny, nt, cw, bd, sc) serve this model?  (sc: the Ruiz-scaled variants exist for the
// block-diagonal BASELINE shapes; other scaled models run on the runtime-shape kernels)
inline bool shape_serves(const ModelDev &M, int cw_model, int a, int nx, int ny, int nt, int cw, int bd, int sc) {
  return M.a == a && M.nx == nx && M.ny == ny && M.nt == nt && cw_model <= cw && (bd == 0 || M.bd != 0) &&
         sc == (M.scaled ? 1 : 0);
}

inline int max_c_width(const ModelDev &M) {
  int wdt = 0;
  for (int c = 0; c < 3; c++) {
    if (M.sc[c].C.width > wdt) wdt = M.sc[c].C.width;
    if (M.sc[c].Ct.width > wdt) wdt = M.sc[c].Ct.width;
  }
  return wdt;
}

}  // namespace rldl
