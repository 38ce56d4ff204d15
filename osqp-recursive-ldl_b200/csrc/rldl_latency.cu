// rldl_latency.cu — latency-mode ADMM kernel (sm_100a): one QP instance per CTA.
//
// The streaming kernel (rldl_stream.cu) is built for throughput: thousands of instances in
// flight, one per lane group, each walking its 2(N+1) stage steps strictly in sequence
// (~1 us per step).  When few instances are left — the tail of a large batch, a small shard on
// one of 8 GPUs, or a batch of one behind the scalar osqp_solve — that chain is all that
// matters.  This kernel shortens it by an order of magnitude:
//
//  * the whole factor of the instance (its N+1 stage records) is loaded into shared memory
//    ONCE by bulk async copies (TMA) and reused for every iteration: no HBM traffic per
//    iteration except the iterates;
//  * the block substitution (reference lin_sys/direct/qdldl/qdldl_interface.c:559-585 over the
//    factor of src/recursive_ldl.c:1139-1318) is re-associated so that everything that does not
//    depend on the neighbouring stage runs for all stages in parallel (one lane group per
//    stage), and only an nx x nx mat-vec per stage stays on the critical path:
//        forward :  sp_k = c_k + T_k sp_{k-1},      c_k = Sinv_k (r^d_k - F_k Xinv_k r^v_k)
//        backward:  vn_k = d_k + T_k' vn_{k+1},     T_k = Sinv_k F_k Xinv_k[:, 0:nx]
//    (T_k is formed once per instance from the resident factor; the same matrix serves both
//    sweeps by symmetry).  Phases per iteration:
//        A  (parallel)  rhs build, a_k = Xinv_k r^v_k, c_k
//        B  (scan)      sp_k                                   — one group, operands by broadcast LDS.128
//        C  (parallel)  e_k = a_k + Xinv_k (F_k' sp_k - J' sp_{k-1}),  d_k = e_k[0:nx]
//        D  (scan)      vn_k
//        E  (parallel)  nu^d_k, v_k = e_k + Xinv_k F_k' Sinv_k vn_{k+1}, ADMM step
//    (reference src/auxil.c:164-228, src/proj.c:4-14);
//    The two recurrences run over a block tree (pairs of stages by default, deeper with the "lat_levels" knob): one
//    group walks the chain of the widest nodes, the other groups fill in the stages behind it as soon as a per-position
//    ready flag says their input is final.  With two or four lane groups per warp a warp owns an item (a pair of stages) and
//    its groups take the two stages side by side ("lat_lockstep", default); timeline of one iteration and the
//    variants measured on the way: profiles/r02_latency_kernel_timeline.txt, DESIGN.md section 3;
//  * residuals / termination / rho adaptation (src/auxil.c:243-789) by one group every
//    check_termination iterations, with the same device functions as the other kernels.
//
// Instances are claimed from the same device counter and with the same resume semantics as
// the streaming kernel, so the host can hand the tail of a streaming launch to this kernel.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "rldl_device_fn.cuh"
#include "rldl_kernels.h"
#include "rldl_stream_common.cuh"

namespace rldl {

constexpr int kLatThreads = 384;

// nx x (nx+2) blocks of global scratch per CTA: T_k and T_k' per stage, the products of the block tree above
// them (levels 1.., both forms) and the chain blocks in scan order (both forms)
constexpr int kLatMaxLevels = 6;
__host__ __device__ inline int lat_tree_blocks(int Nmax) { return Nmax + 2 * kLatMaxLevels + 2; }
__host__ __device__ inline int lat_chain_blocks(int Nmax) { return Nmax / 2 + kLatMaxLevels + 4; }
__host__ __device__ inline int lat_t_blocks(int Nmax) {
  return 2 * Nmax + 2 * lat_tree_blocks(Nmax) + 2 * lat_chain_blocks(Nmax);
}
// first block of level lv >= 1 inside the tree area (level lv holds the nodes j = 1 .. (N-1) >> lv)
__host__ __device__ inline int lat_level_base(int lv, int Nmax) {
  int b = 0;
  for (int i = 1; i < lv; i++) b += (Nmax >> i) + 1;
  return b;
}

template <class SH>
struct LatLayout {  // shared-memory layout (doubles) of k_admm_cta for horizon N
  static constexpr int NG = (kLatThreads / 32) * SH::GPW;
  // T_k ring of the scan phases: NS slots of CH stage blocks; a block is nx rows of TS = nx+2 doubles
  // (row stride chosen so that the per-lane LDS.128 of a row are bank-conflict free)
  static constexpr int CH = 5, NS = 3, TS = SH::NX + 2, TBLK = SH::NX * TS;
  int recs, va, vc, vd, scr, bar, ctl, tring, tbar, qs, red, flg, end;
  // ints behind `flg`: ready flags of the two recurrences, the kind of every position, the chain nodes
  __host__ __device__ static int flag_ints(int N) { return 3 * (N + 2) + 3 * (N / 2 + kLatMaxLevels + 6); }
  __host__ __device__ explicit LatLayout(int N) {
    recs = SH::M_END;
    va = recs + (N + 1) * SH::RS;
    vc = va + (N + 1) * SH::AE;
    vd = vc + (N + 1) * SH::NXE;
    scr = vd + (N + 1) * SH::NXE;
    bar = scr + NG * SH::SC;
    ctl = bar + 2;
    tring = ctl + 4;  // ctl: 4 ints (instance, leave flag, number of chain nodes, spare) padded to 16 bytes
    tbar = tring + NS * CH * TBLK;
    qs = tbar + NS + (NS & 1);
    red = qs + 3 * SH::A * SH::A;  // dense cost blocks padded to A x A (residuals)
    flg = red + NG * 10 + (kLatThreads / 32) * 4;  // per-group partial residuals + per-warp certificate partials
    end = flg + (flag_ints(N) + 1) / 2;
  }
};

#ifdef RLDL_LAT_PROFILE
__device__ long long g_lat_tr[64][12];
#endif
template <class SH>
__global__ void __launch_bounds__(kLatThreads, 1) k_admm_cta(ModelDev M, BatchDev B, SettingsDev S, int warm_start,
                                                              int resume, double *Tg, int claim_mode, int stream_ctas) {
  constexpr int A = SH::A, NX = SH::NX, NY = SH::NY, NT = SH::NT, CW = SH::CW, G = SH::G, NYT = SH::NYT,
                ROWS = SH::ROWS, RS = SH::RS, BQ = SH::B, AE = SH::AE, NXE = SH::NXE, NG = LatLayout<SH>::NG;
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int gl = lane & (G - 1), gw = lane / G;
  const unsigned gm = (G == 32) ? FULL : (((1u << G) - 1u) << (gw * G));
  const int grp = wid * SH::GPW + gw;
  const int N = B.N, n = B.n, m = B.m;
  const LatLayout<SH> L(N);

  // ---- CTA-shared model block (same layout as the streaming kernel) ----
  double *Fd = smem + SH::M_FD, *FdT = smem + SH::M_FDT, *Cv = smem + SH::M_CV, *Ctv = smem + SH::M_CTV;
  int *Ci = reinterpret_cast<int *>(smem + SH::M_CI), *Cti = reinterpret_cast<int *>(smem + SH::M_CTI);
  for (int e = threadIdx.x; e < SH::M_END; e += blockDim.x) smem[e] = 0.0;
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
  for (int t = threadIdx.x; t < 3 * ROWS; t += blockDim.x) {
    const int c = t / ROWS, r = t - c * ROWS;
    const Ell &E = M.sc[c].C;
    if (r < E.rows)
      for (int e = 0; e < E.width && e < CW; e++) {
        Cv[t * CW + e] = M.val[E.off + r * E.width + e];
        Ci[t * CW + e] = M.idx[E.off + r * E.width + e];
      }
  }
  for (int t = threadIdx.x; t < 3 * AE; t += blockDim.x) {
    const int c = t / AE, r = t - c * AE;
    const Ell &E = M.sc[c].Ct;
    if (r < E.rows)
      for (int e = 0; e < E.width && e < CW; e++) {
        Ctv[t * CW + e] = M.val[E.off + r * E.width + e];
        Cti[t * CW + e] = M.idx[E.off + r * E.width + e];
      }
  }
  double *Qs = smem + L.qs, *red = smem + L.red;
  for (int t = threadIdx.x; t < 3 * A * A; t += blockDim.x) {
    const int c = t / (A * A), e = t - c * A * A, j = e / A, i = e - j * A;
    const int dimq = (c == 2) ? NX : A;
    Qs[t] = (i < dimq && j < dimq) ? M.val[M.sc[c].Q + j * dimq + i] : 0.0;
  }
  double *recs = smem + L.recs, *va = smem + L.va, *vc = smem + L.vc, *vd = smem + L.vd;
  double *sc = smem + L.scr + (size_t)grp * SH::SC;
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem + L.bar);
  volatile int *ctl = reinterpret_cast<volatile int *>(smem + L.ctl);
  // ready flags (see the block tree below): rf[k] == iter once sp_k is final, rb[k] == iter once vn_k is final;
  // kf[k]: bit 0 = k ends a chain node, bit 1 = k starts one; ch_lo / ch_hi: the chain nodes in forward order
  volatile int *rf = reinterpret_cast<volatile int *>(smem + L.flg), *rb = rf + (N + 2);
  int *kf = const_cast<int *>(rb) + (N + 2), *ch_lo = kf + (N + 2), *ch_hi = ch_lo + (N / 2 + kLatMaxLevels + 6),
      *ch_lv = ch_hi + (N / 2 + kLatMaxLevels + 6);
  auto wait_flag = [&](volatile int *p, int want) {
    while (*p != want) __nanosleep(64);  // back off: the scan warp shares its scheduler with pollers
    __threadfence_block();
  };
  // groups 2.. work behind the scans (group 1 shares warp 0 with the scanning group and stays out of its way)
  constexpr int NGW = NG - 2;
  double *w = sc + SH::O_W, *tv2 = sc + SH::O_TV2, *yvs = sc + SH::O_YV, *td2 = sc + SH::O_TD2;
  double *vn2 = sc + SH::O_VN2, *vcur = sc + SH::O_VC, *nudv = sc + SH::O_NUD, *gg2 = sc + SH::O_GG2;
  Scratch rs;  // residual routines (group 0 only) alias that group's scratch
  rs.v = sc; rs.w = sc + SH::VM; rs.vn = sc + 2 * SH::VM; rs.td = sc + 3 * SH::VM; rs.sp = sc + 4 * SH::VM;
  rs.g = rs.yv = rs.nud = nullptr;
  constexpr int CH = LatLayout<SH>::CH, NS = LatLayout<SH>::NS, TS = LatLayout<SH>::TS, TBLK = LatLayout<SH>::TBLK;
  static_assert(NX % 2 == 0, "latency-mode kernel: nx must be even (16-byte rows of T)");
  double *tring = smem + L.tring;
  uint64_t *tbars = reinterpret_cast<uint64_t *>(smem + L.tbar);
  // ---- the block tree over the stages 1..N-1 of the two recurrences ----
  // A node (lv, j) is the aligned run of 2^lv stages ending at j 2^lv; its operator M = T_end ... T_start is formed
  // once per instance.  The sequential chain (group 0) runs over the widest nodes only: as many of level LV as fit,
  // then one node per set bit of what is left.  Inside a chain node the aggregates go up the tree before the scan
  // (levels 2.. in CTA-wide steps; level 1 by the owner of the pair) and the states come down behind it, one hop per
  // level, each by the group that owns the pair the position belongs to.  LV = 1 is the pairs-only chain.
  const int LV = min(max(B.lat_levels, 1), kLatMaxLevels);
  if (threadIdx.x == 0) {
    ctl[2] = 0; ctl[3] = 0;
    for (int k = 0; k < N + 2; k++) kf[k] = 0;
    int q = 0, pos = 1, rem = N - 1;
    for (int lv = LV; lv >= 0; lv--)
      while (rem >= (1 << lv)) {
        q++;
        ch_lo[q] = pos; ch_hi[q] = pos + (1 << lv) - 1; ch_lv[q] = lv;
        kf[ch_hi[q]] |= 1; kf[ch_lo[q]] |= 2;
        pos += 1 << lv; rem -= 1 << lv;
      }
    ctl[2] = q;
    mbar_init(bar, 1);
    for (int d = 0; d < NS; d++) mbar_init(&tbars[d], 1);
  }
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const uint32_t bar_s = smem_u32(bar), recs_s = smem_u32(recs), tring_s = smem_u32(tring), tbars_s = smem_u32(tbars);
  long long twait_ = 0;
  uint32_t phase = 0, tphase = 0;  // tphase: one parity bit per ring slot (group 0 only)
  auto lds128 = [](uint32_t a) -> double2 {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
  };
  auto lds64 = [](uint32_t a) -> double {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
  };
  auto sts64 = [](uint32_t a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); };
  auto t_wait = [&](int s) {
    mbar_wait(tbars_s + 8u * s, (tphase >> s) & 1u);
    tphase ^= 1u << s;
  };

  Wraps<SH> wr;
  wr.init(gl);
  const int (&wrapS)[SH::WNS] = wr.S;

  const double sigma = S.sigma, alpha = S.alpha, oma = 1.0 - S.alpha;
  // T_k for phase B (TB[k][i*TS + j] = T_k(i,j)) and phase D (TD[k][i*TS + j] = T_k(j,i)) in global memory
  // (2*Nmax*TBLK doubles per CTA, L2-resident), written once per instance by this CTA and streamed back
  // through the shared-memory ring by bulk async copies, CH stages per copy
  double *TB = Tg + (size_t)blockIdx.x * lat_t_blocks(B.Nmax) * TBLK, *TD = TB + (size_t)B.Nmax * TBLK;
  // Blocked scans over the block tree above.  With the aggregates
  //     C(node) = c_end + T_end (c_end-1 + ...),           kept in place at vc[end],
  //     D(node) = d_start + T_start' (d_start+1 + ...),    kept in place at vd[start],
  // the sequential chain is  sp_end = C(node) + M(node) sp_start-1  (forward) and
  // vn_start = D(node) + M(node)' vn_end+1  (backward) over the NC chain nodes.  Behind it, the end of a left child
  // (forward: sp_e = C(left) + M(left) sp_start-1) and the start of a right child (backward:
  // vn_o = D(right) + M(right)' vn_end+1) follow level by level, down to the pairs; the stage in the middle of a
  // pair is filled in by the group that owns the pair (sp_2j-1 = c_2j-1 + T_2j-1 sp_2j-2, vn_2j = d_2j + T_2j' vn_2j+1).
  // LB / LD: the operators of the levels 1.. (row-major / transposed); CB[q] / CD[q], q = 1..NC: those of the chain
  // nodes, copied out in scan order for the bulk-copy ring.
  double *LB = TD + (size_t)B.Nmax * TBLK, *LD = LB + (size_t)lat_tree_blocks(B.Nmax) * TBLK;
  double *CB = LD + (size_t)lat_tree_blocks(B.Nmax) * TBLK, *CD = CB + (size_t)lat_chain_blocks(B.Nmax) * TBLK;
  const int J = (N - 1) / 2, LEFT = (N - 1) & 1, NC = ctl[2], NI = J + 2;
  // operator of node (lv, j), row-major (mb) or transposed (md); level 0 is T_j itself
  auto mb = [&](int lv, int j) -> double * {
    return lv == 0 ? TB + (size_t)j * TBLK : LB + (size_t)(lat_level_base(lv, B.Nmax) + j) * TBLK;
  };
  auto md = [&](int lv, int j) -> double * {
    return lv == 0 ? TD + (size_t)j * TBLK : LD + (size_t)(lat_level_base(lv, B.Nmax) + j) * TBLK;
  };
  // row gl of an nx x nx block in global memory (L2) times a vector in shared memory
  auto tmv = [&](const double *blk, const double *v) -> double {
    const double2 *r = reinterpret_cast<const double2 *>(blk + (gl < NX ? gl : 0) * TS);
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int j = 0; j < NX / 2; j++) {
      const double2 t = __ldcg(r + j);
      acc[(2 * j) & 3] = fma(t.x, v[2 * j], acc[(2 * j) & 3]);
      acc[(2 * j + 1) & 3] = fma(t.y, v[2 * j + 1], acc[(2 * j + 1) & 3]);
    }
    return (acc[0] + acc[1]) + (acc[2] + acc[3]);
  };
  // the same in two halves: the row is fetched from L2 before a flag is polled, the product follows the flag
  auto ld_rows = [&](const double *blk, double2 (&rowv)[NX / 2]) {
    const double2 *r = reinterpret_cast<const double2 *>(blk + (gl < NX ? gl : 0) * TS);
#pragma unroll
    for (int j = 0; j < NX / 2; j++) rowv[j] = __ldcg(r + j);
  };
  auto dot_rows = [&](const double2 (&rowv)[NX / 2], const double *v) -> double {
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int j = 0; j < NX / 2; j++) {
      acc[(2 * j) & 3] = fma(rowv[j].x, v[2 * j], acc[(2 * j) & 3]);
      acc[(2 * j + 1) & 3] = fma(rowv[j].y, v[2 * j + 1], acc[(2 * j + 1) & 3]);
    }
    return (acc[0] + acc[1]) + (acc[2] + acc[3]);
  };
  // One scan (group 0) over the chain: step q = 1..NC (fwd, vec = vc) writes sp at the end of chain node q from sp
  // in front of it; step q = NC..1 (bwd, vec = vd) writes vn at the start of node q from vn behind it.  The source
  // vector is re-read from shared memory by broadcast LDS.128; one __syncwarp per step.
  auto scan = [&](const bool fwd, const int iter_) {
    const double *Tsrc = fwd ? CB : CD;
    double *vec = fwd ? vc : vd;
    const int nblk = NC, nchunk = (nblk + CH - 1) / CH;
    auto issue_chunk = [&](int c) {  // leader: chunk c of the scan order -> ring slot c % NS
      const int t0 = c * CH, cnt = min(CH, nblk - t0);
      const int p_lo = fwd ? 1 + t0 : (NC - t0) - (cnt - 1);
      const int s = c % NS;
      RLDL_ASSERT(cnt >= 1 && cnt <= CH && p_lo >= 1 && p_lo + cnt - 1 <= NC && NC < lat_chain_blocks(B.Nmax));
      mbar_expect_tx(tbars_s + 8u * s, (uint32_t)(cnt * TBLK * 8));
      bulk_g2s(tring_s + 8u * (uint32_t)(s * CH * TBLK), Tsrc + (size_t)p_lo * TBLK, (uint32_t)(cnt * TBLK * 8),
               tbars_s + 8u * s);
    };
    if (gl == 0)
      for (int c = 0; c < NS && c < nchunk; c++) issue_chunk(c);
    volatile int *ready = fwd ? rf : rb;  // published per chain step; the groups working behind the scan poll it
    // 32-bit shared addresses: nothing but the mat-vec itself is on the chain
    const int row = (gl < NX) ? gl : 0;
    const uint32_t vec_s = smem_u32(vec);
    const int tstep = fwd ? TBLK * 8 : -TBLK * 8;
    int s = 0, t = 0;
    int dst_n = 0, src_n = 0;
    if (NC > 0) {
      dst_n = fwd ? ch_hi[1] : ch_lo[NC];
      src_n = fwd ? ch_lo[1] - 1 : ch_hi[NC] + 1;
    }
    for (int c = 0; c < nchunk; c++) {
#ifdef RLDL_LAT_PROFILE
      const long long w0_ = clock64();
#endif
      t_wait(s);
#ifdef RLDL_LAT_PROFILE
      twait_ += clock64() - w0_;
#endif
      const int cnt = min(CH, nblk - c * CH);
      uint32_t ta = tring_s + 8u * (uint32_t)((s * CH + (fwd ? 0 : cnt - 1)) * TBLK + row * TS);
      for (int r = 0; r < cnt; r++, t++) {
        const int dst = dst_n, src = src_n;
        {  // the next step's positions: off the chain
          const int q = fwd ? min(t + 2, NC) : max(NC - t - 1, 1);
          dst_n = fwd ? ch_hi[q] : ch_lo[q];
          src_n = fwd ? ch_lo[q] - 1 : ch_hi[q] + 1;
        }
        const uint32_t pd = vec_s + 8u * (uint32_t)(dst * NXE), ps = vec_s + 8u * (uint32_t)(src * NXE);
        // (loading the next step's T rows ahead of the STS below was measured slower: the loads sit in the
        // in-order LSU queue in front of the store the next step is waiting for)
        double2 tv[NX / 2], vv[NX / 2];
#pragma unroll
        for (int j = 0; j < NX / 2; j++) tv[j] = lds128(ta + 16u * j);
        double cur = 0.0;
        if (gl < NX) cur = lds64(pd + 8u * gl);
#pragma unroll
        for (int j = 0; j < NX / 2; j++) vv[j] = lds128(ps + 16u * j);
        double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int j = 0; j < NX / 2; j++) {
          acc[(2 * j) & 3] = fma(tv[j].x, vv[j].x, acc[(2 * j) & 3]);
          acc[(2 * j + 1) & 3] = fma(tv[j].y, vv[j].y, acc[(2 * j + 1) & 3]);
        }
        if (gl < NX) sts64(pd + 8u * gl, cur + ((acc[0] + acc[1]) + (acc[2] + acc[3])));
        __syncwarp(gm);
        // publish (after the __syncwarp: the stores of one warp reach shared memory in program order)
        if (gl == 0) ready[dst] = iter_;
        ta += tstep;
      }
      if (gl == 0 && c + NS < nchunk) issue_chunk(c + NS);
      s = (s + 1 == NS) ? 0 : s + 1;
    }
  };
  // work items of the groups behind the scans: item 0 = stage 0, item j = 1..J = the pair (2j-1, 2j),
  // item J+1 = the leftover stage (if any) and the terminal stage
  auto item_lo = [&](int it) -> int { return it == 0 ? 0 : (it <= J ? 2 * it - 1 : 2 * J + 1); };
  auto item_hi = [&](int it) -> int { return it == 0 ? 0 : (it <= J ? 2 * it : N); };
  // Lockstep variant (two lane groups per warp, tuning knob "lat_lockstep"): a WARP owns an item and its two lane
  // groups take the two stages of the item side by side — the upper stage by group 0 of the warp, the lower one by
  // group 1 — so the groups of a warp are released together and never run diverged, and the item released last
  // costs one stage, not two.  Warp 0 keeps the scan.
  // (With four lane groups per warp the groups 2 and 3 of a worker warp stay idle: the items of small shapes are
  // fewer than the warps anyway.)
  const bool LK = (SH::GPW >= 2) && B.lat_lockstep != 0;
  constexpr int NWK = kLatThreads / 32 - 1;
  auto half_stage = [&](int it) -> int {
    if (gw >= 2) return -1;
    if (it == 0) return gw == 0 ? 0 : -1;
    if (it <= J) return gw == 0 ? 2 * it : 2 * it - 1;
    return gw == 0 ? N : (LEFT ? N - 1 : -1);
  };

  for (;;) {
    // ---------------- claim an instance ----------------
    // claim_mode 1: the tail of a streaming launch — first the instances handed over to the latency-mode lane that
    //   nobody took, then every suspended instance, most-unconverged first (the longest remaining chains start first).
    // claim_mode 2: a CTA of the concurrent lane, launched beside the streaming kernel: it waits for hand-overs
    //   (kSchedLWant announces it, the streaming groups answer through ring kSchedLQ); once the streaming kernel has
    //   hit its tail it behaves like mode 1.  A CTA that became resident before every streaming CTA had started
    //   would keep an SM from the streaming kernel for the whole launch: it leaves at once.
    // otherwise: the fresh counter, with the same resume rules as the streaming kernel.
    if (threadIdx.x == 0 && claim_mode) {
      int i = -1;
      bool from_ring = true;
      if (claim_mode == 2 && ld_volatile(B.sched + kSchedStarted) < stream_ctas && !ld_volatile(B.sched + kSchedTailFlag)) {
        i = -1;
        atomicAdd(B.sched + kSchedStats + 1, 1);
      } else {
        if (sched_ring_nonempty(B, kSchedLQ)) i = sched_pop(B, kSchedLQ);
        if (i < 0 && claim_mode == 2 && !ld_volatile(B.sched + kSchedTailFlag)) {
          atomicAdd(B.sched + kSchedLWant, 1);
          unsigned long long t_wait0;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_wait0));
          for (int spins = 1;; spins++) {
            if (sched_ring_nonempty(B, kSchedLQ) && (i = sched_pop(B, kSchedLQ)) >= 0) break;
            if (ld_volatile(B.sched + kSchedUnfinished) <= 0 || ld_volatile(B.sched + kSchedTailFlag)) break;
            if ((spins & 255) == 0) {  // safety net: never wait for another kernel for more than a second
              unsigned long long now;
              asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
              if (now - t_wait0 > 1000000000ull) break;
            }
            __nanosleep(2000);
            if ((spins & 7) == 0) {  // nobody qualified for a while: lower the hand-over threshold by one bucket
              const int t = ld_volatile(B.sched + kSchedTop);
              if (t < kSchedBuckets - 3) atomicCAS(B.sched + kSchedTop, t, t + 1);
            }
          }
          if (i < 0 || !from_ring) {  // leaving without a hand-over: take the announcement back if it is still there
            int w = ld_volatile(B.sched + kSchedLWant);
            while (w > 0 && atomicCAS(B.sched + kSchedLWant, w, w - 1) != w) w = ld_volatile(B.sched + kSchedLWant);
          }
        }
        if (i < 0 && (claim_mode == 1 || ld_volatile(B.sched + kSchedTailFlag))) {
          if (sched_ring_nonempty(B, kSchedLQ)) i = sched_pop(B, kSchedLQ);
          if (i < 0) { i = sched_pop_best(B); from_ring = false; }
        }
      }
      if (i >= 0) __threadfence();  // order the loads of the iterates after the hand-over
      if (claim_mode == 2) atomicAdd(B.sched + kSchedStats + (i >= 0 ? (from_ring ? 2 : 3) : 5), 1);
      ctl[0] = i;
    } else if (threadIdx.x == 0) {
      int i = -1;
      for (;;) {
        i = atomicAdd(B.sched, 1);
        if (i >= B.batch) { i = -1; break; }
        if (B.fact_flag[i]) {  // failed factorisation: report non-convex, like a diverged solve
          if (!resume || B.status[i] == -10) {
            B.status[i] = -7; B.iter[i] = 0; B.obj[i] = kOsqpNan;
            atomicSub(B.sched + kSchedUnfinished, 1);
            deliver_to_host(B, i, 0, 1, n, m, -7, 0, kOsqpNan, 0.0, 0.0, 0.0);
          }
          continue;
        }
        if (resume && __ldcg(B.status + i) != -10) continue;
        break;
      }
      ctl[0] = i;
    }
    __syncthreads();
    const int inst = ctl[0];
    if (inst < 0) break;
    RLDL_ASSERT(inst < B.batch);
    RLDL_ASSERT((long long)(N + 1) * RS <= B.rec_stride && L.end * 8 <= 227 * 1024);
    const size_t vo = (size_t)inst * B.nmax, ro = (size_t)inst * B.mmax;
    double *x = B.x + vo, *z = B.z + ro, *y = B.y + ro, *dx = B.dx + vo, *dy = B.dy + ro;
    const double *recb = B.rec + (size_t)inst * B.rec_stride;
    const double rho_in = __ldcg(B.rho + inst);
    RhoClasses rcl;
    rcl.set(rho_in);
    // ---------------- the whole factor of the instance -> shared memory (TMA bulk copies) ----------------
    if (threadIdx.x == 0) mbar_expect_tx(bar_s, (uint32_t)((N + 1) * RS * 8));
    __syncthreads();
    for (int k = threadIdx.x; k <= N; k += blockDim.x)
      bulk_g2s(recs_s + 8u * (uint32_t)(k * RS), recb + (size_t)k * RS, RS * 8, bar_s);
    int iter;
    const bool fresh = !resume || false;
    if (fresh) {
      if (!warm_start) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) x[i] = 0.0;
        for (int r = threadIdx.x; r < m; r += blockDim.x) { z[r] = 0.0; y[r] = 0.0; }
      }
      if (threadIdx.x == 0) B.status[inst] = -10;
      iter = 0;
    } else {
      iter = (int)__ldcg(B.iter + inst);
    }
    mbar_wait(bar_s, phase);
    phase ^= 1u;
    __syncthreads();
    for (int k = threadIdx.x; k < 2 * (N + 2); k += blockDim.x) rf[k] = -1;  // rf and rb: nothing is ready
    // ---------------- T_k = Sinv_k F_k Xinv_k[:, 0:nx], one group per stage ----------------
    for (int k = grp; k < N; k += NG) {
      const double *sl_ = recs + k * RS;
      const double *Fk = Fd + (k == 0 ? 0 : A * NX);
      for (int j = 0; j < NX; j++) {
        if (gl < A) yvs[gl] = xinv_state_col<SH>(sl_, gl, j);  // column j of Xinv_k
        __syncwarp(gm);
        if (gl < NX) {
          const double wv = dense_mv<A, NX>(Fk + gl, yvs);
          td2[gl] = wv; td2[gl + NX] = wv;
        }
        __syncwarp(gm);
        if (gl < NX) {
          const double t = dmv<NX, SH::WNS>(sl_ + SH::SOFF, sl_ + SH::SOFF + gl, td2 + gl, wrapS);
          TB[(size_t)k * TBLK + gl * TS + j] = t;
          TD[(size_t)k * TBLK + j * TS + gl] = t;
        }
        __syncwarp(gm);
      }
    }
    __threadfence();
    __syncthreads();
    // ---------------- the operators of the block tree, level by level: M(lv, j) = M(lv-1, 2j) M(lv-1, 2j-1),
    // one group per node ----------------
    for (int lv = 1; lv <= LV; lv++) {
      const int cnt = (N - 1) >> lv;
      for (int j = 1 + grp; j <= cnt; j += NG) {
        const double *T2 = mb(lv - 1, 2 * j);       // rows of the left factor
        const double *T1t = md(lv - 1, 2 * j - 1);  // rows of the right factor's transpose = its columns
        double *ob = mb(lv, j), *od = md(lv, j);
        if (gl < NX) {
          double rowv[NX];
#pragma unroll
          for (int mm = 0; mm < NX; mm++) rowv[mm] = __ldcg(T2 + gl * TS + mm);
#pragma unroll 1
          for (int c = 0; c < NX; c++) {
            double v = 0.0;
#pragma unroll
            for (int mm = 0; mm < NX; mm++) v = fma(rowv[mm], __ldcg(T1t + c * TS + mm), v);
            ob[gl * TS + c] = v;
            od[c * TS + gl] = v;
          }
        }
      }
      __threadfence();
      __syncthreads();
    }
    // the chain nodes, copied out in scan order
    for (int q = 1 + grp; q <= NC; q += NG) {
      const int lv = ch_lv[q], j = ch_hi[q] >> lv;
      const double *sb = mb(lv, j), *sd = md(lv, j);
      for (int e = gl; e < TBLK; e += G) {
        CB[(size_t)q * TBLK + e] = __ldcg(sb + e);
        CD[(size_t)q * TBLK + e] = __ldcg(sd + e);
      }
    }
    __threadfence();  // the chain blocks are read back by bulk async copies (async proxy)
    asm volatile("fence.proxy.async;" ::: "memory");
    __syncthreads();
    // ===== phase A for one stage: rhs, a_k = Xinv_k r^v_k, c_k = Sinv_k (r^d_k - F_k a_k), from the iterates of
    // the stage held by the lanes that own them (loaded from HBM for the first iteration of an instance, handed
    // over in registers by phase E afterwards) =====
    auto stage_A = [&](int k, double xv, const double (&zp)[NYT], const double (&yp)[NYT], double zd, double yd) {
      const bool term = (k == N);
      const int cls = term ? 2 : (k == 0 ? 0 : 1);
      const int dim = term ? NX : A, np = term ? NT : NY;
      const int ilo = (k == 0) ? NX : 0;
      const double *sl_ = recs + k * RS;
      double g = 0.0;
      if (gl >= ilo && gl < dim) g = rec_sigma<SH>(sl_, gl, sigma) * xv - sl_[gl];
#pragma unroll
      for (int t = 0; t < NYT; t++) {
        const int rr = gl + G * t;
        if (rr < np) w[rr] = rec_rho<SH>(sl_, rr, rcl) * zp[t] - yp[t];
      }
      __syncwarp(gm);
      double tt = 0.0;
      if (gl < dim) tt = g + ell_row<CW>(Ctv + cls * AE * CW, Cti + cls * AE * CW, gl, w);
      if (gl < A) v2_store<SH>(tv2, gl, tt);
      __syncwarp(gm);
      if (gl < A) {
        const double av = xinv_mv<SH>(sl_, gl, tv2, wr);
        va[k * AE + gl] = av;
        yvs[gl] = av;
      }
      __syncwarp(gm);
      if (!term) {
        if (gl < NX) {
          const double tdv = (zd - yd * rec_rinv<SH>(sl_, NY + gl, rcl)) -
                             dense_mv<A, NX>(Fd + (k == 0 ? 0 : A * NX) + gl, yvs);
          td2[gl] = tdv; td2[gl + NX] = tdv;
        }
        __syncwarp(gm);
        if (gl < NX) vc[k * NXE + gl] = dmv<NX, SH::WNS>(sl_ + SH::SOFF, sl_ + SH::SOFF + gl, td2 + gl, wrapS);
      }
      __syncwarp(gm);
    };
    // forward pre-step of a pair: c_2j += T_2j c_2j-1 (the chain then needs sp_2j-2 only)
    auto pre_fwd = [&](int it) {
      if (it >= 1 && it <= J) {
        if (gl < NX) vc[2 * it * NXE + gl] += tmv(TB + (size_t)(2 * it) * TBLK, vc + (2 * it - 1) * NXE);
        __syncwarp(gm);
      }
    };
    // phase C for one stage (its sp_k and sp_k-1 are final)
    auto stage_C = [&](int k) {
      const bool term = (k == N);
      const double *sl_ = recs + k * RS;
      if (gl < A) {
        double gg = 0.0;
        if (!term) gg = dense_mv<NX, A>(FdT + (k == 0 ? 0 : A * NX) + gl, vc + k * NXE);
        if (k > 0 && gl < NX) gg -= vc[(k - 1) * NXE + gl];
        v2_store<SH>(gg2, gl, gg);
      }
      __syncwarp(gm);
      if (gl < A) {
        const double e = va[k * AE + gl] + xinv_mv<SH>(sl_, gl, gg2, wr);
        va[k * AE + gl] = e;
        if (gl < NX) vd[k * NXE + gl] = e;
      }
      __syncwarp(gm);
    };
    // phase A for one stage from the iterates in HBM (first iteration of an instance)
    auto stage_A_from_hbm = [&](int k) {
      const bool term = (k == N);
      const int dim = term ? NX : A, np = term ? NT : NY;
      const int xoff = k * A - NX, roff = k * BQ, ilo = (k == 0) ? NX : 0;
      const double xv = (gl >= ilo && gl < dim) ? __ldcg(x + xoff + gl) : 0.0;
      double zp[NYT], yp[NYT], zd = 0.0, yd = 0.0;
#pragma unroll
      for (int t = 0; t < NYT; t++) {
        const int rr = gl + G * t;
        zp[t] = yp[t] = 0.0;
        if (rr < np) { zp[t] = __ldcg(z + roff + rr); yp[t] = __ldcg(y + roff + rr); }
      }
      if (!term && gl < NX) { zd = __ldcg(z + roff + NY + gl); yd = __ldcg(y + roff + NY + gl); }
      stage_A(k, xv, zp, yp, zd, yd);
    };
    // first iteration: phase A from the iterates in HBM, by the groups that will own the items in phase E
    if (LK) {
      if (wid >= 1)
        for (int it = wid - 1; it < NI; it += NWK) {
          const int k = half_stage(it);
          if (k >= 0) stage_A_from_hbm(k);
          __syncwarp(FULL);
          if (gw == 0) pre_fwd(it);
          __syncwarp(FULL);
        }
    } else if (grp >= 2) {
      for (int it = grp - 2; it < NI; it += NGW) {
        for (int k = item_lo(it); k <= item_hi(it); k++) stage_A_from_hbm(k);
        pre_fwd(it);
      }
    }
    __syncthreads();
    // ---------------- ADMM iterations (osqp.c:354-519) ----------------
#ifdef RLDL_LAT_PROFILE
    long long pt[7] = {0, 0, 0, 0, 0, 0, 0}, pc = 0;  // [0] / [6]: aggregates before the forward / backward scan
    unsigned long long gt0_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt0_));
    const int iter0_ = iter;
    auto gtime_ = []() -> long long { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return (long long)t; };
#define PT(i) do { const long long now_ = gtime_(); pt[i] += now_ - pc; pc = now_; } while (0)
#define TR(slot) do { if (blockIdx.x == 0 && iter == 101 && gl == 0) g_lat_tr[grp][slot] = gtime_(); } while (0)
#else
#define PT(i)
#define TR(slot)
#endif
    for (;;) {
      iter++;
#ifdef RLDL_LAT_PROFILE
      pc = gtime_();
      if (blockIdx.x == 0 && iter == 101 && gl == 0) for (int e_ = 0; e_ < 12; e_++) g_lat_tr[grp][e_] = 0;
#endif
      TR(0);
      const bool can_check = S.check_termination && (iter % S.check_termination == 0);
      const bool can_adapt = S.adaptive_rho && S.adaptive_rho_interval && (iter % S.adaptive_rho_interval == 0);
      const bool last = iter == S.max_iter;
      const bool time_due = time_check_due(S, iter, can_check);
      const bool store_delta = can_check || last || time_due;
      // ===== phases B and C, pipelined.  B (group 0): sp_k = c_k + T_k sp_{k-1}, in place in vc.
      //       C (groups 2..): e_k = a_k + Xinv_k (F_k' sp_k - J' sp_{k-1}), d_k = e_k[0:nx], as soon as the scan
      //       has published sp_k =====
      // the aggregates of the levels 2.. (level 1 was done by the owners of the pairs: pre_fwd)
      for (int lv = 2; lv <= LV; lv++) {
        const int cnt = (N - 1) >> lv;
        if (cnt == 0) break;
        for (int j = 1 + grp; j <= cnt; j += NG) {
          const int end = j << lv, mid = end - (1 << (lv - 1));
          if (gl < NX) vc[end * NXE + gl] += tmv(mb(lv - 1, 2 * j), vc + mid * NXE);
        }
        __syncthreads();
      }
      PT(0);
      TR(1);
      if (grp == 0) {
        scan(true, iter);
        TR(2);
        PT(1);
      } else if (LK) {
        if (wid >= 1)
          for (int it = wid - 1; it < NI; it += NWK) {
            const int k = half_stage(it);
            if (it >= 1 && it <= J) {
              const int e = 2 * it;
              if (gw == 0) {
                if (!(kf[e] & 1)) {  // sp_e comes down the tree (see the other variant below)
                  const int t = __ffs(e) - 1, src = e - (1 << t);
                  double2 rowv[NX / 2];
                  ld_rows(mb(t, e >> t), rowv);
                  if (src > 0) wait_flag(rf + src, iter);
                  const double add = dot_rows(rowv, vc + src * NXE);
                  if (gl < NX) vc[e * NXE + gl] += add;
                  __syncwarp(gm);
                  if (gl == 0) { __threadfence_block(); rf[e] = iter; }
                } else {
                  wait_flag(rf + e, iter);  // sp_2j from the chain
                }
              } else if (gw == 1) {  // the middle stage: sp_2j-1 = c_2j-1 + T_2j-1 sp_2j-2
                if (e > 2) wait_flag(rf + (e - 2), iter);
                if (gl < NX) vc[(e - 1) * NXE + gl] += tmv(TB + (size_t)(e - 1) * TBLK, vc + (e - 2) * NXE);
              }
              __syncwarp(FULL);
            } else if (it == J + 1) {
              if (2 * J > 0) wait_flag(rf + 2 * J, iter);
              if (LEFT) wait_flag(rf + (N - 1), iter);
            }
            if (it == wid - 1) TR(3);
            if (k >= 0) stage_C(k);
            if (it >= 1 && it <= J) {  // backward pre-step: d_2j-1 += T_2j-1' d_2j
              __syncwarp(FULL);
              if (gw == 1) {
                if (gl < NX) vd[(2 * it - 1) * NXE + gl] += tmv(TD + (size_t)(2 * it - 1) * TBLK, vd + 2 * it * NXE);
              }
              __syncwarp(FULL);
            }
            if (it == wid - 1) TR(4);
          }
      } else if (grp >= 2) {
        for (int it = grp - 2; it < NI; it += NGW) {
          if (it >= 1 && it <= J) {
            const int e = 2 * it;
            if (!(kf[e] & 1)) {  // sp_e comes down the tree: the end of a left child of width 2^t
              const int t = __ffs(e) - 1, src = e - (1 << t);
              const double2 *r = reinterpret_cast<const double2 *>(mb(t, e >> t) + (gl < NX ? gl : 0) * TS);
              double2 rowv[NX / 2];
#pragma unroll
              for (int j = 0; j < NX / 2; j++) rowv[j] = __ldcg(r + j);  // in flight while the flag is polled
              if (src > 0) wait_flag(rf + src, iter);
              const double *v = vc + src * NXE;
              double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
              for (int j = 0; j < NX / 2; j++) {
                acc[(2 * j) & 3] = fma(rowv[j].x, v[2 * j], acc[(2 * j) & 3]);
                acc[(2 * j + 1) & 3] = fma(rowv[j].y, v[2 * j + 1], acc[(2 * j + 1) & 3]);
              }
              if (gl < NX) vc[e * NXE + gl] += (acc[0] + acc[1]) + (acc[2] + acc[3]);
              __syncwarp(gm);
              if (gl == 0) { __threadfence_block(); rf[e] = iter; }
            }
            if (it == grp - 2) TR(2);
            // fill the middle stage: sp_2j-1 = c_2j-1 + T_2j-1 sp_2j-2
            if (e > 2) wait_flag(rf + (e - 2), iter);
            if (gl < NX) vc[(e - 1) * NXE + gl] += tmv(TB + (size_t)(e - 1) * TBLK, vc + (e - 2) * NXE);
            __syncwarp(gm);
            if (it == grp - 2) TR(3);
          } else if (it == J + 1) {  // leftover and terminal stage: sp_N-1 (and sp_N-2 in front of a leftover stage)
            if (2 * J > 0) wait_flag(rf + 2 * J, iter);
            if (LEFT) wait_flag(rf + (N - 1), iter);
          }
          for (int k = item_lo(it); k <= item_hi(it); k++) {
            if (it >= 1 && it <= J && k == 2 * it && (kf[k] & 1)) wait_flag(rf + k, iter);  // sp_2j from the chain
            stage_C(k);
          }
          if (it >= 1 && it <= J) {  // backward pre-step: d_2j-1 += T_2j-1' d_2j
            if (gl < NX) vd[(2 * it - 1) * NXE + gl] += tmv(TD + (size_t)(2 * it - 1) * TBLK, vd + 2 * it * NXE);
            __syncwarp(gm);
          }
          if (it == grp - 2) TR(4);
        }
      }
      __syncthreads();
      TR(5);
      PT(2);
      // ===== phases D and E, pipelined.  D (group 0): vn_k = d_k + T_k' vn_{k+1}, k = N-1 .. 1, in place in vd
      //       (vn_N = d_N; stage 0 has no state) =====
      // (phase E below runs behind this scan: groups 2.. take their stages in descending order and start on
      // stage k once vn_{k+1} is published)
      for (int lv = 2; lv <= LV; lv++) {  // the aggregates of the levels 2.., kept at the start of every node
        const int cnt = (N - 1) >> lv;
        if (cnt == 0) break;
        for (int j = 1 + grp; j <= cnt; j += NG) {
          const int start = ((j - 1) << lv) + 1, rs_ = start + (1 << (lv - 1));
          if (gl < NX) vd[start * NXE + gl] += tmv(md(lv - 1, 2 * j - 1), vd + rs_ * NXE);
        }
        __syncthreads();
      }
      PT(6);
      TR(6);
      if (grp == 0) {
        scan(false, iter);
        TR(7);
        PT(3);
      }
      // ===== phase E: v_k = e_k + Xinv_k F_k' (Sinv_k vn_{k+1}), nu^d_k, ADMM step (auxil.c:188-228) =====
      // phases E and A (of the next iteration) for one stage whose vn_k+1 is final
      auto stage_EA = [&](int k) {
        const bool term = (k == N);
        const int cls = term ? 2 : (k == 0 ? 0 : 1);
        const int dim = term ? NX : A, np = term ? NT : NY;
        const int xoff = k * A - NX, roff = k * BQ, ilo = (k == 0) ? NX : 0;
        const double *sl_ = recs + k * RS;
        const double xo = (gl >= ilo && gl < dim) ? __ldcg(x + xoff + gl) : 0.0;
        double zp[NYT], yp[NYT], zd = 0.0, yd = 0.0;
#pragma unroll
        for (int t = 0; t < NYT; t++) {
          const int rr = gl + G * t;
          zp[t] = yp[t] = 0.0;
          if (rr < np) { zp[t] = __ldcg(z + roff + rr); yp[t] = __ldcg(y + roff + rr); }
        }
        if (!term && gl < NX) { zd = __ldcg(z + roff + NY + gl); yd = __ldcg(y + roff + NY + gl); }
        double vv = 0.0, nud = 0.0;
        if (term) {
          if (gl < A) vv = va[k * AE + gl];
        } else {
          if (gl < NX) {
            const double v1 = vd[(k + 1) * NXE + gl];
            vn2[gl] = v1; vn2[gl + NX] = v1;
          }
          __syncwarp(gm);
          if (gl < NX) {
            const double s = dmv<NX, SH::WNS>(sl_ + SH::SOFF, sl_ + SH::SOFF + gl, vn2 + gl, wrapS);
            nud = -(vc[k * NXE + gl] + s);
            nudv[gl] = s;
          }
          __syncwarp(gm);
          if (gl < A) {
            const double ggv = dense_mv<NX, A>(FdT + (k == 0 ? 0 : A * NX) + gl, nudv);
            v2_store<SH>(gg2, gl, ggv);
          }
          __syncwarp(gm);
          if (gl < A) vv = va[k * AE + gl] + xinv_mv<SH>(sl_, gl, gg2, wr);
        }
        if (gl < A) vcur[gl] = vv;
        __syncwarp(gm);
#pragma unroll
        for (int t = 0; t < NYT; t++) {
          const int rr = gl + G * t;
          if (rr < np) {
            const int ri = roff + rr;
            const double zt = ell_row<CW>(Cv + cls * ROWS * CW, Ci + cls * ROWS * CW, rr, vcur);
            double rho, rinv;
            rec_rho_pair<SH>(sl_, rr, rcl, rho, rinv);
            const double zo = zp[t], yo = yp[t];
            const double zh = alpha * zt + oma * zo;
            const double zn = fmin(fmax(zh + yo * rinv, sl_[SH::LO + rr]), sl_[SH::UP + rr]);
            const double d = rho * (zh - zn);
            z[ri] = zn; y[ri] = yo + d;
            zp[t] = zn; yp[t] = yo + d;  // handed to phase A of the next iteration
            if (store_delta) dy[ri] = d;
          }
        }
        if (!term && gl < NX) {
          const int ri = roff + NY + gl;
          double rho, rinv;
          rec_rho_pair<SH>(sl_, NY + gl, rcl, rho, rinv);
          const double zt = (zd - yd * rinv) + nud * rinv;
          const double zh = alpha * zt + oma * zd;
          const double zn = fmin(fmax(zh + yd * rinv, sl_[SH::LO + NY + gl]), sl_[SH::UP + NY + gl]);
          const double d = rho * (zh - zn);
          z[ri] = zn; y[ri] = yd + d;
          if (store_delta) dy[ri] = d;
          zd = zn; yd = yd + d;
        }
        double xnew = 0.0;
        if (gl >= ilo && gl < dim) {
          const double xn = alpha * vv + oma * xo;
          xnew = xn;
          x[xoff + gl] = xn;
          if (store_delta) dx[xoff + gl] = xn - xo;
        }
        __syncwarp(gm);
        // phase A of the next iteration for this stage, straight from the registers (its inputs are stage-local;
        // va[k] / vc[k] have no reader left in this iteration)
        stage_A(k, xnew, zp, yp, zd, yd);
      };
      if (LK) {
        for (int it = (wid >= 1 && wid - 1 < NI) ? (wid - 1) + ((NI - 1 - (wid - 1)) / NWK) * NWK : -1; it >= 0; it -= NWK) {
          const int k = half_stage(it);
          if (it >= 1 && it <= J) {
            const int o = 2 * it - 1;
            if (gw == 1) {
              if (!(kf[o] & 2)) {  // vn_o comes down the tree (see the other variant below)
                const int t = __ffs(o - 1) - 1, src = o + (1 << t);
                double2 rowv[NX / 2];
                ld_rows(md(t, ((o - 1) >> t) + 1), rowv);
                if (src < N) wait_flag(rb + src, iter);
                const double add = dot_rows(rowv, vd + src * NXE);
                if (gl < NX) vd[o * NXE + gl] += add;
                __syncwarp(gm);
                if (gl == 0) { __threadfence_block(); rb[o] = iter; }
              }
              // the middle stage: vn_2j = d_2j + T_2j' vn_2j+1
              if (o + 2 < N) wait_flag(rb + (o + 2), iter);
              if (gl < NX) vd[2 * it * NXE + gl] += tmv(TD + (size_t)(2 * it) * TBLK, vd + (2 * it + 1) * NXE);
            } else {
              if (o + 2 < N) wait_flag(rb + (o + 2), iter);
            }
            __syncwarp(FULL);
          } else if (it == 0) {
            if (N > 1) wait_flag(rb + 1, iter);
          }
          if (it == wid - 1) TR(8);
          if (k >= 0) stage_EA(k);
          __syncwarp(FULL);
          if (gw == 0) pre_fwd(it);  // c_2j += T_2j c_2j-1
          __syncwarp(FULL);
          if (it == wid - 1) TR(9);
        }
      }
      for (int it = (!LK && grp >= 2 && grp - 2 < NI) ? (grp - 2) + ((NI - 1 - (grp - 2)) / NGW) * NGW : -1; it >= 0; it -= NGW) {
       if (it >= 1 && it <= J) {
         const int o = 2 * it - 1;
         if (!(kf[o] & 2)) {  // vn_o comes down the tree: the start of a right child of width 2^t
           const int t = __ffs(o - 1) - 1, src = o + (1 << t);
           const double2 *r = reinterpret_cast<const double2 *>(md(t, ((o - 1) >> t) + 1) + (gl < NX ? gl : 0) * TS);
           double2 rowv[NX / 2];
#pragma unroll
           for (int j = 0; j < NX / 2; j++) rowv[j] = __ldcg(r + j);  // in flight while the flag is polled
           if (src < N) wait_flag(rb + src, iter);
           const double *v = vd + src * NXE;
           double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
           for (int j = 0; j < NX / 2; j++) {
             acc[(2 * j) & 3] = fma(rowv[j].x, v[2 * j], acc[(2 * j) & 3]);
             acc[(2 * j + 1) & 3] = fma(rowv[j].y, v[2 * j + 1], acc[(2 * j + 1) & 3]);
           }
           if (gl < NX) vd[o * NXE + gl] += (acc[0] + acc[1]) + (acc[2] + acc[3]);
           __syncwarp(gm);
           if (gl == 0) { __threadfence_block(); rb[o] = iter; }
         }
         if (it == grp - 2) TR(7);
         // fill the middle stage: vn_2j = d_2j + T_2j' vn_2j+1
         if (o + 2 < N) wait_flag(rb + (o + 2), iter);
         if (gl < NX) vd[2 * it * NXE + gl] += tmv(TD + (size_t)(2 * it) * TBLK, vd + (2 * it + 1) * NXE);
         __syncwarp(gm);
         if (it == grp - 2) TR(8);
       } else if (it == 0) {  // stage 0 reads vn_1: the end of the chain
         if (N > 1) wait_flag(rb + 1, iter);
       }
       for (int k = item_hi(it); k >= item_lo(it); k--) stage_EA(k);
       pre_fwd(it);
       if (it == grp - 2) TR(9);
      }
      if (!(can_check || can_adapt || last || time_due)) {
        __syncthreads();  // iterates of this iteration visible to every group of the CTA
        TR(10);
#ifdef RLDL_LAT_PROFILE
        if (blockIdx.x == 0 && iter == 101) {
          __threadfence(); __syncthreads();
          if (threadIdx.x == 0) {
            const long long t0_ = g_lat_tr[0][0];
            for (int g_ = 0; g_ < NG; g_++) {
              printf("trace grp %2d:", g_);
              for (int e_ = 0; e_ < 11; e_++) printf(" %6lld", g_lat_tr[g_][e_] ? g_lat_tr[g_][e_] - t0_ : -1);
              printf("\n");
            }
          }
        }
#endif
        PT(4);
        continue;
      }
      __syncthreads();
      PT(4);
      // ---------------- residuals (update_info, auxil.c:567-632), one group per stage ----------------
      {
        double pri = 0, nz = 0, nAx = 0, dua = 0, nq = 0, nAty = 0, nPx = 0, oq = 0, oP = 0;
        double *rv_ = sc, *rw_ = sc + AE, *rtd = rw_ + ROWS, *rvn = rtd + NXE, *rsp = rvn + NXE;
        for (int k = grp; k <= N; k += NG) {
          const bool term = (k == N);
          const int cls = term ? 2 : (k == 0 ? 0 : 1);
          const int dim = term ? NX : A, np = term ? NT : NY;
          const int xoff = k * A - NX, roff = k * BQ, ilo = (k == 0) ? NX : 0;
          const double *sl_ = recs + k * RS;
          const double xv = (gl >= ilo && gl < dim) ? __ldcg(x + xoff + gl) : 0.0;
          if (gl < A) rv_[gl] = xv;
#pragma unroll
          for (int t = 0; t < NYT; t++) {
            const int rr = gl + G * t;
            if (rr < np) rw_[rr] = __ldcg(y + roff + rr);
          }
          if (gl < NX) {
            rtd[gl] = term ? 0.0 : __ldcg(y + roff + NY + gl);
            rvn[gl] = term ? 0.0 : __ldcg(x + xoff + A + gl);
            rsp[gl] = (k > 0) ? __ldcg(y + roff - NX + gl) : 0.0;  // dynamics rows of stage k-1
          }
          __syncwarp(gm);
#pragma unroll
          for (int t = 0; t < NYT; t++) {
            const int rr = gl + G * t;
            if (rr < np) {
              const double ax = ell_row<CW>(Cv + cls * ROWS * CW, Ci + cls * ROWS * CW, rr, rv_), zz = __ldcg(z + roff + rr);
              pri = fmax(pri, fabs(ax - zz)); nz = fmax(nz, fabs(zz)); nAx = fmax(nAx, fabs(ax));
            }
          }
          if (!term && gl < NX) {
            const double ax = dense_mv<A, NX>(Fd + (k == 0 ? 0 : A * NX) + gl, rv_) - rvn[gl], zz = __ldcg(z + roff + NY + gl);
            pri = fmax(pri, fabs(ax - zz)); nz = fmax(nz, fabs(zz)); nAx = fmax(nAx, fabs(ax));
          }
          if (gl >= ilo && gl < dim) {
            const double px = dense_mv<A, A>(Qs + cls * A * A + gl, rv_);
            double aty = ell_row<CW>(Ctv + cls * AE * CW, Cti + cls * AE * CW, gl, rw_);
            if (!term) aty += dense_mv<NX, A>(FdT + (k == 0 ? 0 : A * NX) + gl, rtd);
            if (k > 0 && gl < NX) aty -= rsp[gl];
            const double qv = sl_[gl];
            dua = fmax(dua, fabs((qv + px) + aty)); nq = fmax(nq, fabs(qv));
            nAty = fmax(nAty, fabs(aty)); nPx = fmax(nPx, fabs(px));
            oq = fma(qv, xv, oq); oP = fma(xv, px, oP);
          }
          __syncwarp(gm);
        }
        pri = grp_max<G>(pri, gm); nz = grp_max<G>(nz, gm); nAx = grp_max<G>(nAx, gm); dua = grp_max<G>(dua, gm);
        nq = grp_max<G>(nq, gm); nAty = grp_max<G>(nAty, gm); nPx = grp_max<G>(nPx, gm);
        oq = grp_sum<G>(oq, gm); oP = grp_sum<G>(oP, gm);
        if (gl == 0) {
          double *o = red + grp * 10;
          o[0] = pri; o[1] = nz; o[2] = nAx; o[3] = dua; o[4] = nq; o[5] = nAty; o[6] = nPx; o[7] = oq; o[8] = oP;
        }
      }
      // cheap parts of the infeasibility certificates (auxil.c:364-515) over the flat vectors, whole CTA:
      // projected delta_y with its norm and support-function value; norm of delta_x and q'delta_x
      if (store_delta) {
        const double *lg = B.l + ro, *ug = B.u + ro, *qg = B.q + vo;
        const double *sEg = SH::RZ ? B.sE + ro : nullptr;  // Ruiz-scaled: finiteness is tested on the scaled bounds
        double ny_ = 0.0, lhs_ = 0.0, nx_ = 0.0, qdx_ = 0.0;
        for (int r = threadIdx.x; r < m; r += blockDim.x) {
          double d = __ldcg(dy + r);
          const double lv = __ldcg(lg + r), uv = __ldcg(ug + r);
          const double e = SH::RZ ? sEg[r] : 1.0;
          if (e * uv > kInfBound) d = (e * lv < -kInfBound) ? 0.0 : fmin(d, 0.0);
          else if (e * lv < -kInfBound) d = fmax(d, 0.0);
          dy[r] = d;
          ny_ = fmax(ny_, fabs(d));
          lhs_ += uv * fmax(d, 0.0) + lv * fmin(d, 0.0);
        }
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
          const double d = __ldcg(dx + i);
          nx_ = fmax(nx_, fabs(d));
          qdx_ = fma(__ldcg(qg + i), d, qdx_);
        }
        ny_ = grp_max<32>(ny_, FULL); nx_ = grp_max<32>(nx_, FULL);
        lhs_ = grp_sum<32>(lhs_, FULL); qdx_ = grp_sum<32>(qdx_, FULL);
        if (lane == 0) {
          double *o = red + NG * 10 + wid * 4;
          o[0] = ny_; o[1] = lhs_; o[2] = nx_; o[3] = qdx_;
        }
      }
      __syncthreads();
      // ---------------- epilogue by group 0: termination, rho adaptation ----------------
      if (grp == 0) {
        bool done = false, rho_changed = false;
        int status = -10;
        double obj = 0.0, rho_new = rho_in;
        Resid R = {0, 0, 0, 0, 0, 0, 0, 0};
        const double cs = SH::RZ ? B.sC[inst] : 1.0;
        {
          double a[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
          for (int g2 = gl; g2 < NG; g2 += G) {
#pragma unroll
            for (int e = 0; e < 7; e++) a[e] = fmax(a[e], red[g2 * 10 + e]);
            a[7] += red[g2 * 10 + 7]; a[8] += red[g2 * 10 + 8];
          }
#pragma unroll
          for (int e = 0; e < 7; e++) a[e] = grp_max<G>(a[e], gm);
          a[7] = grp_sum<G>(a[7], gm); a[8] = grp_sum<G>(a[8], gm);
          R.pri = a[0]; R.nz = a[1]; R.nAx = a[2]; R.dua = a[3]; R.nq = a[4]; R.nAty = a[5]; R.nPx = a[6];
          R.obj = 0.5 * a[8] + a[7];
        }
        double cny = 0.0, clhs = 0.0, cnx = 0.0, cqdx = 0.0;
        if (store_delta) {
          for (int w2 = gl; w2 < kLatThreads / 32; w2 += G) {
            const double *o = red + NG * 10 + w2 * 4;
            cny = fmax(cny, o[0]); clhs += o[1]; cnx = fmax(cnx, o[2]); cqdx += o[3];
          }
          cny = grp_max<G>(cny, gm); cnx = grp_max<G>(cnx, gm);
          clhs = grp_sum<G>(clhs, gm); cqdx = grp_sum<G>(cqdx, gm);
        }
        // check_termination (auxil.c:684-789) with the certificates pre-filtered by the CTA-wide partials: the
        // full routines (a stage sweep each) only run when the cheap necessary conditions hold
        auto check_term = [&](int approximate) -> int {
          if (R.pri > kInfty || R.dua > kInfty) { status = -7; obj = kOsqpNan; return 1; }
          double ea = S.eps_abs, er = S.eps_rel, epi = S.eps_prim_inf, edi = S.eps_dual_inf;
          if (approximate) { ea *= 10; er *= 10; epi *= 10; edi *= 10; }
          int prim_ok = 0, dual_ok = 0, prim_inf = 0, dual_inf = 0;
          if (m == 0) prim_ok = 1;
          else if (R.pri < ea + er * fmax(R.nz, R.nAx)) prim_ok = 1;
          else if (cs * cny > epi && clhs < -epi * cny) prim_inf = primal_infeasible<G>(M, B, inst, gl, gm, rs, epi, m);
          if (R.dua < ea + er * fmax(fmax(R.nq, R.nAty), R.nPx)) dual_ok = 1;
          else if (cnx > edi && cqdx < -edi * cnx) dual_inf = dual_infeasible<G>(M, B, inst, gl, gm, rs, edi, n);
          if (prim_ok && dual_ok) { status = approximate ? 2 : 1; return 1; }
          if (prim_inf) { status = approximate ? 3 : -3; obj = kInfty; return 1; }
          if (dual_inf) { status = approximate ? 4 : -4; obj = -kInfty; return 1; }
          return 0;
        };
        if (can_check) {
          obj = R.obj;
          if (check_term(0)) done = true;
        }
        // the rho estimate reads the norms of the SCALED problem (auxil.c:13-55); the stage-parallel residuals above are
        // the unscaled ones, so the Ruiz-scaled variants take one sweep of the shared routine where the estimate is used
        Resid Rs = R;
        if (!done && can_adapt) {
          if (!can_check) { obj = R.obj; }
          if constexpr (SH::RZ) compute_residuals<G>(M, B, inst, gl, gm, rs, &Rs);
          const double rn = rho_estimate(Rs, rho_in);
          if (rn > rho_in * S.adaptive_rho_tolerance || rn < rho_in / S.adaptive_rho_tolerance) {
            rho_changed = true;
            rho_new = fmin(fmax(rn, kRhoMin), kRhoMax);
          }
        }
        bool timed_out = false;
        if (!done && time_due && !last) {  // osqp.c:389-405
          int up = gl == 0 ? (int)time_is_up(B, S) : 0;
          timed_out = __shfl_sync(gm, up, 0, G) != 0;
        }
        if (!done && (last || timed_out)) {
          if (!can_check) {
            obj = R.obj;
            check_term(0);
          }
          if (status == -10) {  // osqp.c:562-576
            if (!check_term(1)) status = timed_out ? -6 : -2;
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
          if constexpr (SH::RZ) {
            if (has_sol) compute_residuals<G>(M, B, inst, gl, gm, rs, &Rs);
          }
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
            trace_finish(B, inst, 2);
          }
        } else if (rho_changed) {
          // parked for the batched re-factorisation, resumed afterwards
          if (gl == 0) { B.iter[inst] = iter; B.status[inst] = kStatusParkedRho; }
        }
        if (gl == 0) ctl[1] = done ? 2 : (rho_changed ? 1 : 0);
      }
      __syncthreads();
      PT(5);
      if (ctl[1]) break;
    }
    if (ctl[1] == 2 && B.h_status)  // fused solve-and-deliver: the whole CTA copies the finished instance out
      deliver_to_host(B, inst, threadIdx.x, blockDim.x, n, m, __ldcg(B.status + inst), __ldcg(B.iter + inst),
                      __ldcg(B.obj + inst), __ldcg(B.pri + inst), __ldcg(B.dua + inst), __ldcg(B.rho_est + inst));
#ifdef RLDL_LAT_PROFILE
    unsigned long long gt1_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt1_));
    if (threadIdx.x == 0 && blockIdx.x < 3)
      printf("ns/iter %lld  ", (long long)(gt1_ - gt0_) / max(iter - iter0_, 1));
    if (threadIdx.x == 0 && blockIdx.x < 3)
      printf("lat profile inst %d iters %d: ns/iter up %lld B %lld C %lld up %lld D %lld E %lld epi %lld  (T waits in B+D %lld)\n", inst, iter,
             pt[0] / iter, pt[1] / iter, pt[2] / iter, pt[6] / iter, pt[3] / iter, pt[4] / iter, pt[5] / iter, twait_ / iter);
#endif
    __syncthreads();  // ctl is rewritten by the next claim
  }
}

// ------------------------------- launcher --------------------------------
template <class SH>
static size_t latency_smem_t(int N) {
  return (size_t)LatLayout<SH>(N).end * sizeof(double);
}

template <class SH>
static cudaError_t launch_latency_t(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                                    int resume, double *Tg, int ctas, int reset_parked, int claim_mode, int stream_ctas,
                                    cudaStream_t st) {
  const size_t shm = latency_smem_t<SH>(B.N);
  auto kern = k_admm_cta<SH>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm);
  if (e != cudaSuccess) return e;
  // fresh counter; the parked-for-rho counter only if no streaming launch of this round has counted into it
  if (claim_mode != 2) {  // (the concurrent lane shares the scheduler state of the running streaming kernel)
    e = cudaMemsetAsync(B.sched, 0, (reset_parked ? 2 : 1) * sizeof(int), st);
    if (e != cudaSuccess) return e;
  }
  kern<<<(unsigned)ctas, kLatThreads, shm, st>>>(M, B, S, warm_start, resume, Tg, claim_mode, stream_ctas);
  return cudaGetLastError();
}

int latency_t_blocks(int Nmax) { return lat_t_blocks(Nmax); }

size_t latency_smem_bytes(const ModelDev &M, int N) {
  const int cw = max_c_width(M);
#define RLDL_X(a_, nx_, ny_, nt_, cw_, bd_, sc_) \
  if (shape_serves(M, cw, a_, nx_, ny_, nt_, cw_, bd_, sc_)) return latency_smem_t<Shape<a_, nx_, ny_, nt_, cw_, bd_, sc_>>(N);
  RLDL_SHAPES(RLDL_X)
#undef RLDL_X
  return (size_t)-1;
}

cudaError_t launch_admm_latency(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                                int resume, double *Tg, int ctas, int reset_parked, int claim_mode, int stream_ctas,
                                cudaStream_t st) {
  const int cw = max_c_width(M);
#define RLDL_X(a_, nx_, ny_, nt_, cw_, bd_, sc_)                    \
  if (shape_serves(M, cw, a_, nx_, ny_, nt_, cw_, bd_, sc_))         \
    return launch_latency_t<Shape<a_, nx_, ny_, nt_, cw_, bd_, sc_>>(M, B, S, warm_start, resume, Tg, ctas, reset_parked, \
                                                                 claim_mode, stream_ctas, st);
  RLDL_SHAPES(RLDL_X)
#undef RLDL_X
  return cudaErrorInvalidValue;
}

}  // namespace rldl
