// rldl_device.h — device-side data model shared by the kernels and the host API.
//
// Stage-ordered block elimination of the MPC KKT matrix (reference
// src/recursive_ldl.c:1139-1318), re-designed for HBM
// streaming.  Per stage k the KKT rows are
//     v_k  : (Q_k + sigma I) v_k + C_k' nu^p_k + F_k' nu^d_k - J' nu^d_{k-1} = r^v_k
//     p_k  : C_k v_k - diag(1/rho^p_k) nu^p_k                              = r^p_k
//     d_k  : F_k v_k - J v_{k+1} - diag(1/rho^d_k) nu^d_k                  = r^d_k
// (v_k = (x_k,u_k), C_k = path rows, F_k = [A B] dynamics rows, J = [I_nx 0]).
// Pivot order inside a stage: path rows, stage variables, dynamics rows — an
// algebraically equivalent ordering of the reference's (variables, all rows):
//     M_k = Q_k + sigma I + C_k' diag(rho^p_k) C_k
//     X_0 = M_0,  S_k = diag(1/rho^d_k) + F_k X_k^-1 F_k',  X_{k+1} = M_{k+1} + J' S_k^-1 J
// The stored factor is the pair of inverted diagonal pivots per stage,
//     Xinv_k (a x a)   and   Sinv_k (nx x nx),
// so every block substitution is a small dense symmetric mat-vec with no
// sequential dependence inside the block.  Stage 0 carries nx phantom variables
// (identity pivot, zero coupling) so all stages have the same shape.
#pragma once
#include <cstdint>

// In-kernel bounds asserts of the debug build (make debug: -DRLDL_DEBUG); compiled out otherwise.
#ifdef RLDL_DEBUG
#include <cstdio>
#define RLDL_ASSERT(cond)                                                                                  \
  do {                                                                                                     \
    if (!(cond)) {                                                                                         \
      printf("RLDL_ASSERT failed: %s  (%s:%d, block %d thread %d)\n", #cond, __FILE__, __LINE__, (int)blockIdx.x, \
             (int)threadIdx.x);                                                                            \
      __trap();                                                                                            \
    }                                                                                                      \
  } while (0)
#else
#define RLDL_ASSERT(cond) ((void)0)
#endif

namespace rldl {

// Row-wise ELL view of one small shared model matrix inside the model buffer.
struct Ell {
  int rows;
  int width;  // entries per row (padded with idx 0 / val 0)
  int off;    // offset of idx[] in ModelDev::idx and of val[] in ModelDev::val
};

struct StageClass {
  int Q;      // offset (doubles) of the dense symmetric cost block, dim x dim
  Ell C, Ct;  // path rows      (np x dim) and transpose
  Ell F, Ft;  // dynamics rows  (nx x dim) and transpose (unused for the terminal class)
};

// Shared stage model: class 0 = stage 0 (phantom-padded), 1 = interior, 2 = terminal.
struct ModelDev {
  int nx, nu, ny, nt, a, b;
  StageClass sc[3];
  const double *val;
  const int *idx;
  int nval;  // entries in val[] / idx[] (the ELL part is copied to shared memory by the streaming kernels)
  int bd;    // 1: no state-input cross terms in the cost and one-variable path rows => block-diagonal pivots
  int scaled;  // 1: the handle is Ruiz-scaled (settings.scaling > 0): selects the record layout with per-row step sizes
};

struct SettingsDev {
  double sigma, alpha, eps_abs, eps_rel, eps_prim_inf, eps_dual_inf;
  double adaptive_rho_tolerance;
  int max_iter, check_termination, adaptive_rho, adaptive_rho_interval;
  // time_limit (reference src/osqp.c:389-405): nanoseconds of solve time left when the launch sequence of a
  // solve starts (0 = no limit), measured on the device against sched[kSchedT0] (%globaltimer at solve start)
  long long time_limit_ns;
};

// Per-instance arrays are instance-major with the strides below.
struct BatchDev {
  int batch, N, Nmax;
  int n, m;                // current problem sizes (N*a, N*b+nt; N==0: nx, nt)
  int nmax, mmax;          // vector strides
  int xstride, sstride;    // factor strides (doubles): (Nmax+1)*a*a, Nmax*nx*nx
  int ykstride, skstride;  // (Nmax+1)*a, Nmax*nx
  double *q, *l, *u, *rho_vec;
  double *x, *z, *y, *dx, *dy;
  double *Xinv, *Sinv;
  double *yk, *sk;
  double *rho;             // current scalar rho per instance
  long long *iter;
  int *status;
  double *obj, *pri, *dua, *rho_est;
  int *rho_updates;
  int *fact_flag;          // 0 ok, >0 = 1 + stage of the first non-positive pivot
  // Ruiz equilibration of the reference (src/scaling.c:44-156), per instance: sD[nmax] variable scaling D,
  // sE[mmax] constraint scaling E, sC cost scaling c; all null when settings.scaling == 0.  The kernels never
  // form scaled data: the scaled ADMM iteration is the unscaled one with per-variable sigma_i = sigma / (c D_i^2)
  // and per-row rho_j = rho_class(E_j l_j, E_j u_j) E_j^2 / c, which is what rho_vec then holds.
  double *sD, *sE, *sC;
  // Fused solve-and-deliver (rldl_batch_solve_to_host): device-accessible addresses of the caller's PINNED host
  // arrays, or null.  The group / CTA that finishes an instance writes its x, y (OSQP_NAN when the status carries
  // no solution) and info scalars straight to host memory, so the device-to-host transfer of a batch runs behind
  // the rest of the solve instead of after it.  Layout as rldl_batch_get_solution / rldl_batch_get_info.
  double *hx, *hy, *h_obj, *h_pri, *h_dua, *h_rho_est;
  long long *h_iter, *h_status, *h_rho_updates;
  // optional trace (rldl_batch_set_tuning "trace"): per instance, milliseconds from the start of the solve to the
  // moment the instance finished, plus 1e6 x the engine that finished it (1 streaming, 2 latency-mode, 3 runtime-shape)
  double *trace;
  // max(r_p / eps_p, r_d / eps_d) of every instance at its previous termination check in the streaming kernel (0: none
  // yet): with the current ratio it gives the convergence rate and from it the predicted remaining iterations, the
  // urgency the device scheduler sorts by
  double *ratio_prev;
  int lat_lockstep;  // latency-mode kernel: 1 = a warp per item, its two lane groups on the two stages side by side
  int lat_levels;  // latency-mode kernel: levels of the block tree under the sequential chain (1 = pairs only)
  int sched_key;  // 0: urgency = log of the distance from the tolerances (round 1); 1: log of the predicted remaining iterations
  // Streaming layout (rldl_stream.cu): one record per (instance, stage), RS doubles:
  //   [ q_k (qa) | Xinv_k packed (PX) | Sinv_k packed (PS) | row classes (kClsWords) | l_k (bb) | u_k (bb) ]
  // packed = lower triangle, column-major, of the a x a / nx x nx symmetric pivot inverse;
  // every part is padded to an even number of doubles (16-byte bulk-copy granularity).
  // bd = 1 (block-diagonal pivots): the Xinv part holds the nx x nx block (PXX doubles) and then the nu x nu block.
  double *rec;
  int qa, PX, PS, bb, RS;
  int bd, PXX;
  int sigw, clsw;          // doubles of the sigma part (0 unless Ruiz-scaled) and of the row-class / step-size part
  long long rec_stride;    // (Nmax+1)*RS doubles per instance
  // Device work scheduler of the persistent ADMM kernel (rldl_stream.cu):
  //   sched[0] fresh-instance counter, sched[1] instances parked for a rho refactorisation,
  //   sched[2 .. 2+kSchedRings) ring heads, sched[2+kSchedRings .. 2+2*kSchedRings) ring tails (rings 0..31: the
  //   priority buckets; ring kSchedLQ: instances handed to the latency-mode CTAs that run beside the streaming kernel);
  //   sched_slots[b*batch + (index mod batch)] = suspended instance, -1 = empty.
  int *sched;
  int *sched_slots;
};

constexpr int kClsWords = 2;  // 64-bit words of 2-bit row classes per stage record (up to 64 rows per stage)
constexpr int kSchedBuckets = 32;  // priority buckets: floor(2*log2(residual / tolerance)), clamped
// sched[kSchedUnfinished]: instances of the current solve not yet finished (kept across the launches of
// one solve: streaming launch, latency-mode launch, adaptive-rho resume rounds); kSchedCounters = array size
constexpr int kSchedLQ = kSchedBuckets;         // ring of instances handed over to the concurrent latency-mode CTAs
constexpr int kSchedRings = kSchedBuckets + 1;
constexpr int kSchedUnfinished = 2 + 2 * kSchedRings;
constexpr int kSchedT0 = kSchedUnfinished + 2;  // 64-bit %globaltimer value at the start of the current solve
// concurrent latency-mode lane: latency-mode CTAs waiting for an instance, the urgency bucket an instance
// must have to be handed over (decaying maximum), streaming CTAs started, "the streaming kernel has hit its tail / ended"
constexpr int kSchedLWant = kSchedT0 + 2, kSchedTop = kSchedT0 + 3, kSchedStarted = kSchedT0 + 4, kSchedTailFlag = kSchedT0 + 5;
// lane statistics of the last streaming launch: hand-overs, lane CTAs that left early, instances a lane CTA took
// from the hand-over ring / from the buckets, SMs that retired
constexpr int kSchedStats = kSchedT0 + 6;
// 64-bit count of the ADMM iterations (over all instances) the last streaming launch executed itself — the rest of a
// solve's iterations ran in the latency-mode kernel; the roofline of the streaming kernel is quoted on this count
constexpr int kSchedStreamIters = kSchedStats + 6;
constexpr int kSchedCounters = kSchedStreamIters + 2;

constexpr int STATUS_FACTOR_ERROR = -7;  // reported as OSQP_NON_CVX
// device-internal: instance parked by an adaptive-rho update until its factor has been refreshed (never
// visible after a solve: k_set_rho_vec turns it back into OSQP_UNSOLVED before the resume launch)
constexpr int kStatusParkedRho = -11;

}  // namespace rldl
