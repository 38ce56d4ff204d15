// rldl_batch.h — the batch handle behind the C ABI (RLDLBatch == rldl::Batch)
#pragma once
#include <cuda_runtime.h>
#include "../../include/osqp_rldl_abi.h"
#include "rldl_device.h"
#include "rldl_kernels.h"

namespace rldl {

c_int validate_settings(const OSQPSettings *s);

struct Batch {
  int device = 0;
  bool generic = false;  // osqp_setup path: one dense stage (N == 0)
  int kernel_choice = 0;   // 0 auto, 1 force the generic (runtime-shape) ADMM kernel
  // hand-off threshold: unfinished instances at which the latency-mode kernel takes over (0: never; -1: automatic,
  // 2048 for batches of 4096 and more — the suspended instances are resumed most-unconverged first, so an earlier
  // hand-off shortens the chain-bound tail of 8k-16k shards by 7 % and is neutral for larger ones — else 1024)
  int lat_below = -1;
  StreamTuning tune;       // launch knobs of the streaming kernel (rldl_batch_set_tuning)
  // concurrent latency-mode lane: lat_lane CTAs of the latency-mode kernel run beside the streaming kernel on the SMs
  // its CTAs vacate once at most lat_retire_below instances are unfinished; lane_tail (> 0) replaces lat_below as the
  // streaming kernel's stop threshold while the lane runs; lat_lane = 0: off
  int lat_lane = 8, lat_retire_below = 0, lane_tail = 0;  // (0: retire at 6 x lat_below; tail at lat_below)
  double *d_ratio = nullptr;
  double *d_trace = nullptr;  // optional per-instance finish-time trace
  c_int set_trace(bool on);
  c_int get_trace(c_float *t);
  cudaStream_t stream2 = nullptr;  // the lane's stream (lower priority than `stream`)
  cudaEvent_t ev_reset = nullptr, ev_lane = nullptr, ev_s0 = nullptr, ev_s1 = nullptr;
  bool stream_timed = false;
  double last_stream_ms = 0, last_stream_iters = 0;  // the streaming kernel of the last solve alone: time, instance-iterations
  bool streaming = false;  // shapes the streaming kernels cover (rldl_stream.cu)
  bool need_unpacked = false;   // a generic-kernel consumer wants Xinv/Sinv unpacked as well as the records
  bool unpacked_valid = false;  // D.Xinv / D.Sinv match the current factorisation
  bool z_pending = false;       // warm_start_shift moved x: z = A x is formed at the next solve
  bool records_stale = false;   // the generic adaptive-rho kernel re-factorised without refreshing the records
  int nx = 0, nu = 0, ny = 0, nt = 0, a = 0, b = 0, N = 0, Nmax = 0, batch = 0, nmax = 0, mmax = 0;
  OSQPSettings settings;
  ModelDev M;
  BatchDev D;
  SettingsDev S;
  double *d_mval = nullptr;
  int *d_midx = nullptr;
  int *d_changed = nullptr;
  double *d_stage = nullptr;
  double *d_latT = nullptr;  // per-CTA T_k blocks of the latency-mode kernel
  double *d_l_alt = nullptr, *d_u_alt = nullptr;  // landing buffers of validated bound updates
  // polish (settings.polish): second factor, its step sizes, three work vectors per instance, outcome
  double *d_polX = nullptr, *d_polS = nullptr, *d_polrho = nullptr, *d_polb = nullptr, *d_pols = nullptr, *d_polr = nullptr;
  int *d_polstat = nullptr;
  double last_polish_ms = 0;
  c_int polish();
  c_int get_polish(c_int *status_polish);
  size_t stage_cap = 0;
  double *h_pin = nullptr;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  double last_factor_ms = 0, last_solve_ms = 0, factor_bytes = 0;
  long long launches = 0;
  double time_spent_before_solve = 0.0;  // seconds already charged against settings.time_limit (scalar API)
  long long solve_rounds = 0;  // extra ADMM launches caused by adaptive-rho refactorisations

  static c_int create(Batch **out, const csc *Q0, const csc *Qi, const csc *QN, const csc *A0,
                      const csc *Ai, const csc *Aij, const csc *AN, const OSQPSettings *st, c_int Nmax,
                      c_int N, c_int nx, c_int nu, c_int ny, c_int nt, c_int batch, const c_float *q,
                      const c_float *l, const c_float *u, const c_float *rho, bool generic);
  ~Batch();

  bool scaled() const { return settings.scaling > 0; }
  bool scale_stale = false;  // the model changed since the last equilibration
  c_int rescale(bool carry_iterates = false);
  bool have_scaling = false;  // an equilibration has been computed on this handle
  c_int fetch_scaling(int inst, c_float *Dv, c_float *Ev, c_float *c);
  c_int configure_stream();
  size_t rec_cap = 0;       // doubles allocated for D.rec
  bool slots_full = false;  // D.sched_slots sized for the scheduler
  c_int upload_model(const csc *Q0, const csc *Qi, const csc *QN, const csc *A0, const csc *Ai,
                     const csc *Aij, const csc *AN);
  void push_settings();
  void set_sizes();
  c_int ensure_stage(size_t doubles);
  c_int factor(int k0, bool only_flagged);
  // instances whose last stage recursion met a non-positive pivot (the reference's "problem seems to be
  // non-convex", lin_sys/direct/qdldl/qdldl_interface.c LDL_factor); -1 on a CUDA error
  c_int failed_factorisations(c_int *flag_out = nullptr);
  c_int factor_launch(int k0, bool only_flagged, bool mark = false);  // kernels only, no events / sync
  c_int ensure_unpacked();
  c_int retype(int k0, int parts, bool only_parked, bool want_flags, bool all_rows = false);
  bool use_stream() const;
  c_int solve();
  c_int retype_and_refactor();
  c_int reset_info();
  c_int update_bounds(const c_float *l, const c_float *u, bool have_l, bool have_u, bool validate = false);
  c_int partial_update_bounds(c_int start, c_int stop, const c_float *l, const c_float *u);
  c_int update_lin_cost(const c_float *q);
  c_int update_data_refactor(const c_float *q, const c_float *l, const c_float *u);
  cudaEvent_t ev_chunk[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  c_int warm_start(const c_float *x, const c_float *y);
  c_int warm_start_shift();
  c_int update_horizon(c_int Nnew);
  c_int update_rho(const c_float *rho);
  c_int refactor();
  c_int kkt_solve(c_float *rhs);
  c_int kkt_solve_static(const c_float *rhs, c_float *sol);
  c_int solve_to_host(c_float *x, c_float *y, c_int *iter, c_int *status, c_float *obj, c_float *pri, c_float *dua,
                      c_int *rho_updates, c_float *rho_est);
  c_int get_solution(c_float *x, c_float *y);
  c_int get_info(c_int *iter, c_int *status, c_float *obj, c_float *pri, c_float *dua,
                 c_int *rho_updates, c_float *rho_est);
  c_int fetch_vectors(int inst, c_float *x, c_float *y, c_float *z, c_float *dx, c_float *dy,
                      c_float *rho_vec, c_float *rho);
  c_int update_settings(const OSQPSettings *s);
  double admm_bytes_per_iteration() const;
};

}  // namespace rldl
