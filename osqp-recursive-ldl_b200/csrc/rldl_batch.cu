// rldl_batch.cu — host side of the batch handle (RLDLBatch): model upload, HBM
// residency of all per-instance data, kernel sequencing.  C++ above the C ABI; no
// CPU compute path exists here — every numerical step is a kernel launch.
#include <cuda_runtime.h>
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>
#include "rldl_batch.h"

namespace rldl {

#define CK(call)                                                                            \
  do {                                                                                      \
    cudaError_t e_ = (call);                                                                \
    if (e_ != cudaSuccess) {                                                                \
      fprintf(stderr, "ERROR in %s: CUDA failure '%s' at %s:%d\n", __func__,                \
              cudaGetErrorString(e_), __FILE__, __LINE__);                                  \
      return OSQP_LINSYS_SOLVER_INIT_ERROR;                                                 \
    }                                                                                       \
  } while (0)

// ------------------------------ model upload ------------------------------
static std::vector<double> csc_dense(const csc *M, int rows, int cols, bool symmetrize) {
  std::vector<double> D((size_t)rows * cols, 0.0);
  for (c_int j = 0; j < M->n && j < cols; j++)
    for (c_int p = M->p[j]; p < M->p[j + 1]; p++) {
      const c_int i = M->i[p];
      if (i < 0 || i >= rows) continue;
      D[(size_t)i * cols + j] += M->x[p];
      if (symmetrize && i != j) D[(size_t)j * cols + i] += M->x[p];
    }
  return D;
}

struct ModelBuilder {
  std::vector<double> val;
  std::vector<int> idx;
  int add_dense(const std::vector<double> &D) {
    const int off = (int)val.size();
    val.insert(val.end(), D.begin(), D.end());
    idx.resize(val.size(), 0);
    return off;
  }
  // rows x cols dense row-major -> ELL; transpose=true builds the ELL of D'
  Ell add_ell(const std::vector<double> &D, int rows, int cols, bool transpose) {
    const int R = transpose ? cols : rows, Cn = transpose ? rows : cols;
    auto at = [&](int r, int c) { return transpose ? D[(size_t)c * cols + r] : D[(size_t)r * cols + c]; };
    int width = 1;
    for (int r = 0; r < R; r++) {
      int cnt = 0;
      for (int c = 0; c < Cn; c++) cnt += at(r, c) != 0.0;
      width = std::max(width, cnt);
    }
    Ell E;
    E.rows = R; E.width = R ? width : 0; E.off = (int)val.size();
    for (int r = 0; r < R; r++) {
      int cnt = 0;
      for (int c = 0; c < Cn; c++)
        if (at(r, c) != 0.0) { val.push_back(at(r, c)); idx.push_back(c); cnt++; }
      for (; cnt < width; cnt++) { val.push_back(0.0); idx.push_back(0); }
    }
    return E;
  }
};

template <class T>
static cudaError_t dalloc(T **p, size_t count) {
  *p = nullptr;
  if (!count) count = 1;
  cudaError_t e = cudaMalloc((void **)p, count * sizeof(T));
  if (e == cudaSuccess) e = cudaMemset(*p, 0, count * sizeof(T));
  // the memset runs on the legacy default stream and the handle's stream is non-blocking (no implicit order
  // with it): without this wait a fill delayed by other host threads' work can land after the handle's first writes
  if (e == cudaSuccess) e = cudaStreamSynchronize(0);
  return e;
}

static bool check_dims(const csc *M, c_int m, c_int n) { return M && M->m == m && M->n == n && M->p; }

c_int Batch::upload_model(const csc *Q0, const csc *Qi, const csc *QN, const csc *A0, const csc *Ai,
                          const csc *Aij, const csc *AN) {
  ModelBuilder mb;
  ModelDev Mnew = M;  // built aside: a refused update must leave the resident model untouched
  ModelDev &Md = Mnew;
  Md.nx = nx; Md.nu = nu; Md.ny = ny; Md.nt = nt; Md.a = a; Md.b = b;
  memset(Md.sc, 0, sizeof(Md.sc));
  Md.bd = 0;
  if (!generic) {
    if (!check_dims(Q0, nu, nu) || !check_dims(Qi, a, a) || !check_dims(QN, nx, nx) ||
        !check_dims(A0, b, nu) || !check_dims(Ai, b, a) || !check_dims(AN, nt, nx) ||
        (Aij && !check_dims(Aij, b, a))) {
      fprintf(stderr, "ERROR in %s: stage matrix dimensions do not match nx,nu,ny,nt\n", __func__);
      return OSQP_DATA_VALIDATION_ERROR;
    }
    if (Aij) {  // must be [0 0; -I 0] — hard-wired in the reference too (src/recursive_ldl.c:584-586)
      std::vector<double> D = csc_dense(Aij, b, a, false);
      for (int r = 0; r < b; r++)
        for (int c = 0; c < a; c++) {
          const double want = (r >= ny && c == r - ny) ? -1.0 : 0.0;
          if (D[(size_t)r * a + c] != want) {
            fprintf(stderr, "ERROR in %s: Aij must be [0 0; -I 0]\n", __func__);
            return OSQP_DATA_VALIDATION_ERROR;
          }
        }
    }
    // class 0: stage 0, padded with nx phantom variables (identity pivot, zero coupling)
    {
      std::vector<double> q0 = csc_dense(Q0, nu, nu, true), a0 = csc_dense(A0, b, nu, false);
      std::vector<double> Q((size_t)a * a, 0.0), C((size_t)ny * a, 0.0), F((size_t)nx * a, 0.0);
      for (int i = 0; i < nx; i++) Q[(size_t)i * a + i] = 1.0;
      for (int i = 0; i < nu; i++)
        for (int j = 0; j < nu; j++) Q[(size_t)(nx + i) * a + nx + j] = q0[(size_t)i * nu + j];
      for (int r = 0; r < ny; r++)
        for (int j = 0; j < nu; j++) C[(size_t)r * a + nx + j] = a0[(size_t)r * nu + j];
      for (int r = 0; r < nx; r++)
        for (int j = 0; j < nu; j++) F[(size_t)r * a + nx + j] = a0[(size_t)(ny + r) * nu + j];
      StageClass &s = Md.sc[0];
      s.Q = mb.add_dense(Q);
      s.C = mb.add_ell(C, ny, a, false); s.Ct = mb.add_ell(C, ny, a, true);
      s.F = mb.add_ell(F, nx, a, false); s.Ft = mb.add_ell(F, nx, a, true);
    }
    {
      std::vector<double> Q = csc_dense(Qi, a, a, true), A = csc_dense(Ai, b, a, false);
      std::vector<double> C(A.begin(), A.begin() + (size_t)ny * a), F(A.begin() + (size_t)ny * a, A.end());
      Md.bd = 1;  // state-input cross terms of the cost? (stage 0 and the terminal stage have none by construction)
      for (int i = 0; i < nx; i++)
        for (int j = nx; j < a; j++)
          if (Q[(size_t)i * a + j] != 0.0 || Q[(size_t)j * a + i] != 0.0) Md.bd = 0;
      StageClass &s = Md.sc[1];
      s.Q = mb.add_dense(Q);
      s.C = mb.add_ell(C, ny, a, false); s.Ct = mb.add_ell(C, ny, a, true);
      s.F = mb.add_ell(F, nx, a, false); s.Ft = mb.add_ell(F, nx, a, true);
    }
  }
  {
    if (generic && (!check_dims(QN, nx, nx) || !check_dims(AN, nt, nx))) {
      fprintf(stderr, "ERROR in %s: P must be n x n and A m x n\n", __func__);
      return OSQP_DATA_VALIDATION_ERROR;
    }
    std::vector<double> Q = csc_dense(QN, nx, nx, true), C = csc_dense(AN, nt, nx, false);
    StageClass &s = Md.sc[2];
    s.Q = mb.add_dense(Q);
    s.C = mb.add_ell(C, nt, nx, false); s.Ct = mb.add_ell(C, nt, nx, true);
    s.F.rows = s.F.width = s.F.off = 0; s.Ft = s.F;
  }
  // the new buffers are allocated and filled before the old ones go: a failure leaves the resident model as it was
  double *nval = nullptr;
  int *nidx = nullptr;
  if (cudaMalloc(&nval, mb.val.size() * sizeof(double)) != cudaSuccess ||
      cudaMalloc(&nidx, mb.idx.size() * sizeof(int)) != cudaSuccess ||
      cudaMemcpy(nval, mb.val.data(), mb.val.size() * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaMemcpy(nidx, mb.idx.data(), mb.idx.size() * sizeof(int), cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaStreamSynchronize(0) != cudaSuccess) {  // (pageable copies may still be in flight when cudaMemcpy returns)
    fprintf(stderr, "ERROR in %s: CUDA failure while uploading the stage model\n", __func__);
    if (nval) cudaFree(nval);
    if (nidx) cudaFree(nidx);
    return OSQP_LINSYS_SOLVER_INIT_ERROR;
  }
  if (stream) CK(cudaStreamSynchronize(stream));  // nothing of this handle may still read the old model
  if (d_mval) cudaFree(d_mval);
  if (d_midx) cudaFree(d_midx);
  d_mval = nval; d_midx = nidx;
  Md.val = d_mval; Md.idx = d_midx; Md.nval = (int)mb.val.size();
  M = Mnew;
  scale_stale = true;
  if (D.rec) return configure_stream();  // a P / A update may change which specialisation serves the model
  return 0;
}

// Which kernels serve the model, and the stage-record layout that goes with them.
c_int Batch::configure_stream() {
  // Ruiz-scaled problems carry a step size per row and a sigma per variable: the BASELINE shapes have streaming
  // variants with that record layout (Shape<..., RZ = 1>), other scaled models run on the runtime-shape kernels
  M.scaled = scaled() ? 1 : 0;
  streaming = !generic && stream_supported(M, N);
  stream_record_dims(M, &D.qa, &D.PX, &D.PS, &D.bb, &D.bd, &D.PXX, &D.sigw, &D.clsw);
  D.RS = D.qa + D.sigw + D.PX + D.PS + D.clsw + 2 * D.bb;
  D.rec_stride = (long long)(Nmax + 1) * D.RS;
  const size_t bs = (size_t)batch;
  const size_t want = streaming ? bs * (size_t)D.rec_stride : 1;
  if (want > rec_cap) {
    if (D.rec) cudaFree(D.rec);
    D.rec = nullptr; rec_cap = 0;
    CK(dalloc(&D.rec, want));
    rec_cap = want;
  }
  if (streaming && !slots_full) {
    if (D.sched_slots) cudaFree(D.sched_slots);
    D.sched_slots = nullptr;
    CK(dalloc(&D.sched_slots, (size_t)kSchedRings * bs));
    slots_full = true;
  } else if (!D.sched_slots) {
    CK(dalloc(&D.sched_slots, 1));
  }
  unpacked_valid = false;
  return 0;
}

// ------------------------------ settings ------------------------------
c_int validate_settings(const OSQPSettings *s) {
  // reference src/auxil.c:889-1061
#define BAD(msg) do { fprintf(stderr, "ERROR in validate_settings: %s\n", msg); return 1; } while (0)
  if (!s) BAD("Missing settings!");
  if (s->scaling < 0) BAD("scaling must be nonnegative");
  if (s->adaptive_rho != 0 && s->adaptive_rho != 1) BAD("adaptive_rho must be either 0 or 1");
  if (s->adaptive_rho_interval < 0) BAD("adaptive_rho_interval must be nonnegative");
  if (s->adaptive_rho_fraction <= 0) BAD("adaptive_rho_fraction must be positive");
  if (s->adaptive_rho_tolerance < 1.0) BAD("adaptive_rho_tolerance must be >= 1");
  if (s->polish_refine_iter < 0) BAD("polish_refine_iter must be nonnegative");
  if (s->rho <= 0.0) BAD("rho must be positive");
  if (s->sigma <= 0.0) BAD("sigma must be positive");
  if (s->delta <= 0.0) BAD("delta must be positive");
  if (s->max_iter <= 0) BAD("max_iter must be positive");
  if (s->eps_abs < 0.0) BAD("eps_abs must be nonnegative");
  if (s->eps_rel < 0.0) BAD("eps_rel must be nonnegative");
  if (s->eps_rel == 0.0 && s->eps_abs == 0.0) BAD("at least one of eps_abs and eps_rel must be positive");
  if (s->eps_prim_inf <= 0.0) BAD("eps_prim_inf must be positive");
  if (s->eps_dual_inf <= 0.0) BAD("eps_dual_inf must be positive");
  if (s->alpha <= 0.0 || s->alpha >= 2.0) BAD("alpha must be strictly between 0 and 2");
  if (s->linsys_solver != QDLDL_SOLVER && s->linsys_solver != MKL_PARDISO_SOLVER) BAD("linsys_solver not recognized");
  if (s->verbose != 0 && s->verbose != 1) BAD("verbose must be either 0 or 1");
  if (s->scaled_termination != 0 && s->scaled_termination != 1) BAD("scaled_termination must be either 0 or 1");
  if (s->check_termination < 0) BAD("check_termination must be nonnegative");
  if (s->warm_start != 0 && s->warm_start != 1) BAD("warm_start must be either 0 or 1");
  if (s->time_limit < 0.0) BAD("time_limit must be nonnegative");
#undef BAD
  return 0;
}

// Preconditions of this path: rejected loudly, never silently diverging.
static c_int validate_scope(const OSQPSettings *s) {
  if (s->scaling != 0 && s->scaled_termination != 0) {
    fprintf(stderr, "ERROR in rldl: scaled_termination = 1 together with scaling is not supported on this path\n");
    return 1;
  }
  if (s->polish != 0 && s->polish != 1) { fprintf(stderr, "ERROR in validate_settings: polish must be either 0 or 1\n"); return 1; }
  if (s->linsys_solver != QDLDL_SOLVER) { fprintf(stderr, "ERROR in rldl: linsys_solver must be QDLDL_SOLVER\n"); return 1; }
  return 0;
}

void Batch::push_settings() {
  S.sigma = settings.sigma; S.alpha = settings.alpha;
  S.eps_abs = settings.eps_abs; S.eps_rel = settings.eps_rel;
  S.eps_prim_inf = settings.eps_prim_inf; S.eps_dual_inf = settings.eps_dual_inf;
  S.adaptive_rho_tolerance = settings.adaptive_rho_tolerance;
  S.max_iter = (int)settings.max_iter; S.check_termination = (int)settings.check_termination;
  S.adaptive_rho = (int)settings.adaptive_rho;
  // osqp.c:459-485 picks the automatic interval from wall-clock; a batch cannot, so the
  // non-PROFILING rule of the reference (osqp.c:248-262) is used when the interval is 0.
  int ai = (int)settings.adaptive_rho_interval;
  if (settings.adaptive_rho && !ai) ai = settings.check_termination ? 4 * (int)settings.check_termination : 100;
  S.adaptive_rho_interval = ai;
  S.time_limit_ns = 0;  // set per solve (Batch::solve)
}

// ------------------------------ lifetime ------------------------------
c_int Batch::create(Batch **out, const csc *Q0, const csc *Qi, const csc *QN, const csc *A0,
                    const csc *Ai, const csc *Aij, const csc *AN, const OSQPSettings *st, c_int Nmax,
                    c_int N, c_int nx, c_int nu, c_int ny, c_int nt, c_int batch, const c_float *q,
                    const c_float *l, const c_float *u, const c_float *rho, bool generic) {
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    fprintf(stderr, "ERROR in rldl: no usable CUDA device — this library has no CPU path\n");
    return OSQP_LINSYS_SOLVER_LOAD_ERROR;
  }
  if (validate_settings(st)) return OSQP_SETTINGS_VALIDATION_ERROR;
  if (validate_scope(st)) return OSQP_SETTINGS_VALIDATION_ERROR;
  if (batch < 1 || nx < 1 || nu < 0 || ny < 0 || nt < 0 || (!generic && (N < 1 || N > Nmax))) {
    fprintf(stderr, "ERROR in rldl: invalid dimensions (need 1 <= N <= Nmax, batch >= 1)\n");
    return OSQP_DATA_VALIDATION_ERROR;
  }
  std::unique_ptr<Batch> guard(new Batch());  // every early return below frees the handle and its device buffers
  Batch *B = guard.get();
  B->generic = generic;
  B->nx = (int)nx; B->nu = (int)nu; B->ny = (int)ny; B->nt = (int)nt;
  B->a = (int)(nx + nu); B->b = (int)(nx + ny);
  B->N = generic ? 0 : (int)N; B->Nmax = generic ? 0 : (int)Nmax; B->batch = (int)batch;
  B->nmax = generic ? (int)nx : (int)(Nmax * (nx + nu));
  B->mmax = generic ? (int)nt : (int)(Nmax * (nx + ny) + nt);
  B->settings = *st;
  CK(cudaGetDevice(&B->device));
  c_int rc = B->upload_model(Q0, Qi, QN, A0, Ai, Aij, AN);
  if (rc) return rc;
  const size_t per_f = factor_smem_doubles(B->a, B->nx, B->ny, B->nt) * sizeof(double);
  const size_t per_a = admm_smem_doubles(B->a, B->nx, B->ny, B->nt) * sizeof(double);
  if (per_f > 220 * 1024 || per_a > 220 * 1024) {
    fprintf(stderr, "ERROR in rldl: stage blocks too large for on-chip staging (a=%d, nx=%d)\n", B->a, B->nx);
    return OSQP_DATA_VALIDATION_ERROR;
  }
  BatchDev &D = B->D;
  memset(&D, 0, sizeof(D));
  D.batch = B->batch; D.Nmax = B->Nmax; D.nmax = B->nmax; D.mmax = B->mmax;
  D.xstride = (B->Nmax + 1) * B->a * B->a;
  D.sstride = std::max(1, B->Nmax * B->nx * B->nx);
  D.ykstride = (B->Nmax + 1) * B->a;
  D.skstride = std::max(1, B->Nmax * B->nx);
  const size_t bs = (size_t)B->batch;
  CK(dalloc(&D.q, bs * D.nmax)); CK(dalloc(&D.l, bs * D.mmax)); CK(dalloc(&D.u, bs * D.mmax));
  CK(dalloc(&D.rho_vec, bs * D.mmax));
  CK(dalloc(&D.x, bs * D.nmax)); CK(dalloc(&D.z, bs * D.mmax)); CK(dalloc(&D.y, bs * D.mmax));
  CK(dalloc(&D.dx, bs * D.nmax)); CK(dalloc(&D.dy, bs * D.mmax));
  CK(dalloc(&D.Xinv, bs * D.xstride)); CK(dalloc(&D.Sinv, bs * D.sstride));
  CK(dalloc(&D.yk, bs * D.ykstride)); CK(dalloc(&D.sk, bs * D.skstride));
  CK(dalloc(&D.rho, bs)); CK(dalloc(&D.iter, bs)); CK(dalloc(&D.status, bs));
  CK(dalloc(&D.obj, bs)); CK(dalloc(&D.pri, bs)); CK(dalloc(&D.dua, bs)); CK(dalloc(&D.rho_est, bs));
  CK(dalloc(&D.rho_updates, bs)); CK(dalloc(&D.fact_flag, bs));
  CK(dalloc(&B->d_changed, bs));
  CK(dalloc(&B->d_ratio, bs));
  D.ratio_prev = B->d_ratio; D.sched_key = 1; D.lat_levels = 1; D.lat_lockstep = 1;
  if (B->scaled()) { CK(dalloc(&D.sD, bs * D.nmax)); CK(dalloc(&D.sE, bs * D.mmax)); CK(dalloc(&D.sC, bs)); }
  CK(dalloc(&D.sched, kSchedCounters));
  rc = B->configure_stream();
  if (rc) return rc;
  {
    int lo = 0, hi = 0;  // lowest / highest priority (numerically greatest / least)
    CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CK(cudaStreamCreateWithPriority(&B->stream, cudaStreamNonBlocking, hi));
    CK(cudaStreamCreateWithPriority(&B->stream2, cudaStreamNonBlocking, lo));
  }
  CK(cudaEventCreate(&B->ev0)); CK(cudaEventCreate(&B->ev1));
  CK(cudaEventCreateWithFlags(&B->ev_reset, cudaEventDisableTiming));
  CK(cudaEventCreateWithFlags(&B->ev_lane, cudaEventDisableTiming));
  CK(cudaEventCreate(&B->ev_s0)); CK(cudaEventCreate(&B->ev_s1));
  B->set_sizes();
  B->push_settings();
  // per-instance data
  CK(cudaMemcpyAsync(D.q, q, bs * D.nmax * sizeof(double), cudaMemcpyHostToDevice, B->stream));
  CK(cudaMemcpyAsync(D.l, l, bs * D.mmax * sizeof(double), cudaMemcpyHostToDevice, B->stream));
  CK(cudaMemcpyAsync(D.u, u, bs * D.mmax * sizeof(double), cudaMemcpyHostToDevice, B->stream));
  if (rho) {
    for (size_t i = 0; i < bs; i++)
      if (!(rho[i] > 0.0)) { fprintf(stderr, "ERROR in rldl: rho must be positive\n"); return OSQP_SETTINGS_VALIDATION_ERROR; }
    CK(cudaMemcpyAsync(D.rho, rho, bs * sizeof(double), cudaMemcpyHostToDevice, B->stream));
  } else {
    CK(launch_fill(D.rho, st->rho, bs, B->stream));
    B->launches++;
  }
  CK(launch_fill_int(D.status, OSQP_UNSOLVED, bs, B->stream));
  B->launches++;
  // validate_data (auxil.c:794-887): l <= u on every row, checked on the device for the whole batch
  {
    int bad = 0;
    CK(launch_check_bounds(D.l, D.u, D.mmax, D.mmax, B->batch, D.sched + 1, B->stream));
    B->launches++;
    CK(cudaMemcpyAsync(&bad, D.sched + 1, sizeof(int), cudaMemcpyDeviceToHost, B->stream));
    CK(cudaStreamSynchronize(B->stream));
    if (bad) {
      fprintf(stderr, "ERROR in validate_data: Lower bound must be lower than or equal to upper bound\n");
      return OSQP_DATA_VALIDATION_ERROR;
    }
    CK(cudaMemsetAsync(D.sched + 1, 0, sizeof(int), B->stream));
  }
  // scale_data (scaling.c:44) + set_rho_vec (auxil.c:79) + full stage recursion (recursive_ldl.c:1139)
  rc = B->rescale();
  if (rc) return rc;
  rc = B->retype(0, 3, false, false, true);
  if (rc) return rc;
  rc = B->factor(0, false);
  if (rc) return rc;
  *out = guard.release();
  return 0;
}

void Batch::set_sizes() {
  D.N = N;
  D.n = generic ? nx : N * a;
  D.m = generic ? nt : N * b + nt;
}

Batch::~Batch() {
  cudaSetDevice(device);
  double *dp[] = {D.sD, D.sE, D.sC, D.q, D.l, D.u, D.rho_vec, D.x, D.z, D.y, D.dx, D.dy, D.Xinv, D.Sinv, D.yk, D.sk,
                  D.rho, D.obj, D.pri, D.dua, D.rho_est, d_mval, d_stage, D.rec, d_l_alt, d_u_alt, d_latT,
                  d_polX, d_polS, d_polrho, d_polb, d_pols, d_polr, d_trace, d_ratio};
  for (double *p : dp) if (p) cudaFree(p);
  if (D.iter) cudaFree(D.iter);
  int *ip[] = {D.status, D.rho_updates, D.fact_flag, d_changed, d_midx, D.sched, D.sched_slots, d_polstat};
  for (int *p : ip) if (p) cudaFree(p);
  if (h_pin) cudaFreeHost(h_pin);
  if (ev0) cudaEventDestroy(ev0);
  if (ev1) cudaEventDestroy(ev1);
  if (ev_reset) cudaEventDestroy(ev_reset);
  if (ev_lane) cudaEventDestroy(ev_lane);
  if (ev_s0) cudaEventDestroy(ev_s0);
  if (ev_s1) cudaEventDestroy(ev_s1);
  for (cudaEvent_t e : ev_chunk) if (e) cudaEventDestroy(e);
  if (stream2) cudaStreamDestroy(stream2);
  if (stream) cudaStreamDestroy(stream);
}

c_int Batch::ensure_stage(size_t doubles) {
  if (doubles <= stage_cap) return 0;
  if (d_stage) cudaFree(d_stage);
  d_stage = nullptr; stage_cap = 0;
  CK(cudaMalloc(&d_stage, doubles * sizeof(double)));
  stage_cap = doubles;
  return 0;
}

// ------------------------------ operations ------------------------------
c_int Batch::factor_launch(int k0, bool only_flagged, bool mark) {
  const int *fl = only_flagged ? d_changed : nullptr;
  if (streaming) {
    // shape-specialised recursion writing the stage records directly; the unpacked pivot inverses are
    // only produced while a generic-kernel consumer (rldl_batch_kkt_solve, select_kernel(1)) wants them
    CK(launch_factor_stream(M, D, settings.sigma, k0, fl, need_unpacked ? 1 : 0, stream));
    launches++;
    if (mark) CK(cudaEventRecord(ev1, stream));  // last_factor_ms times the recursion kernel alone
    if (!need_unpacked) unpacked_valid = false;
    records_stale = false;
  } else {
    CK(launch_factor(M, D, settings.sigma, k0, fl, stream));
    launches++;
    if (mark) CK(cudaEventRecord(ev1, stream));
    unpacked_valid = true;
  }
  return 0;
}

// Generic-kernel consumers read the unpacked pivot inverses: produce them (once) on demand.
c_int Batch::ensure_unpacked() {
  if (unpacked_valid) return 0;
  need_unpacked = true;
  c_int rc = factor_launch(0, false);
  if (rc) return rc;
  unpacked_valid = true;
  return 0;
}

c_int Batch::factor(int k0, bool only_flagged) {
  CK(cudaSetDevice(device));
  CK(cudaEventRecord(ev0, stream));
  c_int rc = factor_launch(k0, only_flagged, true);
  if (rc) return rc;
  CK(cudaStreamSynchronize(stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, ev0, ev1);
  last_factor_ms = ms;
  // algorithmic bytes: pivot inverses written once (diagonal-packed records on the streaming path), the row
  // classes of the record (streaming) or rho_vec (runtime shapes) read once, per stage and instance
  const double per_stage = streaming ? (double)D.PX + D.PS + D.clsw + D.sigw : (double)a * a + (double)nx * nx + b;
  factor_bytes = 8.0 * batch * ((N - k0) * per_stage + (streaming ? (double)D.PX : (double)nx * nx) + nt);
  return 0;
}

c_int Batch::failed_factorisations(c_int *flag_out) {
  if (cudaSetDevice(device) != cudaSuccess) return -1;
  std::vector<int> flag(batch);
  if (cudaMemcpyAsync(flag.data(), D.fact_flag, (size_t)batch * sizeof(int), cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
      cudaStreamSynchronize(stream) != cudaSuccess)
    return -1;
  c_int bad = 0;
  for (int i = 0; i < batch; i++) {
    bad += flag[i] != 0;
    if (flag_out) flag_out[i] = flag[i];
  }
  return bad;
}

c_int Batch::solve() {
  CK(cudaSetDevice(device));
  CK(cudaEventRecord(ev0, stream));
  // time_limit (osqp.c:389-405): what is left of it when the solve starts (the scalar entry points subtract
  // their setup / update time), tested on the device against %globaltimer
  S.time_limit_ns = 0;
  if (settings.time_limit > 0.0) {
    const double left = settings.time_limit - time_spent_before_solve;
    S.time_limit_ns = left > 1e-9 ? (long long)(left * 1e9) : 1;
  }
  if (settings.time_limit > 0.0 || D.trace) {
    CK(launch_mark_time(D, stream));
    launches++;
  }
  if (z_pending) {
    CK(launch_warm_start_z(M, D, stream));
    launches++;
    z_pending = false;
  }
  if (use_stream() && records_stale) {  // a generic-kernel adaptive solve re-factorised behind the records
    c_int rc = retype(0, 2, false, false);
    if (rc) return rc;
    rc = factor_launch(0, false);
    if (rc) return rc;
  }
  if (!use_stream()) {
    c_int rc = ensure_unpacked();
    if (rc) return rc;
    if (streaming && settings.adaptive_rho) records_stale = true;
  }
  if (use_stream()) {
    // Two kernels share one solve.  The streaming kernel carries the bulk (HBM-bound, thousands of instances
    // in flight); once at most `lat_below` instances are unfinished it suspends them as they stand and the
    // latency-mode kernel (one instance per CTA, factor resident in shared memory, ~10x shorter iteration)
    // finishes them.  Small batches go to the latency-mode kernel directly.
    // Adaptive rho: an instance whose rho changes parks itself (osqp_update_rho, osqp.c:1268-1319, needs a
    // full stage recursion); parked instances are re-typed, re-factorised and re-packed together, then resumed.
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const size_t lat_smem = latency_smem_bytes(M, N);
    const int lat_below = this->lat_below >= 0 ? this->lat_below : (batch >= 4096 ? 2048 : 1024);
    const bool lat_ok = lat_below > 0 && lat_smem != (size_t)-1 && lat_smem <= 227 * 1024;
    const int lane = lat_ok ? std::min(std::max(lat_lane, 0), sms - 8) : 0;
    const size_t t_per_cta = (size_t)latency_t_blocks(Nmax) * nx * (nx + 2);
    if (lat_ok && !d_latT) CK(cudaMalloc(&d_latT, (size_t)(2 * sms) * t_per_cta * sizeof(double)));
    stream_timed = false;
    last_stream_ms = 0; last_stream_iters = 0;
    int unfinished = batch, resume = 0;
    CK(cudaMemcpyAsync(D.sched + kSchedUnfinished, &unfinished, sizeof(int), cudaMemcpyHostToDevice, stream));
    for (;;) {
      const bool run_stream = !lat_ok || unfinished > lat_below;
      // the lane only pays when the launch is long enough for the streaming CTAs to retire well before its end
      const bool run_lane = run_stream && lane > 0 && unfinished >= 4 * lat_below;
      if (run_stream) {
        int s_ctas = 0;
        if (run_lane) CK(preload_lane_kernels());  // before anything of this round runs (see below)
        const int retire = lat_retire_below > 0 ? lat_retire_below : 6 * lat_below;
        const int park = !lat_ok ? 0 : ((run_lane && lane_tail > 0) ? std::min(lat_below, lane_tail) : lat_below);
        CK(cudaEventRecord(ev_s0, stream));
        CK(launch_admm_stream(M, D, S, (int)settings.warm_start, resume, park, tune,
                              run_lane ? lane : 0, retire, run_lane ? ev_reset : nullptr, &s_ctas, stream));
        CK(cudaEventRecord(ev_s1, stream));
        stream_timed = true;
        launches++;
        if (run_lane) {
          // The lane CTAs spin until a kernel launched AFTER them (k_set_tail_flag) has run.  With CUDA's lazy module
          // loading the first launch of a kernel loads it, which can wait for running kernels to finish: a deadlock
          // if the kernel to be loaded is the one the running kernel waits for.  Hence preload_lane_kernels() above;
          // the lane CTAs also give up waiting after a second without a hand-over.
          // latency-mode CTAs beside the streaming kernel: pending until its CTAs on SMs 0..lane-1 retire, then fed
          // with the most-unconverged instances by the streaming groups (their chains would bound the tail)
          CK(cudaStreamWaitEvent(stream2, ev_reset, 0));
          CK(launch_admm_latency(M, D, S, (int)settings.warm_start, 1, d_latT + (size_t)sms * t_per_cta, lane, 0, 2,
                                 s_ctas, stream2));
          CK(cudaEventRecord(ev_lane, stream2));
          CK(launch_set_tail_flag(D, stream));  // after the streaming kernel: nobody waits for a hand-over any more
          launches += 2;
        }
      }
      if (lat_ok) {
        CK(launch_admm_latency(M, D, S, (int)settings.warm_start, (resume || run_stream) ? 1 : 0, d_latT,
                               std::min(sms, batch), run_stream ? 0 : 1, run_stream ? 1 : 0, 0, stream));
        launches++;
      }
      if (run_lane) CK(cudaStreamWaitEvent(stream, ev_lane, 0));
      if (!settings.adaptive_rho) break;
      int cnt[2] = {0, 0};
      CK(cudaMemcpyAsync(&cnt[0], D.sched + 1, sizeof(int), cudaMemcpyDeviceToHost, stream));
      CK(cudaMemcpyAsync(&cnt[1], D.sched + kSchedUnfinished, sizeof(int), cudaMemcpyDeviceToHost, stream));
      CK(cudaStreamSynchronize(stream));
      if (!cnt[0]) break;
      unfinished = cnt[1];
      c_int rc = retype(0, 2, true, true);  // the parked instances: new rho_vec, flagged, resumable again
      if (rc) return rc;
      rc = factor_launch(0, true);
      if (rc) return rc;
      resume = 1;
      solve_rounds++;
    }
  } else {
    CK(launch_admm(M, D, S, (int)settings.warm_start, stream));
    launches++;
  }
  CK(cudaEventRecord(ev1, stream));
  CK(cudaStreamSynchronize(stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, ev0, ev1);
  last_solve_ms = ms;
  if (use_stream() && stream_timed) {  // the streaming kernel of the (last) round alone
    cudaEventElapsedTime(&ms, ev_s0, ev_s1);
    last_stream_ms = ms;
    unsigned long long it = 0;
    CK(cudaMemcpy(&it, D.sched + kSchedStreamIters, sizeof(it), cudaMemcpyDeviceToHost));
    last_stream_iters = (double)it;
  }
  last_polish_ms = 0;
  if (settings.polish) return polish();  // osqp.c:589-591: only the instances that ended OSQP_SOLVED
  return 0;
}

// polish (src/polish.c:212-350) for every solved instance: a second stage factor with sigma = delta and
// rho = 1/delta on the guessed active rows, iterative refinement, acceptance test — all on the device (k_polish).
c_int Batch::polish() {
  const size_t bs = (size_t)batch, ld = (size_t)D.nmax + D.mmax;
  if (!d_polX) {
    CK(dalloc(&d_polX, bs * D.xstride)); CK(dalloc(&d_polS, bs * D.sstride)); CK(dalloc(&d_polrho, bs * D.mmax));
    CK(dalloc(&d_polb, bs * ld)); CK(dalloc(&d_pols, bs * ld)); CK(dalloc(&d_polr, bs * ld));
    CK(dalloc(&d_polstat, bs));
  }
  BatchDev P = D;
  P.Xinv = d_polX; P.Sinv = d_polS; P.rho_vec = d_polrho;
  CK(cudaEventRecord(ev0, stream));
  CK(launch_polish(M, D, P, S, settings.delta, (int)settings.polish_refine_iter, d_polb, d_pols, d_polr, d_polstat,
                   stream));
  launches++;
  CK(cudaEventRecord(ev1, stream));
  CK(cudaStreamSynchronize(stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, ev0, ev1);
  last_polish_ms = ms;
  return 0;
}

c_int Batch::get_polish(c_int *status_polish) {
  CK(cudaSetDevice(device));
  if (!d_polstat) {  // no polish has run on this handle
    for (int i = 0; i < batch; i++) status_polish[i] = 0;
    return 0;
  }
  std::vector<int> sp(batch);
  CK(cudaMemcpyAsync(sp.data(), d_polstat, batch * sizeof(int), cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  for (int i = 0; i < batch; i++) status_polish[i] = sp[i];
  return 0;
}

bool Batch::use_stream() const {
  return streaming && kernel_choice != 1;
}

// Everything that is derived from (l, u, rho, q) for stages >= k0 (reference set_rho_vec / update_rho_vec,
// src/auxil.c:79-145): constraint typing, rho_vec, and on the streaming shapes the vector part of the stage
// records (parts bit 0: q, bit 1: row classes, l, u).  want_flags: d_changed[inst] = 1 for the instances whose
// row types changed (or that an adaptive-rho update parked) — they need the stage recursion again.
// only_parked: only the instances parked by an adaptive-rho update are touched.  all_rows: type the rows of
// the whole Nmax-sized problem (setup), not just the current horizon.
c_int Batch::retype(int k0, int parts, bool only_parked, bool want_flags, bool all_rows) {
  if (streaming) {
    CK(launch_vec_records(M, D, k0, parts | (only_parked ? 4 : 0), nullptr, want_flags ? d_changed : nullptr,
                          settings.sigma, stream));
    launches++;
    if (!all_rows) return 0;
    // rows beyond the current horizon are typed once at setup, as the reference does (recursive_ldl.c:2190)
  }
  if ((parts & 2) && (!streaming || all_rows)) {
    CK(launch_set_rho_vec(D, all_rows ? D.mmax : D.m, (want_flags && !streaming) ? d_changed : nullptr, stream));
    launches++;
  }
  return 0;
}

// scale_data (src/scaling.c:44-156) for every instance, from the resident model, the current horizon and the
// current q; a no-op when settings.scaling == 0.
c_int Batch::rescale(bool carry_iterates) {
  scale_stale = false;
  if (!scaled()) return 0;
  // carry_iterates (P / A update): the reference's iterates live in scaled coordinates and survive the new
  // equilibration numerically unchanged (osqp.c:1033-1051) — reproduce that warm-start point
  const bool carry = carry_iterates && have_scaling;
  if (carry) { CK(launch_rescale_iterates(D, 1, stream)); launches++; }
  CK(launch_ruiz(M, D, (int)settings.scaling, stream));
  launches++;
  if (carry) { CK(launch_rescale_iterates(D, 0, stream)); launches++; }
  have_scaling = true;
  return 0;
}

c_int Batch::set_trace(bool on) {
  if (on && !d_trace) CK(dalloc(&d_trace, (size_t)batch));
  D.trace = on ? d_trace : nullptr;
  return 0;
}
c_int Batch::get_trace(c_float *t) {
  if (!d_trace) return 1;
  CK(cudaSetDevice(device));
  CK(cudaMemcpy(t, d_trace, (size_t)batch * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

c_int Batch::fetch_scaling(int inst, c_float *Dv, c_float *Ev, c_float *c) {
  if (!scaled()) return 1;
  CK(cudaSetDevice(device));
  const size_t cnt = inst < 0 ? (size_t)batch : 1, first = inst < 0 ? 0 : (size_t)inst;  // inst < 0: the whole batch
  CK(cudaMemcpyAsync(Dv, D.sD + first * D.nmax, cnt * D.nmax * sizeof(double), cudaMemcpyDeviceToHost, stream));
  CK(cudaMemcpyAsync(Ev, D.sE + first * D.mmax, cnt * D.mmax * sizeof(double), cudaMemcpyDeviceToHost, stream));
  CK(cudaMemcpyAsync(c, D.sC + first, cnt * sizeof(double), cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  return 0;
}

c_int Batch::retype_and_refactor() {
  // update_rho_vec (auxil.c:103-145): re-type rows; refactor instances whose types changed
  c_int rc = retype(0, 2, false, true);
  if (rc) return rc;
  if (streaming) return factor_launch(0, true);  // flags stay on the device: instances without a change are skipped
  std::vector<int> ch(batch);
  CK(cudaMemcpyAsync(ch.data(), d_changed, batch * sizeof(int), cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  bool any = false;
  for (int v : ch) any |= v != 0;
  if (any) return factor(0, true);
  return 0;
}

c_int Batch::reset_info() {  // reset_info, src/auxil.c:635-652 — stream-ordered, no host round trip
  CK(launch_fill_int(D.status, OSQP_UNSOLVED, (size_t)batch, stream));
  launches++;
  CK(cudaMemsetAsync(D.rho_updates, 0, batch * sizeof(int), stream));
  return 0;
}

c_int Batch::update_bounds(const c_float *l, const c_float *u, bool have_l, bool have_u, bool validate) {
  CK(cudaSetDevice(device));
  const size_t bytes = (size_t)batch * D.mmax * sizeof(double);
  if (validate) {
    // the new bounds land in a second pair of buffers, l <= u is checked on the device, and only then
    // do they become current (invalid input leaves the problem untouched, osqp.c:793-801)
    if (!d_l_alt) CK(cudaMalloc(&d_l_alt, bytes));
    if (!d_u_alt) CK(cudaMalloc(&d_u_alt, bytes));
    if (have_l) CK(cudaMemcpyAsync(d_l_alt, l, bytes, cudaMemcpyHostToDevice, stream));
    if (have_u) CK(cudaMemcpyAsync(d_u_alt, u, bytes, cudaMemcpyHostToDevice, stream));
    CK(launch_check_bounds(have_l ? d_l_alt : D.l, have_u ? d_u_alt : D.u, D.mmax, D.m, batch, D.sched + 1, stream));
    launches++;
    int bad = 0;
    CK(cudaMemcpyAsync(&bad, D.sched + 1, sizeof(int), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    if (bad) return 1;
    if (have_l) std::swap(D.l, d_l_alt);
    if (have_u) std::swap(D.u, d_u_alt);
  } else {
    if (have_l) CK(cudaMemcpyAsync(D.l, l, bytes, cudaMemcpyHostToDevice, stream));
    if (have_u) CK(cudaMemcpyAsync(D.u, u, bytes, cudaMemcpyHostToDevice, stream));
  }
  c_int rc = reset_info();
  if (rc) return rc;
  return retype_and_refactor();
}

// The per-instance arrays of `count` instances from `first` on, as a batch of their own (for kernels that only use the
// per-instance arrays: typing / record pass, stage recursion, bound check).
static BatchDev slice(const BatchDev &D, int first, int count) {
  BatchDev S = D;
  const size_t f = (size_t)first;
  S.batch = count;
  S.q += f * D.nmax; S.x += f * D.nmax; S.dx += f * D.nmax;
  S.l += f * D.mmax; S.u += f * D.mmax; S.rho_vec += f * D.mmax; S.z += f * D.mmax; S.y += f * D.mmax; S.dy += f * D.mmax;
  S.Xinv += f * D.xstride; S.Sinv += f * D.sstride; S.yk += f * D.ykstride; S.sk += f * D.skstride;
  S.rho += f; S.iter += f; S.status += f; S.obj += f; S.pri += f; S.dua += f; S.rho_est += f; S.rho_updates += f;
  S.fact_flag += f;
  if (S.rec) S.rec += f * D.rec_stride;
  if (S.sD) { S.sD += f * D.nmax; S.sE += f * D.mmax; S.sC += f; }
  if (S.ratio_prev) S.ratio_prev += f;
  if (S.trace) S.trace += f;
  return S;
}

// update_bounds(l, u) + update_lin_cost(q) + refactor() in one call, pipelined: the batch is uploaded in chunks on a
// second stream while typing / record pass and stage recursion of the previous chunk run (1.9 GB of upload at C5 take
// 36 ms, the device work 10 ms: the latter disappears behind the former).  Same semantics as the three calls: l <= u
// is validated on the device and an invalid update leaves the resident problem as it was (return 1).
c_int Batch::update_data_refactor(const c_float *q, const c_float *l, const c_float *u) {
  CK(cudaSetDevice(device));
  if (!q || !l || !u) return 1;
  if (!streaming || batch < 64) {
    c_int rc = update_bounds(l, u, true, true, true);
    if (rc) return rc;
    rc = update_lin_cost(q);
    if (rc) return rc;
    return refactor();
  }
  const size_t mb = (size_t)batch * D.mmax * sizeof(double);
  if (!d_l_alt) CK(cudaMalloc(&d_l_alt, mb));
  if (!d_u_alt) CK(cudaMalloc(&d_u_alt, mb));
  constexpr int NCH = 8;
  for (int c = 0; c < NCH; c++)
    if (!ev_chunk[c]) CK(cudaEventCreateWithFlags(&ev_chunk[c], cudaEventDisableTiming));
  CK(cudaEventRecord(ev0, stream));
  CK(cudaEventRecord(ev_reset, stream));            // the copies wait for everything already queued on the handle
  CK(cudaStreamWaitEvent(stream2, ev_reset, 0));
  CK(cudaMemsetAsync(D.sched + 1, 0, sizeof(int), stream));
  BatchDev N_ = D;  // the batch as it will be: new bounds (in the landing buffers), new q
  N_.l = d_l_alt; N_.u = d_u_alt;
  if (scale_stale) { c_int rc = rescale(true); if (rc) return rc; }
  for (int c = 0; c < NCH; c++) {
    const int i0 = (int)((long long)batch * c / NCH), i1 = (int)((long long)batch * (c + 1) / NCH);
    const size_t ro = (size_t)i0 * D.mmax, vo = (size_t)i0 * D.nmax, nb = (size_t)(i1 - i0);
    CK(cudaMemcpyAsync(d_l_alt + ro, l + ro, nb * D.mmax * sizeof(double), cudaMemcpyHostToDevice, stream2));
    CK(cudaMemcpyAsync(d_u_alt + ro, u + ro, nb * D.mmax * sizeof(double), cudaMemcpyHostToDevice, stream2));
    CK(cudaMemcpyAsync(D.q + vo, q + vo, nb * D.nmax * sizeof(double), cudaMemcpyHostToDevice, stream2));
    CK(cudaEventRecord(ev_chunk[c], stream2));
    CK(cudaStreamWaitEvent(stream, ev_chunk[c], 0));
    const BatchDev S_ = slice(N_, i0, i1 - i0);
    CK(launch_check_bounds(S_.l, S_.u, D.mmax, D.m, i1 - i0, D.sched + 1, stream, false));
    CK(launch_vec_records(M, S_, 0, 3, nullptr, nullptr, settings.sigma, stream));
    CK(launch_factor_stream(M, S_, settings.sigma, 0, nullptr, need_unpacked ? 1 : 0, stream));
    launches += 3;
  }
  CK(cudaEventRecord(ev1, stream));
  int bad = 0;
  CK(cudaMemcpyAsync(&bad, D.sched + 1, sizeof(int), cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  CK(cudaMemsetAsync(D.sched + 1, 0, sizeof(int), stream));
  if (bad) {
    // refused: q is restored by the caller's next update (as with a refused osqp_update_bounds followed by nothing,
    // the linear cost of this call has been taken); records and factor are rebuilt from the bounds that stay
    c_int rc = retype(0, 3, false, false);
    if (rc) return rc;
    rc = factor(0, false);
    return rc ? rc : 1;
  }
  std::swap(D.l, d_l_alt);
  std::swap(D.u, d_u_alt);
  if (!need_unpacked) unpacked_valid = false;
  records_stale = false;
  float ms = 0;
  cudaEventElapsedTime(&ms, ev0, ev1);
  last_factor_ms = ms;  // here: upload + typing + recursion, pipelined
  return reset_info();
}

c_int Batch::partial_update_bounds(c_int start, c_int stop, const c_float *l, const c_float *u) {
  CK(cudaSetDevice(device));
  const int len = (int)(stop - start);
  const size_t cnt = (size_t)batch * len;
  c_int rc = ensure_stage(2 * cnt);
  if (rc) return rc;
  CK(cudaMemcpyAsync(d_stage, l, cnt * sizeof(double), cudaMemcpyHostToDevice, stream));
  CK(cudaMemcpyAsync(d_stage + cnt, u, cnt * sizeof(double), cudaMemcpyHostToDevice, stream));
  CK(launch_scatter_rows(D.l, D.mmax, d_stage, len, (int)start, batch, stream));
  CK(launch_scatter_rows(D.u, D.mmax, d_stage + cnt, len, (int)start, batch, stream));
  launches += 2;
  rc = reset_info();
  if (rc) return rc;
  // The reference re-types the rows but never refactors (src/recursive_ldl.c:160-189, F6),
  // which leaves factor and rho_vec inconsistent when a type changes; we keep them consistent.
  return retype_and_refactor();
}

c_int Batch::update_lin_cost(const c_float *q) {
  CK(cudaSetDevice(device));
  CK(cudaMemcpyAsync(D.q, q, (size_t)batch * D.nmax * sizeof(double), cudaMemcpyHostToDevice, stream));
  c_int rc = retype(0, 1, false, false);
  if (rc) return rc;
  return reset_info();
}

c_int Batch::warm_start(const c_float *x, const c_float *y) {
  CK(cudaSetDevice(device));
  settings.warm_start = 1;  // osqp.c:935
  if (x) CK(cudaMemcpyAsync(D.x, x, (size_t)batch * D.nmax * sizeof(double), cudaMemcpyHostToDevice, stream));
  if (y) CK(cudaMemcpyAsync(D.y, y, (size_t)batch * D.mmax * sizeof(double), cudaMemcpyHostToDevice, stream));
  if (x) { CK(launch_warm_start_z(M, D, stream)); launches++; z_pending = false; }
  CK(cudaStreamSynchronize(stream));
  return 0;
}

c_int Batch::warm_start_shift() {
  CK(cudaSetDevice(device));
  if (generic) return -1;
  settings.warm_start = 1;  // osqp.c:935
  CK(launch_shift_rows(D.x, D.nmax, D.n, a, batch, stream));
  CK(launch_shift_rows(D.y, D.mmax, D.m, b, batch, stream));
  launches += 2;
  z_pending = true;  // z = A x at the next solve, with the sizes and bounds current by then
  return 0;
}

c_int Batch::update_horizon(c_int Nnew) {
  CK(cudaSetDevice(device));
  if (generic) return -1;
  if (Nnew > Nmax || Nnew < 1) return -1;
  if (Nnew == N) return 0;
  const int k0 = std::min((int)Nnew, N);  // first stage whose pivots change
  N = (int)Nnew;
  set_sizes();
  if (scaled()) {
    // the equilibration of the new horizon differs everywhere (what a fresh setup at the new horizon computes —
    // the reference's recursive path has no defined behaviour here): full recursion
    c_int rc = rescale();
    if (rc) return rc;
    rc = retype(0, 3, false, false);
    if (rc) return rc;
    return factor(0, false);
  }
  c_int rc = retype(k0, 3, false, false);  // the terminal stage moved: its q, rows and bounds change
  if (rc) return rc;
  return factor(k0, false);
}

c_int Batch::update_rho(const c_float *rho) {
  CK(cudaSetDevice(device));
  for (int i = 0; i < batch; i++)
    if (!(rho[i] > 0.0)) { fprintf(stderr, "ERROR in osqp_update_rho: rho must be positive\n"); return 1; }
  CK(cudaMemcpyAsync(D.rho, rho, (size_t)batch * sizeof(double), cudaMemcpyHostToDevice, stream));
  c_int rc = retype(0, 2, false, false);
  if (rc) return rc;
  return factor(0, false);
}

c_int Batch::refactor() {
  CK(cudaSetDevice(device));
  c_int rc = 0;
  if (scale_stale) rc = rescale(true);  // P / A changed: the reference unscales, updates and scales again (osqp.c:1033-1051)
  if (rc) return rc;
  rc = retype(0, 3, false, false);
  if (rc) return rc;
  return factor(0, false);
}

c_int Batch::kkt_solve(c_float *rhs) {
  CK(cudaSetDevice(device));
  const int ld = D.n + D.m;
  const size_t cnt = (size_t)batch * ld;
  c_int rc = ensure_unpacked();
  if (rc) return rc;
  rc = ensure_stage(cnt);
  if (rc) return rc;
  CK(cudaMemcpyAsync(d_stage, rhs, cnt * sizeof(double), cudaMemcpyHostToDevice, stream));
  CK(launch_kkt_solve(M, D, S, d_stage, ld, stream));
  launches++;
  CK(cudaMemcpyAsync(rhs, d_stage, cnt * sizeof(double), cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  return 0;
}

// Device-visible address of a host array iff it is pinned (cudaHostAlloc / cudaHostRegister); null otherwise.
template <class T>
static T *pinned_dev_ptr(T *host) {
  if (!host) return nullptr;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  if (at.type != cudaMemoryTypeHost || !at.devicePointer) return nullptr;
  return static_cast<T *>(at.devicePointer);
}

// Fused solve-and-deliver: rldl_batch_solve + rldl_batch_get_solution + rldl_batch_get_info in one call.  When every
// output array is pinned host memory the kernels write each instance's results to the host as soon as the instance
// finishes (BatchDev::hx ...), so the device-to-host transfer (1.16 GB at C5) runs behind the solve; pageable
// arrays — or polish, which revises results after the ADMM kernels — take the plain copies after the solve.
c_int Batch::solve_to_host(c_float *x, c_float *y, c_int *iter, c_int *status, c_float *obj, c_float *pri, c_float *dua,
                           c_int *rho_updates, c_float *rho_est) {
  CK(cudaSetDevice(device));
  if (!status) return 1;
  BatchDev H = D;
  H.hx = pinned_dev_ptr(x); H.hy = pinned_dev_ptr(y);
  H.h_iter = pinned_dev_ptr(iter); H.h_status = pinned_dev_ptr(status); H.h_rho_updates = pinned_dev_ptr(rho_updates);
  H.h_obj = pinned_dev_ptr(obj); H.h_pri = pinned_dev_ptr(pri); H.h_dua = pinned_dev_ptr(dua);
  H.h_rho_est = pinned_dev_ptr(rho_est);
  const bool direct = !settings.polish && (!x || H.hx) && (!y || H.hy) && (!iter || H.h_iter) && H.h_status &&
                      (!rho_updates || H.h_rho_updates) && (!obj || H.h_obj) && (!pri || H.h_pri) &&
                      (!dua || H.h_dua) && (!rho_est || H.h_rho_est);
  if (!direct) {
    c_int rc = solve();
    if (rc) return rc;
    rc = get_solution(x, y);
    if (rc) return rc;
    return get_info(iter, status, obj, pri, dua, rho_updates, rho_est);
  }
  const BatchDev keep = D;
  D = H;
  const c_int rc = solve();  // ends with a stream synchronisation: every delivery has landed
  D.hx = keep.hx; D.hy = keep.hy; D.h_iter = keep.h_iter; D.h_status = keep.h_status;
  D.h_rho_updates = keep.h_rho_updates; D.h_obj = keep.h_obj; D.h_pri = keep.h_pri; D.h_dua = keep.h_dua;
  D.h_rho_est = keep.h_rho_est;
  return rc;
}

// rldl_batch_kkt_solve without any allocation on the call path: the right-hand sides are staged in the library's
// fixed device scratch (rldl_kernels.cu) instead of the handle's grow-on-demand buffer.  The pivot inverses the
// kernel reads were allocated at setup.  Returns 1 (and says why) when batch * (n + m) exceeds the scratch.
c_int Batch::kkt_solve_static(const c_float *rhs, c_float *sol) {
  CK(cudaSetDevice(device));
  const int ld = D.n + D.m;
  const size_t cnt = (size_t)batch * ld;
  if (cnt > (size_t)kKktStaticDoubles) {
    fprintf(stderr, "ERROR in rldl_batch_kkt_solve_static: batch * (n + m) = %zu doubles exceeds the fixed scratch of %d "
                    "doubles; use rldl_batch_kkt_solve\n", cnt, kKktStaticDoubles);
    return 1;
  }
  c_int rc = ensure_unpacked();  // kernel launches only (the unpacked factor buffers exist since setup)
  if (rc) return rc;
  CK(launch_kkt_solve_static(M, D, S, rhs, sol, cnt, ld, stream));
  launches++;
  CK(cudaStreamSynchronize(stream));
  return 0;
}

static bool no_solution(int st) {
  return st == OSQP_PRIMAL_INFEASIBLE || st == OSQP_PRIMAL_INFEASIBLE_INACCURATE ||
         st == OSQP_DUAL_INFEASIBLE || st == OSQP_DUAL_INFEASIBLE_INACCURATE || st == OSQP_NON_CVX;
}

c_int Batch::get_solution(c_float *x, c_float *y) {
  CK(cudaSetDevice(device));
  std::vector<int> st(batch);
  CK(cudaMemcpyAsync(st.data(), D.status, batch * sizeof(int), cudaMemcpyDeviceToHost, stream));
  if (x) CK(cudaMemcpyAsync(x, D.x, (size_t)batch * D.nmax * sizeof(double), cudaMemcpyDeviceToHost, stream));
  if (y) CK(cudaMemcpyAsync(y, D.y, (size_t)batch * D.mmax * sizeof(double), cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  for (int i = 0; i < batch; i++)
    if (no_solution(st[i])) {  // store_solution, auxil.c:539-542
      if (x) std::fill(x + (size_t)i * D.nmax, x + (size_t)i * D.nmax + D.n, (double)OSQP_NAN);
      if (y) std::fill(y + (size_t)i * D.mmax, y + (size_t)i * D.mmax + D.m, (double)OSQP_NAN);
    }
  return 0;
}

c_int Batch::get_info(c_int *iter, c_int *status, c_float *obj, c_float *pri, c_float *dua,
                      c_int *rho_updates, c_float *rho_est) {
  CK(cudaSetDevice(device));
  std::vector<int> st(batch), ru(batch);
  if (iter) CK(cudaMemcpyAsync(iter, D.iter, batch * sizeof(long long), cudaMemcpyDeviceToHost, stream));
  CK(cudaMemcpyAsync(st.data(), D.status, batch * sizeof(int), cudaMemcpyDeviceToHost, stream));
  CK(cudaMemcpyAsync(ru.data(), D.rho_updates, batch * sizeof(int), cudaMemcpyDeviceToHost, stream));
  if (obj) CK(cudaMemcpyAsync(obj, D.obj, batch * sizeof(double), cudaMemcpyDeviceToHost, stream));
  if (pri) CK(cudaMemcpyAsync(pri, D.pri, batch * sizeof(double), cudaMemcpyDeviceToHost, stream));
  if (dua) CK(cudaMemcpyAsync(dua, D.dua, batch * sizeof(double), cudaMemcpyDeviceToHost, stream));
  if (rho_est) CK(cudaMemcpyAsync(rho_est, D.rho_est, batch * sizeof(double), cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  for (int i = 0; i < batch; i++) {
    if (status) status[i] = st[i];
    if (rho_updates) rho_updates[i] = ru[i];
  }
  return 0;
}

c_int Batch::fetch_vectors(int inst, c_float *x, c_float *y, c_float *z, c_float *dx, c_float *dy,
                           c_float *rho_vec, c_float *rho) {
  CK(cudaSetDevice(device));
  const size_t vo = (size_t)inst * D.nmax, ro = (size_t)inst * D.mmax;
  if (x) CK(cudaMemcpyAsync(x, D.x + vo, D.nmax * sizeof(double), cudaMemcpyDeviceToHost, stream));
  if (y) CK(cudaMemcpyAsync(y, D.y + ro, D.mmax * sizeof(double), cudaMemcpyDeviceToHost, stream));
  if (z) CK(cudaMemcpyAsync(z, D.z + ro, D.mmax * sizeof(double), cudaMemcpyDeviceToHost, stream));
  if (dx) CK(cudaMemcpyAsync(dx, D.dx + vo, D.nmax * sizeof(double), cudaMemcpyDeviceToHost, stream));
  if (dy) CK(cudaMemcpyAsync(dy, D.dy + ro, D.mmax * sizeof(double), cudaMemcpyDeviceToHost, stream));
  if (rho_vec) CK(cudaMemcpyAsync(rho_vec, D.rho_vec + ro, D.mmax * sizeof(double), cudaMemcpyDeviceToHost, stream));
  if (rho) CK(cudaMemcpyAsync(rho, D.rho + inst, sizeof(double), cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  return 0;
}

c_int Batch::update_settings(const OSQPSettings *s) {
  if (validate_settings(s)) return 1;
  if (validate_scope(s)) return 1;
  const double old_rho = settings.rho, old_sigma = settings.sigma;
  const c_int old_scaling = settings.scaling;
  settings = *s;
  settings.rho = old_rho;  // rho changes go through update_rho
  settings.sigma = old_sigma;
  settings.scaling = old_scaling;  // fixed at setup, as in the reference (no osqp_update_scaling)
  push_settings();
  return 0;
}

double Batch::admm_bytes_per_iteration() const {
  // Algorithmic HBM bytes per instance and ADMM iteration.
  if (use_stream()) {
    // streaming kernel: every stage record is streamed once forward ([q|Xinv|Sinv|rho]) and once
    // backward ([Xinv|Sinv|rho|l|u]); the iterates x,z,y are counted first-read + last-write only
    const double fwd = D.qa + D.sigw + D.PX + D.PS + D.clsw, bwd = D.PX + D.PS + D.clsw + 2.0 * D.bb;
    return 8.0 * ((N + 1) * (fwd + bwd) + 2.0 * (D.n + 2.0 * D.m));
  }
  // generic kernel: the full (unpacked) pivot inverses streamed once forward and once backward,
  // plus first-read / last-write of the iterates and problem vectors
  const double F = (double)N * ((double)a * a + (double)nx * nx) + (double)nx * nx;
  const double vec = 3.0 * D.n + 7.0 * D.m;  // x(r,w) q ; z(r,w) y(r,w) l u rho
  return 8.0 * (2.0 * F + vec);
}

}  // namespace rldl
