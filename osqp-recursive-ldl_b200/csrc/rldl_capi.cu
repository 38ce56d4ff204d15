// rldl_capi.cu — the C ABI (include/osqp_rldl_abi.h): the reference's scalar entry
// points as a batch of one, and the additive rldl_batch_* family.
//
// Host-side bookkeeping only (workspace mirrors, global CSC P/A for API fidelity,
// validation, status strings); all numerical work is delegated to rldl::Batch,
// i.e. to CUDA kernels.  Behaviour cited against the reference tree:
//   osqp_setup_recursive ........ src/recursive_ldl.c:2018-2316
//   osqp_update_recursive ....... src/recursive_ldl.c:1973-2016
//   osqp_partial_update_bounds .. src/recursive_ldl.c:119-201
//   setup_AP_matrices ........... src/recursive_ldl.c:1873-1971
//   osqp_solve .................. src/osqp.c:288-641
//   osqp_update_* ............... src/osqp.c:752-1604
//   error codes ................. include/constants.h:43-51, src/error.c
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "rldl_batch.h"

using rldl::Batch;

namespace {

const char *kErrMsg[] = {"Problem data validation.", "Solver settings validation.",
                         "Linear system solver not available.\nTried to obtain it from shared library.",
                         "Linear system solver initialization.", "KKT matrix factorization.\nThe problem seems to be non-convex.",
                         "Memory allocation.", "Solver workspace not initialized."};

c_int osqp_error(int code) {
  fprintf(stderr, "ERROR: %s\n", kErrMsg[code - 1]);
  return code;
}

double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// osqp_setup fails with OSQP_NONCVX_ERROR when the KKT factorisation does not have n positive pivots
// (src/osqp.c:214-216 via init_linsys_solver, qdldl_interface.c LDL_factor: "The problem seems to be
// non-convex").  Here: a non-positive pivot in the stage recursion.  Frees the handle on failure.
bool noncvx_at_setup(Batch *B) {
  if (B->failed_factorisations() == 0) return false;
  fprintf(stderr, "ERROR in LDL_factor: Error in KKT matrix LDL factorization when computing the nonzero elements. "
                  "The problem seems to be non-convex\n");
  delete B;
  return true;
}

struct PrivSolver {
  struct linsys_solver vt;  // must stay first: work->linsys_solver is cast to LinSysSolver*
  Batch *batch;
  OSQPDataRLDL *rldl;       // borrowed, NULL for osqp_setup workspaces
  OSQPWorkspace *work;
  int kind;                 // 0 osqp_setup, 1 recursive, 2 combine
  double t_mark;
};

PrivSolver *priv(OSQPWorkspace *w) { return reinterpret_cast<PrivSolver *>(w->linsys_solver); }
void mirror_scaling(OSQPWorkspace *w);

void set_status(OSQPInfo *info, c_int v) {
  // update_status, src/auxil.c:654-682
  info->status_val = v;
  const char *s = "";
  switch (v) {
    case OSQP_SOLVED: s = "solved"; break;
    case OSQP_SOLVED_INACCURATE: s = "solved inaccurate"; break;
    case OSQP_PRIMAL_INFEASIBLE: s = "primal infeasible"; break;
    case OSQP_PRIMAL_INFEASIBLE_INACCURATE: s = "primal infeasible inaccurate"; break;
    case OSQP_UNSOLVED: s = "unsolved"; break;
    case OSQP_DUAL_INFEASIBLE: s = "dual infeasible"; break;
    case OSQP_DUAL_INFEASIBLE_INACCURATE: s = "dual infeasible inaccurate"; break;
    case OSQP_MAX_ITER_REACHED: s = "maximum iterations reached"; break;
    case OSQP_TIME_LIMIT_REACHED: s = "run time limit reached"; break;
    case OSQP_SIGINT: s = "interrupted"; break;
    case OSQP_NON_CVX: s = "problem non convex"; break;
  }
  strncpy(info->status, s, sizeof(info->status) - 1);
  info->status[sizeof(info->status) - 1] = 0;
}

void reset_info(OSQPInfo *info) {  // src/auxil.c:635-652
  info->solve_time = 0.0;
  info->polish_time = 0.0;
  set_status(info, OSQP_UNSOLVED);
  info->rho_updates = 0;
}

csc *csc_new(c_int m, c_int n, c_int nzmax) {
  csc *M = (csc *)calloc(1, sizeof(csc));
  M->m = m; M->n = n; M->nzmax = nzmax > 0 ? nzmax : 1; M->nz = -1;
  M->p = (c_int *)calloc(n + 1, sizeof(c_int));
  M->i = (c_int *)calloc(M->nzmax, sizeof(c_int));
  M->x = (c_float *)calloc(M->nzmax, sizeof(c_float));
  return M;
}
void csc_del(csc *M) {
  if (!M) return;
  free(M->p); free(M->i); free(M->x); free(M);
}
csc *csc_dup(const csc *A) {
  csc *M = csc_new(A->m, A->n, A->p[A->n]);
  memcpy(M->p, A->p, sizeof(c_int) * (A->n + 1));
  memcpy(M->i, A->i, sizeof(c_int) * A->p[A->n]);
  memcpy(M->x, A->x, sizeof(c_float) * A->p[A->n]);
  return M;
}

// Global P (upper triangle) and A for horizon N from the stage blocks, allocated for
// Nmax stages — the layout of the reference's setup_AP_matrices / update_AP_matrices.
void assemble_PA(OSQPData *d, const OSQPDataRLDL *r, c_int Nmax, c_int N) {
  const csc *Q0 = r->Q0, *Qi = r->Qi, *QN = r->QN, *A0 = r->A0, *Ai = r->Ai, *Aij = r->Aij, *AN = r->AN;
  const c_int a = r->nx + r->nu, b = r->nx + r->ny;
  const c_int pmax = Nmax * Qi->p[Qi->n] + Q0->p[Q0->n] + QN->p[QN->n];
  const c_int amax = Nmax * (Ai->p[Ai->n] + Aij->p[Aij->n]) + A0->p[A0->n] + AN->p[AN->n];
  csc_del(d->P); csc_del(d->A);
  csc *P = csc_new(N * a, N * a, pmax), *A = csc_new(N * b + r->nt, N * a, amax);
  // column pointer arrays must be able to hold the Nmax-sized problem
  free(P->p); free(A->p);
  P->p = (c_int *)calloc(Nmax * a + 1, sizeof(c_int));
  A->p = (c_int *)calloc(Nmax * a + 1, sizeof(c_int));
  c_int np = 0, na = 0, col = 0;
  auto put = [](csc *M, c_int &nz, const csc *blk, c_int j, c_int roff) {
    for (c_int p = blk->p[j]; p < blk->p[j + 1]; p++) { M->i[nz] = roff + blk->i[p]; M->x[nz++] = blk->x[p]; }
  };
  for (c_int j = 0; j < Q0->n; j++, col++) {  // stage 0: u0
    P->p[col] = np; put(P, np, Q0, j, 0);
    A->p[col] = na; put(A, na, A0, j, 0);
  }
  for (c_int k = 1; k < N; k++) {  // stages 1..N-1: (x_k,u_k)
    const c_int prow = r->nu + (k - 1) * a, arow = (k - 1) * b;
    for (c_int j = 0; j < a; j++, col++) {
      P->p[col] = np; put(P, np, Qi, j, prow);
      A->p[col] = na; put(A, na, Aij, j, arow); put(A, na, Ai, j, arow + b);
    }
  }
  {  // terminal: x_N
    const c_int prow = r->nu + (N - 1) * a, arow = (N - 1) * b;
    for (c_int j = 0; j < QN->n; j++, col++) {
      P->p[col] = np; put(P, np, QN, j, prow);
      A->p[col] = na; put(A, na, Aij, j, arow); put(A, na, AN, j, arow + b);
    }
  }
  P->p[col] = np; A->p[col] = na;
  d->P = P; d->A = A;
}

template <class T>
T *zalloc(size_t n) { return (T *)calloc(n ? n : 1, sizeof(T)); }

void constr_types(OSQPWorkspace *w, c_int mtot) {
  // host mirror of set_rho_vec (src/auxil.c:79-101); the device copy is authoritative
  const double rho = w->settings->rho;
  for (c_int i = 0; i < mtot; i++) {
    const double l = w->data->l[i], u = w->data->u[i];
    if (l < -OSQP_INFTY * MIN_SCALING && u > OSQP_INFTY * MIN_SCALING) { w->constr_type[i] = -1; w->rho_vec[i] = RHO_MIN; }
    else if (u - l < RHO_TOL) { w->constr_type[i] = 1; w->rho_vec[i] = RHO_EQ_OVER_RHO_INEQ * rho; }
    else { w->constr_type[i] = 0; w->rho_vec[i] = rho; }
    w->rho_inv_vec[i] = 1.0 / w->rho_vec[i];
  }
}

// vtable: reference include/types.h:298-319
c_int vt_solve(LinSysSolver *self, c_float *b) { return reinterpret_cast<PrivSolver *>(self)->batch->kkt_solve(b); }
void vt_free(LinSysSolver *self) {
  PrivSolver *p = reinterpret_cast<PrivSolver *>(self);
  delete p->batch;
  free(p);
}
c_int vt_update_rho_vec(LinSysSolver *self, const c_float *) {
  // the authoritative rho_vec lives on the device and is derived from (l,u,rho); re-run the recursion
  return reinterpret_cast<PrivSolver *>(self)->batch->refactor();
}
c_int vt_update_matrices(LinSysSolver *self, const csc *, const csc *) {
  return reinterpret_cast<PrivSolver *>(self)->batch->refactor();
}

c_int alloc_workspace(OSQPWorkspace **workp, const OSQPSettings *settings, c_int nmax, c_int mmax) {
  OSQPWorkspace *w = zalloc<OSQPWorkspace>(1);
  if (!w) return OSQP_MEM_ALLOC_ERROR;
  *workp = w;
  w->data = zalloc<OSQPData>(1);
  w->settings = zalloc<OSQPSettings>(1);
  *w->settings = *settings;
  w->rho_vec = zalloc<c_float>(mmax); w->rho_inv_vec = zalloc<c_float>(mmax);
  w->constr_type = zalloc<c_int>(mmax);
  w->x = zalloc<c_float>(nmax); w->y = zalloc<c_float>(mmax); w->z = zalloc<c_float>(mmax);
  w->xz_tilde = zalloc<c_float>(nmax + mmax);
  w->x_prev = zalloc<c_float>(nmax); w->z_prev = zalloc<c_float>(mmax);
  w->Ax = zalloc<c_float>(mmax); w->Px = zalloc<c_float>(nmax); w->Aty = zalloc<c_float>(nmax);
  w->delta_y = zalloc<c_float>(mmax); w->Atdelta_y = zalloc<c_float>(nmax);
  w->delta_x = zalloc<c_float>(nmax); w->Pdelta_x = zalloc<c_float>(nmax); w->Adelta_x = zalloc<c_float>(mmax);
  w->solution = zalloc<OSQPSolution>(1);
  w->solution->x = zalloc<c_float>(nmax); w->solution->y = zalloc<c_float>(mmax);
  w->info = zalloc<OSQPInfo>(1);
  w->timer = (OSQPTimer *)zalloc<char>(64);
  w->pol = nullptr; w->scaling = nullptr;
  return 0;
}

void finish_setup(OSQPWorkspace *w, PrivSolver *p, double t0) {
  w->linsys_solver = reinterpret_cast<LinSysSolver *>(p);
  p->vt.type = QDLDL_SOLVER; p->vt.solve = vt_solve; p->vt.free = vt_free;
  p->vt.update_matrices = vt_update_matrices; p->vt.update_rho_vec = vt_update_rho_vec;
  p->vt.nthreads = 1;
  p->work = w;
  w->info->status_polish = 0;
  set_status(w->info, OSQP_UNSOLVED);
  w->info->setup_time = now_s() - t0;
  w->first_run = 1; w->clear_update_time = 0; w->rho_update_from_solve = 0; w->summary_printed = 0;
  w->info->rho_updates = 0;
  w->info->rho_estimate = w->settings->rho;
  // non-PROFILING rule of the reference for the automatic interval (see Batch::push_settings)
  if (w->settings->adaptive_rho && !w->settings->adaptive_rho_interval)
    w->settings->adaptive_rho_interval = w->settings->check_termination ? 4 * w->settings->check_termination : 100;
  p->batch->settings.adaptive_rho_interval = w->settings->adaptive_rho_interval;
  p->batch->push_settings();
  mirror_scaling(w);
}

void tic_update(OSQPWorkspace *w) {
  if (w->clear_update_time == 1) { w->clear_update_time = 0; w->info->update_time = 0.0; }
  priv(w)->t_mark = now_s();
}
void toc_update(OSQPWorkspace *w) { w->info->update_time += now_s() - priv(w)->t_mark; }

c_int setup_mpc(OSQPWorkspace **workp, OSQPDataRLDL *r, const OSQPSettings *settings, c_int Nmax,
                c_int N, c_int nx, c_int nu, c_int ny, c_int nt, int kind) {
  const double t0 = now_s();
  if (!workp || !r || !settings) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  if (!r->Q0 || !r->Qi || !r->QN || !r->A0 || !r->Ai || !r->Aij || !r->AN || !r->data.q || !r->data.l || !r->data.u)
    return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  if (N < 1 || N > Nmax) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  const c_int nmax = Nmax * (nx + nu), mmax = Nmax * (nx + ny) + nt;
  if (r->data.n != nmax || r->data.m != mmax) {
    fprintf(stderr, "ERROR in osqp_setup_recursive: data.n/data.m must be sized for Nmax (%lld, %lld)\n", nmax, mmax);
    return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  }
  for (c_int i = 0; i < mmax; i++)
    if (r->data.l[i] > r->data.u[i]) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  Batch *B = nullptr;
  c_int rc = Batch::create(&B, r->Q0, r->Qi, r->QN, r->A0, r->Ai, r->Aij, r->AN, settings, Nmax, N, nx,
                           nu, ny, nt, 1, r->data.q, r->data.l, r->data.u, nullptr, false);
  if (rc) return osqp_error((int)rc);
  if (noncvx_at_setup(B)) return osqp_error(OSQP_NONCVX_ERROR);
  OSQPWorkspace *w = nullptr;
  if (alloc_workspace(&w, settings, nmax, mmax)) { delete B; return osqp_error(OSQP_MEM_ALLOC_ERROR); }
  *workp = w;
  r->Nmax = Nmax; r->N = N; r->nx = nx; r->nu = nu; r->ny = ny; r->nt = nt;
  w->data->q = zalloc<c_float>(nmax); w->data->l = zalloc<c_float>(mmax); w->data->u = zalloc<c_float>(mmax);
  memcpy(w->data->q, r->data.q, sizeof(c_float) * nmax);
  memcpy(w->data->l, r->data.l, sizeof(c_float) * mmax);
  memcpy(w->data->u, r->data.u, sizeof(c_float) * mmax);
  assemble_PA(w->data, r, Nmax, N);
  w->data->n = N * (nx + nu); w->data->m = N * (nx + ny) + nt;
  w->settings->rho = fmin(fmax(settings->rho, RHO_MIN), RHO_MAX);
  constr_types(w, mmax);
  // the reference's CPU bookkeeping members (recursive_ldl.c:2194-2222): kept NULL / minimal
  r->Ytemp = nullptr; r->Uhat = nullptr; r->Vhat = nullptr; r->Yhat = nullptr; r->Lz = nullptr; r->Pz = nullptr;
  r->Dzinv = nullptr; r->Ly = nullptr; r->Dyinv = nullptr; r->P0 = nullptr;
  r->Qmats = (csc **)malloc((Nmax + 1) * sizeof(csc *));
  r->Amats = (csc **)malloc((Nmax + 1) * sizeof(csc *));
  r->Qmats[0] = r->Q0; r->Amats[0] = r->A0;
  for (c_int i = 1; i < N; i++) { r->Qmats[i] = r->Qi; r->Amats[i] = r->Ai; }
  r->Qmats[N] = r->QN; r->Amats[N] = r->AN;
  r->X_even = (csc **)calloc(Nmax, sizeof(csc *));
  r->nnz[0] = (nx * (nx + 1)) / 2; r->nnz[1] = (nx + nu) * (nx + nu) + nx * nx;
  PrivSolver *p = zalloc<PrivSolver>(1);
  p->batch = B; p->rldl = r; p->kind = kind;
  finish_setup(w, p, t0);
  return 0;
}

c_int horizon_update(OSQPWorkspace *work, OSQPDataRLDL *r, c_int N) {
  PrivSolver *p = priv(work);
  c_int rc = p->batch->update_horizon(N);
  if (rc) return rc;
  for (c_int i = r->N; i < N; i++) { r->Amats[i] = r->Ai; r->Qmats[i] = r->Qi; }
  r->Amats[N] = r->AN; r->Qmats[N] = r->QN;
  assemble_PA(work->data, r, r->Nmax, N);
  work->data->n = N * (r->nx + r->nu);
  work->data->m = N * (r->nx + r->ny) + r->nt;
  r->N = N;
  mirror_scaling(work);
  return 0;
}

void sync_settings(OSQPWorkspace *w) {
  Batch *B = priv(w)->batch;
  const double rho = B->settings.rho, sigma = B->settings.sigma;
  const c_int scaling = B->settings.scaling;
  B->settings = *w->settings;
  B->settings.rho = rho; B->settings.sigma = sigma; B->settings.scaling = scaling;
  B->push_settings();
}

// Host mirror of the equilibration (work->scaling, reference include/types.h:44-52): D, E, c of the instance as
// the device computed them.  The data mirrors (work->data) stay unscaled — the kernels never form scaled data.
void mirror_scaling(OSQPWorkspace *w) {
  Batch *B = priv(w)->batch;
  if (!B->scaled()) return;
  if (!w->scaling) {
    w->scaling = zalloc<OSQPScaling>(1);
    w->scaling->D = zalloc<c_float>(B->nmax); w->scaling->Dinv = zalloc<c_float>(B->nmax);
    w->scaling->E = zalloc<c_float>(B->mmax); w->scaling->Einv = zalloc<c_float>(B->mmax);
  }
  OSQPScaling *sc = w->scaling;
  if (B->fetch_scaling(0, sc->D, sc->E, &sc->c)) return;
  sc->cinv = 1.0 / sc->c;
  for (int i = 0; i < B->nmax; i++) sc->Dinv[i] = 1.0 / sc->D[i];
  for (int i = 0; i < B->mmax; i++) sc->Einv[i] = 1.0 / sc->E[i];
}

}  // namespace

extern "C" {

const char *rldl_version(void) { return "rldl_b200 0.1 (OSQP 0.6.0 recursive-LDL ABI)"; }

c_int rldl_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}
c_int rldl_set_device(c_int device) { return cudaSetDevice((int)device) == cudaSuccess ? 0 : 1; }

void osqp_set_default_settings(OSQPSettings *s) {
  // src/util.c:167-219 with include/constants.h:59-114
  s->rho = 0.1; s->sigma = 1e-6; s->scaling = 10;
  s->adaptive_rho = 1; s->adaptive_rho_interval = 0; s->adaptive_rho_tolerance = 5; s->adaptive_rho_fraction = 0.4;
  s->max_iter = 4000; s->eps_abs = 1e-3; s->eps_rel = 1e-3; s->eps_prim_inf = 1e-4; s->eps_dual_inf = 1e-4;
  s->alpha = 1.6; s->linsys_solver = QDLDL_SOLVER;
  s->delta = 1e-6; s->polish = 0; s->polish_refine_iter = 3; s->verbose = 1;
  s->scaled_termination = 0; s->check_termination = 25; s->warm_start = 1; s->time_limit = 0;
}

c_int osqp_setup_recursive(OSQPWorkspace **workp, OSQPDataRLDL *data, const OSQPSettings *settings,
                           c_int Nmax, c_int N, c_int nx, c_int nu, c_int ny, c_int nt) {
  if (data) { data->Nplus = 0; data->Nx = 0; }
  return setup_mpc(workp, data, settings, Nmax, N, nx, nu, ny, nt, 1);
}

c_int osqp_setup_combine_recursive(OSQPWorkspace **workp, OSQPDataRLDL *data, const OSQPSettings *settings,
                                   c_int Nmax, c_int N, c_int Nplus, c_int nx, c_int nu, c_int ny, c_int nt) {
  // head of N stages + tail of Nplus stages (src/recursive_ldl.c:2359-2756).  With dense stage
  // blocks the head/tail split is only a restart point of the same recursion.
  c_int rc = setup_mpc(workp, data, settings, Nmax, N + Nplus, nx, nu, ny, nt, 2);
  if (!rc) { data->Nx = N; data->Nplus = Nplus; }
  return rc;
}

c_int osqp_allocate_combine_recursive(OSQPWorkspace **, OSQPDataRLDL *, const OSQPSettings *, c_int, c_int,
                                      c_int, c_int, c_int) {
  return 0;  // no-op in the reference (src/recursive_ldl.c:2319-2356)
}

c_int osqp_update_X_horizon(OSQPWorkspace *, OSQPDataRLDL *, c_int) {
  return 0;  // empty body in the reference (src/recursive_ldl.c:2757-2759)
}

c_int osqp_update_recursive(OSQPWorkspace *work, OSQPDataRLDL *r, c_int N) {
  if (!work || !r) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if (N > r->Nmax) { printf("ERROR: the new horizon is too large: %lld > %lld\n", N, r->Nmax); return -1; }
  if (N == r->N) return 0;
  if (N < 1) { printf("ERROR: the new horizon is too small: %lld < 1\n", N); return -1; }
  return horizon_update(work, r, N);
}

c_int osqp_update_Z_horizon(OSQPWorkspace *work, OSQPDataRLDL *r, c_int deltaN) {
  if (!work || !r) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  const c_int Nnew = r->N + deltaN;
  if (Nnew < r->Nx || Nnew < 1) {
    printf("ERROR: This N is too small for the recursive/combined MPC. Please setup the problem again. ");
    return -1;
  }
  if (Nnew > r->Nmax) { printf("ERROR: the new horizon is too large: %lld > %lld\n", Nnew, r->Nmax); return -1; }
  if (deltaN == 0) return 0;
  c_int rc = horizon_update(work, r, Nnew);
  if (!rc) r->Nplus += deltaN;
  return rc;
}

c_int osqp_setup(OSQPWorkspace **workp, const OSQPData *data, const OSQPSettings *settings) {
  const double t0 = now_s();
  if (!workp || !data || !settings) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  // validate_data, src/auxil.c:794-887
  if (!data->P || !data->A || !data->q) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  if (data->n <= 0 || data->m < 0) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  if (data->P->m != data->n || data->P->m != data->P->n || data->A->m != data->m || data->A->n != data->n)
    return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  for (c_int j = 0; j < data->n; j++)
    for (c_int p = data->P->p[j]; p < data->P->p[j + 1]; p++)
      if (data->P->i[p] > j) { fprintf(stderr, "ERROR in validate_data: P is not upper triangular\n"); return osqp_error(OSQP_DATA_VALIDATION_ERROR); }
  for (c_int i = 0; i < data->m; i++)
    if (data->l[i] > data->u[i]) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  Batch *B = nullptr;
  c_int rc = Batch::create(&B, nullptr, nullptr, data->P, nullptr, nullptr, nullptr, data->A, settings, 0, 0,
                           data->n, 0, 0, data->m, 1, data->q, data->l, data->u, nullptr, true);
  if (rc) return osqp_error((int)rc);
  if (noncvx_at_setup(B)) return osqp_error(OSQP_NONCVX_ERROR);
  OSQPWorkspace *w = nullptr;
  if (alloc_workspace(&w, settings, data->n, data->m)) { delete B; return osqp_error(OSQP_MEM_ALLOC_ERROR); }
  *workp = w;
  w->data->n = data->n; w->data->m = data->m;
  w->data->P = csc_dup(data->P); w->data->A = csc_dup(data->A);
  w->data->q = zalloc<c_float>(data->n); w->data->l = zalloc<c_float>(data->m); w->data->u = zalloc<c_float>(data->m);
  memcpy(w->data->q, data->q, sizeof(c_float) * data->n);
  memcpy(w->data->l, data->l, sizeof(c_float) * data->m);
  memcpy(w->data->u, data->u, sizeof(c_float) * data->m);
  w->settings->rho = fmin(fmax(settings->rho, RHO_MIN), RHO_MAX);
  constr_types(w, data->m);
  PrivSolver *p = zalloc<PrivSolver>(1);
  p->batch = B; p->rldl = nullptr; p->kind = 0;
  finish_setup(w, p, t0);
  return 0;
}

c_int osqp_solve(OSQPWorkspace *work) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  PrivSolver *p = priv(work);
  Batch *B = p->batch;
  if (work->clear_update_time == 1) work->info->update_time = 0.0;
  const double t0 = now_s();
  sync_settings(work);
  // run time so far, as the reference counts it (osqp.c:391-396): setup time on the first run, update time after
  B->time_spent_before_solve = work->first_run ? work->info->setup_time : work->info->update_time;
  c_int rc = B->solve();
  if (rc) return 1;
  c_int iter = 0, status = 0, ru = 0;
  c_float obj = 0, pri = 0, dua = 0, rest = 0, rho = 0;
  if (B->get_info(&iter, &status, &obj, &pri, &dua, &ru, &rest)) return 1;
  if (B->fetch_vectors(0, work->x, work->y, work->z, work->delta_x, work->delta_y, work->rho_vec, &rho)) return 1;
  OSQPInfo *info = work->info;
  info->iter = iter; set_status(info, status); info->obj_val = obj; info->pri_res = pri; info->dua_res = dua;
  info->rho_updates = ru; info->rho_estimate = rest;
  work->settings->rho = rho; B->settings.rho = rho;
  const c_int n = work->data->n, m = work->data->m;
  for (c_int i = 0; i < m; i++) work->rho_inv_vec[i] = 1.0 / work->rho_vec[i];
  const bool nosol = status == OSQP_PRIMAL_INFEASIBLE || status == OSQP_PRIMAL_INFEASIBLE_INACCURATE ||
                     status == OSQP_DUAL_INFEASIBLE || status == OSQP_DUAL_INFEASIBLE_INACCURATE || status == OSQP_NON_CVX;
  for (c_int i = 0; i < n; i++) work->solution->x[i] = nosol ? OSQP_NAN : work->x[i];
  for (c_int i = 0; i < m; i++) work->solution->y[i] = nosol ? OSQP_NAN : work->y[i];
  // polish (src/polish.c) ran inside the solve iff it is on and the instance ended OSQP_SOLVED (osqp.c:589-591);
  // otherwise status_polish / polish_time keep their previous values, as in the reference
  if (work->settings->polish && status == OSQP_SOLVED) {
    c_int sp = 0;
    if (B->get_polish(&sp)) return 1;
    info->status_polish = sp;
    info->polish_time = 1e-3 * B->last_polish_ms;
  }
  info->solve_time = now_s() - t0 - info->polish_time;
  info->run_time = (work->first_run ? info->setup_time : info->update_time) + info->solve_time + info->polish_time;
  if (work->first_run) work->first_run = 0;
  work->clear_update_time = 1;
  work->rho_update_from_solve = 0;
  return 0;
}

c_int osqp_solve_recursive(OSQPWorkspace *work) { return osqp_solve(work); }
c_int osqp_solve_combine_recursive(OSQPWorkspace *work) { return osqp_solve(work); }

c_int osqp_cleanup(OSQPWorkspace *w) {
  if (!w) return 0;
  if (w->linsys_solver) w->linsys_solver->free(w->linsys_solver);
  if (w->data) {
    csc_del(w->data->P); csc_del(w->data->A);
    free(w->data->q); free(w->data->l); free(w->data->u); free(w->data);
  }
  free(w->rho_vec); free(w->rho_inv_vec); free(w->constr_type);
  free(w->x); free(w->y); free(w->z); free(w->xz_tilde); free(w->x_prev); free(w->z_prev);
  free(w->Ax); free(w->Px); free(w->Aty); free(w->delta_y); free(w->Atdelta_y);
  free(w->delta_x); free(w->Pdelta_x); free(w->Adelta_x);
  if (w->solution) { free(w->solution->x); free(w->solution->y); free(w->solution); }
  if (w->scaling) { free(w->scaling->D); free(w->scaling->Dinv); free(w->scaling->E); free(w->scaling->Einv); free(w->scaling); }
  free(w->info); free(w->timer); free(w->settings);
  free(w);
  return 0;
}

void cleanup_rldl(OSQPDataRLDL *data) {
  // src/recursive_ldl.c:15-59 (the reference also frees the struct itself)
  if (data) {
    free(data->X_even); free(data->Amats); free(data->Qmats);
    free(data);
  }
}

void get_L_dimensions(OSQPWorkspace *work, c_int *n, c_int *m, c_int *nnz) {
  if (!work) return;
  Batch *B = priv(work)->batch;
  *n = work->data->n + work->data->m;
  *m = work->data->n + work->data->m;
  *nnz = (c_int)B->N * ((c_int)B->a * B->a + (c_int)B->nx * B->nx) + (c_int)B->nx * B->nx;
}

c_int osqp_update_lin_cost(OSQPWorkspace *work, const c_float *q_new) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  tic_update(work);
  memcpy(work->data->q, q_new, sizeof(c_float) * work->data->n);
  Batch *B = priv(work)->batch;
  std::vector<c_float> q(B->nmax, 0.0);
  memcpy(q.data(), work->data->q, sizeof(c_float) * B->nmax);
  c_int rc = B->update_lin_cost(q.data());
  reset_info(work->info);
  toc_update(work);
  return rc ? 1 : 0;
}

static c_int push_bounds(OSQPWorkspace *work) {
  Batch *B = priv(work)->batch;
  c_int rc = B->update_bounds(work->data->l, work->data->u, true, true);
  reset_info(work->info);
  constr_types(work, B->mmax);
  return rc;
}

c_int osqp_update_bounds(OSQPWorkspace *work, const c_float *l_new, const c_float *u_new) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  tic_update(work);
  const c_int m = work->data->m;
  for (c_int i = 0; i < m; i++)
    if (l_new[i] > u_new[i]) { fprintf(stderr, "ERROR in osqp_update_bounds: lower bound must be lower than or equal to upper bound\n"); return 1; }
  memcpy(work->data->l, l_new, sizeof(c_float) * m);
  memcpy(work->data->u, u_new, sizeof(c_float) * m);
  c_int rc = push_bounds(work);
  toc_update(work);
  return rc;
}

c_int osqp_update_lower_bound(OSQPWorkspace *work, const c_float *l_new) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  tic_update(work);
  const c_int m = work->data->m;
  for (c_int i = 0; i < m; i++)
    if (l_new[i] > work->data->u[i]) { fprintf(stderr, "ERROR in osqp_update_lower_bound: upper bound must be greater than or equal to lower bound\n"); return 1; }
  memcpy(work->data->l, l_new, sizeof(c_float) * m);
  c_int rc = push_bounds(work);
  toc_update(work);
  return rc;
}

c_int osqp_update_upper_bound(OSQPWorkspace *work, const c_float *u_new) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  tic_update(work);
  const c_int m = work->data->m;
  for (c_int i = 0; i < m; i++)
    if (u_new[i] < work->data->l[i]) { fprintf(stderr, "ERROR in osqp_update_upper_bound: lower bound must be lower than or equal to upper bound\n"); return 1; }
  memcpy(work->data->u, u_new, sizeof(c_float) * m);
  c_int rc = push_bounds(work);
  toc_update(work);
  return rc;
}

c_int osqp_partial_update_bounds(OSQPWorkspace *work, c_int start, c_int stop, const c_float *l_new,
                                 const c_float *u_new) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if (start >= stop) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  Batch *B = priv(work)->batch;
  if (start < 0 || stop > B->mmax) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  tic_update(work);
  for (c_int i = 0; i < stop - start; i++)
    if (l_new[i] > u_new[i]) { printf("lower bound must be lower than or equal to upper bound\n"); return 1; }
  memcpy(work->data->l + start, l_new, sizeof(c_float) * (stop - start));
  memcpy(work->data->u + start, u_new, sizeof(c_float) * (stop - start));
  c_int rc = B->partial_update_bounds(start, stop, l_new, u_new);
  reset_info(work->info);
  constr_types(work, B->mmax);
  toc_update(work);
  return rc;
}

c_int osqp_warm_start(OSQPWorkspace *work, const c_float *x, const c_float *y) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  Batch *B = priv(work)->batch;
  if (!work->settings->warm_start) work->settings->warm_start = 1;
  memcpy(work->x, x, sizeof(c_float) * work->data->n);
  memcpy(work->y, y, sizeof(c_float) * work->data->m);
  return B->warm_start(work->x, work->y) ? 1 : 0;
}
c_int osqp_warm_start_x(OSQPWorkspace *work, const c_float *x) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if (!work->settings->warm_start) work->settings->warm_start = 1;
  memcpy(work->x, x, sizeof(c_float) * work->data->n);
  return priv(work)->batch->warm_start(work->x, nullptr) ? 1 : 0;
}
c_int osqp_warm_start_y(OSQPWorkspace *work, const c_float *y) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if (!work->settings->warm_start) work->settings->warm_start = 1;
  memcpy(work->y, y, sizeof(c_float) * work->data->m);
  return priv(work)->batch->warm_start(nullptr, work->y) ? 1 : 0;
}

c_int osqp_update_rho(OSQPWorkspace *work, c_float rho_new) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if (rho_new <= 0) { fprintf(stderr, "ERROR in osqp_update_rho: rho must be positive\n"); return 1; }
  tic_update(work);
  work->settings->rho = fmin(fmax(rho_new, RHO_MIN), RHO_MAX);
  Batch *B = priv(work)->batch;
  B->settings.rho = work->settings->rho;
  c_float r = work->settings->rho;
  c_int rc = B->update_rho(&r);
  constr_types(work, B->mmax);
  toc_update(work);
  return rc;
}

// P / A value updates (src/osqp.c:999-1266): scatter into the host CSC, re-derive the stage
// model from it and re-run the full recursion.  For stage-structured workspaces every stage
// must keep sharing one block (the reference's recursion reads one Qi/Ai for all stages).
static c_int rebuild_model(OSQPWorkspace *work) {
  PrivSolver *p = priv(work);
  Batch *B = p->batch;
  c_int rc;
  if (p->kind == 0) rc = B->upload_model(nullptr, nullptr, work->data->P, nullptr, nullptr, nullptr, work->data->A);
  else {
    OSQPDataRLDL *r = p->rldl;
    rc = B->upload_model(r->Q0, r->Qi, r->QN, r->A0, r->Ai, r->Aij, r->AN);
  }
  if (rc) return -1;
  return B->refactor() ? -1 : 0;
}

static c_int scatter_vals(csc *M, const c_float *v, const c_int *idx, c_int cnt) {
  const c_int nnz = M->p[M->n];
  if (idx) {
    if (cnt > nnz) return 1;
    for (c_int i = 0; i < cnt; i++) {
      if (idx[i] < 0 || idx[i] >= nnz) return 1;
      M->x[idx[i]] = v[i];
    }
  } else {
    for (c_int i = 0; i < nnz; i++) M->x[i] = v[i];
  }
  return 0;
}

// Re-extract the stage blocks from the updated global matrices (stage-structured workspaces).
// commit = false: only check that the update keeps the stage structure (every interior stage equal to stage 1,
// Aij untouched) — nothing is written; commit = true: copy the new values into the user's stage blocks.
static c_int stage_blocks_from_global(OSQPWorkspace *work, bool commit) {
  PrivSolver *p = priv(work);
  OSQPDataRLDL *r = p->rldl;
  const csc *P = work->data->P, *A = work->data->A;
  const c_int a = r->nx + r->nu, N = r->N;
  c_int pp = 0, ap = 0;
  auto take = [commit](csc *blk, const csc *G, c_int &pos) {
    const c_int nnz = blk->p[blk->n];
    if (commit) memcpy(blk->x, G->x + pos, sizeof(c_float) * nnz);
    pos += nnz;
  };
  // columns are stored stage by stage: [Q0|A0] , then per stage [Qi] and [Aij;Ai] interleaved per column
  take(r->Q0, P, pp); take(r->A0, A, ap);
  bool ok = true;
  c_int p1 = -1;                 // position of stage 1's Qi values in P
  std::vector<c_int> a1(a, -1);  // position of stage 1's Ai column values in A
  for (c_int k = 1; k < N; k++) {
    const c_int nq = r->Qi->p[r->Qi->n];
    if (k == 1) { p1 = pp; take(r->Qi, P, pp); }
    else { ok &= memcmp(P->x + p1, P->x + pp, sizeof(c_float) * nq) == 0; pp += nq; }
    for (c_int j = 0; j < a; j++) {
      const c_int nij = r->Aij->p[j + 1] - r->Aij->p[j], ni = r->Ai->p[j + 1] - r->Ai->p[j];
      for (c_int e = 0; e < nij; e++) ok &= A->x[ap + e] == r->Aij->x[r->Aij->p[j] + e];
      ap += nij;
      if (k == 1) {
        a1[j] = ap;
        if (commit) memcpy(r->Ai->x + r->Ai->p[j], A->x + ap, sizeof(c_float) * ni);
      } else {
        ok &= memcmp(A->x + a1[j], A->x + ap, sizeof(c_float) * ni) == 0;
      }
      ap += ni;
    }
  }
  take(r->QN, P, pp);
  for (c_int j = 0; j < r->nx; j++) {
    const c_int nij = r->Aij->p[j + 1] - r->Aij->p[j], ni = r->AN->p[j + 1] - r->AN->p[j];
    for (c_int e = 0; e < nij; e++) ok &= A->x[ap + e] == r->Aij->x[r->Aij->p[j] + e];
    ap += nij;
    if (commit) memcpy(r->AN->x + r->AN->p[j], A->x + ap, sizeof(c_float) * ni);
    ap += ni;
  }
  return ok ? 0 : 1;
}

static c_int update_PA_common(OSQPWorkspace *work, const c_float *Px, const c_int *Pi, c_int Pn,
                              const c_float *Ax, const c_int *Ai_, c_int An) {
  tic_update(work);
  // a refused update leaves the workspace as it was: the host mirrors are restored from these copies
  csc *P = work->data->P, *A = work->data->A;
  const std::vector<c_float> Pold(P->x, P->x + P->p[P->n]), Aold(A->x, A->x + A->p[A->n]);
  auto refuse = [&](c_int rc) {
    memcpy(P->x, Pold.data(), sizeof(c_float) * Pold.size());
    memcpy(A->x, Aold.data(), sizeof(c_float) * Aold.size());
    toc_update(work);
    return rc;
  };
  if (Px && scatter_vals(P, Px, Pi, Pn)) { fprintf(stderr, "ERROR in osqp_update_P: new number of elements greater than elements in P\n"); return refuse(1); }
  if (Ax && scatter_vals(A, Ax, Ai_, An)) { fprintf(stderr, "ERROR in osqp_update_A: new number of elements greater than elements in A\n"); return refuse(Px ? 2 : 1); }
  if (priv(work)->kind != 0) {
    if (stage_blocks_from_global(work, false)) {
      fprintf(stderr, "ERROR in osqp_update_P_A: on a recursive workspace all interior stages must keep sharing one (Qi, Ai) block and Aij is fixed\n");
      return refuse(1);
    }
    stage_blocks_from_global(work, true);
  }
  c_int rc = rebuild_model(work);
  if (!rc) mirror_scaling(work);
  reset_info(work->info);
  toc_update(work);
  return rc;
}

c_int osqp_update_P(OSQPWorkspace *work, const c_float *Px_new, const c_int *Px_new_idx, c_int P_new_n) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  return update_PA_common(work, Px_new, Px_new_idx, P_new_n, nullptr, nullptr, 0);
}
c_int osqp_update_A(OSQPWorkspace *work, const c_float *Ax_new, const c_int *Ax_new_idx, c_int A_new_n) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  return update_PA_common(work, nullptr, nullptr, 0, Ax_new, Ax_new_idx, A_new_n);
}
c_int osqp_update_P_A(OSQPWorkspace *work, const c_float *Px_new, const c_int *Px_new_idx, c_int P_new_n,
                      const c_float *Ax_new, const c_int *Ax_new_idx, c_int A_new_n) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  return update_PA_common(work, Px_new, Px_new_idx, P_new_n, Ax_new, Ax_new_idx, A_new_n);
}

#define SETTER(name, type, field, cond, msg)                                      \
  c_int osqp_update_##name(OSQPWorkspace *work, type v) {                         \
    if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);                  \
    if (cond) { fprintf(stderr, "ERROR in osqp_update_" #name ": " msg "\n"); return 1; } \
    work->settings->field = v;                                                    \
    sync_settings(work);                                                          \
    return 0;                                                                     \
  }
SETTER(max_iter, c_int, max_iter, v <= 0, "max_iter must be positive")
SETTER(eps_abs, c_float, eps_abs, v < 0., "eps_abs must be nonnegative")
SETTER(eps_rel, c_float, eps_rel, v < 0., "eps_rel must be nonnegative")
SETTER(eps_prim_inf, c_float, eps_prim_inf, v < 0., "eps_prim_inf must be nonnegative")
SETTER(eps_dual_inf, c_float, eps_dual_inf, v < 0., "eps_dual_inf must be nonnegative")
SETTER(alpha, c_float, alpha, (v <= 0. || v >= 2.), "alpha must be between 0 and 2")
SETTER(warm_start, c_int, warm_start, (v != 0 && v != 1), "warm_start should be either 0 or 1")
SETTER(scaled_termination, c_int, scaled_termination, (v != 0 && v != 1), "scaled_termination should be either 0 or 1")
SETTER(check_termination, c_int, check_termination, v < 0, "check_termination should be nonnegative")
SETTER(delta, c_float, delta, v <= 0., "delta must be positive")
SETTER(polish_refine_iter, c_int, polish_refine_iter, v < 0, "polish_refine_iter must be nonnegative")
SETTER(verbose, c_int, verbose, (v != 0 && v != 1), "verbose should be either 0 or 1")
SETTER(time_limit, c_float, time_limit, v < 0., "time_limit must be nonnegative")
#undef SETTER

c_int osqp_update_polish(OSQPWorkspace *work, c_int polish_new) {
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if (polish_new != 0 && polish_new != 1) { fprintf(stderr, "ERROR in osqp_update_polish: polish should be either 0 or 1\n"); return 1; }
  work->settings->polish = polish_new;
  work->info->polish_time = 0.0;
  sync_settings(work);
  return 0;
}

// ------------------------------ batch API ------------------------------
struct RLDLBatch { Batch *b; };

c_int rldl_batch_setup(RLDLBatch **out, const OSQPDataRLDL *model, const OSQPSettings *settings, c_int Nmax,
                       c_int N, c_int nx, c_int nu, c_int ny, c_int nt, c_int batch, const c_float *q,
                       const c_float *l, const c_float *u, const c_float *rho) {
  if (!out || !model || !settings || !q || !l || !u) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  *out = nullptr;
  Batch *B = nullptr;
  c_int rc = Batch::create(&B, model->Q0, model->Qi, model->QN, model->A0, model->Ai, model->Aij, model->AN,
                           settings, Nmax, N, nx, nu, ny, nt, batch, q, l, u, rho, false);
  if (rc) return osqp_error((int)rc);
  RLDLBatch *h = (RLDLBatch *)calloc(1, sizeof(RLDLBatch));
  h->b = B;
  *out = h;
  return 0;
}
#define HB(h) if (!(h) || !(h)->b) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR)
c_int rldl_batch_solve(RLDLBatch *h) { HB(h); return h->b->solve(); }
c_int rldl_batch_solve_to_host(RLDLBatch *h, c_float *x, c_float *y, c_int *iter, c_int *status_val, c_float *obj_val,
                               c_float *pri_res, c_float *dua_res, c_int *rho_updates, c_float *rho_estimate) {
  HB(h);
  return h->b->solve_to_host(x, y, iter, status_val, obj_val, pri_res, dua_res, rho_updates, rho_estimate);
}
c_int rldl_batch_update_model(RLDLBatch *h, const OSQPDataRLDL *model) {
  HB(h);
  if (!model) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  Batch *B = h->b;
  if (B->generic) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  c_int rc = B->upload_model(model->Q0, model->Qi, model->QN, model->A0, model->Ai, model->Aij, model->AN);
  if (rc) return rc;
  rc = B->reset_info();
  if (rc) return rc;
  return B->refactor();
}
c_int rldl_batch_failed_factorisations(RLDLBatch *h, c_int *flag) { HB(h); return h->b->failed_factorisations(flag); }
c_int rldl_batch_update_bounds(RLDLBatch *h, const c_float *l, const c_float *u) {
  HB(h);
  const c_int rc = h->b->update_bounds(l, u, true, true, true);  // l <= u is validated on the device
  if (rc == 1) fprintf(stderr, "ERROR in rldl_batch_update_bounds: lower bound must be lower than or equal to upper bound\n");
  return rc;
}
c_int rldl_batch_partial_update_bounds(RLDLBatch *h, c_int start, c_int stop, const c_float *l, const c_float *u) {
  HB(h);
  if (start >= stop || start < 0 || stop > h->b->mmax) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  const size_t cnt = (size_t)h->b->batch * (stop - start);
  for (size_t i = 0; i < cnt; i++)
    if (l[i] > u[i]) { fprintf(stderr, "lower bound must be lower than or equal to upper bound\n"); return 1; }
  return h->b->partial_update_bounds(start, stop, l, u);
}
c_int rldl_batch_update_lin_cost(RLDLBatch *h, const c_float *q) { HB(h); return h->b->update_lin_cost(q); }
c_int rldl_batch_update_data_refactor(RLDLBatch *h, const c_float *q, const c_float *l, const c_float *u) {
  HB(h);
  const c_int rc = h->b->update_data_refactor(q, l, u);
  if (rc == 1) fprintf(stderr, "ERROR in rldl_batch_update_data_refactor: lower bound must be lower than or equal to upper bound\n");
  return rc;
}
c_int rldl_batch_warm_start(RLDLBatch *h, const c_float *x, const c_float *y) { HB(h); return h->b->warm_start(x, y); }
c_int rldl_batch_warm_start_shift(RLDLBatch *h) { HB(h); return h->b->warm_start_shift(); }
c_int rldl_batch_update_horizon(RLDLBatch *h, c_int N) { HB(h); return h->b->update_horizon(N); }
c_int rldl_batch_update_rho(RLDLBatch *h, const c_float *rho) { HB(h); return h->b->update_rho(rho); }
c_int rldl_batch_refactor(RLDLBatch *h) { HB(h); return h->b->refactor(); }
c_int rldl_batch_kkt_solve(RLDLBatch *h, c_float *rhs) { HB(h); return h->b->kkt_solve(rhs); }
c_int rldl_batch_kkt_solve_static(RLDLBatch *h, const c_float *rhs, c_float *sol) {
  HB(h);
  if (!rhs || !sol) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  return h->b->kkt_solve_static(rhs, sol);
}
c_int rldl_batch_max_static_doubles(void) { return (c_int)rldl::kKktStaticDoubles; }
c_int rldl_batch_get_solution(RLDLBatch *h, c_float *x, c_float *y) { HB(h); return h->b->get_solution(x, y); }
c_int rldl_batch_get_info(RLDLBatch *h, c_int *iter, c_int *status_val, c_float *obj_val, c_float *pri_res,
                          c_float *dua_res, c_int *rho_updates, c_float *rho_estimate) {
  HB(h);
  return h->b->get_info(iter, status_val, obj_val, pri_res, dua_res, rho_updates, rho_estimate);
}
c_int rldl_batch_get_polish(RLDLBatch *h, c_int *status_polish) {
  HB(h);
  if (!status_polish) return 1;
  return h->b->get_polish(status_polish);
}
c_int rldl_batch_get_scaling(RLDLBatch *h, c_float *D, c_float *E, c_float *c) {
  HB(h);
  if (!D || !E || !c) return 1;
  return h->b->fetch_scaling(-1, D, E, c);
}
c_int rldl_batch_update_settings(RLDLBatch *h, const OSQPSettings *s) { HB(h); return h->b->update_settings(s); }
void rldl_batch_dims(const RLDLBatch *h, c_int *n, c_int *m, c_int *N, c_int *batch) {
  if (!h || !h->b) return;
  if (n) *n = h->b->D.n;
  if (m) *m = h->b->D.m;
  if (N) *N = h->b->N;
  if (batch) *batch = h->b->batch;
}
void rldl_batch_last_times(const RLDLBatch *h, double *factor_ms, double *solve_ms) {
  if (!h || !h->b) return;
  if (factor_ms) *factor_ms = h->b->last_factor_ms;
  if (solve_ms) *solve_ms = h->b->last_solve_ms;
}
void rldl_batch_stream_stats(const RLDLBatch *h, double *stream_ms, double *stream_instance_iterations) {
  if (!h || !h->b) return;
  if (stream_ms) *stream_ms = h->b->last_stream_ms;
  if (stream_instance_iterations) *stream_instance_iterations = h->b->last_stream_iters;
}
void rldl_batch_counters(const RLDLBatch *h, double *admm_bytes, double *factor_bytes, c_int *launches) {
  if (!h || !h->b) return;
  if (admm_bytes) *admm_bytes = h->b->admm_bytes_per_iteration();
  if (factor_bytes) *factor_bytes = h->b->factor_bytes;
  if (launches) *launches = h->b->launches;
}
c_int rldl_batch_select_kernel(RLDLBatch *h, c_int which) {
  HB(h);
  if (which == 2) return h->b->streaming ? 0 : 1;  // query: is this shape in the specialisation table?
  if (which < 0 || which > 1) return 1;
  h->b->kernel_choice = (int)which;
  return 0;
}
c_int rldl_batch_set_tuning(RLDLBatch *h, const char *name, c_int value) {
  HB(h);
  if (!name || value < -1) return 1;
  Batch *B = h->b;
  if (!strcmp(name, "lat_below")) B->lat_below = (int)value < 0 ? -1 : (int)value;
  else if (!strcmp(name, "ring_depth")) B->tune.ring_depth = (int)value < 0 ? 0 : (int)value;
  else if (!strcmp(name, "sched_s0")) B->tune.sched_s0 = (int)value;
  else if (!strcmp(name, "ctas_per_sm")) B->tune.ctas_per_sm = (int)value < 0 ? 0 : (int)value;
  else if (!strcmp(name, "l2_hints")) B->tune.l2_hints = (int)value < 0 ? 1 : (int)value;
  else if (!strcmp(name, "trace")) return B->set_trace(value > 0);
  else if (!strcmp(name, "lat_lane")) B->lat_lane = (int)value < 0 ? 8 : (int)value;
  else if (!strcmp(name, "lat_retire_below")) B->lat_retire_below = (int)value < 0 ? 0 : (int)value;
  else if (!strcmp(name, "lane_tail")) B->lane_tail = (int)value < 0 ? 0 : (int)value;
  else if (!strcmp(name, "lat_lockstep")) B->D.lat_lockstep = (int)value < 0 ? 1 : ((int)value ? 1 : 0);
  else if (!strcmp(name, "lat_levels")) B->D.lat_levels = (int)value < 1 ? 1 : (int)value;
  else if (!strcmp(name, "sched_key")) B->D.sched_key = (int)value < 0 ? 1 : ((int)value ? 1 : 0);
  else return 1;
  return 0;
}
// diagnostic: raw copy of the device scheduler state (kSchedCounters ints) — usable while a solve is running
c_int rldl_batch_debug_sched(RLDLBatch *h, int *out, c_int n) {
  HB(h);
  if (n > rldl::kSchedCounters) n = rldl::kSchedCounters;
  static cudaStream_t dbg = nullptr;
  if (!dbg && cudaStreamCreateWithFlags(&dbg, cudaStreamNonBlocking) != cudaSuccess) return 1;
  if (cudaMemcpyAsync(out, h->b->D.sched, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, dbg) != cudaSuccess) return 1;
  return cudaStreamSynchronize(dbg) == cudaSuccess ? 0 : 1;
}
c_int rldl_batch_get_trace(RLDLBatch *h, c_float *t) { HB(h); return t ? h->b->get_trace(t) : 1; }
c_int rldl_batch_lane_stats(RLDLBatch *h, c_int *stats) {
  HB(h);
  int v[6] = {0, 0, 0, 0, 0, 0};
  if (cudaSetDevice(h->b->device) != cudaSuccess ||
      cudaMemcpy(v, h->b->D.sched + rldl::kSchedStats, sizeof(v), cudaMemcpyDeviceToHost) != cudaSuccess) return 1;
  for (int i = 0; i < 6; i++) stats[i] = v[i];
  return 0;
}
void *rldl_batch_stream(const RLDLBatch *h) { return (h && h->b) ? (void *)h->b->stream : nullptr; }
void rldl_batch_free(RLDLBatch *h) {
  if (!h) return;
  delete h->b;
  free(h);
}

}  // extern "C"
