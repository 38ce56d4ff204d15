/* oracle/ref_driver.c — TEST INFRASTRUCTURE ONLY.
 *
 * A thin batch driver around the UNMODIFIED reference library
 * (oracle/_ref/libosqp_ref.so, compiled from /root/reference by oracle/Makefile).
 * It exists so that tests/ and bench.py's cpu_baseline / --impl reference legs can
 * run the reference's own C API (osqp_setup_recursive, osqp_solve_recursive,
 * osqp_update_recursive, osqp_update_bounds, osqp_warm_start … — reference
 * include/recursive_ldl.h:52-74, include/osqp.h:32-419) over many independent MPC
 * instances, one OSQPWorkspace per instance, a pthread pool with dynamic instance scheduling (the
 * reference is re-entrant per workspace, SURVEY §8b "Threading").
 *
 * Nothing here is shipped or measured as the product; the product library
 * (librldl_b200.so) never links or loads this file.
 *
 * path = 0 : reference recursive-LDL path (osqp_setup_recursive / osqp_solve_recursive)
 * path = 1 : reference stock path (osqp_setup / osqp_solve on the P/A assembled by
 *            the reference's own setup_AP_matrices, src/recursive_ldl.c:1873)
 *
 * Known reference defects this driver steps around:
 *   F1  rho updates segfault on a recursive workspace -> refb_update_rho is stock-only.
 *   F5  osqp_cleanup on a recursive workspace frees garbage -> recursive workspaces
 *       are leaked on refb_free (bounded batch sizes only).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>
#include "osqp.h"
#include "recursive_ldl.h"
#include "cs.h"

extern c_int setup_AP_matrices(OSQPData *data, c_int Nmax, c_int N, c_int nx, c_int nu,
                               c_int ny, c_int nt, csc *Q0, csc *Qi, csc *QN, csc *A0,
                               csc *Ai, csc *Aij, csc *AN);

typedef struct {
  c_int nx, nu, ny, nt, N, Nmax, batch, nmax, mmax;
  int path;
  csc *mats[7]; /* Q0,Qi,QN,A0,Ai,Aij,AN (shared by the batch, owned here) */
  OSQPSettings settings;
  OSQPWorkspace **work;
  OSQPDataRLDL **rldl;
  OSQPData **sdata;     /* stock path: per-instance assembled data */
  c_float *q, *l, *u;   /* batch x nmax / mmax, current values (for stock rebuilds) */
  c_float *rho;         /* per-instance initial rho */
  double *setup_time;
} RefBatch;


/* ---- tiny pthread parallel-for (no OpenMP dependency) ---- */
typedef int (*job_fn)(void *ctx, c_int k);
typedef struct { job_fn fn; void *ctx; c_int n; volatile long long next; volatile int bad; } ParFor;
static void *parfor_worker(void *p) {
  ParFor *pf = (ParFor *)p;
  for (;;) {
    long long k = __sync_fetch_and_add(&pf->next, 1);
    if (k >= pf->n) break;
    if (pf->fn(pf->ctx, (c_int)k)) __sync_fetch_and_add(&pf->bad, 1);
  }
  return NULL;
}
static int parfor(job_fn fn, void *ctx, c_int n, int nthreads) {
  ParFor pf = {fn, ctx, n, 0, 0};
  if (nthreads < 1) nthreads = 1;
  if (nthreads > n) nthreads = (int)(n > 0 ? n : 1);
  if (nthreads == 1) { parfor_worker(&pf); return pf.bad; }
  pthread_t *t = (pthread_t *)malloc(sizeof(pthread_t) * nthreads);
  for (int i = 0; i < nthreads; i++) pthread_create(&t[i], NULL, parfor_worker, &pf);
  for (int i = 0; i < nthreads; i++) pthread_join(t[i], NULL);
  free(t);
  return pf.bad;
}

static csc *mk_csc(c_int m, c_int n, const c_int *p, const c_int *i, const c_float *x) {
  c_int nnz = p[n];
  csc *M = csc_spalloc(m, n, nnz > 0 ? nnz : 1, 1, 0);
  memcpy(M->p, p, sizeof(c_int) * (n + 1));
  memcpy(M->i, i, sizeof(c_int) * nnz);
  memcpy(M->x, x, sizeof(c_float) * nnz);
  M->nzmax = nnz;
  return M;
}

static int setup_one(RefBatch *h, c_int k) {
  c_int a = h->nx + h->nu, b = h->nx + h->ny;
  OSQPSettings s = h->settings;
  if (h->rho) s.rho = h->rho[k];
  c_int rc;
  if (h->path == 0) {
    OSQPDataRLDL *d = (OSQPDataRLDL *)calloc(1, sizeof(OSQPDataRLDL));
    d->Nmax = h->Nmax; d->N = h->N; d->nx = h->nx; d->nu = h->nu; d->ny = h->ny; d->nt = h->nt;
    d->Q0 = h->mats[0]; d->Qi = h->mats[1]; d->QN = h->mats[2];
    d->A0 = h->mats[3]; d->Ai = h->mats[4]; d->Aij = h->mats[5]; d->AN = h->mats[6];
    d->data.n = h->nmax; d->data.m = h->mmax;
    d->data.q = h->q + k * h->nmax; d->data.l = h->l + k * h->mmax; d->data.u = h->u + k * h->mmax;
    h->rldl[k] = d;
    rc = osqp_setup_recursive(&h->work[k], d, &s, h->Nmax, h->N, h->nx, h->nu, h->ny, h->nt);
    if (!rc) h->setup_time[k] = h->work[k]->info->setup_time;
  } else {
    OSQPData *d = (OSQPData *)calloc(1, sizeof(OSQPData));
    d->n = h->nmax; d->m = h->mmax; /* setup_AP_matrices allocates with these, then shrinks */
    setup_AP_matrices(d, h->Nmax, h->N, h->nx, h->nu, h->ny, h->nt, h->mats[0], h->mats[1],
                      h->mats[2], h->mats[3], h->mats[4], h->mats[5], h->mats[6]);
    d->n = h->N * a; d->m = h->N * b + h->nt;
    d->P->n = d->P->m = d->n; d->A->n = d->n; d->A->m = d->m; /* current horizon, not Nmax */
    d->P->nzmax = d->P->p[d->n]; d->A->nzmax = d->A->p[d->n];
    d->q = h->q + k * h->nmax; d->l = h->l + k * h->mmax; d->u = h->u + k * h->mmax;
    h->sdata[k] = d;
    rc = osqp_setup(&h->work[k], d, &s);
    if (!rc) h->setup_time[k] = h->work[k]->info->setup_time;
  }
  return (int)rc;
}

typedef struct {
  RefBatch *h;
  const c_float *a, *b;
  c_int i0, i1;
} Job;

static int job_setup(void *c, c_int k) { return setup_one(((Job *)c)->h, k) != 0; }

RefBatch *refb_create(c_int nx, c_int nu, c_int ny, c_int nt, c_int N, c_int Nmax,
                      const c_int *mat_m, const c_int *mat_n, const c_int **mat_p,
                      const c_int **mat_i, const c_float **mat_x,
                      const OSQPSettings *settings, c_int batch, const c_float *q,
                      const c_float *l, const c_float *u, const c_float *rho, int path,
                      int nthreads) {
  RefBatch *h = (RefBatch *)calloc(1, sizeof(RefBatch));
  h->nx = nx; h->nu = nu; h->ny = ny; h->nt = nt; h->N = N; h->Nmax = Nmax; h->batch = batch;
  h->nmax = Nmax * (nx + nu); h->mmax = Nmax * (nx + ny) + nt; h->path = path;
  for (int j = 0; j < 7; j++) h->mats[j] = mk_csc(mat_m[j], mat_n[j], mat_p[j], mat_i[j], mat_x[j]);
  h->settings = *settings;
  h->work = (OSQPWorkspace **)calloc(batch, sizeof(void *));
  h->rldl = (OSQPDataRLDL **)calloc(batch, sizeof(void *));
  h->sdata = (OSQPData **)calloc(batch, sizeof(void *));
  h->q = (c_float *)malloc(sizeof(c_float) * batch * h->nmax);
  h->l = (c_float *)malloc(sizeof(c_float) * batch * h->mmax);
  h->u = (c_float *)malloc(sizeof(c_float) * batch * h->mmax);
  memcpy(h->q, q, sizeof(c_float) * batch * h->nmax);
  memcpy(h->l, l, sizeof(c_float) * batch * h->mmax);
  memcpy(h->u, u, sizeof(c_float) * batch * h->mmax);
  if (rho) { h->rho = (c_float *)malloc(sizeof(c_float) * batch); memcpy(h->rho, rho, sizeof(c_float) * batch); }
  h->setup_time = (double *)calloc(batch, sizeof(double));
  Job j = {h, 0, 0, 0, 0};
  int bad = parfor(job_setup, &j, batch, nthreads);
  if (bad) fprintf(stderr, "refb_create: %d setups failed\n", bad);
  return h;
}

static int job_solve(void *c, c_int k) {
  RefBatch *h = ((Job *)c)->h;
  return (h->path == 0 ? osqp_solve_recursive(h->work[k]) : osqp_solve(h->work[k])) != 0;
}
int refb_solve(RefBatch *h, int nthreads) {
  Job j = {h, 0, 0, 0, 0};
  return parfor(job_solve, &j, h->batch, nthreads);
}

void refb_get(RefBatch *h, c_float *x, c_float *y, c_int *iter, c_int *status, c_float *obj,
              c_float *pri, c_float *dua, c_int *rho_updates, c_float *rho_est,
              double *setup_t, double *solve_t) {
  for (c_int k = 0; k < h->batch; k++) {
    OSQPWorkspace *w = h->work[k];
    c_int n = w->data->n, m = w->data->m;
    if (x) { memset(x + k * h->nmax, 0, sizeof(c_float) * h->nmax); memcpy(x + k * h->nmax, w->solution->x, sizeof(c_float) * n); }
    if (y) { memset(y + k * h->mmax, 0, sizeof(c_float) * h->mmax); memcpy(y + k * h->mmax, w->solution->y, sizeof(c_float) * m); }
    if (iter) iter[k] = w->info->iter;
    if (status) status[k] = w->info->status_val;
    if (obj) obj[k] = w->info->obj_val;
    if (pri) pri[k] = w->info->pri_res;
    if (dua) dua[k] = w->info->dua_res;
    if (rho_updates) rho_updates[k] = w->info->rho_updates;
    if (rho_est) rho_est[k] = w->info->rho_estimate;
    if (setup_t) setup_t[k] = h->setup_time[k];
    if (solve_t) solve_t[k] = w->info->solve_time;
  }
}

/* l,u: batch x mmax; only the first (current) m entries of each row are used. */
static int job_bounds(void *c, c_int k) {
  Job *j = (Job *)c; RefBatch *h = j->h;
  return osqp_update_bounds(h->work[k], j->a + k * h->mmax, j->b + k * h->mmax) != 0;
}
int refb_update_bounds(RefBatch *h, const c_float *l, const c_float *u, int nthreads) {
  memcpy(h->l, l, sizeof(c_float) * h->batch * h->mmax);
  memcpy(h->u, u, sizeof(c_float) * h->batch * h->mmax);
  Job j = {h, l, u, 0, 0};
  return parfor(job_bounds, &j, h->batch, nthreads);
}

/* l,u: batch x (stop-start) */
static int job_pbounds(void *c, c_int k) {
  Job *j = (Job *)c; RefBatch *h = j->h;
  c_int len = j->i1 - j->i0;
  memcpy(h->l + k * h->mmax + j->i0, j->a + k * len, sizeof(c_float) * len);
  memcpy(h->u + k * h->mmax + j->i0, j->b + k * len, sizeof(c_float) * len);
  return osqp_partial_update_bounds(h->work[k], j->i0, j->i1, j->a + k * len, j->b + k * len) != 0;
}
int refb_partial_update_bounds(RefBatch *h, c_int start, c_int stop, const c_float *l,
                               const c_float *u, int nthreads) {
  Job j = {h, l, u, start, stop};
  return parfor(job_pbounds, &j, h->batch, nthreads);
}

static int job_q(void *c, c_int k) {
  Job *j = (Job *)c; RefBatch *h = j->h;
  return osqp_update_lin_cost(h->work[k], j->a + k * h->nmax) != 0;
}
int refb_update_lin_cost(RefBatch *h, const c_float *q, int nthreads) {
  memcpy(h->q, q, sizeof(c_float) * h->batch * h->nmax);
  Job j = {h, q, 0, 0, 0};
  return parfor(job_q, &j, h->batch, nthreads);
}

/* x: batch x nmax, y: batch x mmax (either may be NULL) */
static int job_ws(void *c, c_int k) {
  Job *j = (Job *)c; RefBatch *h = j->h;
  const c_float *x = j->a, *y = j->b;
  if (x && y) return osqp_warm_start(h->work[k], x + k * h->nmax, y + k * h->mmax) != 0;
  if (x) return osqp_warm_start_x(h->work[k], x + k * h->nmax) != 0;
  if (y) return osqp_warm_start_y(h->work[k], y + k * h->mmax) != 0;
  return 0;
}
int refb_warm_start(RefBatch *h, const c_float *x, const c_float *y, int nthreads) {
  Job j = {h, x, y, 0, 0};
  return parfor(job_ws, &j, h->batch, nthreads);
}

static int job_hor(void *c, c_int k) {
  Job *j = (Job *)c; RefBatch *h = j->h;
  if (h->path == 0) return osqp_update_recursive(h->work[k], h->rldl[k], j->i0) != 0;
  OSQPData *d = h->sdata[k];
  osqp_cleanup(h->work[k]);
  csc_spfree(d->P); csc_spfree(d->A); free(d);
  return setup_one(h, k) != 0;
}
int refb_update_horizon(RefBatch *h, c_int Nnew, int nthreads) {
  if (Nnew == h->N) return 0;
  if (h->path == 1) h->N = Nnew; /* stock: rebuild at the new size */
  Job j = {h, 0, 0, Nnew, 0};
  int bad = parfor(job_hor, &j, h->batch, nthreads);
  h->N = Nnew;
  return bad;
}

static int job_rho(void *c, c_int k) {
  Job *j = (Job *)c;
  return osqp_update_rho(j->h->work[k], j->a[k]) != 0;
}
int refb_update_rho(RefBatch *h, const c_float *rho, int nthreads) {
  if (h->path == 0) return -1; /* F1: segfaults in the reference */
  Job j = {h, rho, 0, 0, 0};
  return parfor(job_rho, &j, h->batch, nthreads);
}

/* One KKT solve through the reference's vtable (qdldl_interface.c:559) on instance k.
 * rhs has n+m entries (current sizes) and is overwritten with [x_tilde ; z_tilde]. */
int refb_kkt_solve(RefBatch *h, c_int k, c_float *rhs) {
  OSQPWorkspace *w = h->work[k];
  return (int)w->linsys_solver->solve(w->linsys_solver, rhs);
}

/* Equilibration of every instance as the reference computed it (work->scaling, src/scaling.c:44-156):
 * D batch x nmax, E batch x mmax, c batch.  Returns 1 when scaling is off. */
int refb_get_scaling(RefBatch *h, c_float *D, c_float *E, c_float *c) {
  for (c_int k = 0; k < h->batch; k++) {
    OSQPWorkspace *w = h->work[k];
    if (!w->scaling) return 1;
    memset(D + k * h->nmax, 0, sizeof(c_float) * h->nmax);
    memset(E + k * h->mmax, 0, sizeof(c_float) * h->mmax);
    memcpy(D + k * h->nmax, w->scaling->D, sizeof(c_float) * w->data->n);
    memcpy(E + k * h->mmax, w->scaling->E, sizeof(c_float) * w->data->m);
    c[k] = w->scaling->c;
  }
  return 0;
}

/* polish outcome per instance (info->status_polish, src/polish.c) */
void refb_get_polish(RefBatch *h, c_int *status_polish) {
  for (c_int k = 0; k < h->batch; k++) status_polish[k] = h->work[k]->info->status_polish;
}

/* osqp_update_P_A with complete value arrays (the same for every instance; stock path: the values of the
 * assembled global P (upper triangle) and A in CSC order; recursive workspaces hold the same global CSC). */
static int job_pa(void *c, c_int k) {
  Job *j = (Job *)c; RefBatch *h = j->h;
  OSQPWorkspace *w = h->work[k];
  return osqp_update_P_A(w, j->a, OSQP_NULL, w->data->P->p[w->data->P->n], j->b, OSQP_NULL,
                         w->data->A->p[w->data->A->n]) != 0;
}
int refb_update_PA(RefBatch *h, const c_float *Px, const c_float *Ax, int nthreads) {
  if (h->path == 0) return -1; /* the recursive workspace has no KKT to update (F1) */
  Job j = {h, Px, Ax, 0, 0};
  return parfor(job_pa, &j, h->batch, nthreads);
}
/* nnz of the global P / A of instance 0 and their values (to build update arrays in the same order) */
void refb_get_PA(RefBatch *h, c_int *nnzP, c_int *nnzA, c_float *Px, c_float *Ax) {
  OSQPWorkspace *w = h->work[0];
  *nnzP = w->data->P->p[w->data->P->n]; *nnzA = w->data->A->p[w->data->A->n];
  if (Px) memcpy(Px, w->data->P->x, sizeof(c_float) * *nnzP);
  if (Ax) memcpy(Ax, w->data->A->x, sizeof(c_float) * *nnzA);
}

void refb_dims(RefBatch *h, c_int *n, c_int *m) {
  *n = h->work[0]->data->n; *m = h->work[0]->data->m;
}

void refb_free(RefBatch *h) {
  if (!h) return;
  if (h->path == 1) {
    for (c_int k = 0; k < h->batch; k++) {
      if (h->work[k]) osqp_cleanup(h->work[k]);
      if (h->sdata[k]) { csc_spfree(h->sdata[k]->P); csc_spfree(h->sdata[k]->A); free(h->sdata[k]); }
    }
  } /* path 0: leaked on purpose (F5) */
  for (int j = 0; j < 7; j++) csc_spfree(h->mats[j]);
  free(h->work); free(h->rldl); free(h->sdata);
  free(h->q); free(h->l); free(h->u); free(h->rho); free(h->setup_time);
  free(h);
}

void refb_default_settings(OSQPSettings *s) { osqp_set_default_settings(s); }
int refb_sizeof_settings(void) { return (int)sizeof(OSQPSettings); }
int refb_max_threads(void) { long n = sysconf(_SC_NPROCESSORS_ONLN); return n > 0 ? (int)n : 1; }
