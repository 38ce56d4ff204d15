/* tests/dropin/mpc_user.c — TEST INFRASTRUCTURE.
 *
 * A user program written against the REFERENCE's public headers (include/osqp.h,
 * include/recursive_ldl.h, include/cs.h), the way a user of laperss/osqp-recursive-ldl
 * writes a receding-horizon MPC loop.  tests/dropin/Makefile compiles it ONCE (one object
 * file, reference headers) and links that same object twice:
 *     oracle/_ref/dropin_ref    against the unmodified reference  (libosqp_ref.so)
 *     oracle/_ref/dropin_b200   against the product library       (librldl_b200.so)
 * tests/test_dropin_*.py run both and compare what they print.  With -DRLDL_OWN_HEADER the
 * same source compiles against include/osqp_rldl_abi.h alone (dropin_b200_ownhdr).
 *
 * Entry points taken through the boundary (SURVEY §8b "must-keep set"): csc_matrix,
 * osqp_set_default_settings, osqp_setup_recursive, osqp_solve_recursive,
 * osqp_partial_update_bounds, osqp_warm_start, osqp_update_lin_cost, osqp_update_recursive
 * (shrink and grow), osqp_update_bounds, osqp_update_eps_abs/_rel, get_L_dimensions, and the
 * caller-visible fields work->info->{status_val,status,iter,obj_val,pri_res,dua_res},
 * work->solution->{x,y}, work->data->{n,m}.
 *
 * The instance (mpc_model.h): nx=4, nu=2, N=20 (BASELINE config C1), box constraints on state and
 * input, terminal box, Nmax = N+1, cost blocks stored as full upper triangles.
 * osqp_cleanup is not called: on a recursive workspace it is undefined behaviour in the
 * reference (F5), and this object file must run against it.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef RLDL_OWN_HEADER
#include "osqp_rldl_abi.h"
#else
#include "osqp.h"
#include "recursive_ldl.h"
#include "cs.h"
#endif

#include "mpc_model.h"

static void report(const char *tag, OSQPWorkspace *w) {
  c_int n = w->data->n, m = w->data->m;
  printf("SOLVE %s n %lld m %lld status_val %lld iter %lld obj %.17g pri %.17g dua %.17g status \"%s\"\n", tag,
         (long long)n, (long long)m, (long long)w->info->status_val, (long long)w->info->iter, w->info->obj_val,
         w->info->pri_res, w->info->dua_res, w->info->status);
  printf("X %s", tag);
  for (c_int k = 0; k < n; k++) printf(" %.17g", w->solution->x[k]);
  printf("\nY %s", tag);
  for (c_int k = 0; k < m; k++) printf(" %.17g", w->solution->y[k]);
  printf("\n");
}

int main(void) {
  double A[NX * NX], B[NX * NU]; /* column-major */
  OSQPDataRLDL *d = (OSQPDataRLDL *)calloc(1, sizeof(OSQPDataRLDL));
  build_model(d, A, B);

  /* q, l, u sized for Nmax, bounds pre-filled stage-like up to Nmax */
  const c_int nmax = HNMAX * NA, mmax = HNMAX * NB + NT;
  c_float *q = (c_float *)calloc(nmax, sizeof(c_float));
  c_float *l = (c_float *)calloc(mmax, sizeof(c_float));
  c_float *u = (c_float *)calloc(mmax, sizeof(c_float));
  double x0[NX], dyn0[NX];
  for (int r = 0; r < NX; r++) x0[r] = 8.0 * urand();
  fill_bounds(A, x0, l, u);
  d->data.n = nmax;
  d->data.m = mmax;
  d->data.q = q;
  d->data.l = l;
  d->data.u = u;

  OSQPSettings *s = (OSQPSettings *)malloc(sizeof(OSQPSettings));
  osqp_set_default_settings(s);
  s->scaling = 0; /* precondition of the recursive path */
  s->adaptive_rho = 0;
  s->eps_abs = s->eps_rel = 1e-5;
  s->check_termination = 25;
  s->polish = 0;
  s->verbose = 0;
  s->warm_start = 1;

  OSQPWorkspace *w = NULL;
  c_int rc = osqp_setup_recursive(&w, d, s, HNMAX, HN, NX, NU, NY, NT);
  printf("SETUP rc %lld\n", (long long)rc);
  if (rc) return 10 + (int)rc; /* e.g. 13 = OSQP_LINSYS_SOLVER_LOAD_ERROR: no usable device */

  c_int Kn, Km, Lnnz;
  get_L_dimensions(w, &Kn, &Km, &Lnnz);
  printf("DIMS n %lld m %lld\n", (long long)Kn, (long long)Km);

  rc = osqp_solve_recursive(w);
  report("cold", w);

  /* closed loop: apply u0, move the plant, shift nothing (plain warm start from the last solution) */
  c_float *xs = (c_float *)malloc(sizeof(c_float) * nmax), *ys = (c_float *)malloc(sizeof(c_float) * mmax);
  for (int step = 0; step < 2; step++) {
    double x1[NX];
    for (int r = 0; r < NX; r++) {
      x1[r] = 0.0;
      for (int j = 0; j < NX; j++) x1[r] += A[j * NX + r] * x0[j];
      for (int j = 0; j < NU; j++) x1[r] += B[j * NX + r] * w->solution->x[j];
    }
    memcpy(x0, x1, sizeof(x0));
    for (int r = 0; r < NX; r++) {
      dyn0[r] = 0.0;
      for (int j = 0; j < NX; j++) dyn0[r] -= A[j * NX + r] * x0[j];
    }
    memcpy(xs, w->solution->x, sizeof(c_float) * w->data->n);
    memcpy(ys, w->solution->y, sizeof(c_float) * w->data->m);
    rc |= osqp_partial_update_bounds(w, NY, NY + NX, dyn0, dyn0);
    rc |= osqp_warm_start(w, xs, ys);
    rc |= osqp_solve_recursive(w);
    report(step == 0 ? "step1" : "step2", w);
  }

  /* new linear cost (track a non-zero state reference) */
  for (int k = 1; k <= HNMAX; k++)
    for (int r = 0; r < NX; r++) q[NU + (k - 1) * NA + r] = -0.3 * (r + 1) / NX;
  rc |= osqp_update_lin_cost(w, q);
  rc |= osqp_solve_recursive(w);
  report("lincost", w);

  /* shrink the horizon, solve; grow it back, solve */
  rc |= osqp_update_recursive(w, d, HN - 1);
  rc |= osqp_solve_recursive(w);
  report("shrink", w);
  rc |= osqp_update_recursive(w, d, HN);
  rc |= osqp_solve_recursive(w);
  report("grow", w);

  /* tighter input box on every stage, tighter tolerances */
  memcpy(l, w->data->l, sizeof(c_float) * w->data->m);
  memcpy(u, w->data->u, sizeof(c_float) * w->data->m);
  for (int k = 0; k < HN; k++)
    for (int r = 0; r < NU; r++) {
      l[k * NB + NX + r] = -0.3;
      u[k * NB + NX + r] = 0.3;
    }
  rc |= osqp_update_bounds(w, l, u);
  rc |= osqp_update_eps_abs(w, 1e-7);
  rc |= osqp_update_eps_rel(w, 1e-7);
  rc |= osqp_solve_recursive(w);
  report("tight", w);

  /* an invalid update is refused with the reference's return code and changes nothing */
  l[NX] = 1.0;
  u[NX] = -1.0;
  printf("BADBOUNDS rc %lld\n", (long long)osqp_update_bounds(w, l, u));
  printf("DONE rc %lld\n", (long long)rc);
  return 0;
}
