/* tests/dropin/api_user.c — TEST INFRASTRUCTURE.
 *
 * Second user program compiled ONCE against the reference's own headers and linked against the
 * reference and against librldl_b200.so (see mpc_user.c).  It takes the remaining entry points of
 * the boundary through the same object file:
 *   part A — the stock API on a small generic QP (the reference's own basic_qp test problem,
 *     tests/basic_qp/generate_problem.py): osqp_setup, osqp_solve, osqp_update_lin_cost,
 *     osqp_update_lower_bound / _upper_bound, osqp_warm_start_x / _y, osqp_update_rho,
 *     osqp_update_P_A (new values, same pattern), osqp_update_max_iter, osqp_cleanup;
 *   part B — the combine variant on the C1-shaped MPC of mpc_model.h:
 *     osqp_setup_combine_recursive (head 15 + tail 3), osqp_solve_combine_recursive,
 *     osqp_update_Z_horizon (tail +2);
 *   part C — the reference's own unconstrained / non_cvx / primal_dual_infeasibility test problems;
 *   part D — the settings osqp_set_default_settings leaves (Ruiz scaling = 10, adaptive rho; the automatic
 *     wall-clock interval replaced by an explicit one so both sides are deterministic) and polish = 1 on the
 *     basic_qp problem and on a badly scaled QP; osqp_update_polish; time_limit reached.
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

static void report(const char *tag, OSQPWorkspace *w, c_int rc) {
  c_int n = w->data->n, m = w->data->m;
  printf("SOLVE %s rc %lld n %lld m %lld status_val %lld iter %lld obj %.17g status \"%s\"\n", tag, (long long)rc,
         (long long)n, (long long)m, (long long)w->info->status_val, (long long)w->info->iter, w->info->obj_val,
         w->info->status);
  printf("X %s", tag);
  for (c_int k = 0; k < n; k++) printf(" %.17g", w->solution->x[k]);
  printf("\nY %s", tag);
  for (c_int k = 0; k < m; k++) printf(" %.17g", w->solution->y[k]);
  printf("\n");
}

static int part_a(void) {
  /* minimize 1/2 x'Px + q'x  s.t. l <= Ax <= u ;  P upper triangle */
  c_float P_x[3] = {4.0, 1.0, 2.0};
  c_int P_i[3] = {0, 0, 1}, P_p[3] = {0, 1, 3};
  c_float q[2] = {1.0, 1.0};
  c_float A_x[4] = {1.0, 1.0, 1.0, 1.0};
  c_int A_i[4] = {0, 1, 0, 2}, A_p[3] = {0, 2, 4};
  c_float l[3] = {1.0, 0.0, 0.0}, u[3] = {1.0, 0.7, 0.7};
  OSQPData *data = (OSQPData *)calloc(1, sizeof(OSQPData));
  data->n = 2;
  data->m = 3;
  data->P = csc_matrix(2, 2, 3, P_x, P_i, P_p);
  data->q = q;
  data->A = csc_matrix(3, 2, 4, A_x, A_i, A_p);
  data->l = l;
  data->u = u;
  OSQPSettings *s = (OSQPSettings *)malloc(sizeof(OSQPSettings));
  osqp_set_default_settings(s);
  s->scaling = 0;
  s->adaptive_rho = 0;
  s->eps_abs = s->eps_rel = 1e-6;
  s->check_termination = 5;
  s->verbose = 0;
  s->polish = 0;
  s->warm_start = 1;
  OSQPWorkspace *w = NULL;
  c_int rc = osqp_setup(&w, data, s);
  printf("SETUP generic rc %lld\n", (long long)rc);
  if (rc) return 10 + (int)rc;
  rc = osqp_solve(w);
  report("qp", w, rc);

  c_float q2[2] = {2.5, 3.2};
  rc = osqp_update_lin_cost(w, q2);
  rc |= osqp_solve(w);
  report("qp_lincost", w, rc);

  c_float l2[3] = {0.8, -0.1, -0.2}, u2[3] = {1.6, 0.9, 0.5};
  rc = osqp_update_lower_bound(w, l2);
  rc |= osqp_update_upper_bound(w, u2);
  rc |= osqp_solve(w);
  report("qp_bounds", w, rc);

  c_float xw[2] = {0.3, 0.5}, yw[3] = {-2.0, 0.0, 0.1};
  rc = osqp_warm_start_x(w, xw);
  rc |= osqp_solve(w);
  report("qp_warm_x", w, rc);
  rc = osqp_warm_start_y(w, yw);
  rc |= osqp_solve(w);
  report("qp_warm_y", w, rc);

  rc = osqp_update_rho(w, 0.7);
  rc |= osqp_warm_start(w, xw, yw);
  rc |= osqp_solve(w);
  report("qp_rho", w, rc);

  c_float P_new[3] = {5.0, 1.5, 1.0}, A_new[4] = {1.2, 1.5, 1.1, 0.8};
  rc = osqp_update_P_A(w, P_new, NULL, 3, A_new, NULL, 4);
  rc |= osqp_solve(w);
  report("qp_P_A", w, rc);

  /* iteration limit below the first termination check.  osqp_update_lin_cost comes first because only the
   * data updates reset info->status_val: without one the reference leaves the previous solve's "solved" in
   * place when the loop runs out (src/osqp.c:570 tests for OSQP_UNSOLVED), whatever the residuals are */
  rc = osqp_update_max_iter(w, 3);
  rc |= osqp_update_lin_cost(w, q);
  rc |= osqp_warm_start(w, xw, yw);
  rc |= osqp_solve(w);
  report("qp_max_iter", w, rc);
  printf("BADRHO rc %lld\n", (long long)osqp_update_rho(w, -1.0));
  printf("CLEANUP rc %lld\n", (long long)osqp_cleanup(w));
  return 0;
}

static int part_b(void) {
  double A[NX * NX], B[NX * NU];
  OSQPDataRLDL *d = (OSQPDataRLDL *)calloc(1, sizeof(OSQPDataRLDL));
  build_model(d, A, B);
  const c_int nmax = HNMAX * NA, mmax = HNMAX * NB + NT;
  c_float *q = (c_float *)calloc(nmax, sizeof(c_float));
  c_float *l = (c_float *)calloc(mmax, sizeof(c_float));
  c_float *u = (c_float *)calloc(mmax, sizeof(c_float));
  double x0[NX];
  for (int r = 0; r < NX; r++) x0[r] = 8.0 * urand();
  fill_bounds(A, x0, l, u);
  d->data.n = nmax;
  d->data.m = mmax;
  d->data.q = q;
  d->data.l = l;
  d->data.u = u;
  OSQPSettings *s = (OSQPSettings *)malloc(sizeof(OSQPSettings));
  osqp_set_default_settings(s);
  s->scaling = 0;
  s->adaptive_rho = 0;
  s->eps_abs = s->eps_rel = 1e-5;
  s->check_termination = 25;
  s->verbose = 0;
  s->polish = 0;
  s->warm_start = 1;
  OSQPWorkspace *w = NULL;
  const c_int head = 15, tail = 3;
  d->N = head + tail;
  c_int rc = osqp_setup_combine_recursive(&w, d, s, HNMAX, head, tail, NX, NU, NY, NT);
  printf("SETUP combine rc %lld\n", (long long)rc);
  if (rc) return 10 + (int)rc;
  rc = osqp_solve_combine_recursive(w);
  report("combine", w, rc);
  rc = osqp_update_Z_horizon(w, d, 2);
  rc |= osqp_solve_combine_recursive(w);
  report("combine_grow", w, rc);
  /* no tail shrink here: after osqp_update_Z_horizon(-1) the reference itself ends in "problem non convex"
   * with a garbage objective on this instance (as it does for -2 from 15+5), so there is nothing to compare
   * with; the product's shrink is checked against a fresh setup in tests/test_api_gpu.py */
  return 0;
}


/* part C — the reference's own small test problems (tests/unconstrained, tests/non_cvx,
 * tests/primal_dual_infeasibility generate_problem.py): m = 0, a non-convex P refused at setup with
 * sigma 1e-6 and detected in the loop with sigma 5, optimal / primal infeasible / dual infeasible /
 * primal infeasible variants of one QP */
static OSQPSettings *small_settings(void) {
  OSQPSettings *s = (OSQPSettings *)malloc(sizeof(OSQPSettings));
  osqp_set_default_settings(s);
  s->scaling = 0;
  s->adaptive_rho = 0;
  s->eps_abs = s->eps_rel = 1e-5;
  s->check_termination = 1;
  s->verbose = 0;
  s->polish = 0;
  s->warm_start = 0;
  s->max_iter = 2500;
  return s;
}

static int part_c(void) {
  c_int rc;
  OSQPWorkspace *w = NULL;
  { /* unconstrained: x = -P^-1 q, obj -19.2097520268 */
    c_float P_x[5] = {0.617022, 0.92032449, 0.20011437, 0.50233257, 0.34675589};
    c_int P_i[5] = {0, 1, 2, 3, 4}, P_p[6] = {0, 1, 2, 3, 4, 5};
    c_float q[5] = {-1.10593508, -1.65451545, -2.3634686, 1.13534535, -1.01701414};
    c_int A_p[6] = {0, 0, 0, 0, 0, 0};
    c_float dummy_x[1] = {0.0}, none[1] = {0.0};
    c_int dummy_i[1] = {0};
    OSQPData *data = (OSQPData *)calloc(1, sizeof(OSQPData));
    data->n = 5;
    data->m = 0;
    data->P = csc_matrix(5, 5, 5, P_x, P_i, P_p);
    data->q = q;
    data->A = csc_matrix(0, 5, 0, dummy_x, dummy_i, A_p);
    data->l = none;
    data->u = none;
    rc = osqp_setup(&w, data, small_settings());
    printf("SETUP unconstrained rc %lld\n", (long long)rc);
    if (rc) return 10 + (int)rc;
    rc = osqp_solve(w);
    report("unconstrained", w, rc);
    osqp_cleanup(w);
  }
  { /* non convex */
    c_float P_x[3] = {2.0, 5.0, 1.0};
    c_int P_i[3] = {0, 0, 1}, P_p[3] = {0, 1, 3};
    c_float q[2] = {3.0, 4.0};
    c_float A_x[8] = {-1.0, -1.0, 2.0, 3.0, -1.0, 3.0, 5.0, 4.0};
    c_int A_i[8] = {0, 2, 3, 4, 1, 2, 3, 4}, A_p[3] = {0, 4, 8};
    c_float l[5] = {-1e30, -1e30, -1e30, -1e30, -1e30}, u[5] = {0.0, 0.0, -15.0, 100.0, 80.0};
    OSQPData *data = (OSQPData *)calloc(1, sizeof(OSQPData));
    data->n = 2;
    data->m = 5;
    data->P = csc_matrix(2, 2, 3, P_x, P_i, P_p);
    data->q = q;
    data->A = csc_matrix(5, 2, 8, A_x, A_i, A_p);
    data->l = l;
    data->u = u;
    OSQPSettings *s = small_settings();
    s->sigma = 1e-6;
    w = NULL;
    rc = osqp_setup(&w, data, s);
    printf("SETUP non_cvx_small_sigma rc %lld\n", (long long)rc);
    s->sigma = 5.0;
    rc = osqp_setup(&w, data, s);
    printf("SETUP non_cvx rc %lld\n", (long long)rc);
    if (rc) return 10 + (int)rc;
    rc = osqp_solve(w);
    report("non_cvx", w, rc);
    osqp_cleanup(w);
  }
  for (int v = 1; v <= 4; v++) { /* primal / dual infeasibility */
    c_float P_x[2] = {1.0, 0.0};
    c_int P_i[2] = {0, 1}, P_p[3] = {0, 1, 2};
    c_float q[2] = {1.0, -1.0};
    c_float A12_x[4] = {1.0, 1.0, 1.0, 1.0}, A34_x[3] = {1.0, 1.0, 1.0};
    c_int A12_i[4] = {0, 1, 0, 2}, A12_p[3] = {0, 2, 4}, A34_i[3] = {0, 1, 2}, A34_p[3] = {0, 2, 3};
    c_float l[3] = {0.0, 1.0, 1.0};
    c_float uu[4][3] = {{5.0, 3.0, 3.0}, {0.0, 3.0, 3.0}, {2.0, 3.0, 1e30}, {0.0, 3.0, 1e30}};
    OSQPData *data = (OSQPData *)calloc(1, sizeof(OSQPData));
    data->n = 2;
    data->m = 3;
    data->P = csc_matrix(2, 2, 2, P_x, P_i, P_p);
    data->q = q;
    data->A = v <= 2 ? csc_matrix(3, 2, 4, A12_x, A12_i, A12_p) : csc_matrix(3, 2, 3, A34_x, A34_i, A34_p);
    data->l = l;
    data->u = uu[v - 1];
    w = NULL;
    rc = osqp_setup(&w, data, small_settings());
    printf("SETUP infeasibility%d rc %lld\n", v, (long long)rc);
    if (rc) return 10 + (int)rc;
    rc = osqp_solve(w);
    char tag[32];
    snprintf(tag, sizeof tag, "infeasibility%d", v);
    report(tag, w, rc);
    osqp_cleanup(w);
  }
  return 0;
}

/* part D */
static void report_polish(const char *tag, OSQPWorkspace *w) {
  printf("POLISH %s status_polish %lld\n", tag, (long long)w->info->status_polish);
}

static int part_d(void) {
  c_int rc;
  OSQPWorkspace *w = NULL;
  { /* basic_qp, defaults (scaling 10, adaptive rho, eps 1e-3) + polish */
    c_float P_x[3] = {4.0, 1.0, 2.0};
    c_int P_i[3] = {0, 0, 1}, P_p[3] = {0, 1, 3};
    c_float q[2] = {1.0, 1.0};
    c_float A_x[4] = {1.0, 1.0, 1.0, 1.0};
    c_int A_i[4] = {0, 1, 0, 2}, A_p[3] = {0, 2, 4};
    c_float l[3] = {1.0, 0.0, 0.0}, u[3] = {1.0, 0.7, 0.7};
    OSQPData *data = (OSQPData *)calloc(1, sizeof(OSQPData));
    data->n = 2;
    data->m = 3;
    data->P = csc_matrix(2, 2, 3, P_x, P_i, P_p);
    data->q = q;
    data->A = csc_matrix(3, 2, 4, A_x, A_i, A_p);
    data->l = l;
    data->u = u;
    OSQPSettings *s = (OSQPSettings *)malloc(sizeof(OSQPSettings));
    osqp_set_default_settings(s);
    s->adaptive_rho_interval = 25; /* the default 0 picks the interval from wall clock (src/osqp.c:459-485) */
    s->verbose = 0;
    s->polish = 1;
    rc = osqp_setup(&w, data, s);
    printf("SETUP defaults rc %lld\n", (long long)rc);
    if (rc) return 10 + (int)rc;
    rc = osqp_solve(w);
    report("defaults_polish", w, rc);
    report_polish("defaults_polish", w);
    rc = osqp_update_polish(w, 0);
    c_float q2[2] = {2.5, 3.2};
    rc |= osqp_update_lin_cost(w, q2);
    rc |= osqp_solve(w);
    report("defaults_no_polish", w, rc);
    report_polish("defaults_no_polish", w);
    osqp_cleanup(w);
  }
  { /* badly scaled QP: column / row norms from 1e-2 to 1e2, so D, E, c are far from 1 */
    c_float P_x[6] = {200.0, 3.0, 0.5, 0.02, 0.001, 0.004};
    c_int P_i[6] = {0, 0, 1, 1, 2, 3}, P_p[5] = {0, 1, 3, 5, 6};
    c_float q[4] = {-40.0, 0.3, -0.01, 0.002};
    c_float A_x[9] = {100.0, 1.0, 2.0, 0.5, 1.0, 0.01, 0.02, 0.05, 1.0};
    c_int A_i[9] = {0, 3, 0, 1, 4, 1, 2, 2, 4}, A_p[5] = {0, 2, 5, 7, 9};
    c_float l[5] = {-50.0, -1.0, -0.004, -1e30, 0.2}, u[5] = {80.0, 1.5, 0.004, 0.3, 0.2};
    OSQPData *data = (OSQPData *)calloc(1, sizeof(OSQPData));
    data->n = 4;
    data->m = 5;
    data->P = csc_matrix(4, 4, 6, P_x, P_i, P_p);
    data->q = q;
    data->A = csc_matrix(5, 4, 9, A_x, A_i, A_p);
    data->l = l;
    data->u = u;
    OSQPSettings *s = (OSQPSettings *)malloc(sizeof(OSQPSettings));
    osqp_set_default_settings(s);
    s->adaptive_rho_interval = 25;
    s->eps_abs = s->eps_rel = 1e-6;
    s->verbose = 0;
    w = NULL;
    rc = osqp_setup(&w, data, s);
    printf("SETUP badly_scaled rc %lld\n", (long long)rc);
    if (rc) return 10 + (int)rc;
    rc = osqp_solve(w);
    report("badly_scaled", w, rc);
    c_float l2[5] = {-50.0, -1.0, -0.004, -1e30, 0.1}, u2[5] = {80.0, 0.2, 0.004, 0.3, 0.1};
    rc = osqp_update_bounds(w, l2, u2);
    rc |= osqp_update_polish(w, 1);
    rc |= osqp_solve(w);
    report("badly_scaled_bounds_polish", w, rc);
    report_polish("badly_scaled_bounds_polish", w);
    c_float P_new[6] = {150.0, 2.0, 0.8, 0.01, 0.002, 0.004};
    rc = osqp_update_P(w, P_new, NULL, 6);
    rc |= osqp_update_polish(w, 0);
    rc |= osqp_solve(w);
    report("badly_scaled_P", w, rc);
    osqp_cleanup(w);
  }
  return 0;
}

int main(int argc, char **argv) {
  int which = argc > 1 ? atoi(argv[1]) : 15, rc = 0;
  if (which & 1) rc = part_a();
  if (!rc && (which & 2)) rc = part_b();
  if (!rc && (which & 4)) rc = part_c();
  if (!rc && (which & 8)) rc = part_d();
  printf("DONE rc %d\n", rc);
  return rc;
}
