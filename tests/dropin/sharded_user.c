/* tests/dropin/sharded_user.c — TEST INFRASTRUCTURE.
 *
 * The multi-GPU row of the scope table from a C host: independent MPC instances are cut
 * into contiguous shards, one host thread per shard, thread t on device t % G with its own
 * RLDLBatch handle; no collective anywhere, the "gather" is every thread writing its rows of the
 * caller's x / y / info arrays.  The same batch is solved once more through a single handle on
 * device 0 and the two must agree: identical status and iteration count per instance, x and y to
 * 1e-9 (which ADMM kernel finishes an instance depends on the shard size; they agree to rounding).
 * A closed-loop step (rldl_batch_partial_update_bounds + rldl_batch_warm_start_shift + solve) follows.
 * With more threads than devices several handles share a device and are driven concurrently,
 * which is the thread-safety claim of INTEGRATION.md §2.
 *
 *   sharded_user [batch=2048] [threads=max(2, devices)]     exit 0 = agreement
 */
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <string.h>
#include "osqp_rldl_abi.h"
#include "mpc_model.h"

typedef struct {
  int tid, device;
  c_int i0, i1; /* shard = instances [i0, i1) */
  const OSQPDataRLDL *model;
  const OSQPSettings *settings;
  const c_float *q, *l, *u;     /* whole batch, row-major */
  const c_float *dyn1;          /* [batch][NX] stage-0 dynamics rows of the second solve */
  c_float *x[2], *y[2];         /* outputs of solve 1 and 2, whole batch */
  c_int *iter[2], *status[2];
  double factor_ms, solve_ms[2];
  c_int rc;
} Shard;

static const c_int nmax = HNMAX * NA, mmax = HNMAX * NB + NT;

static void *run_shard(void *arg) {
  Shard *s = (Shard *)arg;
  const c_int nb = s->i1 - s->i0;
  RLDLBatch *b = NULL;
  s->rc = rldl_set_device(s->device);
  if (s->rc) return NULL;
  s->rc = rldl_batch_setup(&b, s->model, s->settings, HNMAX, HN, NX, NU, NY, NT, nb, s->q + s->i0 * nmax,
                           s->l + s->i0 * mmax, s->u + s->i0 * mmax, NULL);
  if (s->rc) return NULL;
  for (int pass = 0; pass < 2 && !s->rc; pass++) {
    if (pass == 1) {
      s->rc |= rldl_batch_partial_update_bounds(b, NY, NY + NX, s->dyn1 + s->i0 * NX, s->dyn1 + s->i0 * NX);
      s->rc |= rldl_batch_warm_start_shift(b);
    }
    s->rc |= rldl_batch_solve(b);
    s->rc |= rldl_batch_get_solution(b, s->x[pass] + s->i0 * nmax, s->y[pass] + s->i0 * mmax);
    s->rc |= rldl_batch_get_info(b, s->iter[pass] + s->i0, s->status[pass] + s->i0, NULL, NULL, NULL, NULL, NULL);
    double f, t;
    rldl_batch_last_times(b, &f, &t);
    if (pass == 0) s->factor_ms = f;
    s->solve_ms[pass] = t;
  }
  rldl_batch_free(b);
  return NULL;
}

int main(int argc, char **argv) {
  const c_int batch = argc > 1 ? atoll(argv[1]) : 2048;
  const int G = (int)rldl_device_count();
  if (G < 1) {
    fprintf(stderr, "sharded_user: no CUDA device (the library has no CPU path)\n");
    return 13;
  }
  int T = argc > 2 ? atoi(argv[2]) : (G > 2 ? G : 2);
  if (T > batch) T = (int)batch;

  double A[NX * NX], B[NX * NU];
  OSQPDataRLDL *d = (OSQPDataRLDL *)calloc(1, sizeof(OSQPDataRLDL));
  build_model(d, A, B);
  OSQPSettings st;
  osqp_set_default_settings(&st);
  st.scaling = 0;
  st.adaptive_rho = 0;
  st.eps_abs = st.eps_rel = 1e-5;
  st.check_termination = 25;
  st.verbose = 0;
  st.warm_start = 1;

  c_float *q = (c_float *)calloc(batch * nmax, sizeof(c_float));
  c_float *l = (c_float *)calloc(batch * mmax, sizeof(c_float));
  c_float *u = (c_float *)calloc(batch * mmax, sizeof(c_float));
  double *x0 = (double *)malloc(sizeof(double) * batch * NX);
  for (c_int i = 0; i < batch; i++) {
    for (int r = 0; r < NX; r++) x0[i * NX + r] = 8.0 * urand();
    fill_bounds(A, x0 + i * NX, l + i * mmax, u + i * mmax);
  }

  /* outputs: [0] = one handle on device 0, [1] = sharded over T threads; two solves each */
  c_float *x[2][2], *y[2][2];
  c_int *iter[2][2], *status[2][2];
  for (int w = 0; w < 2; w++)
    for (int p = 0; p < 2; p++) {
      x[w][p] = (c_float *)calloc(batch * nmax, sizeof(c_float));
      y[w][p] = (c_float *)calloc(batch * mmax, sizeof(c_float));
      iter[w][p] = (c_int *)calloc(batch, sizeof(c_int));
      status[w][p] = (c_int *)calloc(batch, sizeof(c_int));
    }
  c_float *dyn1 = (c_float *)malloc(sizeof(c_float) * batch * NX);

  /* ---- whole batch through one handle; the plant step for the second solve comes from its u0 ---- */
  {
    RLDLBatch *b = NULL;
    c_int rc = rldl_set_device(0);
    rc |= rldl_batch_setup(&b, d, &st, HNMAX, HN, NX, NU, NY, NT, batch, q, l, u, NULL);
    if (rc) { fprintf(stderr, "setup failed rc %lld\n", (long long)rc); return 10 + (int)rc; }
    rc |= rldl_batch_solve(b);
    rc |= rldl_batch_get_solution(b, x[0][0], y[0][0]);
    rc |= rldl_batch_get_info(b, iter[0][0], status[0][0], NULL, NULL, NULL, NULL, NULL);
    for (c_int i = 0; i < batch; i++) { /* plant: x0 <- A x0 + B u0;  stage-0 dynamics rows = -A x0 */
      double xn[NX];
      for (int r = 0; r < NX; r++) {
        xn[r] = 0.0;
        for (int j = 0; j < NX; j++) xn[r] += A[j * NX + r] * x0[i * NX + j];
        for (int j = 0; j < NU; j++) xn[r] += B[j * NX + r] * x[0][0][i * nmax + j];
      }
      for (int r = 0; r < NX; r++) {
        dyn1[i * NX + r] = 0.0;
        for (int j = 0; j < NX; j++) dyn1[i * NX + r] -= A[j * NX + r] * xn[j];
      }
    }
    rc |= rldl_batch_partial_update_bounds(b, NY, NY + NX, dyn1, dyn1);
    rc |= rldl_batch_warm_start_shift(b);
    rc |= rldl_batch_solve(b);
    rc |= rldl_batch_get_solution(b, x[0][1], y[0][1]);
    rc |= rldl_batch_get_info(b, iter[0][1], status[0][1], NULL, NULL, NULL, NULL, NULL);
    rldl_batch_free(b);
    if (rc) { fprintf(stderr, "single-handle run failed rc %lld\n", (long long)rc); return 2; }
  }

  /* ---- the same batch, sharded: T threads, thread t on device t % G ---- */
  Shard *sh = (Shard *)calloc(T, sizeof(Shard));
  pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * T);
  for (int t = 0; t < T; t++) {
    Shard s = {t, t % G, batch * t / T, batch * (t + 1) / T, d, &st, q, l, u, dyn1, {x[1][0], x[1][1]},
               {y[1][0], y[1][1]}, {iter[1][0], iter[1][1]}, {status[1][0], status[1][1]}, 0, {0, 0}, 0};
    sh[t] = s;
    pthread_create(&th[t], NULL, run_shard, &sh[t]);
  }
  int bad = 0;
  for (int t = 0; t < T; t++) {
    pthread_join(th[t], NULL);
    printf("SHARD %d device %d instances [%lld,%lld) rc %lld factor %.3f ms solve %.3f / %.3f ms\n", t, sh[t].device,
           (long long)sh[t].i0, (long long)sh[t].i1, (long long)sh[t].rc, sh[t].factor_ms, sh[t].solve_ms[0],
           sh[t].solve_ms[1]);
    bad |= sh[t].rc != 0;
  }

  /* ---- agreement ---- */
  for (int p = 0; p < 2; p++) {
    c_int diff_status = 0, diff_iter = 0, solved = 0;
    double ex = 0.0, ey = 0.0, nx_ = 0.0, ny_ = 0.0;
    for (c_int i = 0; i < batch; i++) {
      diff_status += status[0][p][i] != status[1][p][i];
      diff_iter += iter[0][p][i] != iter[1][p][i];
      solved += status[0][p][i] == OSQP_SOLVED;
    }
    const c_int n = HN * NA, m = HN * NB + NT;
    for (c_int i = 0; i < batch; i++) {
      for (c_int k = 0; k < n; k++) {
        ex = fmax(ex, fabs(x[0][p][i * nmax + k] - x[1][p][i * nmax + k]));
        nx_ = fmax(nx_, fabs(x[0][p][i * nmax + k]));
      }
      for (c_int k = 0; k < m; k++) {
        ey = fmax(ey, fabs(y[0][p][i * mmax + k] - y[1][p][i * mmax + k]));
        ny_ = fmax(ny_, fabs(y[0][p][i * mmax + k]));
      }
    }
    printf("AGREE solve %d batch %lld threads %d devices %d solved %lld status_diff %lld iter_diff %lld x_rel %.3g y_rel %.3g\n",
           p + 1, (long long)batch, T, G, (long long)solved, (long long)diff_status, (long long)diff_iter,
           ex / fmax(nx_, 1e-12), ey / fmax(ny_, 1e-12));
    for (int t = 0; t < T && (diff_status || diff_iter); t++) { /* which shards, and what they hold */
      c_int ds = 0, di = 0;
      for (c_int i = sh[t].i0; i < sh[t].i1; i++) {
        ds += status[0][p][i] != status[1][p][i];
        di += iter[0][p][i] != iter[1][p][i];
      }
      if (ds || di)
        printf("  shard %d: %lld status / %lld iteration mismatches; first instance: status %lld vs %lld, iter %lld vs %lld\n",
               t, (long long)ds, (long long)di, (long long)status[1][p][sh[t].i0], (long long)status[0][p][sh[t].i0],
               (long long)iter[1][p][sh[t].i0], (long long)iter[0][p][sh[t].i0]);
    }
    /* at least some instances must be plain solves, none may disagree */
    bad |= diff_status != 0 || diff_iter != 0 || ex > 1e-9 * fmax(nx_, 1e-12) || ey > 1e-9 * fmax(ny_, 1e-12) ||
           solved == 0;
  }
  printf("%s\n", bad ? "SHARDED MISMATCH" : "SHARDED OK");
  return bad ? 1 : 0;
}
