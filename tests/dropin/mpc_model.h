/* tests/dropin/mpc_model.h — TEST INFRASTRUCTURE: the synthetic MPC instance shared by the user programs
 * of this directory (mpc_user.c, sharded_user.c).  Include after the solver's headers (the reference's
 * osqp.h + recursive_ldl.h + cs.h, or include/osqp_rldl_abi.h).
 *
 * nx=4, nu=2, N=20 (BASELINE config C1), box constraints on state and input (ny = nx+nu, C = I), terminal
 * box (nt = nx, AN = I), Nmax = N+1 and cost blocks stored as full upper triangles with explicit zeros
 * (where the reference's own recursive path is well defined). */
#ifndef RLDL_TEST_MPC_MODEL_H
#define RLDL_TEST_MPC_MODEL_H
#include <stdlib.h>

#define NX 4
#define NU 2
#define NA (NX + NU)      /* stage variables */
#define NY (NX + NU)      /* path rows */
#define NB (NX + NY)      /* stage constraint rows */
#define NT NX
#define HN 20
#define HNMAX 21
#define BIG 1e30

static unsigned long long lcg_state = 20261104ULL;
static double urand(void) { /* U(-0.5, 0.5) */
  lcg_state = lcg_state * 6364136223846793005ULL + 1442695040888963407ULL;
  return (double)(lcg_state >> 11) / 9007199254740992.0 - 0.5;
}

/* dense column-major (rows x cols) -> CSC: upper != 0 keeps the whole upper triangle with explicit zeros
 * (cost blocks, F4), upper == 0 keeps the non-zero entries only (constraint blocks) */
static csc *dense_to_csc(const double *M, int rows, int cols, int upper) {
  c_int *p = (c_int *)malloc(sizeof(c_int) * (cols + 1));
  c_int *i = (c_int *)malloc(sizeof(c_int) * rows * cols);
  c_float *x = (c_float *)malloc(sizeof(c_float) * rows * cols);
  c_int nnz = 0;
  for (int j = 0; j < cols; j++) {
    p[j] = nnz;
    for (int r = 0; r < rows; r++) {
      if (upper && r > j) break;
      if (!upper && M[j * rows + r] == 0.0) continue;
      i[nnz] = r;
      x[nnz] = M[j * rows + r];
      nnz++;
    }
  }
  p[cols] = nnz;
  return csc_matrix(rows, cols, nnz, x, i, p);
}

/* stage model into d (struct dimensions included: the reference reads them from the struct as well as from
 * the arguments of osqp_setup_recursive); A (NX x NX) and B (NX x NU) column-major for the caller's plant */
static void build_model(OSQPDataRLDL *d, double *A, double *B) {
  double Q[NX * NX], R[NU * NU]; /* column-major, like A (NX x NX) and B (NX x NU) */
  for (int j = 0; j < NX; j++)
    for (int r = 0; r < NX; r++) A[j * NX + r] = (r == j ? 0.9 : 0.0) + 0.1 * urand();
  for (int k = 0; k < NX * NU; k++) B[k] = urand();
  for (int j = 0; j < NX; j++)
    for (int r = 0; r < NX; r++) Q[j * NX + r] = (r == j) ? 1.0 : 0.01 / (1 + abs(r - j));
  for (int j = 0; j < NU; j++)
    for (int r = 0; r < NU; r++) R[j * NU + r] = (r == j) ? 0.1 : 0.01 / (1 + abs(r - j));

  /* stage blocks (recursive_ldl.h:30-36) */
  double Qi[NA * NA] = {0}, A0[NB * NU] = {0}, Ai[NB * NA] = {0}, Aij[NB * NA] = {0}, AN[NT * NX] = {0};
  for (int j = 0; j < NX; j++)
    for (int r = 0; r < NX; r++) Qi[j * NA + r] = Q[j * NX + r];
  for (int j = 0; j < NU; j++)
    for (int r = 0; r < NU; r++) Qi[(NX + j) * NA + NX + r] = R[j * NU + r];
  for (int j = 0; j < NU; j++) {
    A0[j * NB + NX + j] = 1.0;                                         /* u0 box rows */
    for (int r = 0; r < NX; r++) A0[j * NB + NY + r] = B[j * NX + r];  /* B u0 - x1 = -A x0 */
  }
  for (int j = 0; j < NA; j++) {
    Ai[j * NB + j] = 1.0; /* C = I */
    for (int r = 0; r < NX; r++) Ai[j * NB + NY + r] = (j < NX) ? A[j * NX + r] : B[(j - NX) * NX + r];
  }
  for (int j = 0; j < NX; j++) Aij[j * NB + NY + j] = -1.0;
  for (int j = 0; j < NX; j++) AN[j * NT + j] = 1.0;

  /* the reference reads the dimensions from the struct as well as from the arguments */
  d->Nmax = HNMAX; d->N = HN; d->nx = NX; d->nu = NU; d->ny = NY; d->nt = NT;
  d->Q0 = dense_to_csc(R, NU, NU, 1);
  d->Qi = dense_to_csc(Qi, NA, NA, 1);
  d->QN = dense_to_csc(Q, NX, NX, 1);
  d->A0 = dense_to_csc(A0, NB, NU, 0);
  d->Ai = dense_to_csc(Ai, NB, NA, 0);
  d->Aij = dense_to_csc(Aij, NB, NA, 0);
  d->AN = dense_to_csc(AN, NT, NX, 0);
}

/* l, u sized for Nmax (mmax = HNMAX*NB + NT), pre-filled stage-like up to Nmax, stage-0 dynamics rows = -A x0 */
static void fill_bounds(const double *A, const double *x0, c_float *l, c_float *u) {
  for (int k = 0; k < HNMAX; k++) {
    for (int r = 0; r < NX; r++) {
      l[k * NB + r] = k == 0 ? -BIG : -5.0;
      u[k * NB + r] = k == 0 ? BIG : 5.0;
    }
    for (int r = 0; r < NU; r++) {
      l[k * NB + NX + r] = -0.5;
      u[k * NB + NX + r] = 0.5;
    }
  }
  for (int r = 0; r < NT; r++) {
    l[HNMAX * NB + r] = -5.0;
    u[HNMAX * NB + r] = 5.0;
  }
  for (int r = 0; r < NX; r++) {
    double dyn = 0.0;
    for (int j = 0; j < NX; j++) dyn -= A[j * NX + r] * x0[j];
    l[NY + r] = u[NY + r] = dyn;
  }
}

#endif
