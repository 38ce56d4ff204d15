/*
 * osqp_rldl_abi.h — C ABI of librldl_b200.so
 *
 * B200-native (sm_100a, FP64) implementation of the recursive-LDL KKT path of
 * laperss/osqp-recursive-ldl and the OSQP ADMM iteration around it, behind the
 * reference's own C API.  A C program written against the reference's headers
 * (include/osqp.h, include/recursive_ldl.h, include/types.h, include/constants.h)
 * links against librldl_b200.so unchanged: every struct below is layout-identical
 * to the reference's default desktop build (DLONG, double, PRINTING, PROFILING,
 * EMBEDDED undefined — reference CMakeLists.txt:50-84) and every scalar entry
 * point keeps the reference's name, argument meaning and return convention.
 * Each declaration cites the reference interface it replaces.
 *
 * Scalar entry points act as a batch of one.  The throughput path is the additive
 * `rldl_batch_*` family at the bottom: one handle = many independent MPC QP
 * instances sharing one stage model, resident in HBM across calls.
 *
 * No CPU fallback exists: every entry point that computes needs a CUDA device and
 * returns an error (after printing why) when none is usable.
 */
#ifndef OSQP_RLDL_ABI_H
#define OSQP_RLDL_ABI_H

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built with -fvisibility=hidden */
#endif

/* ---- scalar types: reference include/glob_opts.h:73-84 (DLONG on, DFLOAT off) ---- */
typedef long long c_int;
typedef double    c_float;

/* ---- status / error codes: reference include/constants.h:18-51 ---- */
#define OSQP_DUAL_INFEASIBLE_INACCURATE   (4)
#define OSQP_PRIMAL_INFEASIBLE_INACCURATE (3)
#define OSQP_SOLVED_INACCURATE            (2)
#define OSQP_SOLVED                       (1)
#define OSQP_MAX_ITER_REACHED             (-2)
#define OSQP_PRIMAL_INFEASIBLE            (-3)
#define OSQP_DUAL_INFEASIBLE              (-4)
#define OSQP_SIGINT                       (-5)
#define OSQP_TIME_LIMIT_REACHED           (-6)
#define OSQP_NON_CVX                      (-7)
#define OSQP_UNSOLVED                     (-10)

enum linsys_solver_type { QDLDL_SOLVER, MKL_PARDISO_SOLVER };

enum osqp_error_type {
  OSQP_DATA_VALIDATION_ERROR = 1,
  OSQP_SETTINGS_VALIDATION_ERROR,
  OSQP_LINSYS_SOLVER_LOAD_ERROR,
  OSQP_LINSYS_SOLVER_INIT_ERROR,
  OSQP_NONCVX_ERROR,
  OSQP_MEM_ALLOC_ERROR,
  OSQP_WORKSPACE_NOT_INIT_ERROR,
};

/* defaults and limits: reference include/constants.h:59-114 */
#define OSQP_INFTY   ((c_float)1e30)
/* The reference defines OSQP_NAN as (c_float)0x7fc00000UL, i.e. the finite double
 * 2143289344.0, not an IEEE NaN (constants.h:95). "No solution" vectors are filled
 * with exactly that value, so we keep it. */
#define OSQP_NAN     ((c_float)0x7fc00000UL)
#define RHO_MIN              (1e-06)
#define RHO_MAX              (1e06)
#define RHO_EQ_OVER_RHO_INEQ (1e03)
#define RHO_TOL              (1e-04)
#define MIN_SCALING          (1e-04)

/* ---- data structures: reference include/types.h ---- */

/* types.h:21-29 — compressed-sparse-column matrix (nz == -1) or triplets */
typedef struct {
  c_int    nzmax;
  c_int    m;
  c_int    n;
  c_int   *p;
  c_int   *i;
  c_float *x;
  c_int    nz;
} csc;

typedef struct linsys_solver LinSysSolver; /* types.h:35, vtable at types.h:298-319 */
typedef struct OSQP_TIMER OSQPTimer;       /* types.h:40, opaque */

/* types.h:45-52 */
typedef struct {
  c_float  c;
  c_float *D;
  c_float *E;
  c_float  cinv;
  c_float *Dinv;
  c_float *Einv;
} OSQPScaling;

/* types.h:57-60 */
typedef struct {
  c_float *x;
  c_float *y;
} OSQPSolution;

/* types.h:66-91 */
typedef struct {
  c_int   iter;
  char    status[32];
  c_int   status_val;
  c_int   status_polish;
  c_float obj_val;
  c_float pri_res;
  c_float dua_res;
  c_float setup_time;
  c_float solve_time;
  c_float update_time;
  c_float polish_time;
  c_float run_time;
  c_int   rho_updates;
  c_float rho_estimate;
} OSQPInfo;

/* types.h:99-114 (polish is out of scope; the struct is kept for layout only) */
typedef struct {
  csc     *Ared;
  c_int    n_low;
  c_int    n_upp;
  c_int   *A_to_Alow;
  c_int   *A_to_Aupp;
  c_int   *Alow_to_A;
  c_int   *Aupp_to_A;
  c_float *x;
  c_float *z;
  c_float *y;
  c_float  obj_val;
  c_float  pri_res;
  c_float  dua_res;
} OSQPPolish;

/* types.h:125-133 */
typedef struct {
  c_int    n;
  c_int    m;
  csc     *P;
  csc     *A;
  c_float *q;
  c_float *l;
  c_float *u;
} OSQPData;

/* types.h:139-176 */
typedef struct {
  c_float rho;
  c_float sigma;
  c_int   scaling;
  c_int   adaptive_rho;
  c_int   adaptive_rho_interval;
  c_float adaptive_rho_tolerance;
  c_float adaptive_rho_fraction;
  c_int   max_iter;
  c_float eps_abs;
  c_float eps_rel;
  c_float eps_prim_inf;
  c_float eps_dual_inf;
  c_float alpha;
  enum linsys_solver_type linsys_solver;
  c_float delta;
  c_int   polish;
  c_int   polish_refine_iter;
  c_int   verbose;
  c_int   scaled_termination;
  c_int   check_termination;
  c_int   warm_start;
  c_float time_limit;
} OSQPSettings;

/* types.h:182-289.  Host mirror of one instance.  The iterate / residual vectors
 * are refreshed from the device after every solve, so code that peeks at
 * work->x, work->y, work->z, work->delta_x (dual-infeasibility certificate) or
 * work->delta_y (primal-infeasibility certificate) keeps working. */
typedef struct {
  OSQPData     *data;
  LinSysSolver *linsys_solver;
  OSQPPolish   *pol;
  c_float *rho_vec;
  c_float *rho_inv_vec;
  c_int   *constr_type;
  c_float *x;
  c_float *y;
  c_float *z;
  c_float *xz_tilde;
  c_float *x_prev;
  c_float *z_prev;
  c_float *Ax;
  c_float *Px;
  c_float *Aty;
  c_float *delta_y;
  c_float *Atdelta_y;
  c_float *delta_x;
  c_float *Pdelta_x;
  c_float *Adelta_x;
  c_float *D_temp;
  c_float *D_temp_A;
  c_float *E_temp;
  OSQPSettings *settings;
  OSQPScaling  *scaling;
  OSQPSolution *solution;
  OSQPInfo     *info;
  OSQPTimer    *timer;
  c_int first_run;
  c_int clear_update_time;
  c_int rho_update_from_solve;
  c_int summary_printed;
} OSQPWorkspace;

/* types.h:298-319 — the linear-system plug-in seam.  work->linsys_solver points
 * at a private object whose leading members are exactly these; `solve` runs one
 * KKT solve on the device (host vector in, host vector out), `update_rho_vec`
 * re-runs the stage recursion. */
struct linsys_solver {
  enum linsys_solver_type type;
  c_int (*solve)(LinSysSolver *self, c_float *b);
  void  (*free)(LinSysSolver *self);
  c_int (*update_matrices)(LinSysSolver *s, const csc *P, const csc *A);
  c_int (*update_rho_vec)(LinSysSolver *s, const c_float *rho_vec);
  c_int nthreads;
};

/* reference include/recursive_ldl.h:17-50.  Stage-block problem description.
 * Q0 nu x nu, Qi (nx+nu)^2, QN nx x nx upper triangles; A0 (ny+nx) x nu,
 * Ai (ny+nx) x (nx+nu), Aij = [0 0; -I 0], AN nt x nx.  The stage matrices are
 * borrowed for the workspace lifetime, like in the reference.  The members from
 * Amats on are the reference's CPU bookkeeping; this library leaves the ones it
 * does not need NULL. */
typedef struct {
  OSQPData data;
  c_int  Nmax;
  c_int  N;
  c_int  N0;
  c_int  N_recursive_temp;
  c_int  Nx;
  c_int  Nplus;
  c_int  nx;
  c_int  ny;
  c_int  nu;
  c_int  nt;
  c_int  nnz[2];
  csc   *A0;
  csc   *Ai;
  csc   *Aij;
  csc   *AN;
  csc   *Q0;
  csc   *Qi;
  csc   *QN;
  csc  **Amats;
  csc  **Qmats;
  csc  **X_even;
  double *Ytemp;
  csc   *Yhat;
  csc   *Uhat;
  csc   *Vhat;
  csc   *Lz;
  c_float *Dzinv;
  c_int *Pz;
  csc   *Ly;
  c_float *Dyinv;
  c_int *P0;
} OSQPDataRLDL;

/* =====================  reference include/osqp.h  ===================== */

/* osqp.h:32 */
void  osqp_set_default_settings(OSQPSettings *settings);
/* osqp.h:58 — generic QP (no stage structure): handled as a single dense stage on
 * the device; intended for the small problems of the reference's own tests. */
c_int osqp_setup(OSQPWorkspace **workp, const OSQPData *data, const OSQPSettings *settings);
/* osqp.h:78 */
c_int osqp_solve(OSQPWorkspace *work);
/* osqp.h:90 */
c_int osqp_cleanup(OSQPWorkspace *work);
/* osqp.h:114 */
c_int osqp_update_lin_cost(OSQPWorkspace *work, const c_float *q_new);
/* osqp.h:125 */
c_int osqp_update_bounds(OSQPWorkspace *work, const c_float *l_new, const c_float *u_new);
/* osqp.h:136 */
c_int osqp_update_lower_bound(OSQPWorkspace *work, const c_float *l_new);
/* osqp.h:146 */
c_int osqp_update_upper_bound(OSQPWorkspace *work, const c_float *u_new);
/* osqp.h:157 */
c_int osqp_warm_start(OSQPWorkspace *work, const c_float *x, const c_float *y);
/* osqp.h:168 */
c_int osqp_warm_start_x(OSQPWorkspace *work, const c_float *x);
/* osqp.h:178 */
c_int osqp_warm_start_y(OSQPWorkspace *work, const c_float *y);
/* osqp.h:200 */
c_int osqp_update_P(OSQPWorkspace *work, const c_float *Px_new, const c_int *Px_new_idx, c_int P_new_n);
/* osqp.h:221 */
c_int osqp_update_A(OSQPWorkspace *work, const c_float *Ax_new, const c_int *Ax_new_idx, c_int A_new_n);
/* osqp.h:250 */
c_int osqp_update_P_A(OSQPWorkspace *work, const c_float *Px_new, const c_int *Px_new_idx, c_int P_new_n,
                      const c_float *Ax_new, const c_int *Ax_new_idx, c_int A_new_n);
/* osqp.h:264 */
c_int osqp_update_rho(OSQPWorkspace *work, c_float rho_new);
/* osqp.h:284-419 — settings setters */
c_int osqp_update_max_iter(OSQPWorkspace *work, c_int max_iter_new);
c_int osqp_update_eps_abs(OSQPWorkspace *work, c_float eps_abs_new);
c_int osqp_update_eps_rel(OSQPWorkspace *work, c_float eps_rel_new);
c_int osqp_update_eps_prim_inf(OSQPWorkspace *work, c_float eps_prim_inf_new);
c_int osqp_update_eps_dual_inf(OSQPWorkspace *work, c_float eps_dual_inf_new);
c_int osqp_update_alpha(OSQPWorkspace *work, c_float alpha_new);
c_int osqp_update_warm_start(OSQPWorkspace *work, c_int warm_start_new);
c_int osqp_update_scaled_termination(OSQPWorkspace *work, c_int scaled_termination_new);
c_int osqp_update_check_termination(OSQPWorkspace *work, c_int check_termination_new);
c_int osqp_update_delta(OSQPWorkspace *work, c_float delta_new);
c_int osqp_update_polish(OSQPWorkspace *work, c_int polish_new);
c_int osqp_update_polish_refine_iter(OSQPWorkspace *work, c_int polish_refine_iter_new);
c_int osqp_update_verbose(OSQPWorkspace *work, c_int verbose_new);
c_int osqp_update_time_limit(OSQPWorkspace *work, c_float time_limit_new);

/* =====================  reference include/recursive_ldl.h:52-74  ===================== */

/* recursive_ldl.h:66 (src/recursive_ldl.c:2018) — data->data.{n,m,q,l,u} sized for Nmax */
c_int osqp_setup_recursive(OSQPWorkspace **workp, OSQPDataRLDL *data, const OSQPSettings *settings,
                           c_int Nmax, c_int N, c_int nx, c_int nu, c_int ny, c_int nt);
/* recursive_ldl.h:64 (src/recursive_ldl.c:2867) */
c_int osqp_solve_recursive(OSQPWorkspace *work);
/* recursive_ldl.h:60 (src/recursive_ldl.c:1973) — change the horizon to N, re-running
 * only the tail of the stage recursion */
c_int osqp_update_recursive(OSQPWorkspace *work, OSQPDataRLDL *data_rldl, c_int N);
/* recursive_ldl.h:67 (src/recursive_ldl.c:2359) — head of N stages + tail of Nplus stages */
c_int osqp_setup_combine_recursive(OSQPWorkspace **workp, OSQPDataRLDL *data, const OSQPSettings *settings,
                                   c_int Nmax, c_int N, c_int Nplus, c_int nx, c_int nu, c_int ny, c_int nt);
/* recursive_ldl.h:65 (src/recursive_ldl.c:2862) */
c_int osqp_solve_combine_recursive(OSQPWorkspace *work);
/* recursive_ldl.h:62 (src/recursive_ldl.c:2761) — grow/shrink the tail by deltaN stages */
c_int osqp_update_Z_horizon(OSQPWorkspace *work, OSQPDataRLDL *data_rldl, c_int deltaN);
/* recursive_ldl.h:61 (src/recursive_ldl.c:2757) — empty body in the reference; kept as a no-op */
c_int osqp_update_X_horizon(OSQPWorkspace *work, OSQPDataRLDL *data_rldl, c_int deltaN);
/* recursive_ldl.h:68 (src/recursive_ldl.c:2319) — no-op in the reference; kept as a no-op */
c_int osqp_allocate_combine_recursive(OSQPWorkspace **workp, OSQPDataRLDL *data, const OSQPSettings *settings,
                                      c_int Nmax, c_int nx, c_int nu, c_int ny, c_int nt);
/* recursive_ldl.h:73 (src/recursive_ldl.c:119) — overwrite l,u on rows [start,stop) */
c_int osqp_partial_update_bounds(OSQPWorkspace *work, c_int start, c_int stop,
                                 const c_float *l_new, const c_float *u_new);
/* recursive_ldl.h:78 (src/recursive_ldl.c:1340) — KKT dimensions and number of stored
 * factor entries (here: doubles of the dense stage-block factor) */
void  get_L_dimensions(OSQPWorkspace *work, c_int *n, c_int *m, c_int *nnz);
/* recursive_ldl.h:52 (src/recursive_ldl.c:15) */
void  cleanup_rldl(OSQPDataRLDL *data);

/* =====================  reference include/cs.h, include/util.h  ===================== */
/* Host-side data-format helpers a user program builds its inputs with (memory from malloc/calloc,
 * i.e. the desktop build's c_malloc/c_calloc, glob_opts.h:64-66).  No numerical work. */

/* util.h:19 */
const char *osqp_version(void);
/* cs.h:26 — wraps caller-owned x,i,p (no copies) */
csc  *csc_matrix(c_int m, c_int n, c_int nzmax, c_float *x, c_int *i, c_int *p);
/* cs.h:44 — allocates p,i (and x when values != 0); triplet form when triplet != 0 */
csc  *csc_spalloc(c_int m, c_int n, c_int nzmax, c_int values, c_int triplet);
/* cs.h:51 — frees p,i,x and the struct */
void  csc_spfree(csc *A);
/* cs.h:70 */
csc  *copy_csc_mat(const csc *A);
/* cs.h:76 — copy into an already allocated B */
void  prea_copy_csc_mat(const csc *A, csc *B);
/* cs.h:92 — triplet form to CSC; TtoC (optional) receives the position of every triplet */
csc  *triplet_to_csc(const csc *T, c_int *TtoC);
/* cs.h:113 — upper triangle (with diagonal) of a square CSC matrix, newly allocated */
csc  *csc_to_triu(csc *M);

/* =====================  additive batch API (new)  ===================== */
/* One handle = `batch` independent QP instances that share the stage model in
 * `model` (dims + Q0,Qi,QN,A0,Ai,Aij,AN) and differ in q, l, u (and optionally
 * rho).  All instance data and the per-instance stage factors live in HBM for the
 * lifetime of the handle; host arrays are row-major [batch][nmax] / [batch][mmax]
 * with nmax = Nmax*(nx+nu), mmax = Nmax*(nx+ny)+nt, exactly the per-instance
 * vectors the scalar API takes, one row per instance.  Functions return 0 on
 * success and an osqp_error_type (or 1, like the reference's update calls) on
 * invalid input.  Every function acts on the CUDA device that was current when
 * the handle was created. */
typedef struct RLDLBatch RLDLBatch;

c_int rldl_batch_setup(RLDLBatch **out, const OSQPDataRLDL *model, const OSQPSettings *settings,
                       c_int Nmax, c_int N, c_int nx, c_int nu, c_int ny, c_int nt, c_int batch,
                       const c_float *q, const c_float *l, const c_float *u,
                       const c_float *rho /* per-instance initial rho, or NULL */);
/* batched osqp_solve_recursive: per-instance ADMM with on-device termination,
 * infeasibility detection and adaptive-rho refactorisation */
c_int rldl_batch_solve(RLDLBatch *b);
/* batched osqp_update_bounds / osqp_partial_update_bounds / osqp_update_lin_cost */
c_int rldl_batch_update_bounds(RLDLBatch *b, const c_float *l, const c_float *u);
c_int rldl_batch_partial_update_bounds(RLDLBatch *b, c_int start, c_int stop,
                                       const c_float *l /* [batch][stop-start] */,
                                       const c_float *u);
c_int rldl_batch_update_lin_cost(RLDLBatch *b, const c_float *q);
/* rldl_batch_update_bounds(l, u) + rldl_batch_update_lin_cost(q) + rldl_batch_refactor() in one call (the per-step
 * update of a controller that re-poses every instance: reference osqp_update_bounds osqp.c:784, osqp_update_lin_cost
 * :752, then the full stage recursion), pipelined: the arrays are uploaded in chunks on a second stream while the
 * typing / record pass and the recursion of the previous chunk run.  Pinned host arrays make the copies
 * asynchronous.  l <= u is validated on the device; a refused update (return 1) leaves the bounds as they were
 * (the linear cost of the call has been taken). */
c_int rldl_batch_update_data_refactor(RLDLBatch *b, const c_float *q, const c_float *l, const c_float *u);
/* batched osqp_warm_start (x or y may be NULL: osqp_warm_start_y / _x) */
c_int rldl_batch_warm_start(RLDLBatch *b, const c_float *x, const c_float *y);
/* receding-horizon warm start without a host round trip: the resident (x, y) of the last solve are
 * moved one stage towards the origin on the device (x[0:n-a] <- x[a:n], y[0:m-b] <- y[b:m], tails kept —
 * what a reference user does with memmove before osqp_warm_start, osqp.c:929-994); z = A x is formed
 * at the next solve, after any horizon / bound update made in between */
c_int rldl_batch_warm_start_shift(RLDLBatch *b);
/* batched osqp_update_recursive */
c_int rldl_batch_update_horizon(RLDLBatch *b, c_int N);
/* batched osqp_update_rho (per-instance value) */
c_int rldl_batch_update_rho(RLDLBatch *b, const c_float *rho);
/* batched osqp_update_P_A (src/osqp.c:1166-1266) for time-varying / re-linearised MPC: new values for the
 * shared stage matrices Q0,Qi,QN,A0,Ai,Aij,AN (same dimensions as at setup; sparsity patterns may differ),
 * followed by the full stage recursion of every instance.  Resets the per-instance info like the
 * reference's update (status OSQP_UNSOLVED); iterates are kept for warm starts. */
c_int rldl_batch_update_model(RLDLBatch *b, const OSQPDataRLDL *model);
/* re-run the full stage recursion (what osqp_setup_recursive does after allocation) */
c_int rldl_batch_refactor(RLDLBatch *b);
/* one KKT solve per instance: rhs [batch][n+m] (current sizes) is overwritten with
 * [x_tilde ; z_tilde] — the batched form of linsys_solver->solve */
c_int rldl_batch_kkt_solve(RLDLBatch *b, c_float *rhs);
/* Allocation-free KKT solve.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same result as rldl_batch_kkt_solve — rhs [batch][n+m] in, sol [batch][n+m] = [x_tilde ; z_tilde] out (host
 * arrays, may alias) — but the right-hand sides are staged in a fixed 4 MiB device array inside the library
 * instead of a cudaMalloc'ed buffer: the call path contains no cudaMalloc / cudaFree.  Limit: batch * (n + m) <=
 * rldl_batch_max_static_doubles() (524 288); larger calls return 1 with a message and do nothing.  Every call on a
 * device shares the one array (it keeps the last call's contents; nothing in it needs clearing): calls that overlap
 * on the same GPU — other handles, other host threads — race and must be serialised by the caller.  Every GPU that
 * loads the library holds the array until exit. */
c_int rldl_batch_kkt_solve_static(RLDLBatch *b, const c_float *rhs, c_float *sol);
c_int rldl_batch_max_static_doubles(void);
/* Fused solve-and-deliver: rldl_batch_solve + rldl_batch_get_solution + rldl_batch_get_info in one call (array
 * layout as in those two; status_val is required, the others may be NULL).  When every array given is PINNED host
 * memory (cudaHostAlloc / cudaHostRegister), the kernels write each instance's x, y (OSQP_NAN for the statuses
 * without a solution, auxil.c:539-542) and info to the host the moment the instance finishes, so the device-to-host
 * transfer runs behind the rest of the solve instead of after it; pageable arrays, or settings.polish = 1 (polish
 * revises the results after the ADMM kernels), take the ordinary copies after the solve.  Only entries [0, n) of
 * each row of x and [0, m) of y are written. */
c_int rldl_batch_solve_to_host(RLDLBatch *b, c_float *x, c_float *y, c_int *iter, c_int *status_val,
                               c_float *obj_val, c_float *pri_res, c_float *dua_res, c_int *rho_updates,
                               c_float *rho_estimate);
/* gather: x [batch][nmax], y [batch][mmax] (any pointer may be NULL) */
c_int rldl_batch_get_solution(RLDLBatch *b, c_float *x, c_float *y);
c_int rldl_batch_get_info(RLDLBatch *b, c_int *iter, c_int *status_val, c_float *obj_val,
                          c_float *pri_res, c_float *dua_res, c_int *rho_updates,
                          c_float *rho_estimate);
/* number of instances whose last stage recursion met a non-positive pivot (the condition on which the
 * reference's osqp_setup fails with OSQP_NONCVX_ERROR); flag (optional, [batch]) receives 0, or 1 + the
 * stage of the first such pivot.  The scalar setups return OSQP_NONCVX_ERROR themselves; a batch keeps its
 * healthy instances and reports the others as OSQP_NON_CVX at the next solve.  -1 on a device error. */
c_int rldl_batch_failed_factorisations(RLDLBatch *b, c_int *flag);
/* polish outcome of the last solve per instance (settings.polish = 1; reference src/polish.c:212-350,
 * info->status_polish): 1 polished point accepted (x, y, obj_val, pri_res, dua_res are the polished ones),
 * -1 polish rejected or failed, 0 not attempted (the instance did not end OSQP_SOLVED, or polish is off) */
c_int rldl_batch_get_polish(RLDLBatch *b, c_int *status_polish);
/* Ruiz equilibration of every instance (settings.scaling > 0; reference scale_data, src/scaling.c:44-156, and
 * the OSQPScaling members of include/types.h:44-52): D [batch x nmax], E [batch x mmax], c [batch].
 * Returns 1 when scaling is off.  x, y, obj, residuals and every other output are always unscaled. */
c_int rldl_batch_get_scaling(RLDLBatch *b, c_float *D, c_float *E, c_float *c);
/* settings of the whole batch (validated like the reference's osqp_update_* setters) */
c_int rldl_batch_update_settings(RLDLBatch *b, const OSQPSettings *settings);
void  rldl_batch_dims(const RLDLBatch *b, c_int *n, c_int *m, c_int *N, c_int *batch);
/* device time of the last factorisation / ADMM launch in milliseconds (CUDA events) */
void  rldl_batch_last_times(const RLDLBatch *b, double *factor_ms, double *solve_ms);
/* the streaming ADMM kernel of the last solve alone: its device time (CUDA events around the launch) and the ADMM
 * iterations, summed over instances, it executed itself (the rest of the solve's iterations ran in the latency-mode
 * kernel) — the pair the kernel's roofline figure is computed from; 0 / 0 if the solve did not use that kernel */
void  rldl_batch_stream_stats(const RLDLBatch *b, double *stream_ms, double *stream_instance_iterations);
/* algorithmic bytes per instance and ADMM iteration, algorithmic bytes of
 * the last factorisation launch, and kernels launched so far on this handle */
void  rldl_batch_counters(const RLDLBatch *b, double *admm_bytes, double *factor_bytes, c_int *launches);
/* ADMM kernel selection: 0 = automatic (the TMA-streaming kernel when the stage shapes allow
 * it, nx+nu <= 32, and adaptive_rho is off), 1 = always the generic runtime-shape kernel.
 * Both compute the same iteration; the switch exists for cross-checking and profiling.
 * which = 2 changes nothing and returns 0 iff the handle's stage shape has specialised kernels. */
c_int rldl_batch_select_kernel(RLDLBatch *b, c_int which);
/* Launch knobs of this handle (per handle, not process-global; the defaults are the measured optima on B200
 * and no setting can change results beyond rounding — the tests use them to steer which kernel finishes a solve):
 *   "lat_below"   unfinished instances at which the streaming kernel hands its tail to the latency-mode kernel
 *                 (default: 2048 for batches of 4096 and more, else 1024; 0 = streaming kernel only; >= batch =
 *                 latency-mode kernel only)
 *   "ring_depth"  stage records in flight per lane group of the streaming kernel (default 2)
 *   "sched_s0"    iterations a fresh instance runs before it may be pre-empted (default check_termination;
 *                 0 = device priority scheduler off)
 *   "ctas_per_sm" resident CTAs per SM of the streaming kernel (default 4, 3 without the hand-off)
 *   "l2_hints"    bit 0: stage records are streamed with an L2 evict-first policy (default 1)
 *   "lat_lane"    latency-mode CTAs that run BESIDE the streaming kernel towards the end of a launch and take over
 *                 its most-unconverged instances, whose chains would otherwise bound the tail (default 8, 0 = off)
 *   "lat_retire_below"  unfinished instances at which the streaming CTAs on the lane's SMs retire
 *                 (default 6 x lat_below)
 *   "lat_levels"  levels of the block tree under the sequential chain of the latency-mode kernel (default 1 = the
 *                 chain runs over pairs of stages; k = over runs of 2^k stages, with k - 1 more hops behind it;
 *                 1..6; measured neutral on B200)
 *   "lat_lockstep"  1 (default): in the latency-mode kernel a warp owns an item and its two lane groups take the two
 *                 stages of the item side by side (shapes with two or four lane groups per warp, nx + nu <= 16;
 *                 wider shapes ignore it); 0: a lane group per item, its stages one after the other
 * value = -1 restores the default.  Returns 0, or 1 for an unknown name / invalid value. */
c_int rldl_batch_set_tuning(RLDLBatch *b, const char *name, c_int value);
/* statistics of the concurrent latency-mode lane in the last streaming launch (6 values): instances handed over,
 * lane CTAs that left at once (resident before the streaming kernel had started), instances a lane CTA took from
 * the hand-over ring / from the priority buckets, instances suspended by retiring SMs, lane CTA exits */
c_int rldl_batch_lane_stats(RLDLBatch *b, c_int *stats);
/* finish-time trace of the last solve (after rldl_batch_set_tuning(b, "trace", 1)): per instance, milliseconds
 * from the start of the solve to the moment it finished + 1e6 x the engine that finished it (1 streaming kernel,
 * 2 latency-mode kernel).  Returns 1 if tracing was never switched on. */
c_int rldl_batch_get_trace(RLDLBatch *b, c_float *t);
/* diagnostic: raw copy of the first n ints of the device scheduler state (csrc/rldl_device.h, kSched*), readable
 * from another host thread while a solve is running */
c_int rldl_batch_debug_sched(RLDLBatch *b, int *out, c_int n);
/* the CUDA stream (cudaStream_t) every launch and copy of this handle is issued on, so a
 * caller can bracket calls with its own events on the launching stream */
void *rldl_batch_stream(const RLDLBatch *b);
void  rldl_batch_free(RLDLBatch *b);

/* ---------------------------------------------------------------------------------------------------------
 * Multi-device handle (SURVEY §8e inside the library).  No counterpart in the reference, which is single-threaded
 * with one workspace per problem (its users parallelise over workspaces, SURVEY §8b "Threading").
 * The batch is cut into `nshards` contiguous shards (0 = one per device), shard s on devices[s % ndev]
 * (devices = NULL: every visible device); each call below runs its per-shard rldl_batch_* calls on one host thread
 * per shard and returns when all have.  Instances share nothing: there is no collective and no peer-to-peer
 * traffic, every shard writes its own rows of the caller's arrays (layouts as in the rldl_batch_* calls, whole
 * batch).  Pinned host arrays make rldl_multi_solve_to_host deliver from inside the kernels on every device.
 * --------------------------------------------------------------------------------------------------------- */
typedef struct RLDLMulti RLDLMulti;
c_int rldl_multi_setup(RLDLMulti **out, const c_int *devices, c_int ndev, c_int nshards, const OSQPDataRLDL *model,
                       const OSQPSettings *settings, c_int Nmax, c_int N, c_int nx, c_int nu, c_int ny, c_int nt,
                       c_int batch, const c_float *q, const c_float *l, const c_float *u, const c_float *rho);
/* number of shards; first[s] = first instance of shard s, device[s] = its device (either may be NULL) */
c_int rldl_multi_shards(const RLDLMulti *m, c_int *first, c_int *device);
/* the shard's own handle, for the rldl_batch_* calls that have no multi form (tuning, diagnostics) */
RLDLBatch *rldl_multi_shard_handle(RLDLMulti *m, c_int shard);
c_int rldl_multi_solve(RLDLMulti *m);
c_int rldl_multi_solve_to_host(RLDLMulti *m, c_float *x, c_float *y, c_int *iter, c_int *status_val,
                               c_float *obj_val, c_float *pri_res, c_float *dua_res, c_int *rho_updates,
                               c_float *rho_estimate);
c_int rldl_multi_get_solution(RLDLMulti *m, c_float *x, c_float *y);
c_int rldl_multi_get_info(RLDLMulti *m, c_int *iter, c_int *status_val, c_float *obj_val, c_float *pri_res,
                          c_float *dua_res, c_int *rho_updates, c_float *rho_estimate);
c_int rldl_multi_update_bounds(RLDLMulti *m, const c_float *l, const c_float *u);
c_int rldl_multi_partial_update_bounds(RLDLMulti *m, c_int start, c_int stop, const c_float *l, const c_float *u);
c_int rldl_multi_update_lin_cost(RLDLMulti *m, const c_float *q);
c_int rldl_multi_warm_start(RLDLMulti *m, const c_float *x, const c_float *y);
c_int rldl_multi_warm_start_shift(RLDLMulti *m);
c_int rldl_multi_update_horizon(RLDLMulti *m, c_int N);
c_int rldl_multi_update_rho(RLDLMulti *m, const c_float *rho);
c_int rldl_multi_refactor(RLDLMulti *m);
c_int rldl_multi_update_settings(RLDLMulti *m, const OSQPSettings *settings);
void  rldl_multi_free(RLDLMulti *m);

/* device plumbing */
c_int rldl_device_count(void);
c_int rldl_set_device(c_int device);
const char *rldl_version(void);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* OSQP_RLDL_ABI_H */
