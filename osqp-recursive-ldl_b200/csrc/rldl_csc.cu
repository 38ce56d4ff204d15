// rldl_csc.cu — the host-side CSC helpers of the reference's public cs.h that a user program calls to
// build the inputs of osqp_setup / osqp_setup_recursive (include/cs.h:26-120, src/cs.c) and
// osqp_version (include/util.h:19).  Pure data-format code (allocation, copies, triplet compression,
// upper-triangle extraction): no numerical work of the solve path lives here.  Memory comes from
// malloc/calloc/free, which is what c_malloc/c_calloc/c_free are in the reference's desktop build
// (include/glob_opts.h:64-66), so a matrix made here may be released by the user's c_free and vice versa.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../../include/osqp_rldl_abi.h"

extern "C" {

// util.h:19 — the reference reports OSQP_VERSION ("0.6.0", configure/osqp_configure.h.in)
const char *osqp_version(void) { return "0.6.0"; }

// cs.h:26 — wrap caller-owned arrays (no copies; csc_spfree would free them, the caller normally c_free()s
// only the struct, as examples/osqp_demo.c does)
csc *csc_matrix(c_int m, c_int n, c_int nzmax, c_float *x, c_int *i, c_int *p) {
  csc *M = static_cast<csc *>(malloc(sizeof(csc)));
  if (!M) return nullptr;
  M->m = m;
  M->n = n;
  M->nz = -1;
  M->nzmax = nzmax;
  M->x = x;
  M->i = i;
  M->p = p;
  return M;
}

// cs.h:51
void csc_spfree(csc *A) {
  if (!A) return;
  free(A->p);
  free(A->i);
  free(A->x);
  free(A);
}

// cs.h:44 — triplet != 0: p holds nzmax column indices and nz counts entries; else p holds n+1 pointers
csc *csc_spalloc(c_int m, c_int n, c_int nzmax, c_int values, c_int triplet) {
  csc *A = static_cast<csc *>(calloc(1, sizeof(csc)));
  if (!A) return nullptr;
  if (nzmax < 1) nzmax = 1;
  A->m = m;
  A->n = n;
  A->nzmax = nzmax;
  A->nz = triplet ? 0 : -1;
  A->p = static_cast<c_int *>(malloc(sizeof(c_int) * (size_t)(triplet ? nzmax : n + 1)));
  A->i = static_cast<c_int *>(malloc(sizeof(c_int) * (size_t)nzmax));
  A->x = values ? static_cast<c_float *>(malloc(sizeof(c_float) * (size_t)nzmax)) : nullptr;
  if (!A->p || !A->i || (values && !A->x)) {
    csc_spfree(A);
    return nullptr;
  }
  return A;
}

// cs.h:70
csc *copy_csc_mat(const csc *A) {
  if (!A) return nullptr;
  const c_int nnz = A->p[A->n];
  csc *B = csc_spalloc(A->m, A->n, nnz, 1, 0);
  if (!B) return nullptr;
  memcpy(B->p, A->p, sizeof(c_int) * (size_t)(A->n + 1));
  memcpy(B->i, A->i, sizeof(c_int) * (size_t)nnz);
  memcpy(B->x, A->x, sizeof(c_float) * (size_t)nnz);
  return B;
}

// cs.h:76 — B already allocated with the same pattern sizes
void prea_copy_csc_mat(const csc *A, csc *B) {
  const c_int nnz = A->p[A->n];
  memcpy(B->p, A->p, sizeof(c_int) * (size_t)(A->n + 1));
  memcpy(B->i, A->i, sizeof(c_int) * (size_t)nnz);
  memcpy(B->x, A->x, sizeof(c_float) * (size_t)nnz);
  B->nzmax = A->nzmax;
}

// cs.h:92 — triplets (T->nz entries, column indices in T->p) to CSC, entries of a column kept in triplet
// order, duplicates kept; TtoC[k] (optional) = position of triplet k in the CSC arrays
csc *triplet_to_csc(const csc *T, c_int *TtoC) {
  if (!T || T->nz < 0) return nullptr;
  const c_int n = T->n, nz = T->nz;
  csc *C = csc_spalloc(T->m, n, nz, T->x != nullptr, 0);
  c_int *next = static_cast<c_int *>(calloc((size_t)(n > 0 ? n : 1), sizeof(c_int)));
  if (!C || !next) {
    csc_spfree(C);
    free(next);
    return nullptr;
  }
  for (c_int k = 0; k < nz; k++) next[T->p[k]]++;
  c_int run = 0;
  for (c_int j = 0; j < n; j++) {  // counts -> start offsets
    C->p[j] = run;
    const c_int cnt = next[j];
    next[j] = run;
    run += cnt;
  }
  C->p[n] = run;
  for (c_int k = 0; k < nz; k++) {
    const c_int dst = next[T->p[k]]++;
    C->i[dst] = T->i[k];
    if (C->x) C->x[dst] = T->x[k];
    if (TtoC) TtoC[k] = dst;
  }
  free(next);
  return C;
}

// cs.h:113 — upper triangle (diagonal included) of a square CSC matrix as a new matrix
csc *csc_to_triu(csc *M) {
  if (!M) return nullptr;
  if (M->m != M->n) {
    fprintf(stderr, "ERROR in csc_to_triu: Matrix M not square\n");
    return nullptr;
  }
  const c_int n = M->n;
  c_int keep = 0;
  for (c_int j = 0; j < n; j++)
    for (c_int e = M->p[j]; e < M->p[j + 1]; e++) keep += (M->i[e] <= j);
  csc *U = csc_spalloc(n, n, keep, 1, 0);
  if (!U) {
    fprintf(stderr, "ERROR in csc_to_triu: Upper triangular matrix extraction failed (out of memory)\n");
    return nullptr;
  }
  c_int w = 0;
  for (c_int j = 0; j < n; j++) {
    U->p[j] = w;
    for (c_int e = M->p[j]; e < M->p[j + 1]; e++)
      if (M->i[e] <= j) {
        U->i[w] = M->i[e];
        U->x[w] = M->x[e];
        w++;
      }
  }
  U->p[n] = w;
  U->nzmax = keep;
  return U;
}

}  // extern "C"
