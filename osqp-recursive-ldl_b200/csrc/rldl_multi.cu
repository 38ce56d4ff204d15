// rldl_multi.cu — library-level multi-device handle (RLDLMulti): SURVEY §8e inside the library.
//
// Independent QP instances share nothing, so a batch is cut into contiguous shards, one RLDLBatch handle per shard,
// shard s on device devices[s % ndev]; every entry point runs its per-shard calls on one host thread per shard
// (each thread makes its device current) and returns when all shards have.  There is no collective and no
// peer-to-peer traffic on any path: the "gather" is every shard writing its rows of the caller's host arrays —
// the only cross-GPU traffic north_star (4) allows.  The reference has no counterpart (it is single-threaded,
// one workspace per problem; SURVEY §8b "Threading" is how its users parallelise it).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>
#include "../../include/osqp_rldl_abi.h"

namespace {

struct Shard {
  RLDLBatch *h = nullptr;
  int device = 0;
  c_int i0 = 0, i1 = 0;  // instances [i0, i1) of the whole batch
};

}  // namespace

struct RLDLMulti {
  std::vector<Shard> shards;
  c_int batch = 0, nmax = 0, mmax = 0;
};

namespace {

// run fn(shard) for every shard on its own host thread with its device current; returns the first non-zero code
template <class F>
c_int for_shards(RLDLMulti *m, F fn) {
  const size_t S = m->shards.size();
  std::vector<c_int> rc(S, 0);
  auto body = [&](size_t s) {
    if (rldl_set_device(m->shards[s].device)) { rc[s] = OSQP_LINSYS_SOLVER_INIT_ERROR; return; }
    rc[s] = fn(m->shards[s]);
  };
  if (S == 1) {
    body(0);
  } else {
    std::vector<std::thread> th;
    th.reserve(S);
    for (size_t s = 0; s < S; s++) th.emplace_back(body, s);
    for (auto &t : th) t.join();
  }
  for (c_int r : rc)
    if (r) return r;
  return 0;
}

template <class T>
T *row(T *base, c_int i, c_int stride) { return base ? base + (size_t)i * stride : nullptr; }

}  // namespace

extern "C" {

c_int rldl_multi_setup(RLDLMulti **out, const c_int *devices, c_int ndev, c_int nshards, const OSQPDataRLDL *model,
                       const OSQPSettings *settings, c_int Nmax, c_int N, c_int nx, c_int nu, c_int ny, c_int nt,
                       c_int batch, const c_float *q, const c_float *l, const c_float *u, const c_float *rho) {
  if (!out) return OSQP_DATA_VALIDATION_ERROR;
  *out = nullptr;
  const c_int have = rldl_device_count();
  if (have < 1) {
    fprintf(stderr, "ERROR in rldl_multi_setup: no usable CUDA device — this library has no CPU path\n");
    return OSQP_LINSYS_SOLVER_LOAD_ERROR;
  }
  if (ndev <= 0 || !devices) ndev = have;  // default: every visible device
  if (nshards <= 0) nshards = ndev;
  if (batch < 1 || nshards > batch) {
    fprintf(stderr, "ERROR in rldl_multi_setup: need 1 <= shards <= batch\n");
    return OSQP_DATA_VALIDATION_ERROR;
  }
  RLDLMulti *m = new RLDLMulti();
  m->batch = batch;
  m->nmax = Nmax * (nx + nu);
  m->mmax = Nmax * (nx + ny) + nt;
  m->shards.resize((size_t)nshards);
  for (c_int s = 0; s < nshards; s++) {
    Shard &sh = m->shards[(size_t)s];
    sh.device = devices ? (int)devices[s % ndev] : (int)(s % ndev);
    sh.i0 = s * batch / nshards;
    sh.i1 = (s + 1) * batch / nshards;
    if (sh.device < 0 || sh.device >= have) {
      fprintf(stderr, "ERROR in rldl_multi_setup: device %d is not visible (%lld devices)\n", sh.device, (long long)have);
      delete m;
      return OSQP_DATA_VALIDATION_ERROR;
    }
  }
  const c_int rc = for_shards(m, [&](Shard &sh) {
    return rldl_batch_setup(&sh.h, model, settings, Nmax, N, nx, nu, ny, nt, sh.i1 - sh.i0, row(q, sh.i0, m->nmax),
                            row(l, sh.i0, m->mmax), row(u, sh.i0, m->mmax), row(rho, sh.i0, 1));
  });
  if (rc) {
    rldl_multi_free(m);
    return rc;
  }
  *out = m;
  return 0;
}

#define HM(m) if (!(m)) return OSQP_WORKSPACE_NOT_INIT_ERROR

c_int rldl_multi_shards(const RLDLMulti *m, c_int *first, c_int *device) {
  if (!m) return 0;
  for (size_t s = 0; s < m->shards.size(); s++) {
    if (first) first[s] = m->shards[s].i0;
    if (device) device[s] = m->shards[s].device;
  }
  return (c_int)m->shards.size();
}

RLDLBatch *rldl_multi_shard_handle(RLDLMulti *m, c_int shard) {
  return (m && shard >= 0 && shard < (c_int)m->shards.size()) ? m->shards[(size_t)shard].h : nullptr;
}

c_int rldl_multi_solve(RLDLMulti *m) {
  HM(m);
  return for_shards(m, [](Shard &sh) { return rldl_batch_solve(sh.h); });
}

c_int rldl_multi_solve_to_host(RLDLMulti *m, c_float *x, c_float *y, c_int *iter, c_int *status_val, c_float *obj_val,
                               c_float *pri_res, c_float *dua_res, c_int *rho_updates, c_float *rho_estimate) {
  HM(m);
  return for_shards(m, [&](Shard &sh) {
    return rldl_batch_solve_to_host(sh.h, row(x, sh.i0, m->nmax), row(y, sh.i0, m->mmax), row(iter, sh.i0, 1),
                                    row(status_val, sh.i0, 1), row(obj_val, sh.i0, 1), row(pri_res, sh.i0, 1),
                                    row(dua_res, sh.i0, 1), row(rho_updates, sh.i0, 1), row(rho_estimate, sh.i0, 1));
  });
}

c_int rldl_multi_get_solution(RLDLMulti *m, c_float *x, c_float *y) {
  HM(m);
  return for_shards(m, [&](Shard &sh) {
    return rldl_batch_get_solution(sh.h, row(x, sh.i0, m->nmax), row(y, sh.i0, m->mmax));
  });
}

c_int rldl_multi_get_info(RLDLMulti *m, c_int *iter, c_int *status_val, c_float *obj_val, c_float *pri_res,
                          c_float *dua_res, c_int *rho_updates, c_float *rho_estimate) {
  HM(m);
  return for_shards(m, [&](Shard &sh) {
    return rldl_batch_get_info(sh.h, row(iter, sh.i0, 1), row(status_val, sh.i0, 1), row(obj_val, sh.i0, 1),
                               row(pri_res, sh.i0, 1), row(dua_res, sh.i0, 1), row(rho_updates, sh.i0, 1),
                               row(rho_estimate, sh.i0, 1));
  });
}

c_int rldl_multi_update_bounds(RLDLMulti *m, const c_float *l, const c_float *u) {
  HM(m);
  return for_shards(m, [&](Shard &sh) {
    return rldl_batch_update_bounds(sh.h, row(l, sh.i0, m->mmax), row(u, sh.i0, m->mmax));
  });
}

c_int rldl_multi_partial_update_bounds(RLDLMulti *m, c_int start, c_int stop, const c_float *l, const c_float *u) {
  HM(m);
  return for_shards(m, [&](Shard &sh) {
    return rldl_batch_partial_update_bounds(sh.h, start, stop, row(l, sh.i0, stop - start), row(u, sh.i0, stop - start));
  });
}

c_int rldl_multi_update_lin_cost(RLDLMulti *m, const c_float *q) {
  HM(m);
  return for_shards(m, [&](Shard &sh) { return rldl_batch_update_lin_cost(sh.h, row(q, sh.i0, m->nmax)); });
}

c_int rldl_multi_warm_start(RLDLMulti *m, const c_float *x, const c_float *y) {
  HM(m);
  return for_shards(m, [&](Shard &sh) {
    return rldl_batch_warm_start(sh.h, row(x, sh.i0, m->nmax), row(y, sh.i0, m->mmax));
  });
}

c_int rldl_multi_warm_start_shift(RLDLMulti *m) {
  HM(m);
  return for_shards(m, [](Shard &sh) { return rldl_batch_warm_start_shift(sh.h); });
}

c_int rldl_multi_update_horizon(RLDLMulti *m, c_int N) {
  HM(m);
  return for_shards(m, [&](Shard &sh) { return rldl_batch_update_horizon(sh.h, N); });
}

c_int rldl_multi_update_rho(RLDLMulti *m, const c_float *rho) {
  HM(m);
  return for_shards(m, [&](Shard &sh) { return rldl_batch_update_rho(sh.h, row(rho, sh.i0, 1)); });
}

c_int rldl_multi_refactor(RLDLMulti *m) {
  HM(m);
  return for_shards(m, [](Shard &sh) { return rldl_batch_refactor(sh.h); });
}

c_int rldl_multi_update_settings(RLDLMulti *m, const OSQPSettings *s) {
  HM(m);
  return for_shards(m, [&](Shard &sh) { return rldl_batch_update_settings(sh.h, s); });
}

void rldl_multi_free(RLDLMulti *m) {
  if (!m) return;
  for (Shard &sh : m->shards)
    if (sh.h) {
      rldl_set_device(sh.device);
      rldl_batch_free(sh.h);
    }
  delete m;
}

}  // extern "C"
