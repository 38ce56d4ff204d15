// rldl_kernels.h — host-callable launchers of the kernels in rldl_kernels.cu
#pragma once
#include <cuda_runtime.h>
#include "rldl_device.h"

namespace rldl {
size_t factor_smem_doubles(int a, int nx, int ny, int nt);
size_t admm_smem_doubles(int a, int nx, int ny, int nt);
cudaError_t launch_rescale_iterates(const BatchDev &B, int to_scaled, cudaStream_t st);
cudaError_t launch_ruiz(const ModelDev &M, const BatchDev &B, int iters, cudaStream_t st);
cudaError_t launch_set_rho_vec(const BatchDev &B, int m, int *changed, cudaStream_t st);
cudaError_t launch_factor(const ModelDev &M, const BatchDev &B, double sigma, int k0,
                          const int *only_flagged, cudaStream_t st);
cudaError_t launch_admm(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                        cudaStream_t st);
cudaError_t launch_polish(const ModelDev &M, const BatchDev &B, const BatchDev &P, const SettingsDev &S, double delta,
                          int refine_iter, double *pb, double *ps, double *pr, int *status_polish, cudaStream_t st);
// allocation-free variant: right-hand sides staged in a fixed file-scope device array of kKktStaticDoubles doubles
constexpr int kKktStaticDoubles = 512 * 1024;  // 4 MiB per GPU that loads the library
cudaError_t launch_kkt_solve_static(const ModelDev &M, const BatchDev &B, const SettingsDev &S, const double *rhs_in,
                                    double *sol_out, size_t count, int ld, cudaStream_t st);
cudaError_t launch_kkt_solve(const ModelDev &M, const BatchDev &B, const SettingsDev &S, double *rhs,
                             int ld, cudaStream_t st);
cudaError_t launch_warm_start_z(const ModelDev &M, const BatchDev &B, cudaStream_t st);
cudaError_t launch_check_bounds(const double *l, const double *u, int ld, int m, int batch, int *bad,
                                cudaStream_t st, bool reset = true);
cudaError_t launch_shift_rows(double *v, int ld, int len, int sh, int batch, cudaStream_t st);
cudaError_t launch_fill(double *p, double v, size_t n, cudaStream_t st);
cudaError_t launch_mark_time(const BatchDev &B, cudaStream_t st);
cudaError_t launch_fill_int(int *p, int v, size_t n, cudaStream_t st);
cudaError_t launch_scatter_rows(double *dst, int ld, const double *src, int len, int start, int batch,
                                cudaStream_t st);
// streaming path (rldl_stream.cu)
bool stream_supported(const ModelDev &M, int N);
void stream_record_dims(const ModelDev &M, int *qa, int *PX, int *PS, int *bb, int *bd, int *PXX, int *sigw,
                        int *clsw);
// vector part of the stage records (parts bit 0: q; bit 1: row classes, rho_vec, l, u + change flags)
cudaError_t launch_vec_records(const ModelDev &M, const BatchDev &B, int k0, int parts, const int *only_flagged,
                               int *changed, double sigma, cudaStream_t st);
// launch knobs of the streaming kernel (rldl_batch_set_tuning; 0 / -1 = the measured default)
struct StreamTuning {
  int ring_depth = 0;   // stage records in flight per lane group (default 2)
  int sched_s0 = -1;    // iterations a fresh instance runs before it may yield (default check_termination; 0 = scheduler off)
  int ctas_per_sm = 0;  // resident CTAs per SM (default 4 with the latency-mode hand-off, 3 without)
  int l2_hints = 1;     // bit 0: stage records evict_first; bit 1: forward intermediates evict_last / evict_first
};
// k_l > 0: the CTAs on SMs 0 .. k_l-1 retire once at most retire_below instances are unfinished (concurrent
// latency-mode lane); reset_done (optional) is recorded after the scheduler reset and before the kernel; grid_out
// (optional) receives the number of CTAs launched
cudaError_t launch_admm_stream(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                               int resume, int park_below, const StreamTuning &tune, int k_l, int retire_below,
                               cudaEvent_t reset_done, int *grid_out, cudaStream_t st);
cudaError_t launch_set_tail_flag(const BatchDev &B, cudaStream_t st);
cudaError_t preload_lane_kernels();  // forces the (lazy) load of the kernels launched while lane CTAs may be spinning
// latency-mode kernel (rldl_latency.cu): one instance per CTA, factor resident in shared memory
int latency_t_blocks(int Nmax);  // nx x (nx+2) blocks of global scratch per CTA of the latency-mode kernel
size_t latency_smem_bytes(const ModelDev &M, int N);  // (size_t)-1 if the shape is not in the table
cudaError_t launch_admm_latency(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                                int resume, double *Tg, int ctas, int reset_parked, int claim_mode, int stream_ctas,
                                cudaStream_t st);
cudaError_t launch_factor_stream(const ModelDev &M, const BatchDev &B, double sigma, int k0,
                                 const int *only_flagged, int write_unpacked, cudaStream_t st);
}  // namespace rldl
