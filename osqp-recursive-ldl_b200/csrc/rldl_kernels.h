// rldl_kernels.h — host-callable launchers of the kernels in rldl_kernels.cu
#pragma once
#include <cuda_runtime.h>
#include "rldl_device.h"

namespace rldl {
size_t factor_smem_doubles(int a, int nx, int ny, int nt);
size_t admm_smem_doubles(int a, int nx, int ny, int nt);
cudaError_t launch_set_rho_vec(const BatchDev &B, int m, int *changed, cudaStream_t st);
cudaError_t launch_factor(const ModelDev &M, const BatchDev &B, double sigma, int k0,
                          const int *only_flagged, cudaStream_t st);
cudaError_t launch_admm(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                        cudaStream_t st);
cudaError_t launch_kkt_solve(const ModelDev &M, const BatchDev &B, const SettingsDev &S, double *rhs,
                             int ld, cudaStream_t st);
cudaError_t launch_warm_start_z(const ModelDev &M, const BatchDev &B, cudaStream_t st);
cudaError_t launch_check_bounds(const double *l, const double *u, int ld, int m, int batch, int *bad,
                                cudaStream_t st);
cudaError_t launch_shift_rows(double *v, int ld, int len, int sh, int batch, cudaStream_t st);
cudaError_t launch_fill(double *p, double v, size_t n, cudaStream_t st);
cudaError_t launch_scatter_rows(double *dst, int ld, const double *src, int len, int start, int batch,
                                cudaStream_t st);
// streaming path (rldl_stream.cu)
bool stream_supported(const ModelDev &M, int N);
void stream_record_dims(const ModelDev &M, int *qa, int *PX, int *PS, int *bb, int *bd, int *PXX);
cudaError_t launch_pack_records(const ModelDev &M, const BatchDev &B, int k0, int parts, const int *only_flagged,
                                cudaStream_t st);
cudaError_t launch_admm_stream(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                               int resume, int park_below, cudaStream_t st);
// latency-mode kernel (rldl_latency.cu): one instance per CTA, factor resident in shared memory
int latency_t_blocks(int Nmax);  // nx x (nx+2) blocks of global scratch per CTA of the latency-mode kernel
size_t latency_smem_bytes(const ModelDev &M, int N);  // (size_t)-1 if the shape is not in the table
cudaError_t launch_admm_latency(const ModelDev &M, const BatchDev &B, const SettingsDev &S, int warm_start,
                                int resume, double *Tg, int ctas, int reset_parked, cudaStream_t st);
cudaError_t launch_factor_stream(const ModelDev &M, const BatchDev &B, double sigma, int k0,
                                 const int *only_flagged, int write_unpacked, cudaStream_t st);
}  // namespace rldl
