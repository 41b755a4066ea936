// Routes launch_frontend / frontend_set_smem to the per-frame-size translation units of frontend_kernel.cu
// (compiled with -DMMB_FRONTEND_N=... -DMMB_FRONTEND_WCOS=...; madmom_b200/build.py).  Empty unless MMB_SPLIT_FRONTEND is defined, so that
// compiling every .cu file on its own without flags (single-unit frontend_kernel.cu) still gives one definition.
#include "frontend_params.h"

#if defined(MMB_SPLIT_FRONTEND)
namespace mmb {

cudaError_t launch_frontend_1024(int, const FrontendParams&, int, int, cudaStream_t);
cudaError_t launch_frontend_2048(int, const FrontendParams&, int, int, cudaStream_t);
cudaError_t launch_frontend_4096(int, const FrontendParams&, int, int, cudaStream_t);
cudaError_t launch_frontend_2048_wcos(int, const FrontendParams&, int, int, cudaStream_t);
cudaError_t launch_frontend_4096_wcos(int, const FrontendParams&, int, int, cudaStream_t);
cudaError_t frontend_set_smem_1024(int, int);
cudaError_t frontend_set_smem_2048(int, int);
cudaError_t frontend_set_smem_4096(int, int);
cudaError_t frontend_set_smem_2048_wcos(int, int);
cudaError_t frontend_set_smem_4096_wcos(int, int);

cudaError_t launch_frontend(int N, int sample_bytes, const FrontendParams& p, int grid, int smem_bytes,
                            cudaStream_t stream) {
  const bool wcos = p.win_cos != 0;      // cosine-sum window evaluated in the kernel (N = 2048, 4096)
  if (N == 1024 && !wcos) return launch_frontend_1024(sample_bytes, p, grid, smem_bytes, stream);
  if (N == 2048) return wcos ? launch_frontend_2048_wcos(sample_bytes, p, grid, smem_bytes, stream)
                             : launch_frontend_2048(sample_bytes, p, grid, smem_bytes, stream);
  if (N == 4096) return wcos ? launch_frontend_4096_wcos(sample_bytes, p, grid, smem_bytes, stream)
                             : launch_frontend_4096(sample_bytes, p, grid, smem_bytes, stream);
  return cudaErrorInvalidValue;
}

cudaError_t frontend_set_smem(int N, int sample_bytes, int smem_bytes) {
  if (N == 1024) return frontend_set_smem_1024(sample_bytes, smem_bytes);
  if (N != 2048 && N != 4096) return cudaErrorInvalidValue;
  cudaError_t e = N == 2048 ? frontend_set_smem_2048(sample_bytes, smem_bytes)
                            : frontend_set_smem_4096(sample_bytes, smem_bytes);
  if (e != cudaSuccess) return e;
  return N == 2048 ? frontend_set_smem_2048_wcos(sample_bytes, smem_bytes)
                   : frontend_set_smem_4096_wcos(sample_bytes, smem_bytes);
}

}  // namespace mmb
#endif
