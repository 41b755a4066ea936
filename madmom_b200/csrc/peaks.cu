// Calibration microbenchmarks behind the C ABI: the FP32 FMA peak of the device (the binding roof of the
// front end) measured the same way MEASURED_PEAKS.json measures HBM and bf16: a
// dependent-chain FFMA kernel over all SMs, best of several launches, CUDA events.
#include <cuda_runtime.h>
#include "../../include/madmom_b200.h"

namespace {
template <int ILP>
__global__ void __launch_bounds__(256) fma_chain(float* out, int iters, float a, float b) {
  float acc[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) acc[i] = (float)(threadIdx.x + i);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) acc[i] = fmaf(acc[i], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += acc[i];
  if (s == 123.456f) out[0] = s;     // never true; keeps the chains alive
}
}  // namespace

extern "C" MMB_API int mmb_measure_fp32_peak(int32_t device, double* tflops) {
  if (!tflops) return MMB_E_INVALID;
  *tflops = 0.0;
  if (cudaSetDevice(device) != cudaSuccess) return MMB_E_CUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return MMB_E_CUDA;
  float* d = nullptr;
  if (cudaMalloc(&d, 4) != cudaSuccess) return MMB_E_CUDA;
  constexpr int ILP = 8;
  const int iters = 1 << 15;
  const int grid = prop.multiProcessorCount * 8;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best = 0.0;
  for (int rep = 0; rep < 6; ++rep) {
    cudaEventRecord(e0);
    fma_chain<ILP><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(d); return MMB_E_CUDA; }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double flops = 2.0 * ILP * (double)iters * 256.0 * grid;
    const double tf = flops / (ms * 1e-3) / 1e12;
    if (rep > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(d);
  *tflops = best;
  return MMB_OK;
}
