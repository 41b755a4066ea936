// Calibration microbenchmarks behind the C ABI: the FP32 FMA peak of the device (the binding roof of the
// front end) measured the same way MEASURED_PEAKS.json measures HBM and bf16: a
// dependent-chain FFMA kernel over all SMs, best of several launches, CUDA events.
#include <cuda_runtime.h>
#include <cmath>
#include <vector>
#include "../../include/madmom_b200.h"
#include "frame_fft.cuh"

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

// ---- instruction-rate probes used to build the issue-slot roofline of the front end --------
// which: 0 FFMA with constant-bank operands (same as mmb_measure_fp32_peak), 1 FFMA with three register
// operands, 2 FADD register-register, 3 FMUL+FADD pairs (butterfly-like), 4 LDS.64 + 2 FFMA (shared-memory fed)
namespace {
template <int WHICH>
__global__ void __launch_bounds__(256) rate_probe(float* out, int iters, float a, float b) {
  __shared__ float2 sm[512];
  constexpr int ILP = 8;
  float acc[ILP], x[ILP], y[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) {
    acc[i] = (float)(threadIdx.x + i); x[i] = a + 1e-7f * (float)(i + threadIdx.x); y[i] = b * (float)(i + 1);
  }
  sm[threadIdx.x] = make_float2(a, b); sm[threadIdx.x + 256] = make_float2(b, a);
  __syncthreads();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      if (WHICH == 1) acc[i] = fmaf(acc[i], x[i], y[i]);
      else if (WHICH == 2) acc[i] = acc[i] + x[i];
      else if (WHICH == 3) { const float t = acc[i] * x[i]; acc[i] = t + y[(i + 1) % ILP]; }
      else if (WHICH == 4) {
        const float2 w = sm[(threadIdx.x + 32 * i + it) & 511];
        acc[i] = fmaf(acc[i], w.x, w.y); acc[i] = fmaf(acc[i], x[i], w.x);
      } else if (WHICH == 5) {           // LDS.32 only (conflict-free)
        acc[i] += reinterpret_cast<const float*>(sm)[(threadIdx.x + 32 * i + it) & 1023];
      } else if (WHICH == 6) {           // SHFL only
        acc[i] = __shfl_sync(0xffffffffu, acc[i], (threadIdx.x + i + 1) & 31) + x[i];
      } else if (WHICH == 8 || WHICH == 9 || WHICH == 10) {
        // packed FP32 pairs (sm_100 add/fma.rn.f32x2 -> FADD2 / FFMA2): two lanes of work per instruction
        unsigned long long p, q, r;
        asm("mov.b64 %0, {%1, %2};" : "=l"(p) : "f"(acc[i]), "f"(y[i]));
        asm("mov.b64 %0, {%1, %2};" : "=l"(q) : "f"(x[i]), "f"(x[(i + 1) % ILP]));
        if (WHICH == 8) asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(p), "l"(q));
        else asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(p), "l"(q), "l"(q));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[i]), "=f"(y[i]) : "l"(r));
        if (WHICH == 10) acc[i] += reinterpret_cast<const float*>(sm)[(threadIdx.x + 32 * i + it) & 1023];
      } else if (WHICH == 11) {          // LDS.64, conflict-free: 2 wavefronts per instruction
        const float2 w = sm[(threadIdx.x + 32 * i + it) & 511];
        acc[i] += w.x + w.y;
      } else if (WHICH == 12) {          // LDS.128, conflict-free: 4 wavefronts per instruction
        const float4 w = reinterpret_cast<const float4*>(sm)[(threadIdx.x + 32 * i + it) & 255];
        acc[i] += (w.x + w.y) + (w.z + w.w);
      } else if (WHICH == 13) {          // STS.64 then LDS.64 of another slot (the FFT exchange pattern)
        const unsigned a0 = (unsigned)__cvta_generic_to_shared(&sm[(threadIdx.x + 32 * i) & 511]);
        const unsigned a1 = (unsigned)__cvta_generic_to_shared(&sm[(threadIdx.x + 32 * i + 7) & 511]);
        float wx, wy;
        asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a0), "f"(acc[i]), "f"(x[i]) : "memory");
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(wx), "=f"(wy) : "r"(a1) : "memory");
        acc[i] += wx + wy;
      } else if (WHICH == 14) {          // STS.64 only
        const unsigned a0 = (unsigned)__cvta_generic_to_shared(&sm[(threadIdx.x + 32 * i + it) & 511]);
        asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a0), "f"(acc[i]), "f"(x[i]) : "memory");
        acc[i] += x[i];
      } else if (WHICH == 7) {           // one LDS.32 + one SHFL per step
        acc[i] = __shfl_sync(0xffffffffu, acc[i], (threadIdx.x + i + 1) & 31) +
                 reinterpret_cast<const float*>(sm)[(threadIdx.x + 32 * i + it) & 1023];
      }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += acc[i];
  if (s == 123.456f) out[0] = s;
}
}  // namespace

// result: warp-level FP instructions per clock per SM at the current clock estimate is left to the caller;
// *ginst_per_s receives 1e9 thread-level FP instructions per second (whole device).
extern "C" MMB_API int mmb_rate_probe(int32_t device, int32_t which, double* ginst_per_s) {
  if (!ginst_per_s || which < 1 || which > 14) return MMB_E_INVALID;
  if (cudaSetDevice(device) != cudaSuccess) return MMB_E_CUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return MMB_E_CUDA;
  float* d = nullptr;
  if (cudaMalloc(&d, 4) != cudaSuccess) return MMB_E_CUDA;
  const int iters = 1 << 14, grid = prop.multiProcessorCount * 8;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best = 0.0;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    if (which == 1) rate_probe<1><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 2) rate_probe<2><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 3) rate_probe<3><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 4) rate_probe<4><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 5) rate_probe<5><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 6) rate_probe<6><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 7) rate_probe<7><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 8) rate_probe<8><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 9) rate_probe<9><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 10) rate_probe<10><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 11) rate_probe<11><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 12) rate_probe<12><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else if (which == 13) rate_probe<13><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    else rate_probe<14><<<grid, 256>>>(d, iters, 1.0000001f, 1e-9f);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(d); return MMB_E_CUDA; }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double per_iter = (which == 3 || which == 4) ? 2.0 : 1.0;      // FP instructions per chain step
    const double n = per_iter * 8.0 * (double)iters * 256.0 * grid;
    const double r = n / (ms * 1e-3) / 1e9;
    if (rep > 0 && r > best) best = r;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(d);
  *ginst_per_s = best;
  return MMB_OK;
}

// ---- FFT-only ceiling of the fused kernel --------------------------------------------------------------
// The same warp-level 2048-point real FFT the front end runs per frame (core_pass1 / core_pass2 / publish / split
// with magnitudes into the warp's scratch), 16 warps per SM, inputs made up in registers: no PCM staging, no
// window, no filterbank, no difference, no global stores.  Its frame rate is the ceiling the fused kernel is
// judged against (bench.py: roofline.fft_only).
namespace {
__global__ void __launch_bounds__(512, 1) fft_only_kernel(const float2* __restrict__ tw1, const float2* __restrict__ tw2k,
                                                           int iters, float* out) {
  using namespace mmb;
  extern __shared__ __align__(16) unsigned char smem[];
  float2* s_tw1 = reinterpret_cast<float2*>(smem);
  float2* s_tw2k = reinterpret_cast<float2*>(smem + 8192);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned char* my = smem + 8192 + 4096 + warp * (8448 + 16);
  float2* ex = reinterpret_cast<float2*>(my);
  float* mag = reinterpret_cast<float*>(my);
  for (int i = tid; i < 1024; i += blockDim.x) s_tw1[i] = tw1[i];
  for (int i = tid; i < 512; i += blockDim.x) s_tw2k[i] = tw2k[i];
  __syncthreads();
  float acc = 0.0f;
  const float a = 1e-3f * (float)(lane + 1), b = 1e-4f * (float)(warp + 1);
  for (int it = 0; it < iters; ++it) {
    float2 v[32];
    const float t = (float)it;
#pragma unroll
    for (int m = 0; m < 32; ++m) v[m] = make_float2(fmaf(a, (float)m, t), fmaf(b, (float)(m * m), a));
    core_pass1(lane, v, ex, s_tw1);
    __syncwarp();
    core_pass2_load(lane, v, ex);
    __syncwarp();
    core_pass2_fft(v);
    core_publish_upper(lane, v, ex, true);
    __syncwarp();
    BinSinkT<false> o;
    o.mag = mag;
    split_2048(lane, v, ex, s_tw2k, o);
    __syncwarp();
    acc += mag[lane + 32 * (it & 31)];
    __syncwarp();
  }
  if (acc == 123.456f) out[0] = acc;
}
}  // namespace

extern "C" MMB_API int mmb_fft_probe(int32_t device, int32_t frame_size, double* frames_per_s) {
  if (!frames_per_s) return MMB_E_INVALID;
  *frames_per_s = 0.0;
  if (frame_size != 2048) return MMB_E_UNSUPPORTED;
  if (cudaSetDevice(device) != cudaSuccess) return MMB_E_CUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return MMB_E_CUDA;
  const double PI = 3.14159265358979323846264338327950288;
  std::vector<float2> tw1(1024), tw2k(512);
  for (int k1 = 0; k1 < 32; ++k1)
    for (int l = 0; l < 32; ++l) {
      const double ang = -2.0 * PI * (double)((l * k1) % 1024) / 1024.0;
      tw1[mmb::tw1_index(k1, l)] = make_float2((float)std::cos(ang), (float)std::sin(ang));
    }
  for (int k = 0; k < 512; ++k) {
    const double ang = -2.0 * PI * (double)k / 2048.0;
    tw2k[k] = make_float2((float)std::cos(ang), (float)std::sin(ang));
  }
  float2 *d1 = nullptr, *d2 = nullptr;
  float* dout = nullptr;
  const int smem = 8192 + 4096 + 16 * (8448 + 16);
  int rc = MMB_OK;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (cudaMalloc(&d1, 1024 * sizeof(float2)) != cudaSuccess || cudaMalloc(&d2, 512 * sizeof(float2)) != cudaSuccess ||
      cudaMalloc(&dout, 4) != cudaSuccess ||
      cudaMemcpy(d1, tw1.data(), 1024 * sizeof(float2), cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaMemcpy(d2, tw2k.data(), 512 * sizeof(float2), cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaFuncSetAttribute(fft_only_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
    rc = MMB_E_CUDA;
  } else {
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 2048, grid = prop.multiProcessorCount;
    double best = 0.0;
    for (int rep = 0; rep < 4 && rc == MMB_OK; ++rep) {
      cudaEventRecord(e0);
      fft_only_kernel<<<grid, 512, smem>>>(d1, d2, iters, dout);
      cudaEventRecord(e1);
      if (cudaEventSynchronize(e1) != cudaSuccess) { rc = MMB_E_CUDA; break; }
      float ms = 0.f;
      cudaEventElapsedTime(&ms, e0, e1);
      const double fps = (double)iters * 16.0 * grid / (ms * 1e-3);
      if (rep > 0 && fps > best) best = fps;
    }
    *frames_per_s = best;
  }
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  cudaFree(d1); cudaFree(d2); cudaFree(dout);
  return rc;
}

// ---- would warp specialisation help?  (probe for DESIGN.md, round 2) ------------------------------------
// The fused kernel keeps 16 warps per SM (128 registers x 512 threads) and its shared-memory pipe at 75 %.  This
// probe runs the same per-frame work in two shapes: mode 0, 16 warps that each do everything (staged-PCM + window
// loads, the warp FFT, split, magnitudes, a filterbank-like pass over the magnitudes: 17 LDS.128 + 68 LDS.32 + 34
// FFMA2 + a dozen stores + log / difference-like arithmetic); mode 1, 12 FFT warps that raise their register budget
// with setmaxnreg and 12 light warps that lower theirs and take the filterbank-like part (no hand-off
// synchronisation: a throughput probe of the mix, not a working pipeline).
namespace {
__device__ __forceinline__ void probe_heavy(int lane, int it, const short* pcm, const float* win, float2* ex,
                                            float* mag, const float2* s_tw1, const float2* s_tw2k) {
  using namespace mmb;
  float2 v[32];
  const short* x = pcm + ((it * 441) & 4095);
#pragma unroll
  for (int m = 0; m < 32; ++m) {
    const int n = lane + 32 * m;
    const float2 w = *reinterpret_cast<const float2*>(win + 2 * n);
    v[m] = mul2(make_float2((float)x[2 * n], (float)x[2 * n + 1]), w);
  }
  core_pass1(lane, v, ex, s_tw1);
  __syncwarp();
  core_pass2_load(lane, v, ex);
  __syncwarp();
  core_pass2_fft(v);
  core_publish_upper(lane, v, ex, true);
  __syncwarp();
  BinSinkT<false> o;
  o.mag = mag;
  split_2048(lane, v, ex, s_tw2k, o);
  __syncwarp();
}

__device__ __forceinline__ float probe_light(int lane, int it, const float* mag, const float* wts, float* part,
                                             float* ring) {
  float2 acc = make_float2(0.0f, 0.0f);
  const float4* w = reinterpret_cast<const float4*>(wts) + lane;
  const float* m0 = mag + ((lane * 7 + it) & 255);
#pragma unroll 1
  for (int g = 0; g < 17; ++g) {                       // 17 x (LDS.128 + 4 LDS.32 + 2 FFMA2)
    const float4 wv = w[32 * g];
    const float* a = m0 + 4 * g;
    acc = mmb::fma2(make_float2(a[0], a[1]), make_float2(wv.x, wv.y), acc);
    acc = mmb::fma2(make_float2(a[2], a[3]), make_float2(wv.z, wv.w), acc);
    if ((g & 3) == 3) part[(g >> 2) * 96 + lane] = acc.x + acc.y;
  }
  __syncwarp();
  float sum = 0.0f;
#pragma unroll 1
  for (int j = 0; j < 3; ++j) {                        // epilogue + difference-like: 3 x 32 bands
    const int b = lane + 32 * j;
    const float s4 = part[b] + part[96 + b] + part[192 + b] + part[288 + b];
    const float lg = __log2f(s4 * s4 + 1.0f) * 0.30103f;
    const float prev = ring[96 + b];
    ring[b] = lg;
    sum += fmaxf(lg - prev, 0.0f);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  __syncwarp();
  return sum;
}

template <int MODE>
__global__ void __launch_bounds__(MODE == 0 ? 512 : 768, 1)
pipeline_probe_kernel(const float2* __restrict__ tw1, const float2* __restrict__ tw2k, int iters, float* out) {
  extern __shared__ __align__(16) unsigned char smem[];
  constexpr int HEAVY = MODE == 0 ? 16 : 12;
  float2* s_tw1 = reinterpret_cast<float2*>(smem);                       // 8 KB
  float2* s_tw2k = reinterpret_cast<float2*>(smem + 8192);               // 4 KB
  float* s_win = reinterpret_cast<float*>(smem + 12288);                 // 8 KB
  short* s_pcm = reinterpret_cast<short*>(smem + 20480);                 // 17 KB
  float* s_wts = reinterpret_cast<float*>(smem + 37888);                 // 17 x 32 x 16 B = 8.5 KB
  unsigned char* tiles = smem + 47104;                                   // HEAVY x 8464
  unsigned char* lightmem = tiles + HEAVY * 8464;                        // per light / monolithic warp: 2 KB
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < 1024; i += blockDim.x) s_tw1[i] = tw1[i];
  for (int i = tid; i < 512; i += blockDim.x) s_tw2k[i] = tw2k[i];
  for (int i = tid; i < 2048; i += blockDim.x) s_win[i] = 1e-4f * (float)(i & 255);
  for (int i = tid; i < 8704; i += blockDim.x) s_pcm[i] = (short)(i * 37);
  for (int i = tid; i < 2176; i += blockDim.x) s_wts[i] = 1e-3f * (float)(i & 63);
  for (int i = tid; i < 24 * 512; i += blockDim.x) reinterpret_cast<float*>(lightmem)[i] = 0.0f;
  __syncthreads();
  float acc = 0.0f;
  if (MODE == 0) {
    float2* ex = reinterpret_cast<float2*>(tiles + warp * 8464);
    float* mag = reinterpret_cast<float*>(ex);
    float* part = reinterpret_cast<float*>(lightmem + warp * 2048);
    for (int it = 0; it < iters; ++it) {
      probe_heavy(lane, it, s_pcm, s_win, ex, mag, s_tw1, s_tw2k);
      acc += probe_light(lane, it, mag, s_wts, part, part + 384 - 96);
    }
  } else {
    if (warp < HEAVY) {
      asm volatile("setmaxnreg.inc.sync.aligned.u32 136;");
      float2* ex = reinterpret_cast<float2*>(tiles + warp * 8464);
      float* mag = reinterpret_cast<float*>(ex);
      for (int it = 0; it < iters; ++it) {
        probe_heavy(lane, it, s_pcm, s_win, ex, mag, s_tw1, s_tw2k);
        acc += mag[lane];
      }
    } else {
      asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
      const int lw = warp - HEAVY;
      const float* mag = reinterpret_cast<const float*>(tiles + lw * 8464);
      float* part = reinterpret_cast<float*>(lightmem + lw * 2048);
      for (int it = 0; it < iters; ++it) acc += probe_light(lane, it, mag, s_wts, part, part + 384 - 96);
    }
  }
  if (acc == 123.456f) out[0] = acc;
}
}  // namespace

extern "C" MMB_API int mmb_pipeline_probe(int32_t device, int32_t mode, double* frames_per_s) {
  if (!frames_per_s || (mode != 0 && mode != 1)) return MMB_E_INVALID;
  *frames_per_s = 0.0;
  if (cudaSetDevice(device) != cudaSuccess) return MMB_E_CUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return MMB_E_CUDA;
  const double PI = 3.14159265358979323846264338327950288;
  std::vector<float2> tw1(1024), tw2k(512);
  for (int k1 = 0; k1 < 32; ++k1)
    for (int l = 0; l < 32; ++l) {
      const double ang = -2.0 * PI * (double)((l * k1) % 1024) / 1024.0;
      tw1[mmb::tw1_index(k1, l)] = make_float2((float)std::cos(ang), (float)std::sin(ang));
    }
  for (int k = 0; k < 512; ++k) {
    const double ang = -2.0 * PI * (double)k / 2048.0;
    tw2k[k] = make_float2((float)std::cos(ang), (float)std::sin(ang));
  }
  float2 *d1 = nullptr, *d2 = nullptr;
  float* dout = nullptr;
  const int heavy = mode == 0 ? 16 : 12;
  const int smem = 47104 + heavy * 8464 + 24 * 2048;
  int rc = MMB_OK;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (cudaMalloc(&d1, 1024 * sizeof(float2)) != cudaSuccess || cudaMalloc(&d2, 512 * sizeof(float2)) != cudaSuccess ||
      cudaMalloc(&dout, 4) != cudaSuccess ||
      cudaMemcpy(d1, tw1.data(), 1024 * sizeof(float2), cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaMemcpy(d2, tw2k.data(), 512 * sizeof(float2), cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaFuncSetAttribute(pipeline_probe_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess ||
      cudaFuncSetAttribute(pipeline_probe_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
    rc = MMB_E_CUDA;
  } else {
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 2048, grid = prop.multiProcessorCount;
    double best = 0.0;
    for (int rep = 0; rep < 4 && rc == MMB_OK; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) pipeline_probe_kernel<0><<<grid, 512, smem>>>(d1, d2, iters, dout);
      else pipeline_probe_kernel<1><<<grid, 768, smem>>>(d1, d2, iters, dout);
      cudaEventRecord(e1);
      if (cudaEventSynchronize(e1) != cudaSuccess || cudaGetLastError() != cudaSuccess) { rc = MMB_E_CUDA; break; }
      float ms = 0.f;
      cudaEventElapsedTime(&ms, e0, e1);
      const double fps = (double)iters * heavy * grid / (ms * 1e-3);
      if (rep > 0 && fps > best) best = fps;
    }
    *frames_per_s = best;
  }
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  cudaFree(d1); cudaFree(d2); cudaFree(dout);
  return rc;
}
