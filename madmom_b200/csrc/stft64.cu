// Double-precision STFT for the launches that hand the raw spectrum to the caller (outputs `stft` / `spec`).
//
// The reference multiplies frame and window in float64, runs a float64 FFT and only then rounds the bins to
// complex64 (madmom/audio/stft.py:105-131, :26); np.abs of those complex64 bins is the Spectrogram
// (madmom/audio/spectrogram.py:122-141).  Consumers of the raw bins -- phase_deviation, rectified_complex_domain,
// modified_kullback_leibler of an unfiltered spectrogram (madmom/features/onsets.py:320-581) -- take phases,
// ratios and signs of bins that sit 70 dB below the loudest one, where an FP32 FFT is off by more than the bin
// itself.  So this path computes like the reference: float64 products, float64 FFT, one rounding to float32 at
// the store.  B200 runs FP64 at half the FP32 rate, and the path is bound by its 8 KB per frame of output
// anyway.  The fused band-feature kernel (frontend_kernel.cu) stays FP32: its outputs pass a filterbank and a
// logarithm and meet the tolerance with room to spare.
//
// One CTA per chunk of frames (same unit / chunk tables as the fused kernel); per frame: even/odd packing into an
// M = N/2 point complex sequence in shared memory, radix-4 DIF passes (two radix-2 stages per pass, twiddles
// W_N^k from a float64 table), a last radix-2 stage when log2 M is odd, real-FFT split straight out of the
// bit-reversed positions.
#include <stdint.h>
#include <cuda_runtime.h>

#include "frontend_params.h"

namespace mmb {

namespace {

__device__ __forceinline__ double2 cadd(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ double2 csub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ double2 cmuld(double2 a, double2 w) {
  return make_double2(a.x * w.x - a.y * w.y, a.x * w.y + a.y * w.x);
}

template <int BITS>
__device__ __forceinline__ int brev(int k) { return (int)(__brev((unsigned)k) >> (32 - BITS)); }

__host__ __device__ constexpr int ilog2(int n) { return n <= 1 ? 0 : 1 + ilog2(n >> 1); }

// slot of bin k in the staged output row: one pad slot per 32 bins, so that bins 32 apart (neighbouring lanes of the
// bit-reversed walk) fall into different banks
__device__ __forceinline__ int out_slot(int k) { return k + (k >> 5); }

// Offsets of the compact twiddle copies behind the natural table of M = N / 2 entries (built in capi.cu): for every
// radix-4 pass, in order, H first-stage and H / 2 second-stage twiddles; then M / 2 split twiddles.
template <int N, int H>
__host__ __device__ constexpr int pass_table_offset() {
  int off = N / 2;
  for (int h = N / 4; h > H; h /= 4) off += h + h / 2;
  return off;
}
template <int N>
__host__ __device__ constexpr int split_table_offset() {
  int off = N / 2;
  for (int h = N / 4; h >= 2; h /= 4) off += h + h / 2;
  return off;
}

// One radix-4 DIF pass with first-stage half-size H (second stage H / 2), then the next one; a last radix-2 stage
// when log2 M is odd.  H is a template parameter so that q / (H / 2) is a shift; the twiddles of a pass are
// contiguous: W_2H^m at tw_a[m], W_H^m at tw_b[m].
template <int N, int H>
__device__ __forceinline__ void radix4_passes(double2* zs, const double2* __restrict__ tw, int tid, int nthreads) {
  constexpr int M = N / 2;
  if constexpr (H >= 2) {
    constexpr int Q2 = H / 2;
    const double2* __restrict__ tw_a = tw + pass_table_offset<N, H>();
    const double2* __restrict__ tw_b = tw_a + H;
    for (int q = tid; q < M / 4; q += nthreads) {
      const int pidx = q & (Q2 - 1);
      const int i0 = ((q / Q2) * 2 * H) + pidx;
      const double2 a0 = zs[i0], a1 = zs[i0 + Q2], a2 = zs[i0 + H], a3 = zs[i0 + H + Q2];
      // first stage: (a0, a2) twiddle W_{2H}^pidx, (a1, a3) twiddle W_{2H}^(pidx + H/2)
      const double2 w0 = tw_a[pidx], w1 = tw_a[pidx + Q2];
      const double2 b0 = cadd(a0, a2), b2 = cmuld(csub(a0, a2), w0);
      const double2 b1 = cadd(a1, a3), b3 = cmuld(csub(a1, a3), w1);
      // second stage (half-size H/2): (b0, b1) and (b2, b3), twiddle W_H^pidx
      const double2 w2 = tw_b[pidx];
      zs[i0] = cadd(b0, b1);
      zs[i0 + Q2] = cmuld(csub(b0, b1), w2);
      zs[i0 + H] = cadd(b2, b3);
      zs[i0 + H + Q2] = cmuld(csub(b2, b3), w2);
    }
    __syncthreads();
    radix4_passes<N, H / 4>(zs, tw, tid, nthreads);
  } else if constexpr (H == 1) {                              // log2 M odd: one radix-2 stage is left (no twiddle)
    for (int q = tid; q < M / 2; q += nthreads) {
      const double2 a0 = zs[2 * q], a1 = zs[2 * q + 1];
      zs[2 * q] = cadd(a0, a1);
      zs[2 * q + 1] = csub(a0, a1);
    }
    __syncthreads();
  }
}

template <int N, typename S>
__global__ void __launch_bounds__(256, 4) stft64_kernel(const Stft64Params p) {
  constexpr int M = N / 2, LOG_M = ilog2(M);
  extern __shared__ double2 stft64_smem[];
  double2* zs = stft64_smem;                                   // [M] the complex sequence of one frame
  float2* so = reinterpret_cast<float2*>(zs + M);              // [stft64_out_slots(M)] staged complex64 bins (out_slot)
  const int tid = threadIdx.x, nthreads = blockDim.x;
  const int chunk = (int)blockIdx.x;
  // unit of this chunk (binary search in the prefix sums of the units' chunk counts)
  int lo = 0, hi = p.n_tracks;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (p.chunk_prefix[mid] <= chunk) lo = mid; else hi = mid;
  }
  const Track tr = p.tracks[lo];
  const long long a_loc = (long long)(chunk - p.chunk_prefix[lo]) * p.chunk_frames;
  const long long b_loc = min(a_loc + (long long)p.chunk_frames, (long long)tr.frame_count);
  const S* src = reinterpret_cast<const S*>(p.pcm) + tr.pcm_offset;

  for (long long f = a_loc; f < b_loc; ++f) {
    const long long g = tr.frame_first + f;
    // int(index * hop_size) - frame_size // 2 - origin in float64, truncation (signal.py:917-919)
    const long long s0 = __double2ll_rz((double)g * p.hop) - (long long)(N / 2) - (long long)p.origin - tr.sample_base;
    // ---- frame * fft_window in float64; circular shift swaps the halves (stft.py:106-124)
    for (int n = tid; n < M; n += nthreads) {
      double v[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int j = 2 * n + h;                              // position in the FFT input
        const int i = p.circular_shift ? (j + N / 2) & (N - 1) : j;      // sample of the frame that sits there
        const long long li = s0 + i;
        double x = 0.0;
        if (li >= 0 && li < tr.pcm_len) x = (double)src[li];
        v[h] = x * p.win[i];
      }
      zs[n] = make_double2(v[0], v[1]);
    }
    __syncthreads();
    // ---- radix-4 DIF passes: stages with half-sizes h and h / 2 on the points i, i + h/2, i + h, i + 3h/2
    //      (compile-time h: the index arithmetic is shifts and masks)
    radix4_passes<N, M / 2>(zs, p.tw, tid, nthreads);
    // ---- real-FFT split, Z[k] sits at the bit-reversed position:
    //      X[k] = (Z[k] + conj Z[M-k]) / 2 + W_N^k (Z[k] - conj Z[M-k]) / 2i
    const long long row = tr.out_row + f;
    // A thread takes the pair (k, M - k): both positions are read once, and W_N^(M-k) = -conj(W_N^k).  The pairs are
    // walked in bit-reversed order, k = brev(2 p'), so that a warp reads Z from consecutive even positions and their
    // partners from one aligned block in reverse (walking k itself puts all 32 lanes on the same banks: measured 32
    // wavefronts per load).  The bins go to a padded shared row and leave in order with coalesced stores below.
    for (int pp = tid; pp < M / 2; pp += nthreads) {
      const int pos = 2 * pp;
      const int k = brev<LOG_M>(pos);                         // 0 <= k < M / 2
      const int km = (M - k) & (M - 1);
      const double2 zk = zs[pos];
      const double2 zm = zs[brev<LOG_M>(km)];
      const double2 w = p.tw[split_table_offset<N>() + pp];   // W_N^k
      // X[k]: a = (zk + conj zm) / 2, b = (zk - conj zm) / 2i
      const double ar = 0.5 * (zk.x + zm.x), ai = 0.5 * (zk.y - zm.y);
      const double br = 0.5 * (zk.y + zm.y), bi = -0.5 * (zk.x - zm.x);
      const double re = ar + (br * w.x - bi * w.y);
      double im = ai + (br * w.y + bi * w.x);
      // X[0] is real.  The sign of its zero imaginary part decides between +pi and -pi in np.angle of a negative
      // sum: scipy.fftpack, the backend of the pinned oracle, returns -0.0 (pyfftw, which the reference's docstrings
      // were made with, +0.0); follow the oracle.
      if (k == 0) im = -0.0;
      so[out_slot(k)] = make_float2((float)re, (float)im);    // the reference's complex64 store
      if (k == 0) {
        so[out_slot(M)] = make_float2((float)(zk.x - zk.y), 0.0f);       // Nyquist bin (emitted on request)
      } else {
        // X[M-k] with the roles of zk and zm swapped and the twiddle (-w.x, w.y)
        const double ar2 = 0.5 * (zm.x + zk.x), ai2 = 0.5 * (zm.y - zk.y);
        const double br2 = 0.5 * (zm.y + zk.y), bi2 = -0.5 * (zm.x - zk.x);
        const double wx = -w.x, wy = w.y;
        so[out_slot(km)] = make_float2((float)(ar2 + (br2 * wx - bi2 * wy)), (float)(ai2 + (br2 * wy + bi2 * wx)));
      }
    }
    if (tid == 0) {                                           // k = M / 2 pairs with itself (position 1)
      const double2 z = zs[1];
      const double2 w = p.tw[M / 2];
      const double br = z.y, bi = 0.0;                        // a = (z.x, 0), b = (z.y, 0) for zk = zm
      so[out_slot(M / 2)] = make_float2((float)(z.x + (br * w.x - bi * w.y)), (float)(0.0 + (br * w.y + bi * w.x)));
    }
    __syncthreads();
    const int out_bins = M + (p.include_nyquist ? 1 : 0);
    for (int k = tid; k < out_bins; k += nthreads) {
      const float2 x = so[out_slot(k)];
      if (p.out_stft) p.out_stft[row * p.ld_stft + k] = x;
      if (p.out_spec) {
        // np.abs of the complex64 bin, float32 (spectrogram.py:22-38)
        float m = (float)sqrt((double)x.x * (double)x.x + (double)x.y * (double)x.y);
        if (p.spec_log_mode != 0) {                           // LogarithmicSpectrogram of an unfiltered one (:527-534)
          const float y = __fadd_rn(__fmul_rn(m, p.mul), p.add);
          m = p.spec_log_mode == 1 ? (float)log10((double)y) : p.spec_log_mode == 3 ? (float)log((double)y)
            : p.spec_log_mode == 4 ? (float)log2((double)y) : y;
        }
        p.out_spec[row * p.ld_spec + k] = m;
      }
    }
    __syncthreads();
  }
}

constexpr int stft64_out_slots(int M) { return M + 1 + (M >> 5) + 1; }

template <int N, typename S>
cudaError_t launch_n(const Stft64Params& p, int grid, cudaStream_t stream) {
  constexpr size_t smem = (size_t)(N / 2) * sizeof(double2) + (size_t)stft64_out_slots(N / 2) * sizeof(float2);
  if (smem > 48 * 1024) {                                     // N = 4096: 49.7 KB, above the default limit
    static const cudaError_t attr = cudaFuncSetAttribute(stft64_kernel<N, S>,
                                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (attr != cudaSuccess) return attr;
  }
  stft64_kernel<N, S><<<grid, 256, smem, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_stft64(int N, int sample_bytes, const Stft64Params& p, int grid, cudaStream_t stream) {
  if (sample_bytes == 2) {
    if (N == 1024) return launch_n<1024, int16_t>(p, grid, stream);
    if (N == 2048) return launch_n<2048, int16_t>(p, grid, stream);
    if (N == 4096) return launch_n<4096, int16_t>(p, grid, stream);
  } else if (sample_bytes == 4) {
    if (N == 1024) return launch_n<1024, float>(p, grid, stream);
    if (N == 2048) return launch_n<2048, float>(p, grid, stream);
    if (N == 4096) return launch_n<4096, float>(p, grid, stream);
  }
  return cudaErrorInvalidValue;
}

}  // namespace mmb
