// Fused spectral front end for sm_100a: framing + window + real FFT + |.| + banded triangular filterbank +
// log10(mul*x+add) + frequency max-filter + lagged (positive) difference + flux row-sum, one pass, only the
// requested feature matrices reach HBM.
//
// Reference path being replaced (CPJKU/madmom): FramedSignal/signal_frame (audio/signal.py:860-962,
// 1103-1197) -> stft() (audio/stft.py:49-133) -> Spectrogram (audio/spectrogram.py:122-141) ->
// FilteredSpectrogram np.dot (:382) -> LogarithmicSpectrogram (:527-534) -> SpectrogramDifference
// (:976-993) -> spectral_flux/superflux row sum (features/onsets.py:163-240).
//
// Work decomposition
//   unit   = track (or frame range of a long signal), split into chunks of `chunk_frames` output frames
//   CTA    = persistent, strides over chunks; a chunk is walked in rounds of FPR consecutive frames
//   warp   = one frame per round (two for N=1024): register-resident 32x32 FFT (fft_core.cuh)
//   round  = stage PCM span -> [A] per-warp FFT + magnitudes -> [B] CTA-wide filterbank + log into a
//            32-row ring -> [C] per-warp max-filter / difference / flux and coalesced row stores
// A chunk recomputes its `diff_frames` lead-in frames instead of exchanging them with its neighbour.
#include "frame_fft.cuh"
#include "frontend_params.h"

namespace mmb {

__device__ __forceinline__ long long frame_start_sample(long long g, double hop, int N, int origin) {
  // int(index * hop_size) - frame_size // 2 - origin, float64 product, truncation (signal.py:917-919)
  return __double2ll_rz((double)g * hop) - (long long)(N / 2) - (long long)origin;
}

__device__ __forceinline__ float fast_log10(float y) { return __log2f(y) * 0.30102999566398119521f; }

template <int N, typename S>
__global__ void __launch_bounds__(warps_per_cta(N) * 32, 1) frontend_kernel(const FrontendParams p) {
  constexpr int kWarps = warps_per_cta(N);
  constexpr int kThreads = kWarps * 32;
  extern __shared__ __align__(16) unsigned char smem[];
  constexpr int FPR = frames_per_round(N);
  constexpr int K = N / 2;
  const SmemLayout L = smem_layout(N, p.n_bands, p.nnz, p.span_cap, (int)sizeof(S));
  float* s_win = reinterpret_cast<float*>(smem + L.win);
  float2* s_tw1 = reinterpret_cast<float2*>(smem + L.tw1);
  float2* s_tw2k = reinterpret_cast<float2*>(smem + L.tw2k);
  float2* s_tw4k = reinterpret_cast<float2*>(smem + L.tw4k);
  int* s_bstart = reinterpret_cast<int*>(smem + L.fb_meta);
  int* s_blen = s_bstart + p.n_bands;
  int* s_bwoff = s_blen + p.n_bands;
  float* s_w = reinterpret_cast<float*>(smem + L.fb_w);
  float* s_ring = reinterpret_cast<float*>(smem + L.ring);
  S* s_pcm = reinterpret_cast<S*>(smem + L.pcm);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned char* my_scratch = smem + L.scratch + warp * warp_scratch_bytes(N);
  float2* ex = reinterpret_cast<float2*>(my_scratch);
  float* my_mag = (N == 4096) ? reinterpret_cast<float*>(my_scratch + 8448) : reinterpret_cast<float*>(my_scratch);

  // ---- plan constants -> shared memory (once per persistent CTA)
  for (int i = tid; i < N; i += kThreads) s_win[i] = p.win[i];
  for (int i = tid; i < 1024; i += kThreads) s_tw1[i] = p.tw1[i];
  if (N != 1024) for (int i = tid; i < 1024; i += kThreads) s_tw2k[i] = p.tw2k[i];
  if (N == 4096) for (int i = tid; i < 512; i += kThreads) s_tw4k[i] = p.tw4k[i];
  for (int i = tid; i < p.n_bands; i += kThreads) {
    s_bstart[i] = p.band_start[i]; s_blen[i] = p.band_len[i]; s_bwoff[i] = p.band_woff[i];
  }
  for (int i = tid; i < p.nnz; i += kThreads) s_w[i] = p.weights[i];
  __syncthreads();

  const S* __restrict__ pcm = reinterpret_cast<const S*>(p.pcm);
  const int nb = p.n_bands;
  const bool want_diff = (p.out_diff != nullptr) || (p.out_odf != nullptr);
  const int lead = want_diff ? p.diff_frames : 0;
  const float sgn_odd = p.circular_shift ? -1.0f : 1.0f;

  for (int chunk = blockIdx.x; chunk < p.total_chunks; chunk += gridDim.x) {
    // ---- locate the unit this chunk belongs to (binary search over the prefix table)
    int lo = 0, hi = p.n_tracks;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (p.chunk_prefix[mid] <= chunk) lo = mid; else hi = mid;
    }
    const Track tr = p.tracks[lo];
    const long long a_loc = (long long)(chunk - p.chunk_prefix[lo]) * p.chunk_frames;
    const long long b_loc = min(a_loc + (long long)p.chunk_frames, (long long)tr.frame_count);
    const long long ga = tr.frame_first + a_loc, gb = tr.frame_first + b_loc;   // global frame range
    const long long g_begin = max(ga - (long long)lead, 0LL);

    for (long long g0 = g_begin; g0 < gb; g0 += FPR) {
      const int nfr = (int)min((long long)FPR, gb - g0);
      // ---- stage the PCM span of this round (zero outside the unit's local samples)
      const long long s_begin = frame_start_sample(g0, p.hop, N, p.origin);
      const long long s_end = frame_start_sample(g0 + nfr - 1, p.hop, N, p.origin) + N;
      const int span = (int)(s_end - s_begin);
      {
        const long long l0 = s_begin - tr.sample_base;       // local index of span[0]
        const S* src = pcm + tr.pcm_offset;
        for (int i = tid; i < span; i += kThreads) {
          const long long li = l0 + i;
          S v = (S)0;
          if (li >= 0 && li < tr.pcm_len) v = src[li];
          s_pcm[i] = v;
        }
      }
      __syncthreads();

      // ================= phase A: per-warp FFT -> magnitudes in warp scratch (+ optional stft/spec rows)
      {
        float2 v[32];
        if (N == 1024) {
          const int fa = 2 * warp, fb = 2 * warp + 1;
          if (fa < nfr) {
            const long long gA = g0 + fa, gB = g0 + fb;
            const bool haveB = fb < nfr;
            const S* xa = s_pcm + (int)(frame_start_sample(gA, p.hop, N, p.origin) - s_begin);
            const S* xb = haveB ? s_pcm + (int)(frame_start_sample(gB, p.hop, N, p.origin) - s_begin) : nullptr;
            load_frames_1024<S>(lane, v, xa, xb, s_win);
            core_pass1(lane, v, ex, s_tw1);
            __syncwarp();
            core_pass2_load(lane, v, ex);
            __syncwarp();
            core_pass2_fft(v);
            core_publish_upper(lane, v, ex, true);
            __syncwarp();
            BinSink oa, ob;
            const long long rowA = tr.out_row + (gA - tr.frame_first), rowB = rowA + 1;
            oa.mag = my_mag; ob.mag = my_mag + 512; oa.sgn_odd = ob.sgn_odd = sgn_odd;
            oa.log_mode = ob.log_mode = p.spec_log_mode; oa.mul = ob.mul = p.mul; oa.add = ob.add = p.add;
            oa.stft_row = (p.out_stft && gA >= ga) ? p.out_stft + rowA * p.ld_stft : nullptr;
            ob.stft_row = (p.out_stft && gB >= ga) ? p.out_stft + rowB * p.ld_stft : nullptr;
            oa.spec_row = (p.out_spec && gA >= ga) ? p.out_spec + rowA * p.ld_spec : nullptr;
            ob.spec_row = (p.out_spec && gB >= ga) ? p.out_spec + rowB * p.ld_spec : nullptr;
            split_1024x2(lane, v, ex, oa, ob, haveB);
          }
        } else if (warp < nfr) {
          const long long g = g0 + warp;
          const S* x = s_pcm + (int)(frame_start_sample(g, p.hop, N, p.origin) - s_begin);
          BinSink o;
          const long long row = tr.out_row + (g - tr.frame_first);
          o.mag = my_mag; o.sgn_odd = sgn_odd; o.log_mode = p.spec_log_mode; o.mul = p.mul; o.add = p.add;
          o.stft_row = (p.out_stft && g >= ga) ? p.out_stft + row * p.ld_stft : nullptr;
          o.spec_row = (p.out_spec && g >= ga) ? p.out_spec + row * p.ld_spec : nullptr;
          if (N == 2048) {
            load_frame_2048<S>(lane, v, x, s_win);
            core_pass1(lane, v, ex, s_tw1);
            __syncwarp();
            core_pass2_load(lane, v, ex);
            __syncwarp();
            core_pass2_fft(v);
            core_publish_upper(lane, v, ex, true);
            __syncwarp();
            split_2048(lane, v, ex, s_tw2k, o);
          } else {
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
              load_frame_4096<S>(lane, c, v, x, s_win, s_tw2k);
              core_pass1(lane, v, ex, s_tw1);
              __syncwarp();
              core_pass2_load(lane, v, ex);
              __syncwarp();
              core_pass2_fft(v);
              core_publish_upper(lane, v, ex, c == 0);
              __syncwarp();
              split_4096(lane, c, v, ex, s_tw2k, s_tw4k, o);
              __syncwarp();
            }
          }
        }
      }
      if (nb == 0) { __syncthreads(); continue; }
      __syncthreads();

      // ================= phase B: banded filterbank (FP32 FMA) + log into the ring
      for (int item = tid; item < FPR * nb; item += kThreads) {
        const int fi = item % FPR, b = item / FPR;
        if (fi >= nfr) continue;
        const float* mg = (N == 1024)
            ? reinterpret_cast<const float*>(smem + L.scratch + (fi >> 1) * warp_scratch_bytes(N)) + (fi & 1) * 512
            : reinterpret_cast<const float*>(smem + L.scratch + fi * warp_scratch_bytes(N) + (N == 4096 ? 8448 : 0));
        const int st = s_bstart[b], ln = s_blen[b];
        const float* w = s_w + s_bwoff[b];
        float acc = 0.0f;
        for (int i = 0; i < ln; ++i) acc = fmaf(mg[st + i], w[i], acc);
        const long long g = g0 + fi;
        if (p.out_filt && g >= ga) p.out_filt[(tr.out_row + (g - tr.frame_first)) * p.ld_filt + b] = acc;
        const float y = p.log_mode != 0 ? log_scaled(acc, p.log_mode, p.mul, p.add) : acc;
        s_ring[(int)(g & (kRing - 1)) * nb + b] = y;
      }
      __syncthreads();

      // ================= phase C: max-filter + lagged difference + flux, coalesced row stores
      for (int fi = warp; fi < nfr; fi += kWarps) {
        const long long g = g0 + fi;
        if (g < ga) continue;                                   // lead-in frame: ring only
        const long long row = tr.out_row + (g - tr.frame_first);
        const float* cur = s_ring + (int)(g & (kRing - 1)) * nb;
        if (p.out_log) {
          float* o = p.out_log + row * p.ld_log;
          for (int b = lane; b < nb; b += 32) o[b] = cur[b];
        }
        if (want_diff) {
          const bool has_prev = (g - p.diff_frames) >= 0;       // first d rows of a signal are zero
          const float* prev = s_ring + (int)((g - p.diff_frames) & (kRing - 1)) * nb;
          float sum = 0.0f;
          const int m = p.max_bins, half = m / 2;
          for (int b = lane; b < nb; b += 32) {
            float dv = 0.0f;
            if (has_prev) {
              float ref;
              if (m > 1) {                                       // scipy maximum_filter, mode='reflect'
                ref = -INFINITY;
                for (int j = 0; j < m; ++j) {
                  int q = b - half + j;
                  q = q < 0 ? -q - 1 : (q >= nb ? 2 * nb - q - 1 : q);
                  ref = fmaxf(ref, prev[q]);
                }
              } else {
                ref = prev[b];
              }
              dv = cur[b] - ref;
              if (p.positive) dv = fmaxf(dv, 0.0f);
            }
            if (p.out_diff) p.out_diff[row * p.ld_diff + b] = dv;
            sum += dv;
          }
          if (p.out_odf) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            if (lane == 0) p.out_odf[row] = sum;
          }
        }
      }
      __syncthreads();
    }
  }
}

template <int N, typename S>
static cudaError_t launch_one(const FrontendParams& p, int grid, int smem_bytes, cudaStream_t stream) {
  cudaError_t e = cudaFuncSetAttribute(frontend_kernel<N, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  if (e != cudaSuccess) return e;
  frontend_kernel<N, S><<<grid, warps_per_cta(N) * 32, smem_bytes, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_frontend(int N, int sample_bytes, const FrontendParams& p, int grid, int smem_bytes,
                            cudaStream_t stream) {
  if (sample_bytes == 2) {
    if (N == 1024) return launch_one<1024, int16_t>(p, grid, smem_bytes, stream);
    if (N == 2048) return launch_one<2048, int16_t>(p, grid, smem_bytes, stream);
    if (N == 4096) return launch_one<4096, int16_t>(p, grid, smem_bytes, stream);
  } else if (sample_bytes == 4) {
    if (N == 1024) return launch_one<1024, float>(p, grid, smem_bytes, stream);
    if (N == 2048) return launch_one<2048, float>(p, grid, smem_bytes, stream);
    if (N == 4096) return launch_one<4096, float>(p, grid, smem_bytes, stream);
  }
  return cudaErrorInvalidValue;
}

template <int N, typename S>
static cudaError_t occ_one(int smem_bytes, int* out) {
  cudaError_t e = cudaFuncSetAttribute(frontend_kernel<N, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  if (e != cudaSuccess) return e;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(out, frontend_kernel<N, S>, warps_per_cta(N) * 32, smem_bytes);
}

cudaError_t frontend_max_ctas_per_sm(int N, int sample_bytes, int smem_bytes, int* out) {
  if (sample_bytes == 2) {
    if (N == 1024) return occ_one<1024, int16_t>(smem_bytes, out);
    if (N == 2048) return occ_one<2048, int16_t>(smem_bytes, out);
    if (N == 4096) return occ_one<4096, int16_t>(smem_bytes, out);
  } else if (sample_bytes == 4) {
    if (N == 1024) return occ_one<1024, float>(smem_bytes, out);
    if (N == 2048) return occ_one<2048, float>(smem_bytes, out);
    if (N == 4096) return occ_one<4096, float>(smem_bytes, out);
  }
  return cudaErrorInvalidValue;
}

}  // namespace mmb
