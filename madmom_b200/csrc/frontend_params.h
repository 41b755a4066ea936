// Kernel parameter block shared by frontend_kernel.cu and capi.cu (internal, not part of the C ABI).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace mmb {

// One unit of independent work: a track (batch mode) or a frame range of a long signal (shard mode).
// Mirrors mmb_track in include/madmom_b200.h.
struct Track {
  int64_t pcm_offset;    // element offset of local sample 0 inside the PCM buffer
  int64_t pcm_len;       // number of local samples; anything outside reads as zero (signal.py:928-962)
  int64_t sample_base;   // global sample index of local sample 0 (0 for whole tracks)
  int64_t frame_first;   // global index of the first output frame
  int64_t frame_count;   // number of output rows
  int64_t out_row;       // first output row of this unit
};

struct FrontendParams {
  const void* pcm;               // int16 or float32 samples
  const Track* tracks;
  const int32_t* chunk_prefix;   // [n_tracks + 1] prefix sum of chunks per track
  int32_t n_tracks;
  int32_t total_chunks;
  int32_t chunk_frames;          // output frames per chunk
  double hop;                    // frame i starts at (int64)(i * hop) - N/2 - origin   (signal.py:917-919)
  int32_t origin;
  int32_t span_cap;              // capacity (samples) of the staged PCM span in shared memory
  // plan constants (device)
  const float* win;              // N floats, fft_window * 0.5
  const float2* tw1;             // 1024: W1024^(l*k1) at [k1*32 + l]
  const float2* tw2k;            // 1024: W2048^k
  const float2* tw4k;            // 512 : W4096^(2j+1)
  int32_t n_bands;               // 0 -> no filterbank stage
  int32_t nnz;
  const int32_t* band_start;
  const int32_t* band_len;
  const int32_t* band_woff;
  const float* weights;
  float mul, add;
  int32_t log_mode;              // 0 none, 1 log10, 2 mul/add only, 3 ln, 4 log2
  int32_t diff_frames;           // d >= 1 when a difference is requested
  int32_t max_bins;              // frequency max-filter width (<=1: none)
  int32_t positive;
  int32_t circular_shift;
  int32_t spec_log_mode;         // log mode applied to the spec output (plans without a filterbank)
  // outputs (device, row-major, leading dimensions in elements; null -> not produced)
  float2* out_stft; int64_t ld_stft;
  float* out_spec;  int64_t ld_spec;
  float* out_filt;  int64_t ld_filt;
  float* out_log;   int64_t ld_log;
  float* out_diff;  int64_t ld_diff;
  float* out_odf;
};

// warps per CTA: one CTA per SM; 16 warps (4 per scheduler) where the per-warp scratch allows it
__host__ __device__ constexpr int warps_per_cta(int N) { return N == 2048 ? 16 : 8; }
constexpr int kRing = 32;          // rows of the log-spectrogram ring (>= frames per round + max diff_frames)
constexpr int kMaxDiffFrames = 16;

// frames one CTA round produces
__host__ __device__ constexpr int frames_per_round(int N) { return N == 1024 ? 2 * warps_per_cta(N) : warps_per_cta(N); }
// per-warp scratch (bytes): exchange tile (+ magnitudes aliased in its lower 4 KB; separate 8 KB for 4096)
__host__ __device__ constexpr int warp_scratch_bytes(int N) { return (N == 4096 ? 8448 + 8192 : 8448) + 16; }

struct SmemLayout {
  int win, tw1, tw2k, tw4k, fb_meta, fb_w, ring, scratch, pcm, total;
};

__host__ __device__ inline int mmb_align16(int x) { return (x + 15) & ~15; }

__host__ __device__ inline SmemLayout smem_layout(int N, int n_bands, int nnz, int span_cap, int sample_bytes) {
  SmemLayout L;
  int o = 0;
  L.win = o;     o += mmb_align16(N * 4);
  L.tw1 = o;     o += 1024 * 8;
  L.tw2k = o;    o += (N == 1024 ? 0 : 1024 * 8);
  L.tw4k = o;    o += (N == 4096 ? 512 * 8 : 0);
  L.fb_meta = o; o += mmb_align16(n_bands * 3 * 4);
  L.fb_w = o;    o += mmb_align16(nnz * 4);
  L.ring = o;    o += mmb_align16(n_bands > 0 ? kRing * n_bands * 4 : 0);
  L.scratch = o; o += warps_per_cta(N) * warp_scratch_bytes(N);
  L.pcm = o;     o += mmb_align16(span_cap * sample_bytes);
  L.total = o;
  return L;
}

cudaError_t launch_frontend(int N, int sample_bytes, const FrontendParams& p, int grid, int smem_bytes,
                            cudaStream_t stream);
cudaError_t frontend_max_ctas_per_sm(int N, int sample_bytes, int smem_bytes, int* out);

}  // namespace mmb
