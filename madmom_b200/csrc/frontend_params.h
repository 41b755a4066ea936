// Kernel parameter block shared by frontend_kernel.cu and capi.cu (internal, not part of the C ABI).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "work_list.h"      // Track, Chunk

namespace mmb {

// Per-warp filterbank schedule (built on the host, capi.cu): every band is cut into pieces; in pass p lane l
// accumulates   sum_i mag[start[p][l] + i] * w[p][i][l],  i < iters[p]   (weights zero-padded to iters[p], stored
// as float4 [p][i / 4][l][i % 4]; the
// starts of a pass are distinct modulo 32, so the reads are free of bank conflicts) and stores the partial sum at
// part[pid[p][l]] = part[q * b_pad + band]; the epilogue adds the pieces q = 0 .. max_q-1 of each band in order.
struct FbPass {
  int32_t iters;   // loop length of this pass (length of its longest piece)
  int32_t woff;    // offset (floats) of w[p][0][0] inside the weight table
};

// byte offsets of the kernel's shared-memory regions (computed once per plan on the host, smem_layout below)
struct SmemLayout {
  int ctl, win, tw1, tw2k, tw4k, fb_pass, fb_w, fb_start, fb_pid, ring, scratch, part, pcm, total;
  int ctl_stride;        // bytes of control block per warp group
  int ring_stride;       // bytes of log-spectrogram ring per warp group
  int part_stride;       // bytes of partial sums per warp; 0: they live in the upper half of the warp's exchange tile
  int pcm_stride;        // bytes of staged PCM span per warp group
};

constexpr int kWinCosFloats = 260;   // cosine-sum window constants: 2 x 32 float4 per-lane factors, then a0 (258 used)
constexpr int kMaxExtraDst = 7;      // destinations of one launch besides the local one (8 GPUs of a node)

struct FrontendParams {
  const void* pcm;               // int16 or float32 samples
  int64_t pcm_extent;            // elements of the PCM buffer that may be read (max over units of offset + len)
  int32_t bulk_allowed;          // PCM base is 16-byte aligned: rounds may be staged with cp.async.bulk
  const Track* tracks;
  const Chunk* chunks;           // [total_chunks] the work list of the launch; warp group v of V takes v, v + V, ...
  int32_t n_tracks;
  int32_t total_chunks;
  double hop;                    // frame i starts at (int64)(i * hop) - N/2 - origin   (signal.py:917-919)
  int32_t origin;
  int32_t span_cap;              // capacity (samples) of one staged PCM span in shared memory
  int32_t warps;                 // warps per CTA (blockDim.x / 32)
  int32_t group_warps;           // warps per group; a group walks its own chunks in rounds of frames_per_round frames
  SmemLayout L;
  // plan constants (device)
  const float* win;              // N floats, fft_window * 0.5; win_cos: the constants of frame_fft.cuh's *_wcos loads
  int32_t win_cos;               // the window is a symmetric cosine sum, evaluated in the kernel (N = 2048, 4096)
  const float2* tw1;             // 1024: W1024^(l*k1) at [tw1_index(k1, l)] (fft_core.cuh)
  const float2* tw2k;            // 1024: W2048^k (N = 2048 uses the first 512)
  const float2* tw4k;            // 512 : W4096^(2j+1)
  int32_t n_bands;               // 0 -> no filterbank stage
  int32_t n_passes, max_q, b_pad, w_floats;
  const FbPass* fb_pass;         // [n_passes]
  const float* fb_w;             // [w_floats]
  const int32_t* fb_start;       // [n_passes * 32]
  const int32_t* fb_pid;         // [n_passes * 32] q * b_pad + band, max_q * b_pad for idle lanes
  float mul, add;
  int32_t log_mode;              // 0 none, 1 log10, 2 mul/add only, 3 ln, 4 log2
  int32_t diff_frames;           // d >= 1 when a difference is requested
  int32_t max_bins;              // frequency max-filter width (<=1: none)
  int32_t positive;
  int32_t circular_shift;
  int32_t spec_log_mode;         // log mode applied to the spec output (plans without a filterbank)
  int32_t ring_rows;             // rows of the log-spectrogram ring: 2 * frames_per_round + diff_frames
  // outputs (device, row-major, leading dimensions in elements; null -> not produced)
  float2* out_stft; int64_t ld_stft;
  float* out_spec;  int64_t ld_spec;
  float* out_filt;  int64_t ld_filt;
  float* out_log;   int64_t ld_log;
  float* out_diff;  int64_t ld_diff;
  float* out_odf;
  float* out_lgd;   int64_t ld_lgd;   // |local group delay| / pi (ComplexFlux), N = 2048 only
  // fused all-gather (mmb_run_gather): the rows of log / diff / odf are also stored to n_extra further destinations
  // (peer memory of the other ranks over NVLink), same leading dimensions as the local matrices
  int32_t n_extra;
  int32_t multicast;             // out_log / out_diff / out_odf are NVLS multicast addresses: stored with multimem.st
  int32_t gather_vec;            // gathered rows may be stored 16 bytes per lane: bands, leading dimensions and all
                                 // destination pointers are multiples of four floats / 16 bytes (checked on the host)
  float* x_log[kMaxExtraDst];
  float* x_diff[kMaxExtraDst];
  float* x_odf[kMaxExtraDst];
};

constexpr int kMaxWarps = 16;
constexpr int kMaxDiffFrames = 16;

// frames one round of a warp group produces
// (N = 1024: two frames per warp, packed into one complex FFT; N = 4096: two warps per frame, one per half of the
// radix-2 pre-stage)
__host__ __device__ constexpr int frames_per_round(int N, int group_warps) {
  return N == 1024 ? 2 * group_warps : (N == 4096 ? group_warps / 2 : group_warps);
}
// per-warp scratch (bytes): exchange tile with the magnitudes aliased in its lower 4 KB.  N = 4096: the two warps of
// a frame keep the even bins (warp 2j) and the odd bins (warp 2j + 1) in their own tiles; the scratch is 16 floats
// longer so that the two arrays lie 16 banks apart and even / odd bins of neighbouring lanes fall into different
// banks.
// the exchange tile above the 4 KB of magnitudes is idle between the split and the next frame's FFT: the partial
// sums of the filterbank may live there when that buys another warp (they must then be cleared per frame)
constexpr int kPartAliasOffset = 4096, kPartAliasBytes = 8448 - 4096;
__host__ __device__ constexpr int warp_scratch_bytes(int N) { return 8448 + (N == 4096 ? 64 : 16); }
// warps that share one set of filterbank partial sums
__host__ __device__ constexpr int part_sharers(int N) { return N == 4096 ? 2 : 1; }

// one descriptor per round of a group, written by one lane, read by the group (4 slots, see frontend_kernel.cu)
struct RoundDesc {
  long long g0;          // first (global) frame of the round
  long long ga;          // first output frame of the chunk (frames below are lead-in)
  long long row0;        // output row of global frame g is row0 + g
  long long s_begin;     // global sample index of the first staged sample
  long long l0;          // local sample index of the first staged sample
  long long pcm_offset, pcm_len, sample_base;
  int nfr;               // frames in this round (0: no more work)
  int span;              // staged samples
  int bulk;              // staged by cp.async.bulk (else by the threads, with zero fill)
  int shift;             // elements between the start of the staging buffer and sample s_begin
  int slot0;             // ring slot of frame g0
  int chunk, unit, pad;
  long long gb;          // end of the chunk's output range
};

// control block of one warp group
struct GroupCtl {
  RoundDesc desc[4];         // rounds r-1 (its phase C may be pending), r, r+1 (being written) live at the same time
  unsigned long long mbar;   // phase r completes when the descriptor of round r is written and its PCM span staged
  unsigned long long ring_bar;   // phase r completes when every warp of the group has written its ring rows of round r
  int loaded;                // warps of the group that have read their frames of the current round
  int described;             // last round whose descriptor is complete
};

__host__ __device__ inline int mmb_align16(int x) { return (x + 15) & ~15; }

// `ring_rows`, `span_cap` are per group
inline SmemLayout smem_layout(int N, int warps, int group_warps, int n_bands, int n_passes, int part_floats,
                              int w_floats, int ring_rows, int span_cap, int sample_bytes, bool alias_part) {
  SmemLayout L;
  const int groups = warps / group_warps;
  int o = 0;
  L.ctl_stride = mmb_align16((int)sizeof(GroupCtl));
  L.ctl = o;         o += groups * L.ctl_stride;
  L.win = o;         o += mmb_align16(N * 4);
  L.tw1 = o;         o += 1024 * 8;
  L.tw2k = o;        o += (N == 1024 ? 0 : (N == 2048 ? 512 * 8 : 1024 * 8));
  L.tw4k = o;        o += (N == 4096 ? 512 * 8 : 0);
  L.fb_pass = o;     o += mmb_align16(n_passes * (int)sizeof(FbPass));
  L.fb_w = o;        o += mmb_align16(w_floats * 4);
  L.fb_start = o;    o += mmb_align16(n_passes * 32 * 4);
  L.fb_pid = o;      o += mmb_align16(n_passes * 32 * 4);
  L.ring_stride = mmb_align16(n_bands > 0 ? ring_rows * n_bands * 4 : 0);
  L.ring = o;        o += groups * L.ring_stride;
  L.scratch = o;     o += warps * warp_scratch_bytes(N);
  L.part_stride = mmb_align16((part_floats + 1) * 4 * (N == 1024 ? 2 : 1));
  // (a pair of warps of a 4096-point frame shares one private set: clearing an aliased one would race with the
  //  partner's epilogue of the frame before)
  if (alias_part && N != 4096 && L.part_stride <= kPartAliasBytes) L.part_stride = 0;
  L.part = o;        o += n_bands > 0 ? (warps / part_sharers(N)) * L.part_stride : 0;
  L.pcm_stride = mmb_align16(span_cap * sample_bytes);
  L.pcm = o;         o += groups * L.pcm_stride;
  L.total = o;
  return L;
}

// parameters of the float64 STFT kernel (stft64.cu): the launches that emit raw bins / magnitudes
struct Stft64Params {
  const void* pcm;
  const Track* tracks;
  const int32_t* chunk_prefix;
  int32_t n_tracks, chunk_frames;
  double hop;
  int32_t origin, circular_shift, include_nyquist, spec_log_mode;
  float mul, add;
  const double* win;             // N doubles: fft_window as the reference multiplies it (float64, unscaled)
  const double2* tw;             // N / 2 doubles2: W_N^k, then the compact per-pass and split copies (capi.cu)
  float2* out_stft; int64_t ld_stft;
  float* out_spec;  int64_t ld_spec;
};

// records the message mmb_last_error() returns on the calling thread; returns `code` (capi.cu)
int set_error(int code, const char* msg);

cudaError_t launch_frontend(int N, int sample_bytes, const FrontendParams& p, int grid, int smem_bytes,
                            cudaStream_t stream);
cudaError_t launch_stft64(int N, int sample_bytes, const Stft64Params& p, int grid, cudaStream_t stream);
cudaError_t frontend_set_smem(int N, int sample_bytes, int smem_bytes);
// interleaved (num_frames, channels) int16 / float32 -> mono of the same type: mean over the channels with the
// reference's rounding (np.mean(...).astype(dtype), signal.py:212), or one channel verbatim (channel >= 0)
cudaError_t launch_remix_mono(const void* in, int sample_bytes, int64_t num_frames, int channels, int channel,
                              void* out, cudaStream_t stream);

}  // namespace mmb
