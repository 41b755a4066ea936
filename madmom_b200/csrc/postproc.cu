// Stages either side of the fused front-end kernel that work on matrices / vectors already in HBM:
//
//   mmb_matrix_diff   SpectrogramDifference of an arbitrary (rows, cols) float32 matrix (used for unfiltered
//                     spectrograms, whose 1024..2048 columns do not fit the fused kernel's shared-memory ring):
//                     scipy maximum_filter(size=(1, m), mode='reflect') on the lagged row, subtraction, optional
//                     half-wave rectification, optional row sums.   madmom/audio/spectrogram.py:976-993,
//                     madmom/features/onsets.py:163-240.
//   mmb_peak_picking  peak_picking() of madmom/features/onsets.py:875-963 on one or many activation functions:
//                     optional Hamming smoothing, moving average (scipy uniform_filter, mode='constant'),
//                     threshold, moving maximum (scipy maximum_filter, mode='constant'), equality test, ordered
//                     compaction of the surviving frame indices.
//
// Both are HBM-bound stencils: every element is read once from DRAM (neighbours come from L1/L2).
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#include <mutex>

#include "../../include/madmom_b200.h"
#include "frontend_params.h"

namespace {

// ---------------------------------------------------------------------------------------------- matrix diff
// one warp per output row; lanes stride over columns
__global__ void __launch_bounds__(256) matrix_diff_kernel(const float* __restrict__ in, int64_t ld_in,
                                                          float* __restrict__ out, int64_t ld_out,
                                                          float* __restrict__ odf, int64_t rows, int32_t cols,
                                                          int64_t seg_rows, int32_t d, int32_t max_bins,
                                                          int32_t positive) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int64_t local = seg_rows > 0 ? row % seg_rows : row;     // frame index inside its own signal
  const bool has_prev = local >= d;
  const float* cur = in + row * ld_in;
  const float* prev = in + (row - d) * ld_in;
  const int half = max_bins / 2;
  float sum = 0.0f;
  for (int c = lane; c < cols; c += 32) {
    float v = 0.0f;
    if (has_prev) {
      float ref;
      if (max_bins > 1) {
        ref = -INFINITY;
        for (int j = 0; j < max_bins; ++j) {
          int q = c - half + j;
          q = q < 0 ? -q - 1 : (q >= cols ? 2 * cols - q - 1 : q);     // 'reflect': (d c b a | a b c d | d c b a)
          ref = fmaxf(ref, prev[q]);
        }
      } else {
        ref = prev[c];
      }
      v = cur[c] - ref;
      if (positive) v = fmaxf(v, 0.0f);
    }
    if (out) out[row * ld_out + c] = v;
    sum += v;
  }
  if (odf) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) odf[row] = sum;
  }
}

// ---------------------------------------------------------------------------------------------- peak picking
struct PeakParams {
  const float* act;
  const int64_t* seg;      // [n_seg + 1] element offsets of the segments, or null (one segment [0, n))
  int64_t n;
  int32_t n_seg;
  double threshold;
  int32_t smooth, pre_avg, post_avg, pre_max, post_max;
  const double* kernel;    // Hamming window of length `smooth` (device), when smooth > 1
  int32_t* peaks;          // [n]   segment s writes its local indices at peaks[seg[s] ...]
  int32_t* counts;         // [n_seg]
};

__device__ __forceinline__ int find_segment(const int64_t* seg, int n_seg, int64_t i) {
  int lo = 0, hi = n_seg;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (seg[mid] <= i) lo = mid; else hi = mid;
  }
  return lo;
}

// activation value after the optional smoothing: np.convolve(x, hamming(k), 'same') evaluated in float64
template <typename T>
__device__ __forceinline__ T smoothed(const PeakParams& p, int64_t s0, int64_t len, int64_t i) {
  if (p.smooth <= 1) return (T)p.act[s0 + i];
  // 'same' output i <-> full output i + (k - 1) / 2;   full[m] = sum_j x[m - j] k[j]
  const int k = p.smooth;
  const int64_t m = i + (k - 1) / 2;
  double acc = 0.0;
  for (int j = 0; j < k; ++j) {
    const int64_t q = m - j;
    if (q >= 0 && q < len) acc += (double)p.act[s0 + q] * p.kernel[j];
  }
  return (T)acc;
}

// detections[i] = act[i] * (act[i] >= mov_avg[i] + threshold)          (onsets.py:927-942)
template <typename T>
__device__ __forceinline__ T detection(const PeakParams& p, int64_t s0, int64_t len, int64_t i) {
  if (i < 0 || i >= len) return (T)0;                                   // mode='constant', cval 0
  const T a = smoothed<T>(p, s0, len, i);
  const int avg_len = p.pre_avg + p.post_avg + 1;
  T limit;
  if (avg_len > 1) {
    // scipy uniform_filter1d: window [i - n/2 - origin, ... + n), origin = floor((pre - post) / 2), zeros outside,
    // accumulated in double, divided by n, stored in the array's dtype
    const int origin = (int)floor((double)(p.pre_avg - p.post_avg) / 2.0);
    const int64_t w0 = i - avg_len / 2 - origin;
    double acc = 0.0;
    for (int j = 0; j < avg_len; ++j) {
      const int64_t q = w0 + j;
      if (q >= 0 && q < len) acc += (double)smoothed<T>(p, s0, len, q);
    }
    const T avg = (T)(acc / (double)avg_len);
    limit = avg + (T)p.threshold;
  } else {
    limit = (T)p.threshold;
  }
  return a >= limit ? a : (T)0;
}

template <typename T>
__global__ void __launch_bounds__(256) peak_flag_kernel(const PeakParams p) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.n) return;
  int64_t s0 = 0, len = p.n;
  if (p.seg) {
    const int s = find_segment(p.seg, p.n_seg, i);
    s0 = p.seg[s]; len = p.seg[s + 1] - s0;
  }
  const int64_t li = i - s0;
  const T det = detection<T>(p, s0, len, li);
  int flag = det != (T)0;
  const int max_len = p.pre_max + p.post_max + 1;
  if (flag && max_len > 1) {
    const int origin = (int)floor((double)(p.pre_max - p.post_max) / 2.0);
    const int64_t w0 = li - max_len / 2 - origin;
    T mx = det;
    for (int j = 0; j < max_len; ++j) mx = fmax(mx, detection<T>(p, s0, len, w0 + j));
    flag = det == mx;
  }
  p.peaks[i] = flag;
}

// ordered in-place compaction: one CTA per segment walks its flags in tiles of blockDim.x
__global__ void __launch_bounds__(1024) peak_compact_kernel(const PeakParams p) {
  __shared__ int warp_tot[32];
  __shared__ int base;
  const int s = blockIdx.x;
  const int64_t s0 = p.seg ? p.seg[s] : 0;
  const int64_t len = (p.seg ? p.seg[s + 1] : p.n) - s0;
  int32_t* buf = p.peaks + s0;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (threadIdx.x == 0) base = 0;
  __syncthreads();
  for (int64_t t0 = 0; t0 < len; t0 += blockDim.x) {
    const int64_t i = t0 + threadIdx.x;
    const int f = (i < len) ? buf[i] : 0;
    const unsigned m = __ballot_sync(0xffffffffu, f != 0);
    if (lane == 0) warp_tot[warp] = __popc(m);
    __syncthreads();                       // every flag of the tile is in a register before anything is written
    int off = base;
    for (int w = 0; w < warp; ++w) off += warp_tot[w];
    if (f) buf[off + __popc(m & ((1u << lane) - 1u))] = (int32_t)i;
    __syncthreads();
    if (threadIdx.x == 0) {
      int tot = 0;
      for (int w = 0; w < nw; ++w) tot += warp_tot[w];
      base += tot;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) p.counts[s] = base;
}

// ---------------------------------------------------------------------------------------------- complex flux
// One warp per frame.  (1) temporal maximum of |lgd| over `tf` frames (scipy maximum_filter size [tf, 1],
// origin 0, mode 'reflect' inside the frame's own signal) into a shared row, (2) per band the minimum over the
// bins [lo, hi) the band touches plus one bin either side, (3) odf = sum_b diff[b] * mask[b].
// madmom/features/onsets.py:282-317.
// One warp per frame, CTAs stride over the rows.  Per frame: (1) temporal maximum of the three neighbouring lgd
// rows into a shared-memory row, coalesced; (2) band minima: every band's bin range is cut into pieces of kCfPiece
// bins (table built once per CTA), one piece per lane and pass, combined with an integer atomicMin in shared memory
// (the values are non-negative floats, so their bit patterns order like integers); (3) sum_b diff[b] * min[b].
constexpr int kCfPiece = 8;
constexpr int kCfMaxPieces = 1024;

__global__ void __launch_bounds__(256) complex_flux_kernel(const float* __restrict__ lgd, int64_t ld_lgd,
                                                           const float* __restrict__ diff, int64_t ld_diff,
                                                           const int32_t* __restrict__ band_lo,
                                                           const int32_t* __restrict__ band_hi, int32_t n_bands,
                                                           int32_t n_bins, int64_t rows, int64_t seg_rows,
                                                           const int64_t* __restrict__ seg, int32_t n_seg, int32_t tf,
                                                           float* __restrict__ odf) {
  extern __shared__ float s_dyn[];                        // [warps][n_bins] rows, [warps][n_bands] minima, tables
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
  float* s_rows = s_dyn;
  int* s_min = reinterpret_cast<int*>(s_rows + (size_t)warps * n_bins);
  int* s_lo = s_min + warps * n_bands;
  int* s_hi = s_lo + n_bands;
  int* s_off = s_hi + n_bands;                            // [n_bands + 1] first piece of each band
  int* s_piece = s_off + n_bands + 1;                     // [kCfMaxPieces] band | piece index << 16
  for (int b = threadIdx.x; b < n_bands; b += blockDim.x) { s_lo[b] = band_lo[b]; s_hi[b] = band_hi[b]; }
  __syncthreads();
  if (threadIdx.x == 0) {
    int o = 0;
    for (int b = 0; b < n_bands; ++b) {
      s_off[b] = o;
      const int len = s_hi[b] - s_lo[b];
      o += len > 0 ? (len + kCfPiece - 1) / kCfPiece : 0;
    }
    s_off[n_bands] = o;
  }
  __syncthreads();
  // band layouts with more pieces than the table holds (heavily overlapping bands) take the serial path below
  const bool piecewise = s_off[n_bands] <= kCfMaxPieces && n_bands <= 0xffff;
  const int n_pieces = piecewise ? s_off[n_bands] : 0;
  if (piecewise)
    for (int b = threadIdx.x; b < n_bands; b += blockDim.x)
      for (int j = 0; j < s_off[b + 1] - s_off[b]; ++j) s_piece[s_off[b] + j] = b | (j << 16);
  __syncthreads();

  float* mx = s_rows + (size_t)warp * n_bins;
  int* mn = s_min + warp * n_bands;
  const int half = tf / 2;
  for (int64_t row = (int64_t)blockIdx.x * warps + warp; row < rows; row += (int64_t)gridDim.x * warps) {
    int64_t s0 = 0, len = rows;
    if (seg) {
      const int s = find_segment(seg, n_seg, row);
      s0 = seg[s]; len = seg[s + 1] - s0;
    } else if (seg_rows > 0) {
      s0 = (row / seg_rows) * seg_rows;
      len = seg_rows < rows - s0 ? seg_rows : rows - s0;     // the last signal may be shorter
    }
    const int64_t t = row - s0;
    // rows of the temporal neighbourhood ('reflect': (b a | a b c))
    int64_t q_prev = t - 1, q_next = t + 1;
    if (q_prev < 0) q_prev = -q_prev - 1;
    if (q_next >= len) q_next = 2 * len - q_next - 1;
    q_prev = q_prev < 0 ? 0 : (q_prev >= len ? len - 1 : q_prev);
    q_next = q_next < 0 ? 0 : (q_next >= len ? len - 1 : q_next);
    const float* r0 = lgd + row * ld_lgd;
    if (tf == 3) {
      const float* ra = lgd + (s0 + q_prev) * ld_lgd;
      const float* rb = lgd + (s0 + q_next) * ld_lgd;
      for (int k = lane; k < n_bins; k += 32) mx[k] = fmaxf(fmaxf(ra[k], r0[k]), rb[k]);
    } else {
      for (int k = lane; k < n_bins; k += 32) {
        float m = r0[k];
        for (int j = -half; j <= half; ++j) {
          if (j == 0 || tf <= 1) continue;
          int64_t q = t + j;
          if (q < 0) q = -q - 1;
          if (q >= len) q = 2 * len - q - 1;
          q = q < 0 ? 0 : (q >= len ? len - 1 : q);
          m = fmaxf(m, lgd[(s0 + q) * ld_lgd + k]);
        }
        mx[k] = m;
      }
    }
    for (int b = lane; b < n_bands; b += 32) mn[b] = 0x7f800000;        // +inf
    __syncwarp();
    for (int pi = lane; pi < n_pieces; pi += 32) {
      const int code = s_piece[pi], b = code & 0xffff, j = code >> 16;
      const int lo = s_lo[b] + j * kCfPiece, hi = min(lo + kCfPiece, s_hi[b]);
      float m = mx[lo];
      for (int k = lo + 1; k < hi; ++k) m = fminf(m, mx[k]);
      atomicMin(&mn[b], __float_as_int(m));
    }
    if (!piecewise)
      for (int b = lane; b < n_bands; b += 32) {
        float m = INFINITY;
        for (int k = s_lo[b]; k < s_hi[b]; ++k) m = fminf(m, mx[k]);
        mn[b] = __float_as_int(m);
      }
    __syncwarp();
    float sum = 0.0f;
    for (int b = lane; b < n_bands; b += 32) {
      const float m = s_hi[b] > s_lo[b] ? __int_as_float(mn[b]) : 0.0f;
      sum = __fmaf_rn(diff[row * ld_diff + b], m, sum);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) odf[row] = sum;
    __syncwarp();                                          // the row buffers are reused by the next frame
  }
}

// Streaming variant for temporal_filter = 3 and n_bins = 128 NV: a warp walks a contiguous run of rows and keeps
// the previous, the current and the next lgd row in registers (float4 chunks, NV per lane), so every row is read
// from HBM once instead of three times, with the row after next already in flight while the band minima of the
// current row are taken.  The temporal maximum goes to a shared-memory row; lane l owns the bands l, l + 32, ...
// and walks their bin ranges itself (neighbouring bands have similar widths, so the uniform trip count of a
// group of 32 bands wastes little), which needs no atomics and no table of minima.
template <int NV>
__device__ __forceinline__ void cf_load_row(float4 (&r)[NV], const float* __restrict__ row, int lane) {
  const float4* p = reinterpret_cast<const float4*>(row);
#pragma unroll
  for (int j = 0; j < NV; ++j) r[j] = __ldg(p + lane + 32 * j);
}
__device__ __forceinline__ float4 fmax4(float4 a, float4 b) {
  return make_float4(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z), fmaxf(a.w, b.w));
}

template <int NV>
__global__ void __launch_bounds__(256, 2) complex_flux_stream_kernel(
    const float* __restrict__ lgd, int64_t ld_lgd, const float* __restrict__ diff, int64_t ld_diff,
    const int32_t* __restrict__ band_lo, const int32_t* __restrict__ band_hi, int32_t n_bands, int64_t rows,
    int64_t seg_rows, const int64_t* __restrict__ seg, int32_t n_seg, int64_t run, float* __restrict__ odf) {
  extern __shared__ float s_dyn[];                        // [warps][128 NV] maxima, band tables
  constexpr int n_bins = 128 * NV;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
  int* s_lo = reinterpret_cast<int*>(s_dyn + (size_t)warps * n_bins);
  int* s_len = s_lo + n_bands;
  for (int b = threadIdx.x; b < n_bands; b += blockDim.x) {
    const int lo = band_lo[b], hi = band_hi[b];
    s_lo[b] = lo; s_len[b] = hi > lo ? hi - lo : 0;
  }
  __syncthreads();
  float* mx = s_dyn + (size_t)warp * n_bins;
  const int64_t first = ((int64_t)blockIdx.x * warps + warp) * run;
  if (first >= rows) return;
  const int64_t stop = first + run < rows ? first + run : rows;

  // [sig_lo, sig_hi): rows of the signal the current row belongs to
  int64_t sig_lo = 0, sig_hi = rows;
  int si = 0;
  if (seg) {
    si = find_segment(seg, n_seg, first);
    sig_lo = seg[si]; sig_hi = seg[si + 1];
  } else if (seg_rows > 0) {
    sig_lo = (first / seg_rows) * seg_rows; sig_hi = sig_lo + seg_rows;
  }
  float4 P[NV], A[NV], B[NV];
  cf_load_row<NV>(A, lgd + first * ld_lgd, lane);
  if (first > sig_lo) cf_load_row<NV>(P, lgd + (first - 1) * ld_lgd, lane);
  else {
#pragma unroll
    for (int j = 0; j < NV; ++j) P[j] = A[j];
  }
  if (first + 1 < rows) cf_load_row<NV>(B, lgd + (first + 1) * ld_lgd, lane);

  for (int64_t row = first; row < stop; ++row) {
    if (row >= sig_hi) {                                  // first row of the next signal
      if (seg) {
        do { ++si; } while (seg[si + 1] <= row);          // empty signals have no rows
        sig_lo = seg[si]; sig_hi = seg[si + 1];
      } else {
        sig_lo = sig_hi; sig_hi = sig_lo + seg_rows;
      }
    }
    // 'reflect': the row itself stands in for its neighbour; the last signal may be shorter than seg_rows
    const bool last = row + 1 >= sig_hi || row + 1 >= rows;
    float4* m4 = reinterpret_cast<float4*>(mx);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      float4 m = fmax4(P[j], A[j]);
      if (!last) m = fmax4(m, B[j]);
      m4[lane + 32 * j] = m;
    }
    // shift the window and put the row after next in flight
#pragma unroll
    for (int j = 0; j < NV; ++j) { P[j] = last ? B[j] : A[j]; A[j] = B[j]; }
    if (row + 2 < rows && row + 1 < stop) cf_load_row<NV>(B, lgd + (row + 2) * ld_lgd, lane);
    __syncwarp();
    float sum = 0.0f;
    const float* drow = diff + row * ld_diff;
    for (int b0 = 0; b0 < n_bands; b0 += 32) {
      const int b = b0 + lane;
      const bool have = b < n_bands;
      const int lo = have ? s_lo[b] : 0, len = have ? s_len[b] : 0;
      const float d = have ? drow[b] : 0.0f;
      const int glen = __reduce_max_sync(0xffffffffu, len);
      float m = INFINITY;
      const int top = lo + (len > 0 ? len - 1 : 0);
      for (int k = 0; k < glen; ++k) m = fminf(m, mx[min(lo + k, top)]);
      sum = __fmaf_rn(d, len > 0 ? m : 0.0f, sum);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) odf[row] = sum;
    __syncwarp();                                         // the maxima are overwritten by the next row
  }
}

// Cover variant (round 2), the one mmb_complex_flux launches for temporal_filter = 3 and 512 / 1024 bins: the
// streaming kernel above with the serial walk over a band's bins (up to 47 dependent shared-memory reads per band of
// the 24-bands-per-octave bank, 93 loop iterations per row) replaced by a hierarchical minimum.
//   * A lane's float4 of maxima gives the minimum of its aligned 4-bin block for free, two shuffles give the aligned
//     16-bin block; maxima and block minima go to the warp's shared row S = [bins | 4-blocks | 16-blocks].
//   * Every band's bin range [lo, hi) is covered once, when the CTA starts, by the largest aligned blocks that fit
//     (at most 3 + 3 + n16 + 3 + 3 entries; 3 - 14 for that bank) and stored as 16-bit byte offsets into S, kCfcCover
//     per band, unused entries repeating the last one (the minimum does not mind), 48 bytes per band so that the
//     lanes' 16-byte table reads fall into different banks.  Lane l owns the bands l, l + 32, ... as before; a group
//     of 32 bands makes as many reads as its longest cover has entries (5 + 5 + 8 + 10 + 14 = 42 for that bank).
//   * Same summation order as the other two kernels: results are bit-identical
//     (tests/test_gpu_processors.py::test_complex_flux_kernel_ragged_signals compares the three).
// A band whose cover does not fit (wider than about 200 bins) makes the CTA fall back to the serial walk.
// A CTA-per-run variant (thread = four bins, three rows in 12 registers, 32 warps per SM, two __syncthreads per
// row) was built first and measured at 3.5 - 5.3 ms per 1.536 M rows against 2.6 ms for the streaming kernel: ncu
// shows 7.9 of 8 warps per scheduler waiting at the barriers (5 of 8 warps take the band minima, 1 of 8 adds the row
// up) -- the warp-synchronous layout is the better one (profiles/r2a_sensitivity.txt).
constexpr int kCfcCover = 24;                   // cover entries per band (48 bytes)
constexpr int kCfcMaxGroups = 8;                // groups of 32 bands: at most 256 bands

template <int NV, int NG>                        // NV float4 per lane and row, NG >= groups of 32 bands
__global__ void __launch_bounds__(256, 2) complex_flux_cover_kernel(
    const float* __restrict__ lgd, int64_t ld_lgd, const float* __restrict__ diff, int64_t ld_diff,
    const int32_t* __restrict__ band_lo, const int32_t* __restrict__ band_hi, int32_t n_bands, int64_t rows,
    int64_t seg_rows, const int64_t* __restrict__ seg, int32_t n_seg, int64_t run, float* __restrict__ odf) {
  extern __shared__ float4 s_dyn4[];            // [warps][S], cover table [groups * 32][24] u16, group lengths, flag
  constexpr int n_bins = 128 * NV, o4 = n_bins, o16 = o4 + n_bins / 4, s_words = o16 + n_bins / 16;
  float* s_dyn = reinterpret_cast<float*>(s_dyn4);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
  const int groups = (n_bands + 31) >> 5;
  uint16_t* s_cover = reinterpret_cast<uint16_t*>(s_dyn + (size_t)warps * s_words);
  int* s_glen = reinterpret_cast<int*>(s_cover + (size_t)groups * 32 * kCfcCover);     // [groups] longest cover
  int* s_lo = s_glen + groups;                                                          // serial fallback only
  int* s_len = s_lo + n_bands;
  int* s_flag = s_len + n_bands;
  for (int g = threadIdx.x; g < groups; g += blockDim.x) s_glen[g] = 0;
  if (threadIdx.x == 0) s_flag[0] = 0;
  __syncthreads();
  for (int b = threadIdx.x; b < groups * 32; b += blockDim.x) {
    const int lo = b < n_bands ? band_lo[b] : 0, hi = b < n_bands ? band_hi[b] : 0;
    if (b < n_bands) { s_lo[b] = lo; s_len[b] = hi > lo ? hi - lo : 0; }
    uint16_t* c = s_cover + (size_t)b * kCfcCover;
    int n = 0, last = hi > lo ? lo : 0;
    for (int k = lo; k < hi; ++n) {
      int off, step;
      if ((k & 15) == 0 && k + 16 <= hi) { off = o16 + (k >> 4); step = 16; }
      else if ((k & 3) == 0 && k + 4 <= hi) { off = o4 + (k >> 2); step = 4; }
      else { off = k; step = 1; }
      if (n < kCfcCover) c[n] = (uint16_t)(4 * off);
      last = off;
      k += step;
    }
    if (n > kCfcCover) s_flag[0] = 1;
    for (int j = n; j < kCfcCover; ++j) c[j] = (uint16_t)(4 * last);
    atomicMax(&s_glen[b >> 5], n < kCfcCover ? n : kCfcCover);
  }
  __syncthreads();
  const bool serial = s_flag[0] != 0;
  float* mx = s_dyn + (size_t)warp * s_words;
  const int64_t first = ((int64_t)blockIdx.x * warps + warp) * run;
  if (first >= rows) return;
  const int64_t stop = first + run < rows ? first + run : rows;

  // [sig_lo, sig_hi): rows of the signal the current row belongs to
  int64_t sig_lo = 0, sig_hi = rows;
  int si = 0;
  if (seg) {
    si = find_segment(seg, n_seg, first);
    sig_lo = seg[si]; sig_hi = seg[si + 1];
  } else if (seg_rows > 0) {
    sig_lo = (first / seg_rows) * seg_rows; sig_hi = sig_lo + seg_rows;
  }
  float4 P[NV], A[NV], B[NV];
  cf_load_row<NV>(A, lgd + first * ld_lgd, lane);
  if (first > sig_lo) cf_load_row<NV>(P, lgd + (first - 1) * ld_lgd, lane);
  else {
#pragma unroll
    for (int j = 0; j < NV; ++j) P[j] = A[j];
  }
  if (first + 1 < rows) cf_load_row<NV>(B, lgd + (first + 1) * ld_lgd, lane);
  float4* m4 = reinterpret_cast<float4*>(mx);
  const char* mxb = reinterpret_cast<const char*>(mx);       // the cover table holds byte offsets

  for (int64_t row = first; row < stop; ++row) {
    if (row >= sig_hi) {                                  // first row of the next signal
      if (seg) {
        do { ++si; } while (seg[si + 1] <= row);          // empty signals have no rows
        sig_lo = seg[si]; sig_hi = seg[si + 1];
      } else {
        sig_lo = sig_hi; sig_hi = sig_lo + seg_rows;
      }
    }
    // prefetches hold no registers: the lgd row three ahead goes to L2 (one 128-byte line per lane), the diff row
    // two ahead to L1 (the band loop below would otherwise wait for DRAM once per group of 32 bands)
    if (row + 3 < rows && row + 2 < stop && lane < NV * 4)
      asm volatile("prefetch.global.L2 [%0];" ::"l"(lgd + (row + 3) * ld_lgd + 32 * lane));
    if (row + 2 < stop && 32 * lane < n_bands)
      asm volatile("prefetch.global.L1 [%0];" ::"l"(diff + (row + 2) * ld_diff + 32 * lane));
    // 'reflect': the row itself stands in for its neighbour; the last signal may be shorter than seg_rows
    const bool last = row + 1 >= sig_hi || row + 1 >= rows;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      float4 m = fmax4(P[j], A[j]);
      if (!last) m = fmax4(m, B[j]);
      m4[lane + 32 * j] = m;
      float q = fminf(fminf(m.x, m.y), fminf(m.z, m.w));
      mx[o4 + lane + 32 * j] = q;
      q = fminf(q, __shfl_xor_sync(0xffffffffu, q, 1));
      q = fminf(q, __shfl_xor_sync(0xffffffffu, q, 2));
      if ((lane & 3) == 0) mx[o16 + (lane >> 2) + 8 * j] = q;
    }
    // shift the window and put the row after next in flight
#pragma unroll
    for (int j = 0; j < NV; ++j) { P[j] = last ? B[j] : A[j]; A[j] = B[j]; }
    if (row + 2 < rows && row + 1 < stop) cf_load_row<NV>(B, lgd + (row + 2) * ld_lgd, lane);
    __syncwarp();
    float sum = 0.0f;
    const float* drow = diff + row * ld_diff;
#pragma unroll
    for (int g = 0; g < NG; ++g) {
      if (32 * g >= n_bands) break;                       // warp-uniform
      const int b = 32 * g + lane;
      const bool have = b < n_bands;
      const float d = have ? drow[b] : 0.0f;
      float m = INFINITY;
      if (!serial) {
        const int glen = s_glen[g];
        const uint4* c = reinterpret_cast<const uint4*>(s_cover + (size_t)b * kCfcCover);
#pragma unroll
        for (int j0 = 0; j0 < kCfcCover; j0 += 8) {
          if (j0 >= glen) break;                          // warp-uniform
          const uint4 o = c[j0 >> 3];
          auto at = [&](uint32_t byte) { return *reinterpret_cast<const float*>(mxb + byte); };
          m = fminf(m, fminf(fminf(at(o.x & 0xffffu), at(o.x >> 16)), fminf(at(o.y & 0xffffu), at(o.y >> 16))));
          if (j0 + 4 < glen)
            m = fminf(m, fminf(fminf(at(o.z & 0xffffu), at(o.z >> 16)), fminf(at(o.w & 0xffffu), at(o.w >> 16))));
        }
        sum = __fmaf_rn(d, have && s_len[b] > 0 ? m : 0.0f, sum);
      } else {
        const int lo = have ? s_lo[b] : 0, len = have ? s_len[b] : 0;
        const int wlen = __reduce_max_sync(0xffffffffu, len);
        const int top = lo + (len > 0 ? len - 1 : 0);
        for (int k = 0; k < wlen; ++k) m = fminf(m, mx[min(lo + k, top)]);
        sum = __fmaf_rn(d, len > 0 ? m : 0.0f, sum);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) odf[row] = sum;
    __syncwarp();                                         // the row buffer is overwritten by the next row
  }
}

// ---------------------------------------------------------------------------------------------- spectral ODFs
// One warp per frame: a reduction over the bins of a magnitude spectrogram and / or a complex STFT with lags
// of one or two frames (madmom/features/onsets.py:102-160, 320-581).  Frames before the start of the row's signal
// count as in the reference (zero rows of the lagged matrices).
__device__ __forceinline__ float cabs2(float2 z) { return sqrtf(z.x * z.x + z.y * z.y); }

// np.angle / atan2(y, x) for the phase-based functions: octant reduction to t = min / max in [0, 1], atan(t) = t P(t^2)
// with the 7-term minimax polynomial of the front end's local-group-delay stage (frame_fft.cuh; 3.2e-7 rad), then the
// reflections.  Exact on the axes (0, +-pi/2, +-pi with the signs of atan2, signed zeros included), which is what the
// pinned bin-0 phases need.  libm's atan2f made the phase kernels compute-bound at 0.17 - 0.4 of the HBM bandwidth.
__device__ __forceinline__ float fast_atan2f(float y, float x) {
  const float ax = fabsf(x), ay = fabsf(y);
  const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
  const float t = mx > 0.0f ? __fdividef(mn, mx) : 0.0f;
  const float u = t * t;
  float p = 0.00681148456f;
  p = fmaf(p, u, -0.0336031997f);
  p = fmaf(p, u, 0.0796223719f);
  p = fmaf(p, u, -0.132332626f);
  p = fmaf(p, u, 0.198077921f);
  p = fmaf(p, u, -0.333173652f);
  p = fmaf(p, u, 0.999996111f);
  float r = p * t;
  if (ay > ax) r = 1.57079632679489661923f - r;
  if (__float_as_int(x) < 0) r = 3.14159265358979323846f - r;       // sign bit: -0.0 counts like atan2's
  return copysignf(r, y);
}

// z times the power of two that brings max(|re|, |im|) into [1, 2): exact, keeps the argument, and lets a product of
// four such factors neither overflow nor underflow (phase_deviation counts a bin at -200 dB like a loud one).  Zero
// counts as 1, i.e. phase 0 like np.angle(0).
__device__ __forceinline__ float2 unit_scaled(float2 z) {
  const float mx = fmaxf(fabsf(z.x), fabsf(z.y));
  if (!(mx > 0.0f)) return make_float2(1.0f, 0.0f);
  const float sc = __int_as_float(0x7f000000 - (__float_as_int(mx) & 0x7f800000));       // 2^-exponent(mx)
  return make_float2(z.x * sc, z.y * sc);
}

// X[t-1] |.| exp(i (2 phi[t-1] - phi[t-2])) without trigonometry: s1 u1^2 conj(u2), u = X / |X| (1 when X = 0)
__device__ __forceinline__ float2 cd_target(float2 x1, float2 x2, bool have2) {
  const float s1 = cabs2(x1);
  if (s1 == 0.0f) return make_float2(0.0f, 0.0f);
  // direction of x1^2 conj(x2) from factors scaled by exact powers of two (no overflow, no underflow; a zero x2 counts
  // as phase 0), normalised with one reciprocal square root, times |x1|: one sqrt + one rsqrt instead of two square
  // roots and four divisions
  const float2 u1 = unit_scaled(x1);
  float2 d = make_float2(u1.x * u1.x - u1.y * u1.y, 2.0f * u1.x * u1.y);              // u1^2
  if (have2) {
    const float2 u2 = unit_scaled(x2);
    d = make_float2(d.x * u2.x + d.y * u2.y, d.y * u2.x - d.x * u2.y);                  // times conj(u2)
  }
  const float g = s1 * rsqrtf(d.x * d.x + d.y * d.y);
  return make_float2(d.x * g, d.y * g);
}

__global__ void __launch_bounds__(256) spectral_odf_kernel(int32_t mode, const float2* __restrict__ stft,
                                                           int64_t ld_stft, const float* __restrict__ spec,
                                                           int64_t ld_spec, int64_t rows, int32_t cols,
                                                           int64_t seg_rows, const int64_t* __restrict__ seg,
                                                           int32_t n_seg, int32_t d, float eps,
                                                           float* __restrict__ out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + warp;
  if (row >= rows) return;
  int64_t s0 = 0;
  if (seg) s0 = seg[find_segment(seg, n_seg, row)];
  else if (seg_rows > 0) s0 = (row / seg_rows) * seg_rows;
  const int64_t t = row - s0;
  float acc = 0.0f, acc2 = 0.0f;
  switch (mode) {
    case MMB_ODF_HFC:
      for (int k = lane; k < cols; k += 32) acc += spec[row * ld_spec + k] * (float)k;
      break;
    case MMB_ODF_SPECTRAL_DIFF:
      if (t >= d)
        for (int k = lane; k < cols; k += 32) {
          const float v = fmaxf(spec[row * ld_spec + k] - spec[(row - d) * ld_spec + k], 0.0f);
          acc += v * v;
        }
      break;
    case MMB_ODF_MKL:
      if (t >= d)
        for (int k = lane; k < cols; k += 32)
          // fast log and division: about 1e-7 per bin against a bound of 1e-4 on the mean of 1024 of them
          acc += __logf(1.0f + __fdividef(spec[row * ld_spec + k], spec[(row - d) * ld_spec + k] + eps));
      break;
    case MMB_ODF_PHASE_DEV:
    case MMB_ODF_WEIGHTED_PHASE_DEV:
    case MMB_ODF_NORM_WEIGHTED_PHASE_DEV:
      for (int k = lane; k < cols; k += 32) {
        const float2 x0 = stft[row * ld_stft + k];
        const float m = cabs2(x0);
        acc2 += m;
        if (t >= 2) {
          // wrap(phi0 - 2 phi1 + phi2) is the argument of x0 conj(x1)^2 x2: one arctangent instead of three and no
          // wrap (only the magnitude of the result is used, so +pi and -pi are the same)
          const float2 u0 = unit_scaled(x0), u1 = unit_scaled(stft[(row - 1) * ld_stft + k]);
          const float2 u2 = unit_scaled(stft[(row - 2) * ld_stft + k]);
          const float2 a = make_float2(u0.x * u1.x + u0.y * u1.y, u0.y * u1.x - u0.x * u1.y);    // u0 conj(u1)
          const float2 b = make_float2(u2.x * u1.x + u2.y * u1.y, u2.y * u1.x - u2.x * u1.y);    // u2 conj(u1)
          const float pd = fast_atan2f(a.x * b.y + a.y * b.x, a.x * b.x - a.y * b.y);
          acc += fabsf(mode == MMB_ODF_PHASE_DEV ? pd : pd * m);
        }
      }
      break;
    case MMB_ODF_COMPLEX_DOMAIN:
    case MMB_ODF_RECTIFIED_COMPLEX_DOMAIN:
      for (int k = lane; k < cols; k += 32) {
        float2 c = stft[row * ld_stft + k];
        const float m0 = cabs2(c);
        float2 x1 = make_float2(0.0f, 0.0f);
        if (t >= 1) {
          x1 = stft[(row - 1) * ld_stft + k];
          const float2 x2 = t >= 2 ? stft[(row - 2) * ld_stft + k] : make_float2(0.0f, 0.0f);
          // frame 0 of a signal has target phase 0: |X[0]| itself
          const float2 tg = t >= 2 ? cd_target(x1, x2, true) : make_float2(cabs2(x1), 0.0f);
          c.x -= tg.x; c.y -= tg.y;
        }
        float v = cabs2(c);
        if (mode == MMB_ODF_RECTIFIED_COMPLEX_DOMAIN) {
          bool rising = false;
          if (t >= d) rising = m0 - cabs2(stft[(row - d) * ld_stft + k]) > 0.0f;
          if (!rising) v = 0.0f;
        }
        acc += v;
      }
      break;
    default: break;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    acc += __shfl_xor_sync(0xffffffffu, acc, o);
    acc2 += __shfl_xor_sync(0xffffffffu, acc2, o);
  }
  if (lane != 0) return;
  const float n = (float)cols;
  float r = acc;
  if (mode == MMB_ODF_HFC || mode == MMB_ODF_MKL || mode == MMB_ODF_PHASE_DEV || mode == MMB_ODF_WEIGHTED_PHASE_DEV)
    r = acc / n;
  else if (mode == MMB_ODF_NORM_WEIGHTED_PHASE_DEV)
    r = (acc / n) / (acc2 / n + eps);
  out[row] = r;
}

}  // namespace

extern "C" MMB_API int mmb_spectral_odf(int32_t mode, const void* stft_dev, int64_t ld_stft, const float* spec_dev,
                                        int64_t ld_spec, int64_t rows, int32_t cols, int64_t rows_per_signal,
                                        const int64_t* seg_offsets_dev, int32_t n_segments, int32_t diff_frames,
                                        double epsilon, float* odf_dev, void* stream) {
  const bool needs_spec = mode == MMB_ODF_HFC || mode == MMB_ODF_SPECTRAL_DIFF || mode == MMB_ODF_MKL;
  if (mode < MMB_ODF_HFC || mode > MMB_ODF_RECTIFIED_COMPLEX_DOMAIN || !odf_dev || rows < 0 || cols <= 0 ||
      (needs_spec && (!spec_dev || ld_spec < cols)) || (!needs_spec && (!stft_dev || ld_stft < cols)) ||
      (seg_offsets_dev && n_segments < 1))
    return MMB_E_INVALID;
  const bool lagged = mode == MMB_ODF_SPECTRAL_DIFF || mode == MMB_ODF_MKL || mode == MMB_ODF_RECTIFIED_COMPLEX_DOMAIN;
  if (lagged && diff_frames < 1) return MMB_E_INVALID;
  if ((mode == MMB_ODF_MKL || mode == MMB_ODF_NORM_WEIGHTED_PHASE_DEV) && !(epsilon > 0.0)) return MMB_E_INVALID;
  if (rows == 0) return MMB_OK;
  const int warps = 8;
  const int64_t grid = (rows + warps - 1) / warps;
  spectral_odf_kernel<<<(unsigned)grid, warps * 32, 0, (cudaStream_t)stream>>>(
      mode, (const float2*)stft_dev, ld_stft, spec_dev, ld_spec, rows, cols, rows_per_signal, seg_offsets_dev,
      n_segments, diff_frames, (float)epsilon, odf_dev);
  return cudaGetLastError() == cudaSuccess ? MMB_OK : MMB_E_CUDA;
}

extern "C" MMB_API int mmb_matrix_diff(const float* in_dev, int64_t ld_in, float* out_dev, int64_t ld_out,
                                       float* odf_dev, int64_t rows, int32_t cols, int64_t rows_per_signal,
                                       int32_t diff_frames, int32_t diff_max_bins, int32_t positive_diffs,
                                       void* stream) {
  if (!in_dev || (!out_dev && !odf_dev) || rows < 0 || cols <= 0 || diff_frames < 1 || ld_in < cols ||
      (out_dev && ld_out < cols))
    return MMB_E_INVALID;
  if (rows == 0) return MMB_OK;
  const int warps = 8;
  const int64_t grid = (rows + warps - 1) / warps;
  matrix_diff_kernel<<<(unsigned)grid, warps * 32, 0, (cudaStream_t)stream>>>(
      in_dev, ld_in, out_dev, ld_out, odf_dev, rows, cols, rows_per_signal, diff_frames, diff_max_bins,
      positive_diffs);
  return cudaGetLastError() == cudaSuccess ? MMB_OK : MMB_E_CUDA;
}

extern "C" MMB_API int mmb_peak_picking(const float* act_dev, const int64_t* seg_offsets_dev, int32_t n_segments,
                                        int64_t n, double threshold, int32_t smooth, const double* smooth_kernel_dev,
                                        int32_t pre_avg, int32_t post_avg, int32_t pre_max, int32_t post_max,
                                        int32_t* peaks_dev, int32_t* counts_dev, void* stream) {
  if (!act_dev || !peaks_dev || !counts_dev || n < 0 || n_segments < 1 || pre_avg < 0 || post_avg < 0 ||
      pre_max < 0 || post_max < 0 || smooth < 0 || (smooth > 1 && !smooth_kernel_dev) ||
      (n_segments > 1 && !seg_offsets_dev))
    return MMB_E_INVALID;
  PeakParams p;
  p.act = act_dev; p.seg = seg_offsets_dev; p.n = n; p.n_seg = n_segments; p.threshold = threshold;
  p.smooth = smooth; p.pre_avg = pre_avg; p.post_avg = post_avg; p.pre_max = pre_max; p.post_max = post_max;
  p.kernel = smooth_kernel_dev; p.peaks = peaks_dev; p.counts = counts_dev;
  cudaStream_t st = (cudaStream_t)stream;
  if (n > 0) {
    const unsigned grid = (unsigned)((n + 255) / 256);
    // with smoothing the activations are float64 from there on (np.convolve with a float64 kernel)
    if (smooth > 1) peak_flag_kernel<double><<<grid, 256, 0, st>>>(p);
    else peak_flag_kernel<float><<<grid, 256, 0, st>>>(p);
  }
  peak_compact_kernel<<<n_segments, 1024, 0, st>>>(p);
  return cudaGetLastError() == cudaSuccess ? MMB_OK : MMB_E_CUDA;
}

// ---------------------------------------------------------------------------------------------- phase / local group delay
// phase()  = np.angle(stft)                                          (madmom/audio/stft.py:136-151)
// local_group_delay() = unwrap(phase) along the bins, lgd[:, k] = up[k] - up[k + 1], last bin 0    (:154-179)
// The difference of neighbouring unwrapped phases needs no scan: up[k] - up[k+1] = -(dd + correction) with
// dd = p[k+1] - p[k] and numpy's unwrap rule  ddmod = mod(dd + pi, 2 pi) - pi,  ddmod = pi where it came out as -pi for
// dd > 0,  correction = ddmod - dd where |dd| >= pi, else 0.
namespace {
// One warp per row, lanes over the bins: the phase of bin k + 1 comes from the next lane by shuffle (lane 31 takes
// the arctangent of its right neighbour itself), loads and stores are coalesced and nothing divides.
__global__ void __launch_bounds__(256) phase_lgd_kernel(const float2* __restrict__ stft, int64_t ld_stft,
                                                        float* __restrict__ phase, int64_t ld_phase,
                                                        float* __restrict__ lgd, int64_t ld_lgd, int64_t rows,
                                                        int32_t cols) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float2* src = stft + row * ld_stft;
  const float pi = 3.14159265358979323846f, two_pi = 6.28318530717958647692f;
  for (int k0 = 0; k0 < cols; k0 += 32) {                    // warp-uniform trip count: the shuffle needs all lanes
    const int k = k0 + lane;
    float p0 = 0.0f;
    if (k < cols) {
      const float2 z = src[k];
      p0 = fast_atan2f(z.y, z.x);
      if (phase) phase[row * ld_phase + k] = p0;
    }
    if (lgd) {
      float pn = __shfl_down_sync(0xffffffffu, p0, 1);
      if (lane == 31 && k + 1 < cols) {
        const float2 z1 = src[k + 1];
        pn = fast_atan2f(z1.y, z1.x);
      }
      if (k < cols) {
        float v = 0.0f;
        if (k + 1 < cols) {
          const float dd = pn - p0;
          // np.mod(dd + pi, 2 pi) for |dd| < 2 pi: one period added or taken off (a - 2 pi is exact for a in
          // [2 pi, 3 pi), and the negative side adds 2 pi like np.mod does), the same values as fmodf gives
          const float a = dd + pi;
          float ddmod = a >= two_pi ? a - two_pi : (a < 0.0f ? a + two_pi : a);
          ddmod -= pi;
          if (ddmod == -pi && dd > 0.0f) ddmod = pi;
          v = fabsf(dd) < pi ? -dd : -ddmod;
        }
        lgd[row * ld_lgd + k] = v;
      }
    }
  }
}
}  // namespace

extern "C" MMB_API int mmb_phase_lgd(const void* stft_dev, int64_t ld_stft, float* phase_dev, int64_t ld_phase,
                                     float* lgd_dev, int64_t ld_lgd, int64_t rows, int32_t cols, void* stream) {
  if (!stft_dev || (!phase_dev && !lgd_dev) || rows < 0 || cols < 1 || ld_stft < cols ||
      (phase_dev && ld_phase < cols) || (lgd_dev && ld_lgd < cols))
    return mmb::set_error(MMB_E_INVALID, "mmb_phase_lgd: invalid argument");
  if (rows == 0) return MMB_OK;
  phase_lgd_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, (cudaStream_t)stream>>>(
      (const float2*)stft_dev, ld_stft, phase_dev, ld_phase, lgd_dev, ld_lgd, rows, cols);
  return cudaGetLastError() == cudaSuccess ? MMB_OK : mmb::set_error(MMB_E_CUDA, "mmb_phase_lgd: kernel launch failed");
}

// ---------------------------------------------------------------------------------------------- no-alloc peak picking
// mmb_peak_picking_static: the variant for callers that may not allocate device memory (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  mmb_peak_picking keeps its per-frame flags in the caller's
// n-element `peaks` buffer and compacts in place; here the flags live in a file-scope device array that every GPU
// loading the library holds until exit (MMB_STATIC_PEAK_MAX_FRAMES int32 = 16 MiB), and the peaks of segment s go
// to a caller buffer of fixed capacity per segment.  Has no counterpart in the reference.
//
// In host code `g_static_peak_flags` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current (static_peak_flags() below).
static __device__ __align__(256) int32_t g_static_peak_flags[MMB_STATIC_PEAK_MAX_FRAMES];

namespace {
// ordered compaction of the flags of one segment into out[s * cap ...]; counts[s] is the number of peaks FOUND
// (it may exceed cap: the caller sees that the segment overflowed; only the first cap indices are stored)
__global__ void __launch_bounds__(1024) peak_compact_out_kernel(const int32_t* __restrict__ flags,
                                                                const int64_t* __restrict__ seg, int64_t n,
                                                                int32_t* __restrict__ out, int32_t cap,
                                                                int32_t* __restrict__ counts) {
  __shared__ int warp_tot[32];
  __shared__ int base;
  const int s = blockIdx.x;
  const int64_t s0 = seg ? seg[s] : 0;
  const int64_t len = (seg ? seg[s + 1] : n) - s0;
  const int32_t* f_in = flags + s0;
  int32_t* dst = out + (int64_t)s * cap;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (threadIdx.x == 0) base = 0;
  __syncthreads();
  for (int64_t t0 = 0; t0 < len; t0 += blockDim.x) {
    const int64_t i = t0 + threadIdx.x;
    const int f = (i < len) ? f_in[i] : 0;
    const unsigned m = __ballot_sync(0xffffffffu, f != 0);
    if (lane == 0) warp_tot[warp] = __popc(m);
    __syncthreads();
    int off = base;
    for (int w = 0; w < warp; ++w) off += warp_tot[w];
    const int at = off + __popc(m & ((1u << lane) - 1u));
    if (f && at < cap) dst[at] = (int32_t)i;
    __syncthreads();
    if (threadIdx.x == 0) {
      int tot = 0;
      for (int w = 0; w < nw; ++w) tot += warp_tot[w];
      base += tot;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) counts[s] = base;
}

// device address of the flag array on the CURRENT device; looked up once per device (no allocation: the array is
// part of the library's device image)
int static_peak_flags(int32_t** out) {
  constexpr int kMaxDevices = 64;
  static int32_t* ptrs[kMaxDevices] = {nullptr};
  static std::mutex mu;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices)
    return mmb::set_error(MMB_E_CUDA, "mmb_peak_picking_static: no current CUDA device");
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_peak_flags);
    if (e != cudaSuccess || !p)
      return mmb::set_error(MMB_E_CUDA, "mmb_peak_picking_static: cudaGetSymbolAddress(g_static_peak_flags) failed");
    ptrs[dev] = static_cast<int32_t*>(p);
  }
  *out = ptrs[dev];
  return MMB_OK;
}
}  // namespace

extern "C" MMB_API int mmb_peak_picking_static(const float* act_dev, const int64_t* seg_offsets_dev,
                                               int32_t n_segments, int64_t n, double threshold, int32_t smooth,
                                               const double* smooth_kernel_dev, int32_t pre_avg, int32_t post_avg,
                                               int32_t pre_max, int32_t post_max, int32_t* peaks_dev,
                                               int32_t peaks_per_segment, int32_t* counts_dev, void* stream) {
  if ((!act_dev && n > 0) || !peaks_dev || !counts_dev || n < 0 || n_segments < 1 || pre_avg < 0 || post_avg < 0 ||
      pre_max < 0 || post_max < 0 || smooth < 0 || (smooth > 1 && !smooth_kernel_dev) ||
      (n_segments > 1 && !seg_offsets_dev) || peaks_per_segment < 1)
    return mmb::set_error(MMB_E_INVALID, "mmb_peak_picking_static: invalid argument");
  if (n > (int64_t)MMB_STATIC_PEAK_MAX_FRAMES)
    return mmb::set_error(MMB_E_INVALID,
                          "mmb_peak_picking_static: more than MMB_STATIC_PEAK_MAX_FRAMES (4194304) frames in one "
                          "call; the static scratch array cannot hold their flags (use mmb_peak_picking)");
  int32_t* flags = nullptr;
  const int rc = static_peak_flags(&flags);
  if (rc != MMB_OK) return rc;
  PeakParams p;
  p.act = act_dev; p.seg = seg_offsets_dev; p.n = n; p.n_seg = n_segments; p.threshold = threshold;
  p.smooth = smooth; p.pre_avg = pre_avg; p.post_avg = post_avg; p.pre_max = pre_max; p.post_max = post_max;
  p.kernel = smooth_kernel_dev; p.peaks = flags; p.counts = counts_dev;
  cudaStream_t st = (cudaStream_t)stream;
  if (n > 0) {
    // the flag kernel writes flags[i] for every i < n, so nothing of the previous call's contents survives in the
    // range this call reads: the array needs no clearing
    const unsigned grid = (unsigned)((n + 255) / 256);
    if (smooth > 1) peak_flag_kernel<double><<<grid, 256, 0, st>>>(p);
    else peak_flag_kernel<float><<<grid, 256, 0, st>>>(p);
  }
  peak_compact_out_kernel<<<n_segments, 1024, 0, st>>>(flags, seg_offsets_dev, n, peaks_dev, peaks_per_segment,
                                                       counts_dev);
  if (cudaGetLastError() != cudaSuccess) return mmb::set_error(MMB_E_CUDA, "mmb_peak_picking_static: kernel launch failed");
  return MMB_OK;
}

extern "C" MMB_API int mmb_complex_flux(const float* lgd_dev, int64_t ld_lgd, const float* diff_dev, int64_t ld_diff,
                                        const int32_t* band_lo_dev, const int32_t* band_hi_dev, int32_t n_bands,
                                        int32_t n_bins, int64_t rows, int64_t rows_per_signal,
                                        const int64_t* seg_offsets_dev, int32_t n_segments, int32_t temporal_filter,
                                        float* odf_dev, void* stream) {
  if (!lgd_dev || !diff_dev || !band_lo_dev || !band_hi_dev || !odf_dev || n_bands <= 0 || n_bins <= 0 || rows < 0 ||
      ld_lgd < n_bins || ld_diff < n_bands || temporal_filter < 0 || (temporal_filter > 1 && temporal_filter % 2 == 0) ||
      (seg_offsets_dev && n_segments < 1))
    return MMB_E_INVALID;
  if (rows == 0) return MMB_OK;
  const int warps = 8;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  // rows read once (streaming kernel) when the rows can be held as float4 registers
  const bool vec_ok = temporal_filter == 3 && (n_bins == 512 || n_bins == 1024) && ld_lgd % 4 == 0 &&
                      (reinterpret_cast<uintptr_t>(lgd_dev) & 15) == 0 && getenv("MMB_CF_GENERIC") == nullptr;
  // the streaming kernel with hierarchical band minima (complex_flux_cover_kernel); MMB_CF_STREAM=1 keeps the serial
  // walk over the bins (the tests compare the two bit for bit)
  if (vec_ok && n_bands <= 32 * kCfcMaxGroups && getenv("MMB_CF_STREAM") == nullptr) {
    const int groups = (n_bands + 31) / 32;
    const size_t s_words = (size_t)n_bins + n_bins / 4 + n_bins / 16;
    const size_t smem = (size_t)warps * s_words * sizeof(float) + (size_t)groups * 32 * kCfcCover * sizeof(uint16_t) +
                        ((size_t)groups + 2 * (size_t)n_bands + 1) * sizeof(int);
    auto kern = n_bins == 1024
                    ? (groups <= 3 ? complex_flux_cover_kernel<8, 3>
                                   : (groups <= 5 ? complex_flux_cover_kernel<8, 5> : complex_flux_cover_kernel<8, 8>))
                    : (groups <= 3 ? complex_flux_cover_kernel<4, 3>
                                   : (groups <= 5 ? complex_flux_cover_kernel<4, 5> : complex_flux_cover_kernel<4, 8>));
    if (smem > 48 * 1024 &&
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
      return MMB_E_CUDA;
    const int64_t total_warps = (int64_t)sms * 2 * warps;          // two CTAs per SM
    const int64_t run = (rows + total_warps - 1) / total_warps;
    const int64_t used = (rows + run - 1) / run;
    const unsigned grid = (unsigned)((used + warps - 1) / warps);
    kern<<<grid, warps * 32, smem, (cudaStream_t)stream>>>(lgd_dev, ld_lgd, diff_dev, ld_diff, band_lo_dev,
                                                           band_hi_dev, n_bands, rows, rows_per_signal,
                                                           seg_offsets_dev, n_segments, run, odf_dev);
    return cudaGetLastError() == cudaSuccess ? MMB_OK : MMB_E_CUDA;
  }
  if (vec_ok) {
    const size_t smem = ((size_t)warps * n_bins + 2 * (size_t)n_bands) * sizeof(float);
    auto kern = n_bins == 1024 ? complex_flux_stream_kernel<8> : complex_flux_stream_kernel<4>;
    if (smem > 48 * 1024 &&
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
      return MMB_E_CUDA;
    const int64_t total_warps = (int64_t)sms * 2 * warps;          // two CTAs per SM
    const int64_t run = (rows + total_warps - 1) / total_warps;
    const int64_t used = (rows + run - 1) / run;
    const unsigned grid = (unsigned)((used + warps - 1) / warps);
    kern<<<grid, warps * 32, smem, (cudaStream_t)stream>>>(lgd_dev, ld_lgd, diff_dev, ld_diff, band_lo_dev,
                                                           band_hi_dev, n_bands, rows, rows_per_signal,
                                                           seg_offsets_dev, n_segments, run, odf_dev);
    return cudaGetLastError() == cudaSuccess ? MMB_OK : MMB_E_CUDA;
  }
  const size_t smem = ((size_t)warps * n_bins + (size_t)warps * n_bands + 3 * (size_t)n_bands + 1 + kCfMaxPieces) *
                      sizeof(float);
  if (smem > 48 * 1024 &&
      cudaFuncSetAttribute(complex_flux_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return MMB_E_CUDA;
  int64_t grid = (rows + warps - 1) / warps;
  if (grid > (int64_t)sms * 4) grid = (int64_t)sms * 4;       // persistent CTAs, a multiple of the SM count
  complex_flux_kernel<<<(unsigned)grid, warps * 32, smem, (cudaStream_t)stream>>>(
      lgd_dev, ld_lgd, diff_dev, ld_diff, band_lo_dev, band_hi_dev, n_bands, n_bins, rows, rows_per_signal,
      seg_offsets_dev, n_segments, temporal_filter, odf_dev);
  return cudaGetLastError() == cudaSuccess ? MMB_OK : MMB_E_CUDA;
}

// ---------------------------------------------------------------------------------------------- PCM ingest
namespace {
// np.mean(signal, axis=-1).astype(signal.dtype) (madmom/audio/signal.py:212): integer samples are averaged
// exactly and truncated towards zero by the cast back; float32 samples are summed in channel order in float32 and
// divided by the channel count.
// stereo int16, four sample frames per thread: one 16-byte load, one 8-byte store, the same integer mean
__global__ void __launch_bounds__(256) remix_stereo16_kernel(const int4* __restrict__ in, int64_t n4,
                                                             int2* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const int4 v = in[i];
    const int w[4] = {v.x, v.y, v.z, v.w};
    int m[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int l = (int)(int16_t)(w[q] & 0xffff), r = w[q] >> 16;      // little endian: left, right (sign-extended)
      m[q] = ((l + r) / 2) & 0xffff;                                     // C division truncates towards zero
    }
    out[i] = make_int2(m[0] | (m[1] << 16), m[2] | (m[3] << 16));
  }
}

template <typename T>
__global__ void __launch_bounds__(256) remix_mono_kernel(const T* __restrict__ in, int64_t n, int channels,
                                                         int channel, T* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const T* f = in + i * channels;
    if (channel >= 0) { out[i] = f[channel]; continue; }
    if (sizeof(T) == 2) {
      int sum = 0;
      for (int c = 0; c < channels; ++c) sum += (int)f[c];
      out[i] = (T)(sum / channels);                    // C division truncates towards zero like astype(int16)
    } else {
      float sum = (float)f[0];
      for (int c = 1; c < channels; ++c) sum += (float)f[c];
      out[i] = (T)(sum / (float)channels);
    }
  }
}
}  // namespace

cudaError_t mmb::launch_remix_mono(const void* in, int sample_bytes, int64_t num_frames, int channels, int channel,
                                   void* out, cudaStream_t stream) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int64_t blocks = (num_frames + 255) / 256;
  if (blocks > (int64_t)sms * 8) blocks = (int64_t)sms * 8;      // grid-stride, a multiple of the SM count
  if (blocks < 1) blocks = 1;
  if (sample_bytes == 2 && channels == 2 && channel < 0 && ((uintptr_t)in & 15) == 0 && ((uintptr_t)out & 7) == 0) {
    const int64_t n4 = num_frames / 4, tail = num_frames - 4 * n4;
    if (n4 > 0) {
      int64_t b4 = (n4 + 255) / 256;
      if (b4 > (int64_t)sms * 8) b4 = (int64_t)sms * 8;
      remix_stereo16_kernel<<<(unsigned)b4, 256, 0, stream>>>((const int4*)in, n4, (int2*)out);
    }
    if (tail > 0)
      remix_mono_kernel<int16_t><<<1, 32, 0, stream>>>((const int16_t*)in + 8 * n4, tail, 2, -1,
                                                      (int16_t*)out + 4 * n4);
  } else if (sample_bytes == 2)
    remix_mono_kernel<int16_t><<<(unsigned)blocks, 256, 0, stream>>>((const int16_t*)in, num_frames, channels, channel,
                                                                    (int16_t*)out);
  else
    remix_mono_kernel<float><<<(unsigned)blocks, 256, 0, stream>>>((const float*)in, num_frames, channels, channel,
                                                                  (float*)out);
  return cudaGetLastError();
}
