/* madmom_b200.h -- C ABI of libmadmom_b200.so, the sm_100a spectral front end behind madmom's Processor API.
 *
 * The reference (CPJKU/madmom) has no FFI on this path: it is numpy/scipy all the way down.  Each entry
 * point therefore replaces a *Python* call site of the reference; INTEGRATION.md shows the ctypes binding a
 * madmom maintainer would add.  What is replaced, per plan field:
 *
 *   framing      FramedSignal.__init__ / signal_frame            madmom/audio/signal.py:1103-1156, 860-962
 *   window+FFT   ShortTimeFourierTransform.__new__ / stft()      madmom/audio/stft.py:317-373, 49-133
 *   magnitude    Spectrogram.__new__ (np.abs)                    madmom/audio/spectrogram.py:122-141
 *   filterbank   FilteredSpectrogram.__new__ (np.dot)            madmom/audio/spectrogram.py:363-390
 *   log          LogarithmicSpectrogram.__new__                  madmom/audio/spectrogram.py:519-544
 *   difference   SpectrogramDifference.__new__ / ...Processor    madmom/audio/spectrogram.py:960-1004, 1085-1139
 *   flux         spectral_flux / superflux (row sum)             madmom/features/onsets.py:163-240
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a negative
 * MMB_E_* code and never throws or aborts; the message of the last failure on the calling thread is
 * available from mmb_last_error().  Device buffers (PCM in, features out) are owned by the caller; the
 * library owns only plan constants (window, twiddles, filterbank) and layout tables.  mmb_run() enqueues
 * work on the caller's stream and does not synchronise.
 */
#ifndef MADMOM_B200_H
#define MADMOM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MMB_VERSION 102 /* 0.1.2: float64 stft / spec outputs, include_nyquist, mmb_run_multi, PCM gain */

#if defined(__GNUC__)
#define MMB_API __attribute__((visibility("default")))
#else
#define MMB_API
#endif

enum {
  MMB_OK = 0,
  MMB_E_INVALID = -1,     /* bad argument (the Python layer raises ValueError / TypeError before this) */
  MMB_E_UNSUPPORTED = -2, /* valid for the reference but not implemented here (NotImplementedError)   */
  MMB_E_CUDA = -3,        /* CUDA runtime failure; message carries cudaGetErrorString                 */
  MMB_E_NOMEM = -4
};

enum { MMB_PCM_I16 = 0, MMB_PCM_F32 = 1 };
enum { MMB_LOG_NONE = 0, MMB_LOG_LOG10 = 1, MMB_LOG_MULADD = 2, MMB_LOG_LN = 3, MMB_LOG_LOG2 = 4 };

typedef struct mmb_plan mmb_plan;     /* immutable after creation, usable from any host thread */
typedef struct mmb_layout mmb_layout; /* device-resident description of a batch of units         */

/* One recipe = one fused kernel configuration. */
typedef struct mmb_config {
  int32_t frame_size;      /* 1024, 2048 or 4096 (FramedSignalProcessor.frame_size); fft_size == frame_size */
  int32_t pcm_dtype;       /* MMB_PCM_I16 / MMB_PCM_F32 (Signal dtype)                                      */
  double hop_size;         /* samples between frames, may be fractional (sample_rate / fps)                 */
  int32_t origin;          /* integer origin, literals already resolved (signal.py:1126-1142)               */
  int32_t circular_shift;  /* STFT circular_shift flag (stft.py:106-119)                                    */
  const double* fft_window;/* frame_size float64 values actually multiplied onto the samples, i.e.
                              window / iinfo(dtype).max for integer PCM (stft.py:331-349).  The band outputs
                              use it in float32; a symmetric cosine sum a0 - a1 cos(2 pi n / (N - 1)) of any
                              scale (np.hanning, np.hamming) is recognised for frame_size 2048 / 4096 and
                              evaluated in the kernel, each value within 3e-7 of the window's peak of the
                              float32 value (features: 4e-6 per band; the stft / spec outputs use float64)    */
  /* filterbank as one contiguous run of non-zero weights per band (Filterbank (K x B) float32 columns) */
  int32_t num_bands;       /* 0: no filterbank stage; only stft / spec outputs are available               */
  const int32_t* band_start; /* [num_bands] first non-zero bin                                             */
  const int32_t* band_len;   /* [num_bands] run length                                                     */
  const float* band_weights; /* concatenated runs                                                          */
  int32_t log_mode;        /* MMB_LOG_*: none / log10(x*mul+add) / x*mul+add / ln(..) / log2(..)            */
  float mul, add;
  int32_t diff_frames;     /* d >= 1 (from _diff_frames, spectrogram.py:834-864), 0: no difference stage   */
  int32_t diff_max_bins;   /* frequency maximum-filter width (<= 1: none), scipy 'reflect' borders         */
  int32_t positive_diffs;
  int32_t chunk_frames;    /* > 0: cut every unit into chunks of this many output frames; 0 = the library  */
                           /* lays out each launch so that the SMs get equal work (results are the same)   */
  int32_t device;          /* CUDA device ordinal                                                          */
  int32_t include_nyquist; /* stft / spec rows carry frame_size/2 + 1 bins (stft.py:93-95); plans without a
                              filterbank only                                                               */
} mmb_config;

/* One independent unit of work: a whole track (sample_base = frame_first = 0, pcm_len = num_samples) or a
 * frame range of a long signal with its PCM halo (FramedSignal.__getitem__(slice), signal.py:1181-1194).
 * Frame g of the unit's signal starts at sample int(g * hop) - frame_size/2 - origin; samples outside
 * [sample_base, sample_base + pcm_len) read as zero. */
typedef struct mmb_track {
  int64_t pcm_offset;  /* element offset of the unit's first local sample in the PCM buffer */
  int64_t pcm_len;     /* local samples                                                     */
  int64_t sample_base; /* global sample index of local sample 0                             */
  int64_t frame_first; /* global index of the first output frame                            */
  int64_t frame_count; /* output rows                                                       */
  int64_t out_row;     /* first output row                                                  */
} mmb_track;

/* Output matrices, row-major, leading dimensions in ELEMENTS; a null pointer disables that output.
 * hstack((spec, diff)) of SpectrogramDifferenceProcessor(stack_diffs=np.hstack) is log = base,
 * diff = base + num_bands, both with ld = row length of the stacked matrix. */
typedef struct mmb_outputs {
  void* stft;  int64_t ld_stft;  /* complex64 (rows, frame_size/2 [+ 1 with include_nyquist]): float64 FFT rounded
                                    once, like the reference (stft.py:26, :105-131)                             */
  float* spec; int64_t ld_spec;  /* magnitudes of those complex64 bins (rows, same width)                        */
  float* filt; int64_t ld_filt;  /* filtered, before log (rows, num_bands)      */
  float* log;  int64_t ld_log;   /* log-filtered spectrogram (rows, num_bands)  */
  float* diff; int64_t ld_diff;  /* (positive, max-filtered) difference         */
  float* odf;                    /* row sums of diff (rows)                     */
  float* lgd;  int64_t ld_lgd;   /* |local group delay| / pi per bin (rows, frame_size/2), frame_size 2048 only:
                                    |unwrap(phase)[k] - unwrap(phase)[k+1]| / pi, last bin 0 (stft.py:154-179,
                                    onsets.py:282-284); input of mmb_complex_flux                              */
} mmb_outputs;

typedef struct mmb_plan_info {
  int32_t frame_size, num_bins, num_bands, chunk_frames;
  int32_t smem_bytes, ctas_per_sm, num_sms, frames_per_round;
  int32_t warps;           /* warps per CTA chosen for this plan (shared-memory budget of one SM) */
} mmb_plan_info;

MMB_API int mmb_version(void);
MMB_API const char* mmb_last_error(void);

/* integer framing helpers (bit-exact with signal.py:1145-1156 and :917-919) */
MMB_API int64_t mmb_num_frames(int64_t num_samples, double hop_size, int32_t end_extend);
MMB_API int64_t mmb_frame_start(int64_t index, int32_t frame_size, double hop_size, int32_t origin);

MMB_API int mmb_plan_create(const mmb_config* cfg, mmb_plan** out);
MMB_API int mmb_plan_destroy(mmb_plan* plan);
MMB_API int mmb_plan_get_info(const mmb_plan* plan, mmb_plan_info* info);

/* Upload the unit table of a batch (host array of n units) and lay out the launch's work list. */
MMB_API int mmb_layout_create(const mmb_plan* plan, const mmb_track* units, int32_t n_units, mmb_layout** out);
MMB_API int mmb_layout_destroy(mmb_layout* layout);

/* Device path: pcm_dev and every non-null output are DEVICE pointers; asynchronous on `stream`
 * (a cudaStream_t passed as void*).  *launches (optional) receives the number of kernels launched. */
MMB_API int mmb_run(const mmb_plan* plan, const mmb_layout* layout, const void* pcm_dev, const mmb_outputs* outs,
            void* stream, int32_t* launches);

/* Device path fused with the all-gather of a sharded long signal (the reference has no counterpart, it
 * replaces mmb_run followed by ncclAllGather): dsts[0] are this GPU's matrices, dsts[1 .. n_dst-1] the SAME matrices
 * of the other ranks mapped into this process (torch symmetric memory, cudaIpcOpenMemHandle, or cudaMalloc memory
 * with peer access enabled), every pointer already offset to this rank's row range.  The kernel stores each row of
 * log / diff / odf to all n_dst destinations as it is produced, so the NVLink transfer overlaps the computation
 * row by row; the caller synchronises the ranks afterwards (barrier) before reading.  Only the band outputs log,
 * diff and odf are allowed; all destinations must use the same leading dimensions.  n_dst <= 8. */
MMB_API int mmb_run_gather(const mmb_plan* plan, const mmb_layout* layout, const void* pcm_dev,
                           const mmb_outputs* dsts, int32_t n_dst, void* stream, int32_t* launches);

/* Same, through an NVLS multicast mapping (NVSwitch): the pointers of `mc` are MULTICAST addresses of the ranks'
 * result matrices (torch symmetric memory: multicast_ptr; cuMulticastCreate / cuMulticastBindMem), offset to this
 * rank's rows.  The kernel stores each row once with multimem.st and the switch delivers it to every GPU of the
 * multicast object, this one included, so a row leaves the GPU once instead of once per peer. */
MMB_API int mmb_run_multicast(const mmb_plan* plan, const mmb_layout* layout, const void* pcm_dev,
                              const mmb_outputs* mc, void* stream, int32_t* launches);

/* Host path (what a madmom call site holding numpy arrays uses): pcm_host and the outputs are HOST
 * pointers (pinned for full speed); the library stages units through its own device buffers with
 * host-to-device copy, kernel and device-to-host copy overlapped on three streams, and returns when the
 * outputs are complete.  `units` is the same host table given to a layout (out_row / pcm_offset global). */
MMB_API int mmb_run_host(const mmb_plan* plan, const mmb_track* units, int32_t n_units, const void* pcm_host,
                 const mmb_outputs* outs_host, int32_t* launches);

/* pinned host memory for mmb_run_host callers without their own allocator */
MMB_API int mmb_host_alloc(void** ptr, int64_t bytes);
MMB_API int mmb_host_free(void* ptr);

/* SpectrogramDifference of a (rows, cols) float32 matrix already on the device (unfiltered spectrograms):
 * out[r, c] = in[r, c] - max_{|j| <= max_bins/2} in[r - d, reflect(c + j)], rows whose lagged row lies before the
 * start of their signal are zero; rows_per_signal > 0: the matrix holds equal-length signals back to back.
 * out_dev and odf_dev (row sums) are optional.  Replaces madmom/audio/spectrogram.py:976-993 and the row sum of
 * madmom/features/onsets.py:163-240. */
MMB_API int mmb_matrix_diff(const float* in_dev, int64_t ld_in, float* out_dev, int64_t ld_out, float* odf_dev,
                            int64_t rows, int32_t cols, int64_t rows_per_signal, int32_t diff_frames,
                            int32_t diff_max_bins, int32_t positive_diffs, void* stream);

/* peak_picking() of madmom/features/onsets.py:875-963 on n_segments activation functions stored back to back in
 * act_dev (segment s = [seg_offsets[s], seg_offsets[s+1]); seg_offsets_dev may be null for one segment).  All
 * lengths are in frames.  smooth > 1: Hamming kernel of that length in smooth_kernel_dev (float64).  On return
 * (stream order) peaks_dev[seg_offsets[s] + j], j < counts_dev[s], holds the ascending local frame indices. */
MMB_API int mmb_peak_picking(const float* act_dev, const int64_t* seg_offsets_dev, int32_t n_segments, int64_t n,
                             double threshold, int32_t smooth, const double* smooth_kernel_dev, int32_t pre_avg,
                             int32_t post_avg, int32_t pre_max, int32_t post_max, int32_t* peaks_dev,
                             int32_t* counts_dev, void* stream);

/* mmb_peak_picking for callers that may not allocate device memory (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same peak definition and arguments as mmb_peak_picking, but the per-frame flags, which mmb_peak_picking keeps in
 * the caller's n-element peaks buffer, live in a static device array inside the library of
 * MMB_STATIC_PEAK_MAX_FRAMES int32 (16 MiB, held by every GPU that loads the library until the process exits), and
 * the output has a fixed capacity per segment: peaks_dev[s * peaks_per_segment + j], j < min(counts_dev[s],
 * peaks_per_segment), holds the ascending local frame indices of segment s; counts_dev[s] is the number of peaks
 * FOUND, so counts_dev[s] > peaks_per_segment tells the caller that the segment overflowed.
 * The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine): it launches two kernels on
 * `stream`.  The device of `stream` must be current; the first call on a device looks the array's address up
 * (cudaGetSymbolAddress), so make one call outside a graph capture first (n = 0 is enough).
 * n > MMB_STATIC_PEAK_MAX_FRAMES returns MMB_E_INVALID with a message.  Every call on a device shares the array; it
 * keeps the previous call's flags, which is harmless because a call rewrites every flag it later reads, but two
 * calls that overlap on the GPU (different streams, no dependency between them) RACE: serialise them. */
#define MMB_STATIC_PEAK_MAX_FRAMES (1 << 22)
MMB_API int mmb_peak_picking_static(const float* act_dev, const int64_t* seg_offsets_dev, int32_t n_segments,
                                    int64_t n, double threshold, int32_t smooth, const double* smooth_kernel_dev,
                                    int32_t pre_avg, int32_t post_avg, int32_t pre_max, int32_t post_max,
                                    int32_t* peaks_dev, int32_t peaks_per_segment, int32_t* counts_dev,
                                    void* stream);

/* complex_flux() of madmom/features/onsets.py:245-317 from the lgd and diff outputs of mmb_run:
 * temporal maximum filter of width temporal_filter (odd, reflect borders inside each signal; <= 1: none) on lgd,
 * mask[b] = min over bins [band_lo[b], band_hi[b]) (the band's non-zero bins plus one either side, clipped),
 * odf[row] = sum_b diff[row, b] * mask[b].  Signals are either equal-length (rows_per_signal > 0), given by
 * seg_offsets_dev[n_segments + 1], or the whole matrix is one signal. */
MMB_API int mmb_complex_flux(const float* lgd_dev, int64_t ld_lgd, const float* diff_dev, int64_t ld_diff,
                             const int32_t* band_lo_dev, const int32_t* band_hi_dev, int32_t n_bands, int32_t n_bins,
                             int64_t rows, int64_t rows_per_signal, const int64_t* seg_offsets_dev, int32_t n_segments,
                             int32_t temporal_filter, float* odf_dev, void* stream);

/* phase() and local_group_delay() of madmom/audio/stft.py:136-179 (the Phase / LocalGroupDelay classes, :556-708) on a
 * complex64 STFT on the device (rows x cols, re / im pairs): phase = angle(stft); lgd[:, k] = unwrap(phase)[k] -
 * unwrap(phase)[k + 1] with numpy's unwrap rule, last column 0.  Either output may be null.  The arctangent is a
 * polynomial one (3.2e-7 rad; exact on the axes with atan2's signs, signed zeros included). */
MMB_API int mmb_phase_lgd(const void* stft_dev, int64_t ld_stft, float* phase_dev, int64_t ld_phase, float* lgd_dev,
                          int64_t ld_lgd, int64_t rows, int32_t cols, void* stream);

/* PCM ingest (madmom/audio/signal.py:170-223 `remix`, called from Signal.__new__ :600-632): interleaved
 * (num_frames, num_channels) int16 (sample_bytes 2) or float32 (4) on the device -> mono of the same type.
 * channel = -1: mean over the channels with the reference's rounding (np.mean(..., axis=-1).astype(dtype): integer
 * samples truncate towards zero); channel >= 0: that channel verbatim. */
MMB_API int mmb_remix_mono(const void* in_dev, int32_t sample_bytes, int64_t num_frames, int32_t num_channels,
                           int32_t channel, void* out_dev, void* stream);

/* mmb_run_host for interleaved multi-channel host PCM: unit offsets / lengths count sample FRAMES; each group is
 * copied to the device interleaved and mixed down there (mmb_remix_mono) before the front end runs. */
MMB_API int mmb_run_host_interleaved(const mmb_plan* plan, const mmb_track* units, int32_t n, const void* pcm_host,
                                     int32_t num_channels, int32_t channel, const mmb_outputs* outs_host,
                                     int32_t* launches);

/* The other spectral onset detection functions of madmom/features/onsets.py, one value per row:
 *   HFC                      high_frequency_content   :102-127   mean_k spec[t,k] * k
 *   SPECTRAL_DIFF            spectral_diff            :130-160   sum_k max(spec[t,k] - spec[t-d,k], 0)^2
 *   MKL                      modified_kullback_leibler :320-362  mean_k log(1 + spec[t,k] / (spec[t-d,k] + eps))
 *   PHASE_DEV                phase_deviation          :390-414   mean_k |wrap(phi[t] - 2 phi[t-1] + phi[t-2])|
 *   WEIGHTED_PHASE_DEV       weighted_phase_deviation :417-446   mean_k |wrap(..) * |X[t,k]||
 *   NORM_WEIGHTED_PHASE_DEV  normalized_weighted_...  :449-479   the above / (mean_k |X[t,k]| + eps)
 *   COMPLEX_DOMAIN           complex_domain           :482-549   sum_k |X[t,k] - |X[t-1,k]| e^{i(2 phi[t-1] - phi[t-2])}|
 *   RECTIFIED_COMPLEX_DOMAIN rectified_complex_domain :552-581   the same over bins whose magnitude rose against t-d
 * The first three read a magnitude matrix `spec_dev` (any spectrogram: raw, filtered, logarithmic), the others the
 * complex64 STFT `stft_dev` (re, im pairs).  Rows before the start of a signal count as zero rows of the lagged
 * terms, as in the reference.  Signals: see mmb_complex_flux.  The three phase-deviation functions take the wrapped
 * second difference as the argument of X[t] conj(X[t-1])^2 X[t-2] (one arctangent per bin, factors scaled by exact
 * powers of two); a bin that is exactly zero counts as phase 0, like np.angle(0). */
enum {
  MMB_ODF_HFC = 1, MMB_ODF_SPECTRAL_DIFF = 2, MMB_ODF_MKL = 3, MMB_ODF_PHASE_DEV = 4,
  MMB_ODF_WEIGHTED_PHASE_DEV = 5, MMB_ODF_NORM_WEIGHTED_PHASE_DEV = 6, MMB_ODF_COMPLEX_DOMAIN = 7,
  MMB_ODF_RECTIFIED_COMPLEX_DOMAIN = 8
};
MMB_API int mmb_spectral_odf(int32_t mode, const void* stft_dev, int64_t ld_stft, const float* spec_dev,
                             int64_t ld_spec, int64_t rows, int32_t cols, int64_t rows_per_signal,
                             const int64_t* seg_offsets_dev, int32_t n_segments, int32_t diff_frames, double epsilon,
                             float* odf_dev, void* stream);

/* Calibration: measured FP32 FMA throughput of `device` in TFLOP/s (dependent FFMA chains on every SM,
 * best of 5, CUDA events).  The front end is FP32-pipe bound; bench.py reports its roofline against this. */
MMB_API int mmb_measure_fp32_peak(int32_t device, double* tflops);

/* Instruction-rate probes behind the issue-slot roofline in DESIGN.md: which = 1 FFMA with three register
 * operands, 2 FADD, 3 FMUL+FADD pairs, 4 LDS.64 + 2 FFMA, 5 LDS.32, 6 SHFL, 7 LDS.32 + SHFL, 8 FADD2 (f32x2),
 * 9 FFMA2 (f32x2), 10 FFMA2 + LDS.32 + FADD, 11 LDS.64, 12 LDS.128, 13 STS.64 + LDS.64, 14 STS.64 (one step of
 * 11 .. 14 counts as one instruction).  *ginst_per_s = 1e9 thread-level FP instructions
 * per second over the whole device. */
MMB_API int mmb_rate_probe(int32_t device, int32_t which, double* ginst_per_s);

/* Ceiling of the fused kernel: frames per second of its warp-level real FFT alone (frame_size 2048: both 32-point
 * register passes, the shared-memory transpose, inter-pass twiddles, real-FFT split and magnitudes, 16 warps per SM;
 * no PCM staging, window, filterbank, difference or stores).  bench.py reports the fused kernel against it. */
MMB_API int mmb_fft_probe(int32_t device, int32_t frame_size, double* frames_per_s);

/* Design probe: the per-frame work of the fused kernel (staged-sample and window loads, warp FFT, split, magnitudes,
 * a filterbank-like pass, log / difference-like arithmetic) as mode 0: 16 warps that each do everything, mode 1: 12
 * FFT warps + 12 light warps with their register budgets moved by setmaxnreg.  Synthetic data, no hand-off
 * synchronisation; frames per second of the mix. */
MMB_API int mmb_pipeline_probe(int32_t device, int32_t mode, double* frames_per_s);

#ifdef __cplusplus
}
#endif
#endif /* MADMOM_B200_H */
