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
//   CTA    = persistent (one per SM); its warps form one group (or several independent ones, group_warps) that
//            strides over chunks and walks a chunk in rounds of FPR consecutive frames
//   round  = the PCM span of the round sits in the group's staging buffer, put there by ONE cp.async.bulk (TMA bulk
//            copy, mbarrier completion) that the last warp to have read the previous span issued; spans that touch
//            the edge of a signal are filled by the threads with zero padding instead
//   warp   = owns one frame per round (two for N = 1024) end to end: register-resident 32x32 FFT
//            (fft_core.cuh) -> magnitudes in its scratch -> its own filterbank pass (FbPass schedule) -> log ->
//            one row of the group's log-spectrogram ring
//   sync   = no CTA barrier in the steady state: an mbarrier counts the rows of a round; a warp waits for it only
//            when it is about to write its row of the NEXT round and then does the lagged difference, flux and
//            output stores of the previous one (warps drift by up to one round)
// A chunk recomputes its `diff_frames` lead-in frames instead of exchanging them with its neighbour.
#include "frame_fft.cuh"
#include "frontend_params.h"

namespace mmb {

__device__ __forceinline__ long long frame_start_sample(long long g, double hop, int N, int origin) {
  // int(index * hop_size) - frame_size // 2 - origin, float64 product, truncation (signal.py:917-919)
  return __double2ll_rz((double)g * hop) - (long long)(N / 2) - (long long)origin;
}

// ---------------------------------------------------------------------------------------------- mbarrier / TMA
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@!P1 bra WAIT_LOOP;\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (UBLKCP); dst, src 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------- round cursor
// One lane walks the rounds of its warp group; `d` is advanced to the round that follows.  Group `vid` of `vstride`
// groups in the grid takes the chunks vid, vid + vstride, ... of the launch's work list (FrontendParams::chunks).
template <int N, typename S>
__device__ void next_round(const FrontendParams& p, RoundDesc& d, bool first, int fpr, int vid, int vstride) {
  const int lead = ((p.out_diff != nullptr) || (p.out_odf != nullptr)) ? p.diff_frames : 0;
  // ring slots rotate from round to round, across chunk changes too: a round takes the slots behind those of the
  // round before it, so the rows a pending phase C still needs are never the ones being written (the ring holds two
  // rounds + d rows) and a new chunk needs neither a drain nor a barrier
  int slot = first ? 0 : d.slot0 + d.nfr;
  if (slot >= p.ring_rows) slot -= p.ring_rows;
  long long g0;
  bool new_chunk = first;
  int chunk = first ? vid : d.chunk;
  if (!first) {
    g0 = d.g0 + d.nfr;
    if (g0 >= d.gb) { chunk += vstride; new_chunk = true; }
  }
  if (new_chunk) {
    // the next entry of the work list that holds frames (holes: the lists of the groups differ in length)
    Chunk c;
    for (;;) {
      if (chunk >= p.total_chunks) { d.nfr = 0; return; }
      c = p.chunks[chunk];
      if (c.frames > 0) break;
      chunk += vstride;
    }
    const int lo = c.unit;
    const Track tr = p.tracks[lo];
    const long long a_loc = c.first;
    const long long b_loc = min(a_loc + (long long)c.frames, (long long)tr.frame_count);
    d.chunk = chunk; d.unit = lo;
    d.ga = tr.frame_first + a_loc; d.gb = tr.frame_first + b_loc;
    d.row0 = tr.out_row - tr.frame_first;
    d.pcm_offset = tr.pcm_offset; d.pcm_len = tr.pcm_len;
    d.sample_base = tr.sample_base;
    g0 = max(d.ga - (long long)lead, 0LL);
    // N = 1024 packs frames (2m, 2m + 1) into one FFT: pairing by the global frame index (one more lead-in frame when
    // needed; the host cuts in front of even frames, capi.cu) keeps a frame's bits independent of how its unit is cut
    if (N == 1024) g0 &= ~1LL;
  }
  d.g0 = g0;
  d.nfr = (int)min((long long)fpr, d.gb - g0);
  d.s_begin = frame_start_sample(g0, p.hop, N, p.origin);
  const long long s_end = frame_start_sample(g0 + d.nfr - 1, p.hop, N, p.origin) + N;
  d.span = (int)(s_end - d.s_begin);
  d.l0 = d.s_begin - d.sample_base;
  d.slot0 = slot;
  // bulk staging: every staged sample is a real sample of the unit and the 16-byte aligned hull stays inside
  // the PCM buffer
  constexpr int A = 16 / (int)sizeof(S);
  const long long abs0 = d.pcm_offset + d.l0;
  const long long e0 = abs0 & ~(long long)(A - 1);
  const long long e1 = (abs0 + d.span + A - 1) & ~(long long)(A - 1);
  d.bulk = p.bulk_allowed && d.l0 >= 0 && d.l0 + d.span <= d.pcm_len && e1 <= p.pcm_extent &&
           (e1 - e0) <= (long long)p.span_cap;
  d.shift = d.bulk ? (int)(abs0 - e0) : 0;
}

// barrier of one warp group (named barrier 1 + group; a single-warp group needs only __syncwarp)
__device__ __forceinline__ void group_sync(int grp, int gthreads) {
  if (gthreads == 32) __syncwarp();
  else asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "r"(gthreads) : "memory");
}

template <typename S>
__device__ __forceinline__ void issue_bulk(const FrontendParams& p, const RoundDesc& d, S* buf, uint64_t* bar) {
  constexpr int A = 16 / (int)sizeof(S);
  const long long abs0 = d.pcm_offset + d.l0;
  const long long e0 = abs0 & ~(long long)(A - 1);
  const long long e1 = (abs0 + d.span + A - 1) & ~(long long)(A - 1);
  const uint32_t bytes = (uint32_t)((e1 - e0) * (long long)sizeof(S));
  fence_proxy_async();                            // earlier generic-proxy accesses of buf are ordered before the copy
  mbar_expect_tx(bar, bytes);
  bulk_g2s(buf, reinterpret_cast<const S*>(p.pcm) + e0, bytes, bar);
}

// Called by every warp of a group once per round, after its last read of the staged span.  The last warp to
// arrive describes the following round and, when that round can be staged by a bulk copy, starts the copy into the
// (now idle) span buffer: it lands while the group is busy with the FFTs of the current round.  Otherwise it just
// completes the barrier phase and the threads fill the span at the start of the round.
template <int N, typename S>
__device__ __forceinline__ void frames_loaded(const FrontendParams& p, GroupCtl* G, int r, S* buf, int lane, int gw,
                                              int fpr, int vid, int vstride) {
  __syncwarp();
  if (lane == 0) {
    __threadfence_block();
    if (atomicAdd(&G->loaded, 1) == gw - 1) {
      G->loaded = 0;
      __threadfence_block();
      RoundDesc* Dn = &G->desc[(r + 1) & 3];
      *Dn = G->desc[r & 3];
      next_round<N, S>(p, *Dn, false, fpr, vid, vstride);
      if (Dn->nfr > 0 && Dn->bulk) issue_bulk<S>(p, *Dn, buf, reinterpret_cast<uint64_t*>(&G->mbar));
      else mbar_arrive(reinterpret_cast<uint64_t*>(&G->mbar));
    }
  }
}

// ---------------------------------------------------------------------------------------------- per-warp stages
// Filterbank of one frame (two for N = 1024) owned by this warp: FbPass schedule over the magnitudes in the
// warp's scratch, then the per-band sum of pieces, log scaling and the ring row.  `mag1` is the second frame of a
// 1024-point pair or null.
// SPLIT: the magnitudes are stored as even bins in `mag0` and odd bins in `mag1` (N = 4096).
// The warp takes the passes ps0, ps0 + ps_step, ...: the two warps of a 4096-point frame share the passes of one
// frame (ps0 = 0 / 1, ps_step = 2) and write into one set of partial sums.
template <int NF, bool SPLIT = false>
__device__ __forceinline__ void warp_filterbank(const FrontendParams& p, int lane, const float* __restrict__ mag0,
                                                const float* __restrict__ mag1, float* part,
                                                const FbPass* __restrict__ s_pass, const float* __restrict__ s_w,
                                                const int* __restrict__ s_start, const int* __restrict__ s_pid,
                                                int ps0 = 0, int ps_step = 1) {
  // the previous frame's epilogue reads of `part` are complete: the warp has passed a __syncwarp since
  const int np1 = p.max_q * p.b_pad + 1;
  if (p.L.part_stride == 0) {          // the slots share memory with the FFT exchange tile: clear them per frame
    const int n4 = (NF * np1 + 3) >> 2;
    for (int i = lane; i < n4; i += 32) reinterpret_cast<float4*>(part)[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    __syncwarp();
  }
  for (int ps = ps0; ps < p.n_passes; ps += ps_step) {
    const FbPass fp = s_pass[ps];
    const int st = s_start[ps * 32 + lane];
    const float4* w = reinterpret_cast<const float4*>(s_w + fp.woff) + lane;
    // SPLIT: taps 0, 2 read bins st, st + 2 (array of st's parity), taps 1, 3 the other array
    const float* m0 = SPLIT ? ((st & 1) ? mag1 + (st >> 1) : mag0 + (st >> 1)) : mag0 + st;
    const float* m1 = SPLIT ? ((st & 1) ? mag0 + ((st + 1) >> 1) : mag1 + (st >> 1)) : ((NF == 2) ? mag1 + st : mag0);
    float2 acc0 = make_float2(0.0f, 0.0f), acc1 = make_float2(0.0f, 0.0f);    // (even taps, odd taps), two frames
    // four taps of this lane per group (pass lengths are multiples of four, capi.cu); groups are taken two at a time
    // (an odd one first) so that the loop overhead is paid once per eight taps; the order of the FMAs is unchanged
    constexpr int MS = SPLIT ? 2 : 4;                // magnitudes a group advances in each array
    auto group = [&](int g) {
      const float4 wv = w[32 * g];                   // taps 4 g .. 4 g + 3 of this lane
      const float2 w01 = make_float2(wv.x, wv.y), w23 = make_float2(wv.z, wv.w);
      const float* a = m0 + MS * g;
      const float* b = m1 + MS * g;
      const float x0 = a[0], x1 = SPLIT ? b[0] : a[1], x2 = SPLIT ? a[1] : a[2], x3 = SPLIT ? b[1] : a[3];
      acc0 = fma2(make_float2(x0, x1), w01, acc0);   // packed pairs: same per-half FMA chains as the scalar form
      acc0 = fma2(make_float2(x2, x3), w23, acc0);
      if (!SPLIT && NF == 2) {
        const float y0 = b[0], y1 = b[1], y2 = b[2], y3 = b[3];
        acc1 = fma2(make_float2(y0, y1), w01, acc1);
        acc1 = fma2(make_float2(y2, y3), w23, acc1);
      }
    };
    const int n4 = fp.iters >> 2;
    if (n4 & 1) {
      group(0);
      w += 32; m0 += MS; if (SPLIT || NF == 2) m1 += MS;
    }
#pragma unroll 1
    for (int k = n4 >> 1; k > 0; --k) {
      group(0);
      group(1);
      w += 64; m0 += 2 * MS; if (SPLIT || NF == 2) m1 += 2 * MS;
    }
    const float a0 = acc0.x, c0 = acc0.y, a1 = acc1.x, c1 = acc1.y;
    const int pid = s_pid[ps * 32 + lane];
    part[pid] = a0 + c0;
    if (NF == 2) part[np1 + pid] = a1 + c1;
  }
  __syncwarp();
}

// rows of the ring / outputs for one frame g handled by this warp
__device__ __forceinline__ void warp_band_epilogue(const FrontendParams& p, int lane, const float* part,
                                                   float* ring_row, long long row, bool emit, int b0 = 0,
                                                   int b_step = 32) {
  const int nb = p.n_bands, bp = p.b_pad, nq = p.max_q;
  for (int b = b0 + lane; b < nb; b += b_step) {
    // pieces are added in order q = 0, 1, ...; slots without a piece stay zero.  The first four loads are issued
    // together (the piece count is warp-uniform), the rare longer tails in a loop.
    const float* pp = part + b;
    const float a0 = pp[0];
    const float a1 = nq > 1 ? pp[bp] : 0.0f;
    const float a2 = nq > 2 ? pp[2 * bp] : 0.0f;
    const float a3 = nq > 3 ? pp[3 * bp] : 0.0f;
    float acc = a0;
    if (nq > 1) acc += a1;
    if (nq > 2) acc += a2;
    if (nq > 3) acc += a3;
    for (int q = 4; q < nq; ++q) acc += pp[q * bp];
    if (emit && p.out_filt) p.out_filt[row * p.ld_filt + b] = acc;
    ring_row[b] = p.log_mode != 0 ? log_scaled(acc, p.log_mode, p.mul, p.add) : acc;
  }
}

// store through an NVLS multicast address: the NVSwitch delivers the value to the same offset of every GPU bound to
// the multicast object (this one included), so a gathered row leaves the SM once instead of once per peer
__device__ __forceinline__ void mc_store(float* a, float v) {
  asm volatile("multimem.st.weak.global.f32 [%0], %1;" ::"l"(a), "f"(v) : "memory");
}

__device__ __forceinline__ void mc_store4(float* a, float4 v) {
  asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// difference value of band b: cur[b] - max over the frequency window of prev (scipy maximum_filter, 'reflect'),
// optionally rectified; 0 for the first d frames of a signal
__device__ __forceinline__ float band_difference(const FrontendParams& p, const float* cur, const float* prev,
                                                 bool has_prev, int b, int nb) {
  if (!has_prev) return 0.0f;
  const int m = p.max_bins;
  float ref;
  if (m == 3) {                                          // the SuperFlux case, 'reflect' duplicates the edge bin
    ref = fmaxf(fmaxf(prev[b > 0 ? b - 1 : 0], prev[b]), prev[b + 1 < nb ? b + 1 : nb - 1]);
  } else if (m > 1) {                                    // scipy maximum_filter, mode='reflect'
    const int half = m / 2;
    ref = -INFINITY;
    for (int j = 0; j < m; ++j) {
      int q = b - half + j;
      q = q < 0 ? -q - 1 : (q >= nb ? 2 * nb - q - 1 : q);
      ref = fmaxf(ref, prev[q]);
    }
  } else {
    ref = prev[b];
  }
  const float dv = cur[b] - ref;
  return p.positive ? fmaxf(dv, 0.0f) : dv;
}

// lagged difference of frame g against the ring row of frame g - d (phase C)
// GATHER: the launch also stores to peer / multicast destinations (mmb_run_gather, mmb_run_multicast); the plain
// launches are a separate instantiation without that code (it cost them 4 % when it was a runtime branch).
// Remote rows go out as 16 bytes per lane when the host found everything aligned (p.gather_vec): a lane then owns
// four consecutive bands, recomputed with the same operations as in the lane = band mapping of the local row and
// of the flux, so every destination receives the same bits; a warp store covers 512 contiguous bytes of the peer's
// row instead of 128.
template <bool GATHER>
__device__ __forceinline__ void warp_difference(const FrontendParams& p, int lane, const float* cur,
                                                const float* prev, bool has_prev, long long row) {
  const int nb = p.n_bands, nx = GATHER ? p.n_extra : 0;
  const bool mc = GATHER && p.multicast != 0;
  const bool vec = GATHER && p.gather_vec != 0;
  if (p.out_log) {
    float* o = p.out_log + row * p.ld_log;
    if (mc && vec) {
      for (int b = 4 * lane; b < nb; b += 128) mc_store4(o + b, *reinterpret_cast<const float4*>(cur + b));
    } else if (mc) {
      for (int b = lane; b < nb; b += 32) mc_store(o + b, cur[b]);
    } else {
      for (int b = lane; b < nb; b += 32) o[b] = cur[b];
    }
    if (vec) {
      for (int b = 4 * lane; b < nb; b += 128) {
        const float4 v = *reinterpret_cast<const float4*>(cur + b);
        for (int t = 0; t < nx; ++t) *reinterpret_cast<float4*>(p.x_log[t] + row * p.ld_log + b) = v;
      }
    } else {
      for (int t = 0; t < nx; ++t) {                     // fused all-gather: the same row into the peers' matrices
        float* ox = p.x_log[t] + row * p.ld_log;
        for (int b = lane; b < nb; b += 32) ox[b] = cur[b];
      }
    }
  }
  if (p.out_diff == nullptr && p.out_odf == nullptr) return;
  float sum = 0.0f;
  for (int b = lane; b < nb; b += 32) {
    const float dv = band_difference(p, cur, prev, has_prev, b, nb);
    if (p.out_diff) {
      if (mc) { if (!vec) mc_store(p.out_diff + row * p.ld_diff + b, dv); }
      else p.out_diff[row * p.ld_diff + b] = dv;
      if (!vec) for (int t = 0; t < nx; ++t) p.x_diff[t][row * p.ld_diff + b] = dv;
    }
    sum += dv;
  }
  if (vec && p.out_diff) {
    for (int b = 4 * lane; b < nb; b += 128) {
      float4 v;
      v.x = band_difference(p, cur, prev, has_prev, b, nb);
      v.y = band_difference(p, cur, prev, has_prev, b + 1, nb);
      v.z = band_difference(p, cur, prev, has_prev, b + 2, nb);
      v.w = band_difference(p, cur, prev, has_prev, b + 3, nb);
      if (mc) mc_store4(p.out_diff + row * p.ld_diff + b, v);
      for (int t = 0; t < nx; ++t) *reinterpret_cast<float4*>(p.x_diff[t] + row * p.ld_diff + b) = v;
    }
  }
  if (p.out_odf) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) {
      if (mc) mc_store(p.out_odf + row, sum);
      else p.out_odf[row] = sum;
      for (int t = 0; t < nx; ++t) p.x_odf[t][row] = sum;
    }
  }
}

// phase C of one round: max-filter + lagged difference + flux of its frames, coalesced row stores
template <bool GATHER>
__device__ __forceinline__ void round_difference(const FrontendParams& p, const RoundDesc* Dc, const float* s_ring,
                                                 int gwarp, int GW, int lane, bool want_diff) {
  const int nb = p.n_bands, R = p.ring_rows, nfr = Dc->nfr, slot0 = Dc->slot0;
  const long long g0 = Dc->g0, ga = Dc->ga, row0 = Dc->row0;
  for (int fi = gwarp; fi < nfr; fi += GW) {
    const long long g = g0 + fi;
    if (g < ga) continue;                                   // lead-in frame: ring only
    int sc = slot0 + fi; while (sc >= R) sc -= R;
    int sp = sc - p.diff_frames; if (sp < 0) sp += R;
    const bool has_prev = want_diff && (g - p.diff_frames) >= 0;       // first d rows of a signal are zero
    warp_difference<GATHER>(p, lane, s_ring + sc * nb, s_ring + sp * nb, has_prev, row0 + g);
  }
}

// ---------------------------------------------------------------------------------------------- the kernel
// SPECOUT: per-bin outputs the launch may write: 0 none, 2 lgd (N = 2048): its own instantiation, so that launches
// that only produce band outputs carry no emitting copy of the split.  (Raw bins and magnitudes -- the `stft` and
// `spec` outputs -- come from the float64 kernel in stft64.cu, not from here.)
// WCOS: the window is a symmetric cosine sum and is evaluated instead of loaded (frame_fft.cuh; N = 2048, 4096)
template <int N, typename S, bool GATHER, int SPECOUT, bool WCOS>
__global__ void __launch_bounds__(kMaxWarps * 32, 1) frontend_kernel(const FrontendParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  constexpr int K = N / 2;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nthreads = blockDim.x, warps = nthreads >> 5;
  const int GW = p.group_warps, gthreads = GW * 32;
  const int grp = warp / GW, gwarp = warp - grp * GW, gtid = tid - grp * gthreads;
  const int vid = (int)blockIdx.x * (warps / GW) + grp, vstride = (int)gridDim.x * (warps / GW);
  const int FPR = frames_per_round(N, GW);
  const SmemLayout& L = p.L;
  GroupCtl* G = reinterpret_cast<GroupCtl*>(smem + L.ctl + grp * L.ctl_stride);
  uint64_t* g_mbar = reinterpret_cast<uint64_t*>(&G->mbar);
  float* s_win = reinterpret_cast<float*>(smem + L.win);
  float2* s_tw1 = reinterpret_cast<float2*>(smem + L.tw1);
  float2* s_tw2k = reinterpret_cast<float2*>(smem + L.tw2k);
  float2* s_tw4k = reinterpret_cast<float2*>(smem + L.tw4k);
  FbPass* s_pass = reinterpret_cast<FbPass*>(smem + L.fb_pass);
  float* s_w = reinterpret_cast<float*>(smem + L.fb_w);
  int* s_start = reinterpret_cast<int*>(smem + L.fb_start);
  int* s_pid = reinterpret_cast<int*>(smem + L.fb_pid);
  float* s_ring = reinterpret_cast<float*>(smem + L.ring + grp * L.ring_stride);
  S* buf = reinterpret_cast<S*>(smem + L.pcm + grp * L.pcm_stride);

  unsigned char* my_scratch = smem + L.scratch + warp * warp_scratch_bytes(N);
  float2* ex = reinterpret_cast<float2*>(my_scratch);
  float* my_mag = reinterpret_cast<float*>(my_scratch);     // N = 4096: even bins in warp 2f, odd bins in warp 2f + 1
  // N = 4096: the magnitude arrays of this warp's pair and the pair's barrier (ids behind the groups' barriers)
  const float* pair_even = reinterpret_cast<const float*>(smem + L.scratch + (warp & ~1) * warp_scratch_bytes(N));
  const float* pair_odd = reinterpret_cast<const float*>(smem + L.scratch + (warp | 1) * warp_scratch_bytes(N));
  const int pair_bar = 1 + warps / GW + (warp >> 1);
  float* my_part = L.part_stride == 0 ? reinterpret_cast<float*>(my_scratch + kPartAliasOffset)
                                      : reinterpret_cast<float*>(smem + L.part + (warp / part_sharers(N)) * L.part_stride);

  // ---- plan constants -> shared memory (once per persistent CTA)
  for (int i = tid; i < (WCOS ? kWinCosFloats : N); i += nthreads) s_win[i] = p.win[i];
  for (int i = tid; i < 1024; i += nthreads) s_tw1[i] = p.tw1[i];
  if (N == 2048) for (int i = tid; i < 512; i += nthreads) s_tw2k[i] = p.tw2k[i];
  if (N == 4096) {
    for (int i = tid; i < 1024; i += nthreads) s_tw2k[i] = p.tw2k[i];
    for (int i = tid; i < 512; i += nthreads) s_tw4k[i] = p.tw4k[i];
  }
  const int nb = p.n_bands;
  if (nb > 0) {
    for (int i = tid; i < p.n_passes; i += nthreads) s_pass[i] = p.fb_pass[i];
    for (int i = tid; i < p.w_floats; i += nthreads) s_w[i] = p.fb_w[i];
    for (int i = tid; i < p.n_passes * 32; i += nthreads) { s_start[i] = p.fb_start[i]; s_pid[i] = p.fb_pid[i]; }
    for (int i = tid; i < (warps / part_sharers(N)) * (L.part_stride / 4); i += nthreads)     // private slots only
      reinterpret_cast<float*>(smem + L.part)[i] = 0.0f;     // part slots without a piece are never written again
  }
  uint64_t* ring_bar = reinterpret_cast<uint64_t*>(&G->ring_bar);
  if (gtid == 0) {
    mbar_init(g_mbar, 1);
    mbar_init(ring_bar, GW);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    G->loaded = 0;
    G->described = 0;
    next_round<N, S>(p, G->desc[0], true, FPR, vid, vstride);
    if (G->desc[0].nfr > 0 && G->desc[0].bulk) issue_bulk<S>(p, G->desc[0], buf, g_mbar);
    else mbar_arrive(g_mbar);
  }
  __syncthreads();

  const bool want_diff = (p.out_diff != nullptr) || (p.out_odf != nullptr);
  const float sgn_odd = p.circular_shift ? -1.0f : 1.0f;
  const int R = p.ring_rows;
  uint32_t pcm_phase = 0, ring_phase = 0;   // parities of the next completions of the group's two mbarriers
  bool pending = false;                     // phase C of the previous round is still to be done by this warp

  // Phase C of round r-1 is deferred until this warp is about to write its ring row of round r: by then every
  // warp of the group has long finished its row of round r-1, so the wait costs nothing and the warps are free
  // to drift apart by up to one round (the ring holds 2 * FPR + d rows).
#define MMB_PENDING_DIFFERENCE()                                                                       \
  if (pending) {                                                                                       \
    mbar_wait(ring_bar, ring_phase);                                                                   \
    ring_phase ^= 1;                                                                                   \
    round_difference<GATHER>(p, &G->desc[(r + 3) & 3], s_ring, gwarp, GW, lane, want_diff);             \
    pending = false;                                                                                   \
  }
#define MMB_ROWS_WRITTEN()                                                                             \
  {                                                                                                    \
    __syncwarp();                                                                                      \
    if (lane == 0) mbar_arrive(ring_bar);                                                              \
    pending = true;                                                                                    \
  }

  int r = 0;
  for (;; ++r) {
    mbar_wait(g_mbar, pcm_phase);           // descriptor of round r is written, its span staged if `bulk`
    pcm_phase ^= 1;
    const RoundDesc* D = &G->desc[r & 3];
    const int nfr = D->nfr;
    if (nfr == 0) break;
    // ---- spans that touch the edge of a signal are filled here with zero padding (all warps have read the
    //      previous span: the descriptor is written by the last of them)
    if (!D->bulk) {
      const long long l0 = D->l0, plen = D->pcm_len;
      const S* src = reinterpret_cast<const S*>(p.pcm) + D->pcm_offset;
      const int span = D->span;
      for (int i = gtid; i < span; i += gthreads) {
        const long long li = l0 + i;
        S v = (S)0;
        if (li >= 0 && li < plen) v = src[li];
        buf[i] = v;
      }
      group_sync(grp, gthreads);
    }
    const S* span0 = buf + D->shift;

    // ================= this warp's frame(s): FFT -> magnitudes -> filterbank -> [phase C of round r-1] -> ring row
    // The round descriptor is read again after the FFT instead of keeping its 64-bit fields in registers across it.
    float2 v[32];
    if (N == 1024) {
      const int fa = 2 * gwarp, fb = 2 * gwarp + 1;
      if (fa >= nfr) {
        frames_loaded<N, S>(p, G, r, buf, lane, GW, FPR, vid, vstride);
        if (nb > 0) { MMB_PENDING_DIFFERENCE(); MMB_ROWS_WRITTEN(); }
      } else {
        const bool haveB = fb < nfr;
        {
          const long long gA = D->g0 + fa, s_begin = D->s_begin;
          const S* xa = span0 + (int)(frame_start_sample(gA, p.hop, N, p.origin) - s_begin);
          const S* xb = haveB ? span0 + (int)(frame_start_sample(gA + 1, p.hop, N, p.origin) - s_begin) : nullptr;
          load_frames_1024<S>(lane, v, xa, xb, s_win);
        }
        // The two frames share one complex FFT, and the split separates them only up to rounding: a frame carries
        // about 1e-7 of its partner's bins.  Between frames of music that is nothing, but a frame that is exactly
        // silent after windowing (zero padding, digital silence, the all-but-one-sample first frame of an online
        // framing under a window whose last tap is 0) must read exactly 0 like in the reference, which transforms
        // every frame on its own.  So: which of the two frames has no non-zero sample?  (Depends on the frame's own
        // samples only: rows stay independent of chunking and sharding.)
        bool silentA, silentB;
        {
          unsigned bitsA = 0, bitsB = 0;
#pragma unroll
          for (int m = 0; m < 32; ++m) { bitsA |= __float_as_uint(v[m].x); bitsB |= __float_as_uint(v[m].y); }
          silentA = __all_sync(0xffffffffu, (bitsA << 1) == 0);         // the sign of a zero does not count
          silentB = __all_sync(0xffffffffu, (bitsB << 1) == 0);         // true when there is no frame B
        }
        frames_loaded<N, S>(p, G, r, buf, lane, GW, FPR, vid, vstride);
        core_pass1(lane, v, ex, s_tw1);
        __syncwarp();
        core_pass2_load(lane, v, ex);
        __syncwarp();
        core_pass2_fft(v);
        core_publish_upper(lane, v, ex, true);
        __syncwarp();
        const long long gA = D->g0 + fa, gB = gA + 1, ga = D->ga;
        const long long rowA = D->row0 + gA, rowB = rowA + 1;
        {
          BinSinkT<false> oa, ob;
          oa.mag = my_mag; ob.mag = my_mag + 512;
          split_1024x2(lane, v, ex, oa, ob, haveB);
        }
        if (nb > 0) {
          __syncwarp();
          if (silentA) for (int i = lane; i < 512; i += 32) my_mag[i] = 0.0f;
          if (silentB) for (int i = lane; i < 512; i += 32) my_mag[512 + i] = 0.0f;
          __syncwarp();
          warp_filterbank<2>(p, lane, my_mag, my_mag + 512, my_part, s_pass, s_w, s_start, s_pid);
          MMB_PENDING_DIFFERENCE();
          int sa = D->slot0 + fa; if (sa >= R) sa -= R;
          warp_band_epilogue(p, lane, my_part, s_ring + sa * nb, rowA, gA >= ga);
          if (haveB) {
            int sb = sa + 1; if (sb >= R) sb -= R;
            warp_band_epilogue(p, lane, my_part + p.max_q * p.b_pad + 1, s_ring + sb * nb, rowB, gB >= ga);
          }
          MMB_ROWS_WRITTEN();
        }
      }
    } else if (N == 4096) {
      // two warps per frame: the radix-2 pre-stage splits the packed 2048-point sequence into the half that gives the
      // even bins (warp 2f, c = 0) and the one that gives the odd bins (warp 2f + 1, c = 1); each is a 1024-point
      // complex FFT of its own.  The pair meets twice in the load phase (above) and twice more: before the filterbank (both magnitude arrays are in place;
      // each warp then takes every other pass) and after it (all partial sums of the frame are written).
      const int f = gwarp >> 1, c = gwarp & 1;
      if (f >= nfr) {
        frames_loaded<N, S>(p, G, r, buf, lane, GW, FPR, vid, vstride);
        if (nb > 0) { MMB_PENDING_DIFFERENCE(); MMB_ROWS_WRITTEN(); }
      } else {
        const S* x = span0 + (int)(frame_start_sample(D->g0 + f, p.hop, N, p.origin) - D->s_begin);
        // each warp loads half of the frame and the pair swaps the pre-stage results (frame_fft.cuh)
        load_frame_4096_half<S, WCOS>(lane, c, v, x, s_win, s_tw2k, ex);
        frames_loaded<N, S>(p, G, r, buf, lane, GW, FPR, vid, vstride);
        asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
        take_frame_4096_half(lane, c, v, reinterpret_cast<const float2*>(c ? pair_even : pair_odd));
        asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");      // before pass 1 overwrites the tiles
        core_pass1(lane, v, ex, s_tw1);
        __syncwarp();
        core_pass2_load(lane, v, ex);
        __syncwarp();
        core_pass2_fft(v);
        core_publish_upper(lane, v, ex, c == 0);
        __syncwarp();
        BinSinkT<false> o;
        o.mag = my_mag;
        split_4096(lane, c, v, ex, s_tw2k, s_tw4k, o);
        if (nb > 0) {
          asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
          warp_filterbank<1, true>(p, lane, pair_even, pair_odd, my_part, s_pass, s_w, s_start, s_pid, c, 2);
          asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
          MMB_PENDING_DIFFERENCE();
          const long long g = D->g0 + f;
          int sl = D->slot0 + f; if (sl >= R) sl -= R;
          warp_band_epilogue(p, lane, my_part, s_ring + sl * nb, D->row0 + g, g >= D->ga, 32 * c, 64);
          MMB_ROWS_WRITTEN();
        }
      }
    } else if (gwarp >= nfr) {
      frames_loaded<N, S>(p, G, r, buf, lane, GW, FPR, vid, vstride);
      if (nb > 0) { MMB_PENDING_DIFFERENCE(); MMB_ROWS_WRITTEN(); }
    } else {
      const S* x = span0 + (int)(frame_start_sample(D->g0 + gwarp, p.hop, N, p.origin) - D->s_begin);
      {
        if (WCOS) load_frame_2048_wcos<S>(lane, v, x, s_win);
        else load_frame_2048<S>(lane, v, x, s_win);
        frames_loaded<N, S>(p, G, r, buf, lane, GW, FPR, vid, vstride);
        core_pass1(lane, v, ex, s_tw1);
        __syncwarp();
        core_pass2_load(lane, v, ex);
        __syncwarp();
        core_pass2_fft(v);
        core_publish_upper(lane, v, ex, true);
        __syncwarp();
        BinSinkT<false> o;
        o.mag = my_mag;
        if (SPECOUT & 2) {
          const long long g = D->g0 + gwarp, row = D->row0 + g;
          if (p.out_lgd && g >= D->ga) split_2048_lgd(lane, v, ex, s_tw2k, o, p.out_lgd + row * p.ld_lgd, sgn_odd);
          else split_2048(lane, v, ex, s_tw2k, o);
        } else {
          split_2048(lane, v, ex, s_tw2k, o);
        }
      }
      if (nb > 0) {
        __syncwarp();
        warp_filterbank<1>(p, lane, my_mag, nullptr, my_part, s_pass, s_w, s_start, s_pid);
        MMB_PENDING_DIFFERENCE();
        const long long g = D->g0 + gwarp;
        int sl = D->slot0 + gwarp; if (sl >= R) sl -= R;
        warp_band_epilogue(p, lane, my_part, s_ring + sl * nb, D->row0 + g, g >= D->ga);
        MMB_ROWS_WRITTEN();
      }
    }
  }
  MMB_PENDING_DIFFERENCE();
#undef MMB_PENDING_DIFFERENCE
#undef MMB_ROWS_WRITTEN
}

// ---------------------------------------------------------------------------------------------- host launchers
template <int N, typename S, bool WCOS>
static cudaError_t launch_one(const FrontendParams& p, int grid, int smem_bytes, cudaStream_t stream) {
  const bool lgd = p.out_lgd != nullptr;
  const int threads = p.warps * 32;
  if (p.n_extra > 0 || p.multicast) frontend_kernel<N, S, true, 0, WCOS><<<grid, threads, smem_bytes, stream>>>(p);
  else if (N == 2048 && lgd) frontend_kernel<N, S, false, (N == 2048 ? 2 : 0), WCOS><<<grid, threads, smem_bytes, stream>>>(p);
  else frontend_kernel<N, S, false, 0, WCOS><<<grid, threads, smem_bytes, stream>>>(p);
  return cudaGetLastError();
}

template <int N, typename S, bool WCOS>
static cudaError_t set_one(int smem_bytes) {
  const cudaFuncAttribute a = cudaFuncAttributeMaxDynamicSharedMemorySize;
  cudaError_t e = cudaFuncSetAttribute(frontend_kernel<N, S, false, 0, WCOS>, a, smem_bytes);
  if (e == cudaSuccess && N == 2048)
    e = cudaFuncSetAttribute(frontend_kernel<N, S, false, (N == 2048 ? 2 : 0), WCOS>, a, smem_bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(frontend_kernel<N, S, true, 0, WCOS>, a, smem_bytes);
  return e;
}

template <int N, bool WCOS>
static cudaError_t launch_n(int sample_bytes, const FrontendParams& p, int grid, int smem_bytes, cudaStream_t stream) {
  if (sample_bytes == 2) return launch_one<N, int16_t, WCOS>(p, grid, smem_bytes, stream);
  if (sample_bytes == 4) return launch_one<N, float, WCOS>(p, grid, smem_bytes, stream);
  return cudaErrorInvalidValue;
}
template <int N, bool WCOS>
static cudaError_t set_n(int sample_bytes, int smem_bytes) {
  if (sample_bytes == 2) return set_one<N, int16_t, WCOS>(smem_bytes);
  if (sample_bytes == 4) return set_one<N, float, WCOS>(smem_bytes);
  return cudaErrorInvalidValue;
}

#if defined(MMB_FRONTEND_N)
// One frame size and one window path per translation unit (madmom_b200/build.py compiles this file once per
// (N, WCOS) in parallel with -DMMB_FRONTEND_N=1024 / 2048 / 4096 -DMMB_FRONTEND_WCOS=0 / 1; frontend_dispatch.cu
// routes launch_frontend / frontend_set_smem to them).
#if !defined(MMB_FRONTEND_WCOS)
#define MMB_FRONTEND_WCOS 0
#endif
#define MMB_CAT3(a, b, c) a##b##c
#define MMB_CAT(a, b, c) MMB_CAT3(a, b, c)
#if MMB_FRONTEND_WCOS
#define MMB_UNIT_SUFFIX _wcos
#else
#define MMB_UNIT_SUFFIX
#endif
cudaError_t MMB_CAT(launch_frontend_, MMB_FRONTEND_N, MMB_UNIT_SUFFIX)(int sample_bytes, const FrontendParams& p,
                                                                        int grid, int smem_bytes,
                                                                        cudaStream_t stream) {
  return launch_n<MMB_FRONTEND_N, MMB_FRONTEND_WCOS != 0>(sample_bytes, p, grid, smem_bytes, stream);
}
cudaError_t MMB_CAT(frontend_set_smem_, MMB_FRONTEND_N, MMB_UNIT_SUFFIX)(int sample_bytes, int smem_bytes) {
  return set_n<MMB_FRONTEND_N, MMB_FRONTEND_WCOS != 0>(sample_bytes, smem_bytes);
}
#else
cudaError_t launch_frontend(int N, int sample_bytes, const FrontendParams& p, int grid, int smem_bytes,
                            cudaStream_t stream) {
  const bool wcos = p.win_cos != 0;
  if (N == 1024 && !wcos) return launch_n<1024, false>(sample_bytes, p, grid, smem_bytes, stream);
  if (N == 2048) return wcos ? launch_n<2048, true>(sample_bytes, p, grid, smem_bytes, stream)
                             : launch_n<2048, false>(sample_bytes, p, grid, smem_bytes, stream);
  if (N == 4096) return wcos ? launch_n<4096, true>(sample_bytes, p, grid, smem_bytes, stream)
                             : launch_n<4096, false>(sample_bytes, p, grid, smem_bytes, stream);
  return cudaErrorInvalidValue;
}

// opt the kernel variants in to `smem_bytes` of dynamic shared memory (once per plan)
cudaError_t frontend_set_smem(int N, int sample_bytes, int smem_bytes) {
  if (N == 1024) return set_n<1024, false>(sample_bytes, smem_bytes);
  cudaError_t e = N == 2048 ? set_n<2048, false>(sample_bytes, smem_bytes)
                : N == 4096 ? set_n<4096, false>(sample_bytes, smem_bytes) : cudaErrorInvalidValue;
  if (e != cudaSuccess) return e;
  return N == 2048 ? set_n<2048, true>(sample_bytes, smem_bytes) : set_n<4096, true>(sample_bytes, smem_bytes);
}
#endif

}  // namespace mmb
