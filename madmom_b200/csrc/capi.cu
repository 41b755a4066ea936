// C ABI of libmadmom_b200.so (see include/madmom_b200.h).  Host-side plan construction: twiddle tables
// are generated in float64 and rounded once to float32; the fft_window is scaled by 0.5 (exact) because
// the kernel's real-FFT split omits that factor.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>    // header-only NVTX v3: ranges around the entry points show up in nsys / ncu timelines

#include "../../include/madmom_b200.h"
#include "frontend_params.h"
#include "work_list.h"
#include "fft_core.cuh"          // tw1_index: layout of the inter-pass twiddle table

using mmb::FrontendParams;
using mmb::Track;

static_assert(sizeof(mmb_track) == sizeof(Track), "mmb_track layout");

static thread_local std::string g_err;

static int fail(int code, const std::string& msg) { g_err = msg; return code; }
// for the entry points that live in other translation units (postproc.cu): same thread-local message
namespace mmb { int set_error(int code, const char* msg) { return fail(code, msg); } }
static int cuda_fail(cudaError_t e, const char* what) {
  return fail(MMB_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return cuda_fail(e_, #x); } while (0)

struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

// The entry points work on the plan's device and leave the caller's current device as they found it.
struct DeviceGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int device) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != device) err = cudaSetDevice(device); else if (err == cudaSuccess) prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
#define ON_DEVICE(d) DeviceGuard guard_(d); if (guard_.err != cudaSuccess) return cuda_fail(guard_.err, "cudaSetDevice")

namespace {
// One stage of the host pipeline of mmb_run_host: its own stream and device buffers, reused from call to call.
struct Slot {
  cudaStream_t stream = nullptr;
  void* d_pcm = nullptr; size_t pcm_cap = 0;
  void* d_raw = nullptr; size_t raw_cap = 0;     // interleaved multi-channel PCM before the down-mix
  void* d_out[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr}; size_t out_cap[7] = {0, 0, 0, 0, 0, 0, 0};
  Track* d_tracks = nullptr; size_t tracks_cap = 0;
  int32_t* d_prefix = nullptr; size_t prefix_cap = 0;
  mmb::Chunk* d_work = nullptr; size_t work_cap = 0;
};

static cudaError_t ensure(void** p, size_t* cap, size_t need) {
  if (need <= *cap) return cudaSuccess;
  if (*p) cudaFree(*p);
  *p = nullptr; *cap = 0;
  cudaError_t e = cudaMalloc(p, need);
  if (e == cudaSuccess) *cap = need;
  return e;
}

struct HostPipe {
  std::mutex mu;
  Slot slots[3];
};

void free_slots(Slot* slots, int n) {
  for (int s = 0; s < n; ++s) {
    if (slots[s].stream) { cudaStreamSynchronize(slots[s].stream); cudaStreamDestroy(slots[s].stream); }
    cudaFree(slots[s].d_pcm); cudaFree(slots[s].d_raw); cudaFree(slots[s].d_tracks); cudaFree(slots[s].d_prefix); cudaFree(slots[s].d_work);
    for (int k = 0; k < 7; ++k) cudaFree(slots[s].d_out[k]);
    slots[s] = Slot();
  }
}
}  // namespace

struct mmb_plan {
  HostPipe* pipe = nullptr;      // buffers of mmb_run_host, kept between calls
  mmb_config cfg;
  int num_sms = 0;
  int nnz = 0;
  int span_cap = 0;
  int sample_bytes = 2;
  int smem_bytes = 0;
  int warps = 0;
  int group_warps = 0;
  mmb::SmemLayout L;
  int ring_rows = 0;
  int chunk_frames = 0;          // output frames per chunk of a launch with plenty of work (upper bound)
  bool chunk_fixed = false;      // set by the caller / MMB_CHUNK_ROUNDS: never adapted to the size of a launch
  int n_passes = 0, max_q = 0, b_pad = 0, w_floats = 0, fb_iters = 0;
  float* d_win = nullptr;        // fft_window * 0.5 as float32, or the constants of a cosine-sum window (win_cos)
  bool win_cos = false;
  double* d_win64 = nullptr;     // fft_window in float64 and W_N^k, k < N/2: the float64 STFT kernel (stft64.cu)
  double2* d_tw64 = nullptr;
  float2 *d_tw1 = nullptr, *d_tw2k = nullptr, *d_tw4k = nullptr;
  mmb::FbPass* d_pass = nullptr;
  float* d_fbw = nullptr;
  int32_t *d_fbstart = nullptr, *d_fbpid = nullptr;
};

struct mmb_layout {
  Track* d_tracks = nullptr;
  int32_t* d_prefix = nullptr;    // float64 STFT kernel: prefix sum of the units' chunks of chunk_frames frames
  int32_t n = 0;
  int32_t total_chunks = 0;
  int32_t chunk_frames = 0;       // (pick_chunk_frames)
  mmb::Chunk* d_work = nullptr;   // fused kernel: its work list (build_work) and the CTAs it is laid out for
  int32_t n_work = 0, work_grid = 0;
  int64_t total_rows = 0;
  int64_t pcm_extent = 0;     // max over units of pcm_offset + pcm_len
};

extern "C" int mmb_version(void) { return MMB_VERSION; }
extern "C" const char* mmb_last_error(void) { return g_err.c_str(); }

extern "C" int64_t mmb_num_frames(int64_t num_samples, double hop_size, int32_t end_extend) {
  // np.ceil(len / hop) ('normal') or np.floor(len / hop + 1) ('extend'), float64 (signal.py:1145-1156)
  const double q = (double)num_samples / hop_size;
  return end_extend ? (int64_t)std::floor(q + 1.0) : (int64_t)std::ceil(q);
}

extern "C" int64_t mmb_frame_start(int64_t index, int32_t frame_size, double hop_size, int32_t origin) {
  return (int64_t)((double)index * hop_size) - (int64_t)(frame_size / 2) - (int64_t)origin;
}

template <typename T>
static cudaError_t upload(T** dst, const std::vector<T>& src) {
  *dst = nullptr;
  if (src.empty()) return cudaSuccess;
  cudaError_t e = cudaMalloc((void**)dst, src.size() * sizeof(T));
  if (e != cudaSuccess) return e;
  return cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice);
}

extern "C" int mmb_plan_destroy(mmb_plan* p) {
  if (!p) return MMB_OK;
  cudaFree(p->d_win); cudaFree(p->d_win64); cudaFree(p->d_tw64); cudaFree(p->d_tw1); cudaFree(p->d_tw2k); cudaFree(p->d_tw4k);
  cudaFree(p->d_pass); cudaFree(p->d_fbw); cudaFree(p->d_fbstart); cudaFree(p->d_fbpid);
  if (p->pipe) { free_slots(p->pipe->slots, 3); delete p->pipe; }
  delete p;
  return MMB_OK;
}

// Per-warp filterbank schedule (frontend_params.h: FbPass).  Every band is cut into pieces of at most P bins.
// A pass gives each of the 32 lanes at most one piece, chosen so that the lanes' first bins are distinct modulo
// 32: lane l reads mag[start_l + i] in iteration i, so a pass is free of shared-memory bank conflicts by
// construction.  A piece may be shifted down by `lead` bins (zero weights in front) to reach a free bank.  The
// longest remaining piece sets the pass length; weights are stored [iteration][lane], zero padded.
// The partial sum of piece q of band b is stored at part[q * b_pad + b]; the epilogue adds q = 0 .. max_q-1.
namespace {
struct FbSchedule {
  std::vector<mmb::FbPass> pass;
  std::vector<float> w;
  std::vector<int32_t> start, pid;
  int max_q = 0, b_pad = 0, iters = 0;
};
struct Piece { int start, len, band, q; const float* w; };

// weight of iteration i of lane `lane`: four consecutive iterations of a lane sit in one float4
inline size_t widx(int off, int i, int lane) { return (size_t)off + ((size_t)(i >> 2) * 32 + lane) * 4 + (i & 3); }

void schedule_with(const mmb_config* cfg, int K, const std::vector<int32_t>& woff, int P, FbSchedule& s) {
  const int B = cfg->num_bands;
  // N = 4096 keeps even and odd bins in two arrays (frontend_params.h): a lane that starts at bin st reads array
  // st & 1 at st / 2 + i and the other one beside it.  Even starts that are distinct modulo 64 put the lanes of a
  // pass on distinct banks of both arrays (distinct starts modulo 32 do not: every load then took two wavefronts); a
  // piece that begins at an odd bin gets one zero tap in front.
  const bool split = cfg->frame_size == 4096 && !std::getenv("MMB_FB_SPLIT_ANY");
  s = FbSchedule();
  s.b_pad = (B + 31) & ~31;
  std::vector<Piece> rem;
  for (int b = 0; b < B; ++b) {
    const int len = cfg->band_len[b];
    const int n = len > 0 ? (len + P - 1) / P : 0;
    s.max_q = std::max(s.max_q, n);
    int at = 0;
    for (int j = 0; j < n; ++j) {
      const int l = len / n + (j < len % n ? 1 : 0);
      rem.push_back(Piece{cfg->band_start[b] + at, l, b, j, cfg->band_weights + woff[b] + at});
      at += l;
    }
  }
  const int dump = s.max_q * s.b_pad;                  // idle lanes write here
  auto need = [split](const Piece& x) { return x.len + (split ? (x.start & 1) : 0); };      // taps incl. a forced lead
  std::stable_sort(rem.begin(), rem.end(), [&](const Piece& x, const Piece& y) { return need(x) > need(y); });
  while (!rem.empty()) {
    const int T = std::min((need(rem[0]) + 3) & ~3, K & ~3);      // pass lengths are multiples of four
    const int off = (int)s.w.size();
    const int ps = (int)s.pass.size();
    s.pass.push_back(mmb::FbPass{T, off});
    s.w.resize(off + (size_t)T * 32, 0.0f);
    s.start.resize((ps + 1) * 32, 0);
    s.pid.resize((ps + 1) * 32, dump);
    bool lane_used[32] = {false};
    int placed = 0;
    std::vector<Piece> next;
    for (size_t i = 0; i < rem.size(); ++i) {
      const Piece& pc = rem[i];
      bool ok = false;
      if (placed < 32) {
        for (int lead = split ? (pc.start & 1) : 0; lead <= T - pc.len && !ok; lead += split ? 2 : 1) {
          const int st = pc.start - lead;
          if (st < 0 || st + T > K) continue;
          const int lane = split ? ((st >> 1) & 31) : (st & 31);
          if (lane_used[lane]) continue;
          lane_used[lane] = true; ++placed; ok = true;
          s.start[ps * 32 + lane] = st;
          s.pid[ps * 32 + lane] = pc.q * s.b_pad + pc.band;
          for (int k = 0; k < pc.len; ++k) s.w[widx(off, k + lead, lane)] = pc.w[k];
        }
        if (!ok && i == 0) {
          // the pass-defining piece cannot stay inside [0, K) without a shift larger than its slack (only when
          // T > K - start): shift it down as far as needed, bank conflicts or not
          const int lead = std::max(0, pc.start + T - K);
          int lane = (pc.start - lead) & 31;
          while (lane_used[lane]) lane = (lane + 1) & 31;
          lane_used[lane] = true; ++placed; ok = true;
          s.start[ps * 32 + lane] = pc.start - lead;
          s.pid[ps * 32 + lane] = pc.q * s.b_pad + pc.band;
          for (int k = 0; k < pc.len; ++k) s.w[widx(off, k + lead, lane)] = pc.w[k];
        }
      }
      if (!ok) next.push_back(pc);
    }
    rem.swap(next);
  }
  s.iters = 0;
  for (const mmb::FbPass& fp : s.pass) s.iters += fp.iters;
}

void build_schedule(const mmb_config* cfg, int K, const std::vector<int32_t>& woff, FbSchedule& best) {
  int forced = 0;
  if (const char* env = std::getenv("MMB_FB_PIECE")) forced = std::atoi(env);
  bool have = false;
  int best_cost = 0;
  for (int P = 6; P <= 96; ++P) {                       // every piece length: a schedule takes microseconds
    if (forced > 0 && P != forced) continue;
    FbSchedule s;
    schedule_with(cfg, K, woff, P, s);
    if (s.max_q > 16) continue;
    // issue slots per frame: ~4 per iteration (tap + magnitude loads, FMA, loop), ~12 per pass, ~3.5 per piece
    // slot and 32 bands in the epilogue; the piece slots also cost shared memory per warp
    const int cost = 8 * s.iters + 24 * (int)s.pass.size() + 7 * s.max_q * (s.b_pad / 32);
    if (!have || cost < best_cost) { best = s; best_cost = cost; have = true; }
  }
  if (!have) schedule_with(cfg, K, woff, 64, best);
}
}  // namespace

// Is `w` (N float64 values) a symmetric cosine sum  a0 - a1 cos(2 pi n / (N - 1))  of any scale, to 1e-9 of its
// peak?  Then `wc` gets the constants the kernel evaluates 0.5 w from (the 0.5 of the real-FFT split, like the table):
// float4 l (l < 32): A = -0.5 a1 cos(t n), B = 0.5 a1 sin(t n) for n = 2 l and 2 l + 1 as (A, A', B, B');
// float4 32 + l: the same for n + 2048 (N = 4096); float 256: 0.5 a0.
static bool cosine_window_constants(const double* w, int N, std::vector<float>& wc) {
  const double PI = 3.14159265358979323846264338327950288;
  const double t = 2.0 * PI / (N - 1);
  // a0 - a1 from n = 0, a0 - a1 cos(t N/2) from the middle
  const double c_mid = std::cos(t * (N / 2));
  const double a1 = (w[N / 2] - w[0]) / (1.0 - c_mid), a0 = w[0] + a1;
  double peak = 0.0, worst = 0.0;
  for (int n = 0; n < N; ++n) {
    peak = std::max(peak, std::fabs(w[n]));
    worst = std::max(worst, std::fabs(w[n] - (a0 - a1 * std::cos(t * n))));
  }
  if (!(peak > 0.0) || !(worst <= 1e-9 * peak) || !(std::fabs(a1) > 1e-6 * peak)) return false;
  wc.assign(mmb::kWinCosFloats, 0.0f);
  for (int half = 0; half < (N == 4096 ? 2 : 1); ++half)
    for (int l = 0; l < 32; ++l)
      for (int odd = 0; odd < 2; ++odd) {
        const double n = 2 * l + odd + 2048 * half;
        wc[(32 * half + l) * 4 + odd] = (float)(-0.5 * a1 * std::cos(t * n));
        wc[(32 * half + l) * 4 + 2 + odd] = (float)(0.5 * a1 * std::sin(t * n));
      }
  wc[256] = (float)(0.5 * a0);
  return true;
}

extern "C" int mmb_plan_create(const mmb_config* cfg, mmb_plan** out) {
  NvtxRange nvtx_("mmb_plan_create");
  if (!cfg || !out) return fail(MMB_E_INVALID, "null argument");
  *out = nullptr;
  const int N = cfg->frame_size;
  if (N != 1024 && N != 2048 && N != 4096)
    return fail(MMB_E_UNSUPPORTED, "frame_size must be 1024, 2048 or 4096");
  if (cfg->pcm_dtype != MMB_PCM_I16 && cfg->pcm_dtype != MMB_PCM_F32)
    return fail(MMB_E_INVALID, "pcm_dtype must be MMB_PCM_I16 or MMB_PCM_F32");
  if (!(cfg->hop_size > 0.0) || !std::isfinite(cfg->hop_size)) return fail(MMB_E_INVALID, "hop_size must be > 0");
  if (!cfg->fft_window) return fail(MMB_E_INVALID, "fft_window is null");
  if (cfg->num_bands < 0) return fail(MMB_E_INVALID, "num_bands < 0");
  if (cfg->num_bands > 0 && (!cfg->band_start || !cfg->band_len || !cfg->band_weights))
    return fail(MMB_E_INVALID, "filterbank arrays are null");
  if (cfg->diff_frames < 0 || cfg->diff_frames > mmb::kMaxDiffFrames)
    return fail(MMB_E_UNSUPPORTED, "diff_frames must be in [0, 16]");
  if (cfg->log_mode < 0 || cfg->log_mode > MMB_LOG_LOG2) return fail(MMB_E_INVALID, "bad log_mode");
  if (cfg->include_nyquist && cfg->num_bands > 0)
    return fail(MMB_E_UNSUPPORTED, "include_nyquist is available for the stft / spec outputs only (no filterbank)");
  if (!(cfg->hop_size < 1e6)) return fail(MMB_E_UNSUPPORTED, "hop_size must be below 1e6 samples");
  const int K = N / 2;
  std::vector<int32_t> woff(cfg->num_bands);
  int nnz = 0;
  for (int b = 0; b < cfg->num_bands; ++b) {
    if (cfg->band_start[b] < 0 || cfg->band_len[b] < 0 || cfg->band_start[b] + cfg->band_len[b] > K)
      return fail(MMB_E_INVALID, "filterbank band outside [0, frame_size/2)");
    woff[b] = nnz;
    nnz += cfg->band_len[b];
  }
  mmb_plan* p = new (std::nothrow) mmb_plan();
  if (!p) return fail(MMB_E_NOMEM, "out of host memory");
  p->pipe = new (std::nothrow) HostPipe();
  if (!p->pipe) { mmb_plan_destroy(p); return fail(MMB_E_NOMEM, "out of host memory"); }
  p->cfg = *cfg;
  p->cfg.fft_window = nullptr; p->cfg.band_start = nullptr; p->cfg.band_len = nullptr; p->cfg.band_weights = nullptr;
  p->nnz = nnz;
  p->sample_bytes = cfg->pcm_dtype == MMB_PCM_I16 ? 2 : 4;
  DeviceGuard guard_(cfg->device);
  cudaError_t e = guard_.err;
  if (e != cudaSuccess) { mmb_plan_destroy(p); return cuda_fail(e, "cudaSetDevice"); }
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, cfg->device);
  if (e != cudaSuccess) { mmb_plan_destroy(p); return cuda_fail(e, "cudaGetDeviceProperties"); }
  p->num_sms = prop.multiProcessorCount;

  // ---- tables (float64 -> float32)
  const double PI = 3.14159265358979323846264338327950288;
  std::vector<float> win(N);
  for (int i = 0; i < N; ++i) win[i] = (float)(0.5 * cfg->fft_window[i]);
  // a symmetric cosine-sum window (np.hanning, np.hamming, any scale) is evaluated in the kernel from a few
  // constants instead of loaded (frame_fft.cuh: load_frame_2048_wcos); MMB_WIN_TABLE=1 keeps the table
  p->win_cos = (N == 2048 || N == 4096) && !std::getenv("MMB_WIN_TABLE") && cosine_window_constants(cfg->fft_window, N, win);
  std::vector<float2> tw1(1024), tw2k(1024), tw4k(512);
  for (int k1 = 0; k1 < 32; ++k1)
    for (int l = 0; l < 32; ++l) {
      const double a = -2.0 * PI * (double)((l * k1) % 1024) / 1024.0;
      tw1[mmb::tw1_index(k1, l)] = make_float2((float)std::cos(a), (float)std::sin(a));
    }
  for (int k = 0; k < 1024; ++k) {
    const double a = -2.0 * PI * (double)k / 2048.0;
    tw2k[k] = make_float2((float)std::cos(a), (float)std::sin(a));
  }
  for (int j = 0; j < 512; ++j) {
    const double a = -2.0 * PI * (double)(2 * j + 1) / 4096.0;
    tw4k[j] = make_float2((float)std::cos(a), (float)std::sin(a));
  }
  std::vector<double> win64(cfg->fft_window, cfg->fft_window + N);
  std::vector<double2> tw64(N / 2);
  for (int k = 0; k < N / 2; ++k) {
    const double a = -2.0 * PI * (double)k / (double)N;
    tw64[k] = make_double2(std::cos(a), std::sin(a));
  }
  // compact copies for stft64.cu, behind the natural table (same doubles, so the same results): per radix-4 pass
  // with first-stage half-size H the H first-stage twiddles W_2H^m and the H / 2 second-stage twiddles W_H^m, each
  // contiguous (a warp's 32 loads are one 512-byte run instead of 32 lines), then the split's twiddles in the
  // bit-reversed order the split walks them, W_N^brev(2 p')
  {
    const int M = N / 2;
    int log_m = 0;
    while ((1 << log_m) < M) ++log_m;
    for (int H = M / 2; H >= 2; H /= 4) {
      for (int m = 0; m < H; ++m) tw64.push_back(tw64[(size_t)m * (N / (2 * H))]);
      for (int m = 0; m < H / 2; ++m) tw64.push_back(tw64[(size_t)m * (N / H)]);
    }
    for (int pp = 0; pp < M / 2; ++pp) {
      unsigned v = (unsigned)(2 * pp), r = 0;
      for (int b = 0; b < log_m; ++b) { r = (r << 1) | (v & 1u); v >>= 1; }
      tw64.push_back(tw64[r]);
    }
  }
  FbSchedule sch;
  build_schedule(cfg, K, woff, sch);
  p->n_passes = (int)sch.pass.size(); p->max_q = sch.max_q; p->b_pad = sch.b_pad; p->w_floats = (int)sch.w.size();
  p->fb_iters = sch.iters;
  if ((e = upload(&p->d_win, win)) != cudaSuccess || (e = upload(&p->d_tw1, tw1)) != cudaSuccess ||
      (e = upload(&p->d_win64, win64)) != cudaSuccess || (e = upload(&p->d_tw64, tw64)) != cudaSuccess ||
      (e = upload(&p->d_tw2k, tw2k)) != cudaSuccess || (e = upload(&p->d_tw4k, tw4k)) != cudaSuccess ||
      (e = upload(&p->d_pass, sch.pass)) != cudaSuccess || (e = upload(&p->d_fbw, sch.w)) != cudaSuccess ||
      (e = upload(&p->d_fbstart, sch.start)) != cudaSuccess || (e = upload(&p->d_fbpid, sch.pid)) != cudaSuccess) {
    mmb_plan_destroy(p);
    return cuda_fail(e, "plan constant upload");
  }
  // ---- warps per CTA: as many as the shared-memory budget of one SM allows (one persistent CTA per SM)
  const int align_elems = 16 / p->sample_bytes;
  //      split into groups of `gw` warps that walk their own chunks (their phases drift apart, which keeps the
  //      FP32 pipes and the shared-memory pipe busy at the same time)
  p->warps = 0;
  int gw_req = 0;                                           // 0: one group of all warps
  if (const char* env = std::getenv("MMB_GROUP_WARPS")) gw_req = std::max(std::atoi(env), 0);
  int wmax = mmb::kMaxWarps;
  if (const char* env = std::getenv("MMB_WARPS")) wmax = std::min(std::max(std::atoi(env), 1), mmb::kMaxWarps);
  for (int w = wmax; w >= 1 && p->warps == 0; --w) {
    const int gw = (gw_req > 0 && gw_req <= w) ? gw_req : w;
    if (w % gw != 0 || (w / gw > 15 && gw > 1)) continue;      // named barriers 1 .. 15, one per group
    // N = 4096: two warps per frame, and their pair barriers take the ids behind the groups' (2 .. 15)
    if (N == 4096 && (gw % 2 != 0 || w / gw + w / 2 > 15)) continue;
    const int fpr = mmb::frames_per_round(N, gw);
    const int span = (((int)std::ceil((fpr - 1) * cfg->hop_size) + 2 + N + align_elems) + 7) & ~7;
    const int rows = 2 * fpr + cfg->diff_frames;
    for (int alias = 0; alias <= 1 && p->warps == 0; ++alias) {      // private partial sums first (no clearing)
      const mmb::SmemLayout L = mmb::smem_layout(N, w, gw, cfg->num_bands, p->n_passes, p->max_q * p->b_pad,
                                                 p->w_floats, rows, span, p->sample_bytes, alias != 0);
      if (L.total <= (int)prop.sharedMemPerBlockOptin) {
        p->warps = w; p->group_warps = gw; p->span_cap = span; p->ring_rows = rows; p->smem_bytes = L.total;
        p->L = L;
      }
    }
  }
  if (p->warps == 0) {
    mmb_plan_destroy(p);
    return fail(MMB_E_UNSUPPORTED, "configuration needs more shared memory than one SM has (hop_size or filterbank too large)");
  }
  e = mmb::frontend_set_smem(N, p->sample_bytes, (int)prop.sharedMemPerBlockOptin);
  if (e != cudaSuccess) { mmb_plan_destroy(p); return cuda_fail(e, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)"); }
  const int fpr = mmb::frames_per_round(N, p->group_warps);
  int rounds = 32;
  if (const char* env = std::getenv("MMB_CHUNK_ROUNDS")) { rounds = std::max(std::atoi(env), 1); p->chunk_fixed = true; }
  if (cfg->chunk_frames > 0) p->chunk_fixed = true;
  p->chunk_frames = cfg->chunk_frames > 0 ? cfg->chunk_frames : std::max(rounds * fpr - cfg->diff_frames, 1);
  // N = 1024: even chunks, so that chunks of a unit with an even first frame start at even frames (the kernel pairs
  // frames 2m, 2m + 1 and a chunk that starts at an odd frame computes one more lead-in frame)
  if (N == 1024) p->chunk_frames = std::max(cfg->chunk_frames > 0 ? (p->chunk_frames + 1) & ~1 : p->chunk_frames & ~1, 2);
  *out = p;
  return MMB_OK;
}

extern "C" int mmb_plan_get_info(const mmb_plan* p, mmb_plan_info* info) {
  if (!p || !info) return fail(MMB_E_INVALID, "null argument");
  info->frame_size = p->cfg.frame_size; info->num_bins = p->cfg.frame_size / 2 + (p->cfg.include_nyquist ? 1 : 0);
  info->num_bands = p->cfg.num_bands;
  info->chunk_frames = p->chunk_frames; info->smem_bytes = p->smem_bytes; info->ctas_per_sm = 1;
  info->num_sms = p->num_sms; info->frames_per_round = mmb::frames_per_round(p->cfg.frame_size, p->group_warps);
  info->warps = p->warps;
  return MMB_OK;
}

// ------------------------------------------------------------------------------------------ layouts
static int validate_units(const mmb_track* u, int32_t n) {
  if (!u || n <= 0) return fail(MMB_E_INVALID, "no units");
  for (int32_t i = 0; i < n; ++i) {
    if (u[i].pcm_offset < 0 || u[i].pcm_len < 0 || u[i].frame_first < 0 || u[i].frame_count < 0 || u[i].out_row < 0)
      return fail(MMB_E_INVALID, "negative field in unit table");
  }
  return MMB_OK;
}

// Output frames per chunk of the float64 STFT kernel (one CTA per chunk) for one launch.  With plenty of work it is
// the plan's chunk (32 rounds).  A launch with fewer frames than that spreads over the SMs (one short file, the
// drop-in use) gets the fewest rounds per chunk that still give every SM at most one chunk, so the latency is that
// of a few frames instead of that of a whole chunk on one or two SMs.  Results do not depend on it.  (The fused
// kernel has its own work list, build_work.)
static int pick_chunk_frames(const mmb_plan* p, const mmb_track* u, int32_t n) {
  if (p->chunk_fixed) return p->chunk_frames;
  int64_t total = 0;
  for (int32_t i = 0; i < n; ++i) total += u[i].frame_count;
  const int fpr = mmb::frames_per_round(p->cfg.frame_size, p->group_warps);
  const int64_t slots = (int64_t)p->num_sms * (p->warps / p->group_warps);
  const int64_t per_slot = (total + slots - 1) / slots;
  const int64_t rounds = std::max<int64_t>((per_slot + p->cfg.diff_frames + fpr - 1) / fpr, 1);
  int64_t chunk = std::max<int64_t>(rounds * fpr - p->cfg.diff_frames, 1);
  if (p->cfg.frame_size == 1024) chunk = std::max<int64_t>(chunk & ~(int64_t)1, 2);      // even, see mmb_plan_create
  return (int)std::min<int64_t>(chunk, p->chunk_frames);
}

static void build_prefix(const mmb_plan* p, const mmb_track* u, int32_t n, int chunk_frames,
                         std::vector<int32_t>& prefix, int64_t* rows) {
  (void)p;
  prefix.assign(n + 1, 0);
  int64_t r = 0;
  for (int32_t i = 0; i < n; ++i) {
    const int64_t c = (u[i].frame_count + chunk_frames - 1) / chunk_frames;
    prefix[i + 1] = prefix[i] + (int32_t)c;
    r += u[i].frame_count;
  }
  *rows = r;
}

// Work list of the fused kernel for one launch (work_list.h).  A plan with a fixed chunk size
// (mmb_plan_config::chunk_frames, MMB_CHUNK_ROUNDS) gets equal chunks in unit order, else every group equal rounds.
static void build_work(const mmb_plan* p, const mmb_track* u, int32_t n, std::vector<mmb::Chunk>& out, int* grid) {
  static_assert(sizeof(mmb_track) == sizeof(Track), "mmb_track and Track are the same record");
  int64_t wave_rounds = 64;
  if (const char* env = std::getenv("MMB_WAVE_ROUNDS")) wave_rounds = std::max(std::atoi(env), 0);      // experiments
  mmb::build_work_list(p->cfg.frame_size, mmb::frames_per_round(p->cfg.frame_size, p->group_warps),
                       p->warps / p->group_warps, p->num_sms, p->cfg.diff_frames, p->chunk_fixed ? p->chunk_frames : 0,
                       wave_rounds,
                       reinterpret_cast<const Track*>(u), n, out, grid);
}

extern "C" int mmb_layout_destroy(mmb_layout* l) {
  if (!l) return MMB_OK;
  cudaFree(l->d_tracks); cudaFree(l->d_prefix); cudaFree(l->d_work);
  delete l;
  return MMB_OK;
}

extern "C" int mmb_layout_create(const mmb_plan* p, const mmb_track* units, int32_t n, mmb_layout** out) {
  if (!p || !out) return fail(MMB_E_INVALID, "null argument");
  *out = nullptr;
  int rc = validate_units(units, n);
  if (rc) return rc;
  ON_DEVICE(p->cfg.device);
  mmb_layout* l = new (std::nothrow) mmb_layout();
  if (!l) return fail(MMB_E_NOMEM, "out of host memory");
  std::vector<int32_t> prefix;
  l->chunk_frames = pick_chunk_frames(p, units, n);
  build_prefix(p, units, n, l->chunk_frames, prefix, &l->total_rows);
  l->n = n;
  l->total_chunks = prefix[n];
  for (int32_t i = 0; i < n; ++i) l->pcm_extent = std::max(l->pcm_extent, units[i].pcm_offset + units[i].pcm_len);
  std::vector<Track> t(n);
  std::memcpy(t.data(), units, sizeof(Track) * n);
  cudaError_t e;
  if ((e = upload(&l->d_tracks, t)) != cudaSuccess || (e = upload(&l->d_prefix, prefix)) != cudaSuccess) {
    mmb_layout_destroy(l);
    return cuda_fail(e, "layout upload");
  }
  std::vector<mmb::Chunk> work;
  int grid = 0;
  build_work(p, units, n, work, &grid);
  l->n_work = (int32_t)work.size(); l->work_grid = grid;
  if (!work.empty() && (e = upload(&l->d_work, work)) != cudaSuccess) {
    mmb_layout_destroy(l);
    return cuda_fail(e, "layout upload");
  }
  *out = l;
  return MMB_OK;
}

// ------------------------------------------------------------------------------------------ run
static int check_outputs(const mmb_plan* p, const mmb_outputs* o) {
  if (!o) return fail(MMB_E_INVALID, "outputs is null");
  const bool band_out = o->filt || o->log || o->diff || o->odf;
  if (!o->stft && !o->spec && !band_out && !o->lgd) return fail(MMB_E_INVALID, "no output requested");
  if (o->lgd && p->cfg.frame_size != 2048)
    return fail(MMB_E_UNSUPPORTED, "the local group delay output is implemented for frame_size 2048 only");
  if (band_out && p->cfg.num_bands == 0)
    return fail(MMB_E_UNSUPPORTED, "filt/log/diff/odf outputs need a filterbank in the plan");
  if ((o->diff || o->odf) && p->cfg.diff_frames < 1)
    return fail(MMB_E_INVALID, "diff/odf outputs need diff_frames >= 1 in the plan");
  return MMB_OK;
}

// Launches of one call: the raw-spectrum outputs (stft, spec) come from the float64 kernel (stft64.cu), everything
// behind the filterbank and the lgd rows from the fused FP32 kernel.  *n_launched counts the kernels started.
// The units of one launch on the device, cut for the two kernels.
struct DeviceWork {
  const Track* tracks; int32_t n;
  const int32_t* prefix; int32_t chunks, chunk_frames;      // float64 STFT kernel: one CTA per chunk
  const mmb::Chunk* work; int32_t n_work, grid;             // fused kernel (build_work)
};

static int launch(const mmb_plan* p, const DeviceWork& w, const void* pcm_dev, int64_t pcm_extent,
                  const mmb_outputs* o, cudaStream_t stream,
                  int* n_launched, const mmb_outputs* extra = nullptr, int n_extra = 0, bool multicast = false) {
  const Track* d_tracks = w.tracks;
  const int32_t n = w.n, chunks = w.chunks, chunk_frames = w.chunk_frames;
  const int32_t* d_prefix = w.prefix;
  if (chunks < 1) return MMB_OK;
  if (o->stft || o->spec) {
    mmb::Stft64Params sp;
    std::memset(&sp, 0, sizeof(sp));
    sp.pcm = pcm_dev; sp.tracks = d_tracks; sp.chunk_prefix = d_prefix; sp.n_tracks = n; sp.chunk_frames = chunk_frames;
    sp.hop = p->cfg.hop_size; sp.origin = p->cfg.origin; sp.circular_shift = p->cfg.circular_shift;
    sp.include_nyquist = p->cfg.include_nyquist;
    sp.spec_log_mode = p->cfg.num_bands == 0 ? p->cfg.log_mode : 0;
    sp.mul = p->cfg.mul; sp.add = p->cfg.add;
    sp.win = p->d_win64; sp.tw = p->d_tw64;
    sp.out_stft = (float2*)o->stft; sp.ld_stft = o->ld_stft; sp.out_spec = o->spec; sp.ld_spec = o->ld_spec;
    cudaError_t e = mmb::launch_stft64(p->cfg.frame_size, p->sample_bytes, sp, chunks, stream);
    if (e != cudaSuccess) return cuda_fail(e, "float64 STFT kernel launch");
    if (n_launched) ++*n_launched;
  }
  const bool band_out = o->filt || o->log || o->diff || o->odf;
  if (!band_out && !o->lgd) return MMB_OK;
  if (w.n_work < 1 || w.grid < 1) return MMB_OK;
  FrontendParams fp;
  std::memset(&fp, 0, sizeof(fp));
  fp.pcm = pcm_dev; fp.tracks = d_tracks; fp.chunks = w.work; fp.n_tracks = n; fp.total_chunks = w.n_work;
  fp.hop = p->cfg.hop_size; fp.origin = p->cfg.origin; fp.span_cap = p->span_cap;
  fp.win = p->d_win; fp.win_cos = p->win_cos ? 1 : 0; fp.tw1 = p->d_tw1; fp.tw2k = p->d_tw2k; fp.tw4k = p->d_tw4k;
  fp.n_bands = band_out ? p->cfg.num_bands : 0;     // skip phases B/C when only the lgd rows are wanted
  fp.n_passes = p->n_passes; fp.max_q = p->max_q; fp.b_pad = p->b_pad; fp.w_floats = p->w_floats;
  fp.fb_pass = p->d_pass; fp.fb_w = p->d_fbw; fp.fb_start = p->d_fbstart; fp.fb_pid = p->d_fbpid;
  fp.pcm_extent = pcm_extent; fp.bulk_allowed = ((uintptr_t)pcm_dev % 16) == 0;
  fp.warps = p->warps; fp.group_warps = p->group_warps; fp.L = p->L; fp.ring_rows = p->ring_rows;
  fp.mul = p->cfg.mul; fp.add = p->cfg.add; fp.log_mode = p->cfg.log_mode;
  fp.diff_frames = p->cfg.diff_frames; fp.max_bins = p->cfg.diff_max_bins; fp.positive = p->cfg.positive_diffs;
  fp.circular_shift = p->cfg.circular_shift;
  fp.out_filt = o->filt; fp.ld_filt = o->ld_filt; fp.out_log = o->log; fp.ld_log = o->ld_log;
  fp.out_diff = o->diff; fp.ld_diff = o->ld_diff; fp.out_odf = o->odf;
  fp.out_lgd = o->lgd; fp.ld_lgd = o->ld_lgd;
  fp.n_extra = n_extra; fp.multicast = multicast ? 1 : 0;
  for (int t = 0; t < n_extra; ++t) { fp.x_log[t] = extra[t].log; fp.x_diff[t] = extra[t].diff; fp.x_odf[t] = extra[t].odf; }
  if (n_extra > 0 || multicast) {
    // 16-byte gathered stores: ring rows (n_bands floats each) and every destination row must start on 16 bytes
    bool ok = fp.n_bands % 4 == 0 && (!o->log || o->ld_log % 4 == 0) && (!o->diff || o->ld_diff % 4 == 0) &&
              p->L.ring % 16 == 0 && p->L.ring_stride % 16 == 0;
    auto aligned = [](const void* q) { return q == nullptr || ((uintptr_t)q % 16) == 0; };
    ok = ok && aligned(o->log) && aligned(o->diff);
    for (int t = 0; t < n_extra; ++t) ok = ok && aligned(extra[t].log) && aligned(extra[t].diff);
    if (const char* env = std::getenv("MMB_GATHER_VEC")) ok = ok && std::atoi(env) != 0;      // experiments
    fp.gather_vec = ok ? 1 : 0;
  }
  // one persistent CTA per SM, fewer for a launch of a few rounds: the work list is laid out for this grid
  cudaError_t e = mmb::launch_frontend(p->cfg.frame_size, p->sample_bytes, fp, w.grid, p->smem_bytes, stream);
  if (e != cudaSuccess) return cuda_fail(e, "frontend kernel launch");
  if (n_launched) ++*n_launched;
  return MMB_OK;
}

static DeviceWork work_of(const mmb_layout* l) {
  return DeviceWork{l->d_tracks, l->n, l->d_prefix, l->total_chunks, l->chunk_frames, l->d_work, l->n_work, l->work_grid};
}

extern "C" int mmb_run(const mmb_plan* p, const mmb_layout* l, const void* pcm_dev, const mmb_outputs* outs,
                       void* stream, int32_t* launches) {
  NvtxRange nvtx_("mmb_run");
  if (launches) *launches = 0;
  if (!p || !l || !pcm_dev) return fail(MMB_E_INVALID, "null argument");
  int rc = check_outputs(p, outs);
  if (rc) return rc;
  ON_DEVICE(p->cfg.device);
  int nl = 0;
  rc = launch(p, work_of(l), pcm_dev, l->pcm_extent, outs,
              (cudaStream_t)stream, &nl);
  if (launches) *launches = nl;
  return rc;
}

extern "C" int mmb_run_gather(const mmb_plan* p, const mmb_layout* l, const void* pcm_dev, const mmb_outputs* dsts,
                              int32_t n_dst, void* stream, int32_t* launches) {
  NvtxRange nvtx_("mmb_run_gather");
  if (launches) *launches = 0;
  if (!p || !l || !pcm_dev || !dsts) return fail(MMB_E_INVALID, "null argument");
  if (n_dst < 1 || n_dst > mmb::kMaxExtraDst + 1) return fail(MMB_E_INVALID, "n_dst must be in [1, 8]");
  int rc = check_outputs(p, &dsts[0]);
  if (rc) return rc;
  if (dsts[0].stft || dsts[0].spec || dsts[0].filt || dsts[0].lgd)
    return fail(MMB_E_UNSUPPORTED, "mmb_run_gather stores the band outputs log, diff and odf only");
  for (int32_t t = 1; t < n_dst; ++t) {
    const mmb_outputs& a = dsts[0];
    const mmb_outputs& b = dsts[t];
    if (b.stft || b.spec || b.filt || b.lgd || !a.log != !b.log || !a.diff != !b.diff || !a.odf != !b.odf ||
        (a.log && a.ld_log != b.ld_log) || (a.diff && a.ld_diff != b.ld_diff))
      return fail(MMB_E_INVALID, "every destination must hold the same outputs with the same leading dimensions");
  }
  ON_DEVICE(p->cfg.device);
  int nl = 0;
  rc = launch(p, work_of(l), pcm_dev, l->pcm_extent, &dsts[0],
              (cudaStream_t)stream, &nl, dsts + 1, n_dst - 1);
  if (launches) *launches = nl;
  return rc;
}

extern "C" int mmb_run_multicast(const mmb_plan* p, const mmb_layout* l, const void* pcm_dev, const mmb_outputs* mc,
                                 void* stream, int32_t* launches) {
  NvtxRange nvtx_("mmb_run_multicast");
  if (launches) *launches = 0;
  if (!p || !l || !pcm_dev || !mc) return fail(MMB_E_INVALID, "null argument");
  int rc = check_outputs(p, mc);
  if (rc) return rc;
  if (mc->stft || mc->spec || mc->filt || mc->lgd)
    return fail(MMB_E_UNSUPPORTED, "mmb_run_multicast stores the band outputs log, diff and odf only");
  ON_DEVICE(p->cfg.device);
  int nl = 0;
  rc = launch(p, work_of(l), pcm_dev, l->pcm_extent, mc,
              (cudaStream_t)stream, &nl, nullptr, 0, true);
  if (launches) *launches = nl;
  return rc;
}

extern "C" int mmb_host_alloc(void** ptr, int64_t bytes) {
  if (!ptr || bytes < 0) return fail(MMB_E_INVALID, "bad argument");
  CU(cudaHostAlloc(ptr, (size_t)bytes, cudaHostAllocDefault));
  return MMB_OK;
}
extern "C" int mmb_host_free(void* ptr) {
  if (ptr) CU(cudaFreeHost(ptr));
  return MMB_OK;
}

// ------------------------------------------------------------------------------------------ host path
// mmb_run_host: units are cut into groups of <= kGroupBytes of PCM; each group goes through one of
// kSlots staging slots (own stream: H2D copy -> kernel -> D2H copies), so the copy engines and the SMs
// work on different groups at the same time.  Requires units ordered by pcm_offset and out_row.
namespace {
constexpr int kSlots = 3;
constexpr int64_t kGroupBytes = 64ll << 20;

}  // namespace

// A unit whose PCM alone exceeds a pipeline group (one long file) would pass through mmb_run_host as a single
// host-to-device copy, kernel and copy back, one after the other.  It is cut into frame ranges of about half a
// group, each with the PCM its frames touch: the lead-in frames of the lagged difference on the left (one more for
// frame_size 1024, whose frames are paired by their global index) and frame_size / 2 on the right -- what
// FramedSignal.__getitem__(slice) means (madmom/audio/signal.py:1181-1194) and what the multi-GPU shards do
// (madmom_b200/parallel.py).  Frame starts use the global frame index, so the rows are bit-identical to the unsplit run.
static void split_long_units(const mmb_plan* p, const mmb_track* u, int32_t n, int64_t group_bytes,
                             std::vector<mmb_track>& out) {
  const int N = p->cfg.frame_size;
  const double hop = p->cfg.hop_size;
  const int64_t sb = p->sample_bytes;
  const int64_t lead = p->cfg.diff_frames + (N == 1024 ? 1 : 0);
  int64_t nf = (int64_t)((double)(group_bytes / 2 / sb) / hop) & ~(int64_t)1;
  if (nf < 64) nf = 64;
  out.clear();
  for (int32_t i = 0; i < n; ++i) {
    const mmb_track& t = u[i];
    if (t.pcm_len * sb <= group_bytes || t.frame_count <= nf) { out.push_back(t); continue; }
    for (int64_t f = 0; f < t.frame_count; f += nf) {
      const int64_t f0 = t.frame_first + f, f1 = std::min(f0 + nf, t.frame_first + t.frame_count);
      const int64_t gl = std::max<int64_t>(f0 - lead, 0);
      const int64_t base = (int64_t)(N / 2) + (int64_t)p->cfg.origin + t.sample_base;
      int64_t lo = (int64_t)((double)gl * hop) - base;
      int64_t hi = (int64_t)((double)(f1 - 1) * hop) - base + N;
      lo = std::min(std::max<int64_t>(lo, 0), t.pcm_len);
      hi = std::min(std::max<int64_t>(hi, lo), t.pcm_len);
      mmb_track s;
      s.pcm_offset = t.pcm_offset + lo; s.pcm_len = hi - lo; s.sample_base = t.sample_base + lo;
      s.frame_first = f0; s.frame_count = f1 - f0; s.out_row = t.out_row + (f0 - t.frame_first);
      out.push_back(s);
    }
  }
}

static int run_host_impl(const mmb_plan* p, const mmb_track* units, int32_t n, const void* pcm_host,
                         int32_t channels, int32_t channel, const mmb_outputs* oh, int32_t* launches) {
  NvtxRange nvtx_("mmb_run_host");
  if (launches) *launches = 0;
  if (!p || !pcm_host) return fail(MMB_E_INVALID, "null argument");
  if (channels < 1 || channel < -1 || channel >= channels) return fail(MMB_E_INVALID, "bad channel count / index");
  int rc = validate_units(units, n);
  if (rc) return rc;
  rc = check_outputs(p, oh);
  if (rc) return rc;
  ON_DEVICE(p->cfg.device);
  for (int32_t i = 1; i < n; ++i)
    if (units[i].pcm_offset < units[i - 1].pcm_offset || units[i].out_row != units[i - 1].out_row + units[i - 1].frame_count)
      return fail(MMB_E_INVALID, "mmb_run_host needs units ordered by pcm_offset with consecutive out_row ranges "
                                 "(out_row[i] == out_row[i-1] + frame_count[i-1])");
  const int K = p->cfg.frame_size / 2 + (p->cfg.include_nyquist ? 1 : 0), B = p->cfg.num_bands;
  const size_t sb = (size_t)p->sample_bytes;
  // widths (elements) and element sizes of the six outputs: stft spec filt log diff odf
  void* hptr[7] = {oh->stft, oh->spec, oh->filt, oh->log, oh->diff, oh->odf, oh->lgd};
  const int64_t hld[7] = {oh->ld_stft, oh->ld_spec, oh->ld_filt, oh->ld_log, oh->ld_diff, 1, oh->ld_lgd};
  const int width[7] = {K, K, B, B, B, 1, K};
  const size_t esz[7] = {8, 4, 4, 4, 4, 4, 4};
  for (int k = 0; k < 7; ++k)
    if (hptr[k] && hld[k] < width[k]) return fail(MMB_E_INVALID, "leading dimension smaller than the row width");

  // the plan's own pipeline buffers when no other thread is using them, else temporary ones
  int64_t group_bytes = kGroupBytes;
  if (const char* env = std::getenv("MMB_HOST_GROUP_MB")) group_bytes = std::max(1, std::atoi(env)) * (1ll << 20);
  std::vector<mmb_track> split;
  split_long_units(p, units, n, group_bytes, split);
  units = split.data(); n = (int32_t)split.size();
  Slot local[kSlots];
  std::unique_lock<std::mutex> lock(p->pipe->mu, std::try_to_lock);
  Slot* slots = lock.owns_lock() ? p->pipe->slots : local;
  int used_slots = 0, n_launch = 0;
  int status = MMB_OK;
  auto cleanup = [&]() {
    if (slots == local) free_slots(local, kSlots);
    else
      for (int s = 0; s < kSlots; ++s)
        if (slots[s].stream) cudaStreamSynchronize(slots[s].stream);
  };
#define CUS(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { status = cuda_fail(e_, #x); goto done; } } while (0)
  {
    std::vector<Track> tab;
    std::vector<int32_t> prefix;
    std::vector<mmb::Chunk> work;
    int32_t i0 = 0;
    int g = 0;
    while (i0 < n) {
      // ---- group [i0, i1)
      int32_t i1 = i0;
      const int64_t pcm0 = units[i0].pcm_offset;
      int64_t pcm1 = pcm0;
      while (i1 < n) {
        const int64_t e1 = units[i1].pcm_offset + units[i1].pcm_len;
        const int64_t np1 = e1 > pcm1 ? e1 : pcm1;
        if (i1 > i0 && (np1 - pcm0) * (int64_t)sb > group_bytes) break;
        pcm1 = np1; ++i1;
      }
      const int64_t row0 = units[i0].out_row;
      const int64_t row1 = units[i1 - 1].out_row + units[i1 - 1].frame_count;
      const int64_t rows = row1 - row0;
      Slot& S = slots[g % kSlots];
      if (!S.stream) { CUS(cudaStreamCreateWithFlags(&S.stream, cudaStreamNonBlocking)); ++used_slots; }
      // the slot's previous group must be fully drained before its buffers are reused
      CUS(cudaStreamSynchronize(S.stream));
      tab.assign(i1 - i0, Track());
      for (int32_t i = i0; i < i1; ++i) {
        std::memcpy(&tab[i - i0], &units[i], sizeof(Track));
        tab[i - i0].pcm_offset -= pcm0;
        tab[i - i0].out_row -= row0;
      }
      int64_t dummy;
      const int chunk_frames = pick_chunk_frames(p, reinterpret_cast<const mmb_track*>(tab.data()), i1 - i0);
      build_prefix(p, reinterpret_cast<const mmb_track*>(tab.data()), i1 - i0, chunk_frames, prefix, &dummy);
      int work_grid = 0;
      build_work(p, reinterpret_cast<const mmb_track*>(tab.data()), i1 - i0, work, &work_grid);
      size_t pcm_bytes = (size_t)(pcm1 - pcm0) * sb;
      if (pcm_bytes == 0) pcm_bytes = sb;
      CUS(ensure(&S.d_pcm, &S.pcm_cap, pcm_bytes));
      CUS(ensure((void**)&S.d_tracks, &S.tracks_cap, tab.size() * sizeof(Track)));
      CUS(ensure((void**)&S.d_prefix, &S.prefix_cap, prefix.size() * sizeof(int32_t)));
      CUS(ensure((void**)&S.d_work, &S.work_cap, std::max<size_t>(work.size(), 1) * sizeof(mmb::Chunk)));
      mmb_outputs od;
      std::memset(&od, 0, sizeof(od));
      for (int k = 0; k < 7; ++k) {
        if (!hptr[k]) continue;
        CUS(ensure(&S.d_out[k], &S.out_cap[k], (size_t)(rows > 0 ? rows : 1) * width[k] * esz[k]));
      }
      od.stft = hptr[0] ? S.d_out[0] : nullptr; od.ld_stft = K;
      od.spec = hptr[1] ? (float*)S.d_out[1] : nullptr; od.ld_spec = K;
      od.filt = hptr[2] ? (float*)S.d_out[2] : nullptr; od.ld_filt = B;
      od.log = hptr[3] ? (float*)S.d_out[3] : nullptr; od.ld_log = B;
      od.diff = hptr[4] ? (float*)S.d_out[4] : nullptr; od.ld_diff = B;
      od.odf = hptr[5] ? (float*)S.d_out[5] : nullptr;
      od.lgd = hptr[6] ? (float*)S.d_out[6] : nullptr; od.ld_lgd = K;
      CUS(cudaMemcpyAsync(S.d_tracks, tab.data(), tab.size() * sizeof(Track), cudaMemcpyHostToDevice, S.stream));
      CUS(cudaMemcpyAsync(S.d_prefix, prefix.data(), prefix.size() * sizeof(int32_t), cudaMemcpyHostToDevice, S.stream));
      if (!work.empty())
        CUS(cudaMemcpyAsync(S.d_work, work.data(), work.size() * sizeof(mmb::Chunk), cudaMemcpyHostToDevice, S.stream));
      if (channels == 1) {
        CUS(cudaMemcpyAsync(S.d_pcm, (const char*)pcm_host + (size_t)pcm0 * sb, (size_t)(pcm1 - pcm0) * sb,
                            cudaMemcpyHostToDevice, S.stream));
      } else {
        // interleaved frames: copy the group, then mix down on the device (remix, signal.py:170-223)
        const size_t raw_bytes = pcm_bytes * (size_t)channels;
        CUS(ensure(&S.d_raw, &S.raw_cap, raw_bytes));
        CUS(cudaMemcpyAsync(S.d_raw, (const char*)pcm_host + (size_t)pcm0 * sb * channels,
                            (size_t)(pcm1 - pcm0) * sb * channels, cudaMemcpyHostToDevice, S.stream));
        if (pcm1 > pcm0) {
          CUS(mmb::launch_remix_mono(S.d_raw, p->sample_bytes, pcm1 - pcm0, channels, channel, S.d_pcm, S.stream));
          ++n_launch;
        }
      }
      // pageable tab/prefix vectors are reused by the next group: the async copies above are staged
      // synchronously by the runtime for pageable memory, so this is safe.
      if (prefix.back() > 0) {
        const DeviceWork dw{S.d_tracks, (int32_t)tab.size(), S.d_prefix, prefix.back(), chunk_frames,
                            S.d_work, (int32_t)work.size(), work_grid};
        status = launch(p, dw, S.d_pcm, pcm1 - pcm0, &od, S.stream, &n_launch);
        if (status) goto done;
        for (int k = 0; k < 7; ++k) {
          if (!hptr[k]) continue;
          char* dst = (char*)hptr[k] + (size_t)row0 * hld[k] * esz[k];
          CUS(cudaMemcpy2DAsync(dst, (size_t)hld[k] * esz[k], S.d_out[k], (size_t)width[k] * esz[k],
                                (size_t)width[k] * esz[k], (size_t)rows, cudaMemcpyDeviceToHost, S.stream));
        }
      }
      i0 = i1; ++g;
    }
    for (int s = 0; s < kSlots; ++s)
      if (slots[s].stream) CUS(cudaStreamSynchronize(slots[s].stream));
  }
done:
#undef CUS
  cleanup();
  (void)used_slots;
  if (launches) *launches = n_launch;
  return status;
}

extern "C" int mmb_run_host(const mmb_plan* p, const mmb_track* units, int32_t n, const void* pcm_host,
                            const mmb_outputs* oh, int32_t* launches) {
  return run_host_impl(p, units, n, pcm_host, 1, -1, oh, launches);
}

extern "C" int mmb_run_host_interleaved(const mmb_plan* p, const mmb_track* units, int32_t n, const void* pcm_host,
                                        int32_t num_channels, int32_t channel, const mmb_outputs* oh,
                                        int32_t* launches) {
  return run_host_impl(p, units, n, pcm_host, num_channels, channel, oh, launches);
}

extern "C" int mmb_remix_mono(const void* in_dev, int32_t sample_bytes, int64_t num_frames, int32_t num_channels,
                              int32_t channel, void* out_dev, void* stream) {
  if (!in_dev || !out_dev || (sample_bytes != 2 && sample_bytes != 4) || num_frames < 0 || num_channels < 1 ||
      channel < -1 || channel >= num_channels)
    return fail(MMB_E_INVALID, "mmb_remix_mono: invalid argument");
  if (num_frames == 0) return MMB_OK;
  cudaError_t e = mmb::launch_remix_mono(in_dev, sample_bytes, num_frames, num_channels, channel, out_dev,
                                         (cudaStream_t)stream);
  return e == cudaSuccess ? MMB_OK : cuda_fail(e, "remix kernel launch");
}
