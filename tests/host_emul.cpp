// CPU emulation of ONE warp of the CUDA front-end kernel: the phase functions of
// madmom_b200/csrc/{fft_core,frame_fft}.cuh are compiled for the host and the 32 lanes are run one
// after the other, phase by phase (each __syncwarp() of the kernel is a phase boundary here).
// Used by tests/test_host_emul.py to validate the index / twiddle algebra without a GPU.
//
//   emul_frame(N, x, win_half, mag, re, im, circ)
//     N = 2048 / 4096: x = N samples (float), outputs K = N/2 bins
//     N = 1024       : x = 2*1024 samples (two frames back to back), outputs 2*512 bins
#include <cmath>
#include <cstring>
#include <vector>
#include "../madmom_b200/csrc/frame_fft.cuh"

using namespace mmb;

static void make_tables(std::vector<float2>& tw1, std::vector<float2>& tw2k, std::vector<float2>& tw4k) {
  const double PI = 3.14159265358979323846;
  tw1.resize(1024); tw2k.resize(1024); tw4k.resize(512);
  for (int k1 = 0; k1 < 32; ++k1)
    for (int l = 0; l < 32; ++l) {
      double a = -2.0 * PI * (double)(l * k1) / 1024.0;
      tw1[tw1_index(k1, l)] = make_float2((float)cos(a), (float)sin(a));
    }
  for (int k = 0; k < 1024; ++k) {   // index q*32 + lane == k
    double a = -2.0 * PI * (double)k / 2048.0;
    tw2k[k] = make_float2((float)cos(a), (float)sin(a));
  }
  for (int j = 0; j < 512; ++j) {
    double a = -2.0 * PI * (double)(2 * j + 1) / 4096.0;
    tw4k[j] = make_float2((float)cos(a), (float)sin(a));
  }
}

struct Warp {
  float2 v[32][32];
  float2 ex[kExElems];
};

static void run_core(Warp& w, const std::vector<float2>& tw1, bool publish_zero) {
  for (int l = 0; l < 32; ++l) core_pass1(l, w.v[l], w.ex, tw1.data());
  for (int l = 0; l < 32; ++l) core_pass2_load(l, w.v[l], w.ex);
  for (int l = 0; l < 32; ++l) core_pass2_fft(w.v[l]);
  for (int l = 0; l < 32; ++l) core_publish_upper(l, w.v[l], w.ex, publish_zero);
}

extern "C" int emul_frame(int N, const float* x, const float* win_half, float* mag, float* re, float* im,
                          int circ) {
  std::vector<float2> tw1, tw2k, tw4k;
  make_tables(tw1, tw2k, tw4k);
  static Warp w;
  const int K = N / 2;
  std::vector<float2> st((N == 1024) ? 1024 : K);
  std::vector<float> mg((N == 1024) ? 1024 : K);
  BinSink out; out.mag = mg.data(); out.stft_row = st.data(); out.spec_row = nullptr;
  out.sgn_odd = circ ? -1.0f : 1.0f;
  if (N == 2048) {
    for (int l = 0; l < 32; ++l) load_frame_2048<float>(l, w.v[l], x, win_half);
    run_core(w, tw1, true);
    for (int l = 0; l < 32; ++l) split_2048(l, w.v[l], w.ex, tw2k.data(), out);
  } else if (N == 1024) {
    for (int l = 0; l < 32; ++l) load_frames_1024<float>(l, w.v[l], x, x + 1024, win_half);
    run_core(w, tw1, true);
    BinSink outB = out; outB.mag = mg.data() + 512; outB.stft_row = st.data() + 512;
    for (int l = 0; l < 32; ++l) split_1024x2(l, w.v[l], w.ex, out, outB, true);
  } else if (N == 4096) {
    for (int c = 0; c < 2; ++c) {
      for (int l = 0; l < 32; ++l) load_frame_4096<float>(l, c, w.v[l], x, win_half, tw2k.data());
      run_core(w, tw1, c == 0);
      // the kernel keeps the magnitudes of the even (c = 0) and odd (c = 1) bins in two arrays
      std::vector<float> half(1025, 0.0f);
      BinSink oh = out; oh.mag = half.data();
      for (int l = 0; l < 32; ++l) split_4096(l, c, w.v[l], w.ex, tw2k.data(), tw4k.data(), oh);
      for (int j = 0; 2 * j + c < K; ++j) mg[2 * j + c] = half[j];
    }
  } else {
    return -1;
  }
  for (size_t i = 0; i < mg.size(); ++i) { mag[i] = mg[i]; re[i] = st[i].x; im[i] = st[i].y; }
  return 0;
}

extern "C" void emul_fft32(const float* re_in, const float* im_in, float* re_out, float* im_out) {
  float2 v[32];
  for (int i = 0; i < 32; ++i) v[i] = make_float2(re_in[i], im_in[i]);
  fft32_dif(v);
  for (int r = 0; r < 32; ++r) { re_out[brev5(r)] = v[r].x; im_out[brev5(r)] = v[r].y; }
}

// ---------------------------------------------------------------------------------------------- work list
// The host-side cut of a launch's units into the fused kernel's work list (madmom_b200/csrc/work_list.h), for
// tests/test_host_logic.py.  units: n x 6 int64 (the mmb_track record).  Returns the number of entries written to
// `out` (3 int64 each: unit, frames, first), -1 when `cap` entries do not hold them; *grid = CTAs of the launch.
#include "../madmom_b200/csrc/work_list.h"

extern "C" long long emul_work_list(int N, int fpr, int groups, int num_sms, long long lead, long long fixed_chunk,
                                    long long wave_rounds, const long long* units, int n, long long* out, long long cap, int* grid) {
  static_assert(sizeof(mmb::Track) == 6 * sizeof(long long), "Track is six int64");
  std::vector<mmb::Chunk> work;
  mmb::build_work_list(N, fpr, groups, num_sms, lead, fixed_chunk, wave_rounds, reinterpret_cast<const mmb::Track*>(units), n, work,
                       grid);
  if ((long long)work.size() > cap) return -1;
  for (size_t k = 0; k < work.size(); ++k) {
    out[3 * k] = work[k].unit; out[3 * k + 1] = work[k].frames; out[3 * k + 2] = work[k].first;
  }
  return (long long)work.size();
}
