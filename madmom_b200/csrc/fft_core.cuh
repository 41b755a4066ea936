// Warp-level 1024-point complex FFT (32 lanes x 32 points in registers) and the per-frame-size
// pre/post-processing that turns it into the real FFTs madmom needs (frame sizes 1024 / 2048 / 4096).
//
// Replaces, per frame, the reference's   fftpack.fft(frame * fft_window)[:N/2]   (madmom/audio/stft.py:105-131)
// followed by np.abs (madmom/audio/spectrogram.py:122-141).
//
// Everything here is written as *phase functions of one lane*: a phase touches only the lane's own
// registers plus warp-shared scratch, and phases are separated by __syncwarp() in the kernel.  The same
// functions compile for the host, where tests/host_emul.cpp runs the 32 lanes of a warp one after the
// other; that is how the index/twiddle algebra is validated on a machine without a GPU.
//
// Decomposition (Cooley-Tukey, n = 32*n1 + n2, k = k1 + 32*k2, lane = n2 in pass 1, lane = k1 in pass 2):
//   pass 1  A_l[k1]   = sum_n1 z[32 n1 + l] W32^(n1 k1)            32-point FFT in registers
//   twiddle A_l[k1]  *= W1024^(l k1)                               table T1 (tw1_index: k1 and k1 + 16 in one LDS.128)
//   xchg    B_j[n2]   = A_n2[j]                                    32x33 float2 transpose through smem
//   pass 2  Z[j+32k2] = sum_n2 B_j[n2] W32^(n2 k2)                 32-point FFT in registers
// The in-register FFT is a fully unrolled radix-2 DIF (natural in, bit-reversed out); the bit reversal is
// pure compile-time register renaming.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define MMB_HD __host__ __device__ __forceinline__
#else
#define MMB_HD inline
#include <cmath>
struct float2 { float x, y; };
static inline float2 make_float2(float x, float y) { float2 r; r.x = x; r.y = y; return r; }
#endif

namespace mmb {

constexpr int kExStride = 33;                 // row stride (float2) of the 32x32 exchange tile
constexpr int kExElems = 32 * kExStride;      // 1056 float2 = 8448 B per warp

MMB_HD constexpr int brev5(int r) {
  return ((r & 1) << 4) | ((r & 2) << 2) | (r & 4) | ((r & 8) >> 2) | ((r & 16) >> 4);
}

// cos/sin(2 pi j / 32) for j = 0..8 (first quadrant incl. end points), rounded from double.
#define MMB_C32_0 1.0f
#define MMB_C32_1 0.98078528040323044913f
#define MMB_C32_2 0.92387953251128675613f
#define MMB_C32_3 0.83146961230254523708f
#define MMB_C32_4 0.70710678118654752440f
#define MMB_C32_5 0.55557023301960222474f
#define MMB_C32_6 0.38268343236508977173f
#define MMB_C32_7 0.19509032201612826785f

MMB_HD constexpr float cos32(int j) {   // cos(2 pi j / 32), j in [0, 16)
  return j == 0 ? 1.0f : j == 1 ? MMB_C32_1 : j == 2 ? MMB_C32_2 : j == 3 ? MMB_C32_3 : j == 4 ? MMB_C32_4
       : j == 5 ? MMB_C32_5 : j == 6 ? MMB_C32_6 : j == 7 ? MMB_C32_7 : j == 8 ? 0.0f
       : j == 9 ? -MMB_C32_7 : j == 10 ? -MMB_C32_6 : j == 11 ? -MMB_C32_5 : j == 12 ? -MMB_C32_4
       : j == 13 ? -MMB_C32_3 : j == 14 ? -MMB_C32_2 : -MMB_C32_1;
}
MMB_HD constexpr float sin32(int j) {   // sin(2 pi j / 32), j in [0, 16)
  return j <= 8 ? cos32(8 - j) : cos32(j - 8);
}

// Packed FP32 pairs (PTX add/sub/mul.rn.f32x2 -> SASS FADD2 / FMUL2, sm_100): both halves of a complex number in
// one instruction.  Each half is the same IEEE operation as the scalar form, so results are bit-identical; what is
// saved are issue slots (measured: -8 % kernel time from the butterflies alone).  The host build (tests/host_emul.cpp)
// uses the scalar form.
MMB_HD float2 add2(float2 a, float2 b) {
#if defined(__CUDA_ARCH__)
  unsigned long long x, y, r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(y) : "f"(b.x), "f"(b.y));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
  float2 o;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(o.x), "=f"(o.y) : "l"(r));
  return o;
#else
  return make_float2(a.x + b.x, a.y + b.y);
#endif
}
MMB_HD float2 sub2(float2 a, float2 b) {
#if defined(__CUDA_ARCH__)
  unsigned long long x, y, r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(y) : "f"(b.x), "f"(b.y));
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
  float2 o;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(o.x), "=f"(o.y) : "l"(r));
  return o;
#else
  return make_float2(a.x - b.x, a.y - b.y);
#endif
}
MMB_HD float2 mul2(float2 a, float2 b) {
#if defined(__CUDA_ARCH__)
  unsigned long long x, y, r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(y) : "f"(b.x), "f"(b.y));
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
  float2 o;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(o.x), "=f"(o.y) : "l"(r));
  return o;
#else
  return make_float2(a.x * b.x, a.y * b.y);
#endif
}

MMB_HD float2 fma2(float2 a, float2 b, float2 c) {      // a * b + c per half, each a fused multiply-add
#if defined(__CUDA_ARCH__)
  unsigned long long x, y, z, r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(y) : "f"(b.x), "f"(b.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(z) : "f"(c.x), "f"(c.y));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(x), "l"(y), "l"(z));
  float2 o;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(o.x), "=f"(o.y) : "l"(r));
  return o;
#else
  return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y));
#endif
}

// One radix-2 DIF butterfly with twiddle W_32^(tw) = exp(-2 pi i tw / 32) on the difference output:
//   (dr + i di)(c - i s) = (dr c + di s) + i (di c - dr s)
// as one packed multiply by the broadcast immediate c and two scalar FMAs with the immediate s.  (The fully
// packed form, FFMA2 with a (s, -s) register pair, was measured slower: three register-pair operands.)
template <int TW>
MMB_HD void bfly(float2& u, float2& w) {
  const float2 sum = add2(u, w), dif = sub2(u, w);
  u = sum;
  if (TW == 0) {                       // * 1
    w = dif;
  } else if (TW == 8) {                // * (-i)
    w = make_float2(dif.y, -dif.x);
  } else {
    constexpr float c = cos32(TW), s = sin32(TW);
    const float2 t = mul2(dif, make_float2(c, c));
    w = make_float2(fmaf(dif.y, s, t.x), fmaf(-dif.x, s, t.y));
  }
}

template <int HALF, int BASE, int J>
struct DifStageInner {
  static MMB_HD void run(float2 (&v)[32]) {
    bfly<J * (16 / HALF)>(v[BASE + J], v[BASE + J + HALF]);
    DifStageInner<HALF, BASE, J + 1>::run(v);
  }
};
template <int HALF, int BASE>
struct DifStageInner<HALF, BASE, HALF> { static MMB_HD void run(float2 (&)[32]) {} };

template <int HALF, int BASE>
struct DifStageBlocks {
  static MMB_HD void run(float2 (&v)[32]) {
    DifStageInner<HALF, BASE, 0>::run(v);
    DifStageBlocks<HALF, BASE + 2 * HALF>::run(v);
  }
};
template <int HALF>
struct DifStageBlocks<HALF, 32> { static MMB_HD void run(float2 (&)[32]) {} };

// In-place 32-point FFT, natural-order input, output X[brev5(r)] in v[r].
MMB_HD void fft32_dif(float2 (&v)[32]) {
  DifStageBlocks<16, 0>::run(v);
  DifStageBlocks<8, 0>::run(v);
  DifStageBlocks<4, 0>::run(v);
  DifStageBlocks<2, 0>::run(v);
  DifStageBlocks<1, 0>::run(v);
}

MMB_HD float2 cmul(float2 a, float2 w) { return make_float2(a.x * w.x - a.y * w.y, a.x * w.y + a.y * w.x); }

// Layout of the inter-pass twiddle table: W1024^(lane k1) and W1024^(lane (k1 + 16)) sit side by side, so that the
// two twiddles of the register pair (r, r + 1) (k1 = brev5(r) < 16 and k1 + 16) arrive with one 16-byte load.
MMB_HD constexpr int tw1_index(int k1, int lane) { return ((k1 & 15) * 32 + lane) * 2 + (k1 >> 4); }

// ---- the four phases of the 1024-point warp FFT -------------------------------------------------------
// phase 1: in-register FFT over n1, inter-pass twiddle, store into the exchange tile (row k1, column lane)
MMB_HD void core_pass1(int lane, float2 (&v)[32], float2* ex, const float2* __restrict__ tw1) {
  fft32_dif(v);
#pragma unroll
  for (int r = 0; r < 32; r += 2) {
    const int k1 = brev5(r);                        // < 16; register r + 1 holds k1 + 16
#if defined(__CUDA_ARCH__)
    const float4 t = reinterpret_cast<const float4*>(tw1)[k1 * 32 + lane];
    const float2 ta = make_float2(t.x, t.y), tb = make_float2(t.z, t.w);
#else
    const float2 ta = tw1[tw1_index(k1, lane)], tb = tw1[tw1_index(k1 + 16, lane)];
#endif
    float2 a = v[r];
    if (k1 != 0) a = cmul(a, ta);
    ex[k1 * kExStride + lane] = a;
    ex[(k1 + 16) * kExStride + lane] = cmul(v[r + 1], tb);
  }
}
// phase 2 (after __syncwarp): load row `lane` of the tile, FFT over n2.  v[r] = Z[lane + 32*brev5(r)].
MMB_HD void core_pass2_load(int lane, float2 (&v)[32], const float2* ex) {
#pragma unroll
  for (int n2 = 0; n2 < 32; ++n2) v[n2] = ex[lane * kExStride + n2];
}
MMB_HD void core_pass2_fft(float2 (&v)[32]) { fft32_dif(v); }

// phase 3 (after __syncwarp; the tile is dead): publish the upper half Z[512..1023] (and Z[0] in slot
// 1024) at flat index k of the same buffer, i.e. bytes [4096, 8200) -- bytes [0, 4096) stay free for the
// magnitudes of this frame.  `partner_of_zero`: index that Z[0] is published at (1024 for the pairings
// (k, 1024-k); unused (-1) for the (j, 1023-j) pairing of the odd half of a 4096-point frame).
MMB_HD void core_publish_upper(int lane, const float2 (&v)[32], float2* zs, bool publish_zero) {
#pragma unroll
  for (int r = 0; r < 32; ++r) {
    const int k2 = brev5(r);
    if (k2 >= 16) zs[lane + 32 * k2] = v[r];
  }
  if (publish_zero && lane == 0) zs[1024] = v[0];
}

MMB_HD float mag_of(float re, float im) {
#if defined(__CUDA_ARCH__)
  float s = re * re + im * im, r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));
  return r;
#else
  return sqrtf(re * re + im * im);
#endif
}

}  // namespace mmb
