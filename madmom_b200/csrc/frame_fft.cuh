// Per-frame-size wrappers around the 1024-point warp FFT core (fft_core.cuh):
//   N = 2048 : one frame per core call, even/odd packing z[n] = x[2n] + i x[2n+1], split with W2048^k
//   N = 1024 : two frames per core call, z[n] = xA[n] + i xB[n], twiddle-free split
//   N = 4096 : one frame per two core calls: radix-2 DIF pre-stage on the packed 2048-point sequence
//              (y0 = z[n] + z[n+1024], y1 = (z[n] - z[n+1024]) W2048^n), each half split with W4096^k
// All variants produce X[k] = sum_n x[n] w[n] exp(-2 pi i k n / N), k = 0 .. N/2-1  (Nyquist dropped, as
// madmom/audio/stft.py:93-95 does by default).  The window passed in is fft_window * 0.5 (exact scaling;
// it absorbs the 1/2 of the real-FFT split), see capi.cu.
//
// Sample type S is int16_t or float (madmom keeps integer PCM un-scaled and scales the window instead,
// madmom/audio/stft.py:331-349).
#pragma once
#include "fft_core.cuh"
#if defined(__CUDACC__)
#include "win_cos_tables.h"
#endif

namespace mmb {

template <typename S>
MMB_HD float to_f32(S s) { return (float)s; }
#if defined(__CUDA_ARCH__)
// int16 -> float through a sign-extending load and the 32-bit conversion (I2FP, FMA-pipe rate) instead of the
// 16-bit one (I2F.S16 runs on the quarter-rate XU pipe and would bound the load phase: 64 per lane and frame)
template <>
__device__ __forceinline__ float to_f32<int16_t>(int16_t s) {
  float f;
  asm("cvt.rn.f32.s32 %0, %1;" : "=f"(f) : "r"((int)s));
  return f;
}
#endif

// ------------------------------------------------------------------------------------------ loads
// x points at the first sample of the frame inside the staged PCM span (any alignment).
template <typename S>
MMB_HD void load_frame_2048(int lane, float2 (&v)[32], const S* x, const float* __restrict__ win) {
#pragma unroll
  for (int m = 0; m < 32; ++m) {
    const int n = lane + 32 * m;
    const float2 w = *reinterpret_cast<const float2*>(win + 2 * n);
    v[m] = mul2(make_float2(to_f32(x[2 * n]), to_f32(x[2 * n + 1])), w);
  }
}

// The same for a symmetric cosine-sum window w[n] = a0 - a1 cos(t n), t = 2 pi / (N - 1) (np.hanning, np.hamming;
// capi.cu recognises them): at n = n0 + 64 m, n0 = 2 lane (+ 1),
//   w = a0 - a1 cos(t n0) cos(64 t m) + a1 sin(t n0) sin(64 t m) = G + A C_m + B S_m
// with the lane's A = -a1 cos(t n0), B = a1 sin(t n0) for its even and its odd sample (`wc`: 32 x float4
// (A_even, A_odd, B_even, B_odd), G at float 256; capi.cu: cosine_window_constants) and C_m, S_m from the constant bank (win_cos_tables.h: uniform-register
// operands).  Two packed FMAs per complex element instead of an 8-byte shared-memory load: the window loads are 64 of the
// 663 shared-memory wavefronts of a frame and that pipe is the busiest unit of the kernel.
// Each window value is within 3e-7 of the window's peak of the float32 table value.
#if defined(__CUDACC__)
template <typename S>
__device__ __forceinline__ void load_frame_2048_wcos(int lane, float2 (&v)[32], const S* x,
                                                     const float* __restrict__ wc) {
  const float4 k = reinterpret_cast<const float4*>(wc)[lane];
  const float G = wc[256];
#pragma unroll
  for (int m = 0; m < 32; ++m) {
    const int n = lane + 32 * m;
    const float C = kWinCos2048[m], Sn = kWinSin2048[m];
    // two packed FMAs: (A, A') C_m + ((B, B') S_m + G), the scalars broadcast (FFMA2 R, R.F32x2, UR.F32, R.F32)
    const float2 w = fma2(make_float2(k.x, k.y), make_float2(C, C),
                          fma2(make_float2(k.z, k.w), make_float2(Sn, Sn), make_float2(G, G)));
    v[m] = mul2(make_float2(to_f32(x[2 * n]), to_f32(x[2 * n + 1])), w);
  }
}
#endif

// two frames (xb may be null -> zero imaginary part)
template <typename S>
MMB_HD void load_frames_1024(int lane, float2 (&v)[32], const S* xa, const S* xb, const float* __restrict__ win) {
#pragma unroll
  for (int m = 0; m < 32; ++m) {
    const int n = lane + 32 * m;
    const float w = win[n];
    v[m] = make_float2(to_f32(xa[n]) * w, xb ? to_f32(xb[n]) * w : 0.0f);
  }
}

// half c (0 or 1) of a 4096-point frame; tw2k[q*32 + lane] = W2048^(lane + 32 q)
template <typename S>
MMB_HD void load_frame_4096(int lane, int c, float2 (&v)[32], const S* x, const float* __restrict__ win,
                            const float2* __restrict__ tw2k) {
#pragma unroll
  for (int m = 0; m < 32; ++m) {
    const int n = lane + 32 * m;
    const float2 w0 = *reinterpret_cast<const float2*>(win + 2 * n);
    const float2 w1 = *reinterpret_cast<const float2*>(win + 2 * n + 2048);
    const float2 z0 = mul2(make_float2(to_f32(x[2 * n]), to_f32(x[2 * n + 1])), w0);
    const float2 z1 = mul2(make_float2(to_f32(x[2 * n + 2048]), to_f32(x[2 * n + 2049])), w1);
    if (c == 0) {
      v[m] = add2(z0, z1);
    } else {
      v[m] = cmul(sub2(z0, z1), tw2k[m * 32 + lane]);
    }
  }
}

// cosine-sum window (see load_frame_2048_wcos): `wc` holds 32 x float4 for the samples 2 n, 2 n + 1, 32 x float4 for
// 2 n + 2048, 2 n + 2049, G at float 256
#if defined(__CUDACC__)
template <typename S>
__device__ __forceinline__ void load_frame_4096_wcos(int lane, int c, float2 (&v)[32], const S* x,
                                                     const float* __restrict__ wc,
                                                     const float2* __restrict__ tw2k) {
  const float4 k0 = reinterpret_cast<const float4*>(wc)[lane];
  const float4 k1 = reinterpret_cast<const float4*>(wc)[32 + lane];
  const float G = wc[256];
#pragma unroll
  for (int m = 0; m < 32; ++m) {
    const int n = lane + 32 * m;
    const float C = kWinCos4096[m], Sn = kWinSin4096[m];
    const float2 cc = make_float2(C, C), ss = make_float2(Sn, Sn), gg = make_float2(G, G);
    const float2 w0 = fma2(make_float2(k0.x, k0.y), cc, fma2(make_float2(k0.z, k0.w), ss, gg));
    const float2 w1 = fma2(make_float2(k1.x, k1.y), cc, fma2(make_float2(k1.z, k1.w), ss, gg));
    const float2 z0 = mul2(make_float2(to_f32(x[2 * n]), to_f32(x[2 * n + 1])), w0);
    const float2 z1 = mul2(make_float2(to_f32(x[2 * n + 2048]), to_f32(x[2 * n + 2049])), w1);
    if (c == 0) {
      v[m] = add2(z0, z1);
    } else {
      v[m] = cmul(sub2(z0, z1), tw2k[m * 32 + lane]);
    }
  }
}
#endif

// N = 4096, the pair's loads shared.  Both halves of the pre-stage need every packed element of the frame,
// z[n] + z[n + 1024] for one warp and (z[n] - z[n + 1024]) W2048^n for the other; instead of both warps loading,
// converting and windowing all 4096 samples, warp c does so for the elements m = 16 c .. 16 c + 15 of every lane
// (n = lane + 32 m), forms BOTH combinations, keeps its own and leaves the partner's in `mine` ([16][32] float2,
// the lower 4 KB of its exchange tile, idle in this phase).  After the pair's barrier take_frame_4096_half() picks
// up the other sixteen from the partner's tile.  Same operations on the same operands as load_frame_4096: the bits do
// not change; each warp has half of the load phase, and the odd warp no longer carries all 32 twiddle products.
#if defined(__CUDACC__)
template <typename S, bool WCOS>
__device__ __forceinline__ void load_frame_4096_half(int lane, int c, float2 (&v)[32], const S* x,
                                                     const float* __restrict__ win,
                                                     const float2* __restrict__ tw2k, float2* mine) {
  float4 k0 = make_float4(0.f, 0.f, 0.f, 0.f), k1 = k0;
  float G = 0.f;
  if (WCOS) {
    k0 = reinterpret_cast<const float4*>(win)[lane];
    k1 = reinterpret_cast<const float4*>(win)[32 + lane];
    G = win[256];
  }
#pragma unroll
  for (int m = 0; m < 32; ++m) {
    if ((m >> 4) != c) continue;                       // warp-uniform: the other half is the partner's
    const int n = lane + 32 * m;
    float2 w0, w1;
    if (WCOS) {
      const float C = kWinCos4096[m], Sn = kWinSin4096[m];
      const float2 cc = make_float2(C, C), ss = make_float2(Sn, Sn), gg = make_float2(G, G);
      w0 = fma2(make_float2(k0.x, k0.y), cc, fma2(make_float2(k0.z, k0.w), ss, gg));
      w1 = fma2(make_float2(k1.x, k1.y), cc, fma2(make_float2(k1.z, k1.w), ss, gg));
    } else {
      w0 = *reinterpret_cast<const float2*>(win + 2 * n);
      w1 = *reinterpret_cast<const float2*>(win + 2 * n + 2048);
    }
    const float2 z0 = mul2(make_float2(to_f32(x[2 * n]), to_f32(x[2 * n + 1])), w0);
    const float2 z1 = mul2(make_float2(to_f32(x[2 * n + 2048]), to_f32(x[2 * n + 2049])), w1);
    const float2 y0 = add2(z0, z1), y1 = cmul(sub2(z0, z1), tw2k[m * 32 + lane]);
    if (m < 16) { v[m] = y0; mine[(m & 15) * 32 + lane] = y1; }      // executed by warp c = 0 only
    else { v[m] = y1; mine[(m & 15) * 32 + lane] = y0; }            // executed by warp c = 1 only
  }
}
__device__ __forceinline__ void take_frame_4096_half(int lane, int c, float2 (&v)[32], const float2* theirs) {
#pragma unroll
  for (int m = 0; m < 32; ++m)
    if ((m >> 4) != c) v[m] = theirs[(m & 15) * 32 + lane];
}
#endif

// log(mul * x + add) with the reference's order of operations and roundings (x *= mul; x += add; log(x),
// madmom/audio/spectrogram.py:527-534).  mode: 1 log10, 2 none (mul/add only), 3 ln, 4 log2.
MMB_HD float log_scaled(float x, int mode, float mul, float add) {
#if defined(__CUDA_ARCH__)
  const float y = __fadd_rn(__fmul_rn(x, mul), add);
  if (mode == 1) return __log2f(y) * 0.30102999566398119521f;
  if (mode == 3) return __log2f(y) * 0.69314718055994530942f;
  if (mode == 4) return __log2f(y);
  return y;
#else
  const float y = x * mul + add;
  return mode == 1 ? log10f(y) : mode == 3 ? logf(y) : mode == 4 ? log2f(y) : y;
#endif
}

// ------------------------------------------------------------------------------------------ splits
// Output sink: magnitudes into warp scratch `mag` (always), and optionally complex bins / magnitudes
// straight to global memory (row pointers, may be null).  `sgn_odd` = -1 when circular_shift is set:
// swapping the two halves of the frame multiplies X[k] by (-1)^k (madmom/audio/stft.py:106-119).
template <bool EMIT>
struct BinSinkT {
  float* mag;        // warp scratch, K floats
  float2* stft_row;  // global, K complex, or null
  float* spec_row;   // global, K floats, or null
  float sgn_odd;     // +1 or -1
  int log_mode;      // != 0: the spec output is log(m * mul + add) (LogarithmicSpectrogram of an unfiltered one)
  float mul, add;
  MMB_HD void put(int k, float re, float im) const { put_at(k, k, re, im); }
  // magnitude slot `slot` of the scratch, bin k of the output rows
  MMB_HD void put_at(int slot, int k, float re, float im) const {
    const float m = mag_of(re, im);
    mag[slot] = m;
    if (!EMIT) return;                 // hot path: magnitudes stay on chip, no per-bin branches
    if (spec_row) spec_row[k] = log_mode ? log_scaled(m, log_mode, mul, add) : m;
    if (stft_row) {
      const float s = (k & 1) ? sgn_odd : 1.0f;
      stft_row[k] = make_float2(re * s, im * s);
    }
  }
};
using BinSink = BinSinkT<true>;
MMB_HD BinSinkT<false> quiet(const BinSink& o) { BinSinkT<false> q; q.mag = o.mag; q.stft_row = nullptr; q.spec_row = nullptr; q.sgn_odd = 1.0f; q.log_mode = 0; q.mul = 1.0f; q.add = 0.0f; return q; }


// N = 2048.  v[r] = Z[lane + 32 brev5(r)], zs[k] valid for k in [512, 1024]  (zs[1024] = Z[0]).
template <class Sink>
MMB_HD void split_2048(int lane, const float2 (&v)[32], const float2* zs, const float2* __restrict__ tw2k,
                       const Sink& out) {
#pragma unroll
  for (int q = 0; q < 16; ++q) {
    const float2 zk = v[brev5(q)];
    const int k = lane + 32 * q, kp = 1024 - k;
    const float2 zp = zs[kp];
    const float2 w = tw2k[q * 32 + lane];
    const float2 sp = add2(zk, zp), sm = sub2(zk, zp);          // (ar, bi), (br, ai)
    const float ar = sp.x, ai = sm.y, br = sm.x, bi = sp.y;
    const float tr = w.x * br - w.y * bi, ti = w.x * bi + w.y * br;
    out.put(k, ar + ti, ai - tr);
    if (kp < 1024) out.put(kp, ar - ti, -(ai + tr));
  }
  if (lane == 0) {                       // k = 512 pairs with itself: X[512] = conj(Z[512])
    const float2 z = v[brev5(16)];
    out.put(512, 2.0f * z.x, -2.0f * z.y);
  }
}

#if defined(__CUDACC__)
// |angle(a * conj(b))| / pi in [0, 1]; `flip` = -1 negates the product (circular shift: X[k] -> (-1)^k X[k]).
__device__ __forceinline__ float abs_angle_over_pi(float2 a, float2 b, float flip) {
  const float re = (a.x * b.x + a.y * b.y) * flip;
  const float im = fabsf(a.y * b.x - a.x * b.y);
  // atan2(im, re) / pi for im >= 0: octant reduction to t = min / max in [0, 1], atan(t) = t P(t^2) with a
  // 7-term minimax polynomial (3.2e-7 rad), then the reflections; 0 for the zero vector like np.angle
  const float ar = fabsf(re);
  const float mx = fmaxf(im, ar), mn = fminf(im, ar);
  float rmx;                                            // the zero vector gives t = 0 (0 times a finite reciprocal)
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rmx) : "f"(fmaxf(mx, 1e-30f)));
  const float t = mn * rmx;
  const float u = t * t;
  float p = 0.00681148456f;
  p = fmaf(p, u, -0.0336031997f);
  p = fmaf(p, u, 0.0796223719f);
  p = fmaf(p, u, -0.132332626f);
  p = fmaf(p, u, 0.198077921f);
  p = fmaf(p, u, -0.333173652f);
  p = fmaf(p, u, 0.999996111f);
  float r = p * t * 0.318309886183790671538f;          // in units of pi, [0, 1/4]
  if (im > ar) r = 0.5f - r;
  if (re < 0.0f) r = 1.0f - r;
  return r;
}

// N = 2048 split that also emits the magnitude of the local group delay (ComplexFlux, madmom/audio/stft.py:
// 154-179 and madmom/features/onsets.py:282-284):  lgd[k] = |unwrap(phase)[k] - unwrap(phase)[k+1]| / pi, last bin
// 0.  The difference of neighbouring unwrapped phases is the principal value of the phase difference, i.e.
// |angle(X[k+1] conj X[k])|, so no scan along the bins is needed.  Bins are walked from high q to low q so that
// the neighbour of lane 31 (bin 32(q+1), owned by lane 0) is already known; neighbours travel by warp shuffle.
template <class Sink>
__device__ __forceinline__ void split_2048_lgd(int lane, const float2 (&v)[32], const float2* zs,
                                               const float2* __restrict__ tw2k, const Sink& out,
                                               float* __restrict__ lgd_row, float flip) {
  const float2 z512 = v[brev5(16)];
  const float2 x512 = make_float2(2.0f * z512.x, -2.0f * z512.y);       // meaningful in lane 0 only
  if (lane == 0) out.put(512, x512.x, x512.y);
  float2 prev_lo = x512, prev_hi = x512;
  const int up = (lane + 1) & 31;
#pragma unroll
  for (int q = 15; q >= 0; --q) {
    const float2 zk = v[brev5(q)];
    const int k = lane + 32 * q, kp = 1024 - k;
    const float2 zp = zs[kp];
    const float2 w = tw2k[q * 32 + lane];
    const float2 sp = add2(zk, zp), sm = sub2(zk, zp);          // (ar, bi), (br, ai)
    const float ar = sp.x, ai = sm.y, br = sm.x, bi = sp.y;
    const float tr = w.x * br - w.y * bi, ti = w.x * bi + w.y * br;
    const float2 xk = make_float2(ar + ti, ai - tr);
    const float2 xp = make_float2(ar - ti, -(ai + tr));                  // X[1024 - k]
    out.put(k, xk.x, xk.y);
    if (kp < 1024) out.put(kp, xp.x, xp.y);
    // lane 0 hands lane 31 its bins of the previous (higher) q
    const float2 give_lo = lane == 0 ? prev_lo : xk;
    const float2 give_hi = lane == 0 ? prev_hi : xp;
    float2 nlo, nhi;
    nlo.x = __shfl_sync(0xffffffffu, give_lo.x, up); nlo.y = __shfl_sync(0xffffffffu, give_lo.y, up);
    nhi.x = __shfl_sync(0xffffffffu, give_hi.x, up); nhi.y = __shfl_sync(0xffffffffu, give_hi.y, up);
    lgd_row[k] = abs_angle_over_pi(nlo, xk, flip);                       // X[k+1] against X[k]
    lgd_row[kp - 1] = (kp == 1024) ? 0.0f : abs_angle_over_pi(xp, nhi, flip);   // X[kp] against X[kp-1]
    prev_lo = xk; prev_hi = xp;
  }
}
#endif

// N = 1024, two frames.  Bins of frame A via outA, frame B via outB (k in [0, 512)).
template <class Sink>
MMB_HD void split_1024x2(int lane, const float2 (&v)[32], const float2* zs, const Sink& outA,
                         const Sink& outB, bool haveB) {
#pragma unroll
  for (int q = 0; q < 16; ++q) {
    const float2 zk = v[brev5(q)];
    const int k = lane + 32 * q;
    const float2 zp = zs[1024 - k];
    const float2 sp = add2(zk, zp), sm = sub2(zk, zp);
    outA.put(k, sp.x, sm.y);                             // (Z[k] + conj Z[-k]) / 2
    if (haveB) outB.put(k, sp.y, -sm.x);                 // (Z[k] - conj Z[-k]) / 2i
  }
}

// N = 4096, half c.  Even bins (c = 0) pair as (j, 1024 - j) with twiddle W4096^(2j) = W2048^j;
// odd bins (c = 1) pair as (j, 1023 - j) with twiddle W4096^(2j+1) = tw4k[q*32 + lane].  The sink's `mag` is the
// half's own array: magnitude of bin 2j + c goes to mag[j].
template <class Sink>
MMB_HD void split_4096(int lane, int c, const float2 (&v)[32], const float2* zs,
                       const float2* __restrict__ tw2k, const float2* __restrict__ tw4k, const Sink& out) {
#pragma unroll
  for (int q = 0; q < 16; ++q) {
    const float2 zk = v[brev5(q)];
    const int j = lane + 32 * q;
    const int jp = (c == 0) ? 1024 - j : 1023 - j;
    const float2 zp = zs[jp];
    const float2 w = (c == 0) ? tw2k[q * 32 + lane] : tw4k[q * 32 + lane];
    const float2 sp = add2(zk, zp), sm = sub2(zk, zp);          // (ar, bi), (br, ai)
    const float ar = sp.x, ai = sm.y, br = sm.x, bi = sp.y;
    const float tr = w.x * br - w.y * bi, ti = w.x * bi + w.y * br;
    const int k = 2 * j + c, kp = 2 * jp + c;
    out.put_at(j, k, ar + ti, ai - tr);
    if (kp < 2048) out.put_at(jp, kp, ar - ti, -(ai + tr));
  }
  if (c == 0 && lane == 0) {             // bin 1024 pairs with itself
    const float2 z = v[brev5(16)];
    out.put_at(512, 1024, 2.0f * z.x, -2.0f * z.y);
  }
}

}  // namespace mmb
