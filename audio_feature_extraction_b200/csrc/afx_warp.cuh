// Warp-level building blocks shared by the warp steps (afx_step.cuh): one warp turns the raw samples of 1, 2 or 4
// consecutive frames (its private TMA tile) into MFCC rows, energies and log-mel extrema without any CTA-wide barrier.
//
//   n_fft = 1024: two frames per step as the real / imaginary lanes of one complex 1024-point FFT
//   n_fft =  512: four frames per step as two complex 512-point FFTs (16 + 16 complex values per lane)
//   n_fft = 2048: one frame per step, even / odd samples packed into one complex 1024-point FFT + a split pass
//
// Replaces the framing / STFT / mel / dB / DCT chain of librosa.feature.mfcc and librosa.feature.rms reached from
// audio_feature_extraction_toolkit/core/feature_extractor.py:127-134,164-168 (reference).
#pragma once
#include "afx_args.cuh"
#include "afx_common.cuh"
#include "fft_gen.cuh"

namespace afx {

// where a streaming step reads: frames f (relative to `origin`) cover origin + f*hop - n_fft/2 + [0, n_fft),
// samples outside [lo, hi) count as zero (librosa centre padding of the trimmed signal)
struct Region { int origin, lo, hi; int raw = 0; };   // raw: the input is already pre-emphasised (no first-sample rule)

// shared-memory tables + per-warp buffers
struct Smem {
  uint64_t* mbar;     // tile barriers, then work-item barriers, one of each per warp
  float2* tw;         // [32][32] W_1024^(row*col) (n_fft 1024, 2048) or [16][32] W_512^(row*col) (n_fft 512)
  float2* ptw;        // n_fft 2048: W_2048^k, k = 0..512
  const float* dct;   // [32][2 n16 + 4] per lane: DCT-II coefficients of band `lane`, then of band 63 - lane (padded stride)
};

// per-(n_fft, hop) geometry of a warp step
// HOP == 0: the hop is a run-time value (any hop <= n_fft / 2, streaming kernel only); sizes are those of the largest hop
template <int NFFT, int HOP>
struct Geo {
  static constexpr int FPW = NFFT == 2048 ? 1 : (NFFT == 1024 ? 2 : 4);   // frames per warp step
  static constexpr int HOPMAX = HOP ? HOP : NFFT / 2;
  static constexpr int SPAN = (FPW - 1) * HOPMAX + NFFT;                  // samples a step spans (at most)
  static constexpr int RAWW = SPAN + (HOP ? 8 : 12);                      // floats of the warp's tile
  static constexpr int NG = SPAN / HOPMAX;                                // hop-sized groups in the span
  static constexpr int G0 = NFFT / (2 * HOPMAX);                          // group holding [fA*hop, (fA+1)*hop)
};

__device__ __forceinline__ float preemph_op(float x, float xm1, float coef) {
  // scipy.signal.lfilter with b=[1,-c], a=[1] in float32 computes fl(x + fl(-c * xm1)); the fused form differs by at
  // most one rounding of the product (<= 2^-24 |c xm1|), far inside every tolerance of the path, and saves an issue slot
  return fmaf(-coef, xm1, x);
}

// pre-emphasised sample at tile offset t (absolute sample gA + t); rp0 = raw + 4.
// FAST: the whole step lies inside [max(lo,1), hi), no checks.
template <bool FAST>
__device__ __forceinline__ float load_y(const float* rp0, int t, int gA, const Region& rg, int L0, float coef) {
  if (FAST) return preemph_op(rp0[t], rp0[t - 1], coef);
  const int g = gA + t;
  float y = 0.f;
  if (g >= rg.lo && g < rg.hi) {
    const float xv = rp0[t];
    if (g == 0) {
      const float x1 = (L0 > 1) ? rp0[t + 1] : 0.f;
      y = rg.raw ? xv : xv + (2.f * xv - x1);   // zi = 2*y[0]-y[1] handed to lfilter as state (librosa.effects.preemphasis)
    } else {
      y = preemph_op(xv, rp0[t - 1], coef);
    }
  }
  return y;
}

// 10 log10(max(1e-10, m)).  The clamped argument is never subnormal, so the bare MUFU.LG2 (lg2.approx.ftz) returns what
// __log2f returns without its subnormal pre-scaling (FSETP / FMUL / FADD around the MUFU): 3 instead of 6 instructions
__device__ __forceinline__ float to_db(float m) {
  float l;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(fmaxf(m, 1e-10f)));
  return 3.0102999566398120f * l;
}

// Run lengths of the mel segment walk.  MelCnt<NFFT, false>: read from the tables (any sample rate).  MelCnt<NFFT, true>:
// the stock pairing of this frame length (22.05 kHz / 1024, 16 kHz / 512, 44.1 kHz / 2048; the host checks that the
// tables it built carry exactly these counts before it picks the variant): compile-time trip counts, the walk is
// fully unrolled and every shared-memory offset becomes an immediate.
template <int NFFT, bool STOCK> struct MelCnt { static constexpr bool kStatic = false; };
template <> struct MelCnt<1024, true> { static constexpr bool kStatic = true; static constexpr int c0 = 2, c1 = 3, c2 = 6, c3 = 14, cz = 2; };
template <> struct MelCnt<512, true>  { static constexpr bool kStatic = true; static constexpr int c0 = 1, c1 = 2, c2 = 3, c3 = 7, cz = 1; };
template <> struct MelCnt<2048, true> { static constexpr bool kStatic = true; static constexpr int c0 = 2, c1 = 4, c2 = 12, c3 = 33, cz = 2; };
template <typename MC> __device__ __forceinline__ constexpr int mel_cnt_static(int q) {
  if constexpr (MC::kStatic) return q == 0 ? MC::c0 : q == 1 ? MC::c1 : q == 2 ? MC::c2 : MC::c3;
  else return -1;
}

// Reduce M (power of two <= 16) per-lane values over the warp with M - 1 + log2(32 / M) shuffles instead of 5 M:
// at each of the first log2(M) steps a lane keeps one half of its values and hands the other half to its partner.
// On return v[0] is the full reduction of value index (lane >> (5 - log2 M)).
template <int M, typename Op>
__device__ __forceinline__ void warp_reduce_scatter(float (&v)[M], int lane, Op op) {
  int dist = 16;
#pragma unroll
  for (int half = M / 2; half >= 1; half >>= 1, dist >>= 1) {
    const bool hi = (lane & dist) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      const float send = hi ? v[i] : v[i + half];
      const float keep = hi ? v[i + half] : v[i];
      v[i] = op(keep, __shfl_xor_sync(0xffffffffu, send, dist));
    }
  }
#pragma unroll
  for (; dist >= 1; dist >>= 1) v[0] = op(v[0], __shfl_xor_sync(0xffffffffu, v[0], dist));
}

}  // namespace afx
