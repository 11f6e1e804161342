// Warp-level building blocks of the fused kernel: one warp turns the raw samples of 1, 2 or 4 consecutive
// frames (its private TMA tile) into MFCC rows, RMS and log-mel extrema without any CTA-wide barrier.
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

constexpr int kPlaneFloats = 32 * 33;

// where a streaming step reads: frames f (relative to `origin`) cover origin + f*hop - n_fft/2 + [0, n_fft),
// samples outside [lo, hi) count as zero (librosa centre padding of the trimmed signal)
struct Region { int origin, lo, hi; int raw = 0; };   // raw: the input is already pre-emphasised (no first-sample rule)

// shared-memory tables + per-warp buffers
struct Smem {
  uint64_t* mbar;     // [16] one TMA barrier per warp
  float2* tw;         // [32][32] W_1024^(row*col) (n_fft 1024, 2048) or [16][32] W_512^(row*col) (n_fft 512)
  float2* ptw;        // n_fft 2048: W_2048^k, k = 0..512
  float* win;         // [NFFT] 0.5 * periodic Hamming
  const float* dct;   // [32][2 n16 + 4] per lane: DCT-II coefficients of band `lane`, then of band 63 - lane (padded stride)
  const float* melw;  // copy of Tables::mel_w (float2 entries)
  float* raw;         // [warps][RAWW] per-warp raw-sample tiles
  float* planes;      // [warps][32*33]
};

struct WarpCtx {
  float* raw;        // this warp's raw-sample tile (TMA destination); position p holds sample gA - 4 + p
  float* plane;      // [32*33] transpose scratch / power spectrum / DCT reduction
  uint64_t* mbar;    // this warp's TMA barrier
  unsigned parity;
  uint64_t pol_in;   // L2 policy of the audio stream (evict_first)
  uint64_t pol_slab; // L2 policy of the scratch slab (evict_last)
};

// per-(n_fft, hop) geometry of a warp step
template <int NFFT, int HOP>
struct Geo {
  static constexpr int FPW = NFFT == 2048 ? 1 : (NFFT == 1024 ? 2 : 4);   // frames per warp step
  static constexpr int SPAN = (FPW - 1) * HOP + NFFT;                     // samples a step spans
  static constexpr int RAWW = SPAN + 8;                                   // floats of the warp's tile
  static constexpr int NG = SPAN / HOP;                                   // hop-sized groups in the span
  static constexpr int G0 = NFFT / (2 * HOP);                             // group holding [fA*hop, (fA+1)*hop)
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

// issue this warp's next tile as soon as the current one sits in registers
__device__ __forceinline__ void prefetch_next(WarpCtx& wc, unsigned bytes, const float* src, float* dst, int lane) {
  __syncwarp();
  if (lane == 0 && bytes) {
    fence_proxy_async();
    mbar_expect_tx(wc.mbar, bytes);
    tma_load_1d_hint(dst, src, bytes, wc.mbar, wc.pol_in);
  }
}

// transpose (lane a, reg b) -> (lane b, reg a) of a 32 x 32 tile of packed complex values through the warp's plane
// (one half at a time: the plane holds 32 x 33 floats), writing the result in bit-reversed register order (the input
// order of the next 32-point DIT)
__device__ __forceinline__ void transpose32_bitrev(c64 (&z)[32], float* plane, int lane) {
  float t[32];
#pragma unroll
  for (int k = 0; k < 32; ++k) plane[k * 33 + lane] = lo2(z[k]);
  __syncwarp();
#pragma unroll
  for (int n = 0; n < 32; ++n) t[n] = plane[lane * 33 + n];
  __syncwarp();
#pragma unroll
  for (int k = 0; k < 32; ++k) plane[k * 33 + lane] = hi2(z[k]);
  __syncwarp();
#pragma unroll
  for (int n = 0; n < 32; ++n) z[bitrev32(n)] = pk2(t[n], plane[lane * 33 + n]);
  __syncwarp();
}

// complex 1024-point FFT of z[32*n1 + lane] held at register bitrev32(n1); result Z[lane + 32*k2] in z[k2]
__device__ __forceinline__ void fft1024_core(c64 (&z)[32], const Smem& sm, float* plane, int lane) {
  fft32_dit2(z);
#pragma unroll
  for (int k1 = 1; k1 < 32; ++k1) {   // twiddle by W_1024^(k1 * lane)
    const float2 w = sm.tw[k1 * 32 + lane];
    z[k1] = cmul2(z[k1], w.x, w.y);
  }
  transpose32_bitrev(z, plane, lane);
  fft32_dit2(z);
}

// two complex 512-point FFTs: signal s in registers [16 s, 16 s + 16) as z_s[32*n1 + lane] at bitrev16(n1);
// result: lane 16 s + k1 holds Z_s[k1 + 16*k2] in z[k2]
__device__ __forceinline__ void fft512x2_core(c64 (&z)[32], const Smem& sm, float* plane, int lane) {
  fft16x2_dit2(z);
#pragma unroll
  for (int j = 0; j < 32; ++j) {      // twiddle by W_512^(k1 * lane), k1 = j & 15
    if ((j & 15) == 0) continue;
    const float2 w = sm.tw[(j & 15) * 32 + lane];
    z[j] = cmul2(z[j], w.x, w.y);
  }
  transpose32_bitrev(z, plane, lane);
  fft32_dit2(z);
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

// Packed variant: NP pairs of frames per bin (power spectra stored as (frame 2p, frame 2p + 1) pairs): one FFMA2 per
// (weight, pair).  d[2 p + h][q].
template <int NP, typename MC, typename LoadP>
__device__ __forceinline__ void mel_db2(const ClipArgs& a, const Smem& sm, LoadP loadp, float (&d)[2 * NP][4], int lane) {
  const float2* mw = reinterpret_cast<const float2*>(sm.melw);
  c64 U[4][NP], D[4][NP];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
#pragma unroll
    for (int f = 0; f < NP; ++f) { U[q][f] = 0ull; D[q][f] = 0ull; }
    const int k0 = __ldg(a.tab.mel_k0 + 32 * q + lane);
    const float2* wq = mw + a.tab.mel_woff[q] + lane;
    auto body = [&](int i) {
      const float2 w = wq[32 * i];
      c64 p[NP];
      loadp(k0 + i, p);
#pragma unroll
      for (int f = 0; f < NP; ++f) { U[q][f] = fma2(p[f], bc2(w.x), U[q][f]); D[q][f] = fma2(p[f], bc2(w.y), D[q][f]); }
    };
    if constexpr (MC::kStatic) {
      constexpr int kCnt[4] = {mel_cnt_static<MC>(0), mel_cnt_static<MC>(1), mel_cnt_static<MC>(2), mel_cnt_static<MC>(3)};
#pragma unroll
      for (int i = 0; i < kCnt[q]; ++i) body(i);
    } else {
      const int cnt = a.tab.mel_cnt[q];
#pragma unroll 2
      for (int i = 0; i < cnt; ++i) body(i);
    }
  }
  c64 u00[NP];   // segment 0: the bins below the first mel point rise into band 0
#pragma unroll
  for (int f = 0; f < NP; ++f) u00[f] = 0ull;
  {
    auto body0 = [&](int i) {
      const float w = mw[a.tab.mel_woff0 + i].x;
      c64 p[NP];
      loadp(a.tab.mel_k00 + i, p);
#pragma unroll
      for (int f = 0; f < NP; ++f) u00[f] = fma2(p[f], bc2(w), u00[f]);
    };
    if constexpr (MC::kStatic) {
#pragma unroll
      for (int i = 0; i < MC::cz; ++i) body0(i);
    } else {
      for (int i = 0; i < a.tab.mel_cnt0; ++i) body0(i);
    }
  }
#pragma unroll
  for (int f = 0; f < NP; ++f) {
    c64 up0 = __shfl_up_sync(0xffffffffu, U[0][f], 1);     // band lane - 1
    c64 up1 = __shfl_down_sync(0xffffffffu, U[1][f], 1);   // band 62 - lane = (q1, lane + 1)
    c64 up2 = __shfl_up_sync(0xffffffffu, U[2][f], 1);     // band 63 + lane = (q2, lane - 1)
    c64 up3 = __shfl_down_sync(0xffffffffu, U[3][f], 1);   // band 126 - lane = (q3, lane + 1)
    if (lane == 0) { up0 = u00[f]; up2 = U[1][f]; }        // band 63 is (q1, lane 0)
    if (lane == 31) { up1 = U[0][f]; up3 = U[2][f]; }      // bands 31 and 95 are (q0, 31) and (q2, 31)
    const c64 m0 = add2(D[0][f], up0), m1 = add2(D[1][f], up1), m2 = add2(D[2][f], up2), m3 = add2(D[3][f], up3);
    d[2 * f][0] = to_db(lo2(m0)); d[2 * f + 1][0] = to_db(hi2(m0));
    d[2 * f][1] = to_db(lo2(m1)); d[2 * f + 1][1] = to_db(hi2(m1));
    d[2 * f][2] = to_db(lo2(m2)); d[2 * f + 1][2] = to_db(hi2(m2));
    d[2 * f][3] = to_db(lo2(m3)); d[2 * f + 1][3] = to_db(hi2(m3));
  }
}

// ---- DCT-II (ortho) of two clamped log-mel rows, 16 coefficients at a time.  A lane holds the mirror pairs
// (lane, 127 - lane) = d[0], d[3] and (63 - lane, 64 + lane) = d[1], d[2]; cos(pi k (2 (127 - b) + 1) / 256) =
// (-1)^k cos(pi k (2 b + 1) / 256), so even coefficients need only the pair sums and odd ones the pair differences:
// half the multiplies.  Lane partials, then a transposed reduction through the plane (lane j sums row j =
// (frame j/16, coefficient j%16) over the 32 lanes' partials); rows `row` (A) and `row + 1` (B, if vB) of the slab
__device__ __forceinline__ void dct_pair(const ClipArgs& a, const Smem& sm, const Slab& sl, const WarpCtx& wc,
                                         const float (&da)[4], const float (&db)[4], int row, bool vB, int lane) {
  float* plane = wc.plane;
  const int stride = 2 * a.n16 + 4;
  // (frame A, frame B) pairs: one FMUL2 + one FFMA2 per coefficient
  const c64 d0 = pk2(da[0], db[0]), d1 = pk2(da[1], db[1]), d2 = pk2(da[2], db[2]), d3 = pk2(da[3], db[3]);
  const c64 e0 = add2(d0, d3), o0 = sub2(d0, d3), e1 = add2(d1, d2), o1 = sub2(d1, d2);
  for (int kg = 0; kg < a.n16; kg += 16) {
    float pa[16], pb[16];
    const float* c0row = sm.dct + lane * stride + kg;   // band lane
    const float* c1row = c0row + a.n16;                 // band 63 - lane
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) {
      const float4 c0 = *reinterpret_cast<const float4*>(c0row + 4 * k4);
      const float4 c1 = *reinterpret_cast<const float4*>(c1row + 4 * k4);
      const c64 p0 = fma2(e0, bc2(c0.x), mul2(e1, bc2(c1.x)));
      const c64 p1 = fma2(o0, bc2(c0.y), mul2(o1, bc2(c1.y)));
      const c64 p2 = fma2(e0, bc2(c0.z), mul2(e1, bc2(c1.z)));
      const c64 p3 = fma2(o0, bc2(c0.w), mul2(o1, bc2(c1.w)));
      pa[4 * k4 + 0] = lo2(p0); pb[4 * k4 + 0] = hi2(p0);
      pa[4 * k4 + 1] = lo2(p1); pb[4 * k4 + 1] = hi2(p1);
      pa[4 * k4 + 2] = lo2(p2); pb[4 * k4 + 2] = hi2(p2);
      pa[4 * k4 + 3] = lo2(p3); pb[4 * k4 + 3] = hi2(p3);
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) { plane[k * 33 + lane] = pa[k]; plane[(16 + k) * 33 + lane] = pb[k]; }
    __syncwarp();
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    const float* prow = plane + lane * 33;
#pragma unroll
    for (int i = 0; i < 32; i += 4) { s0 += prow[i]; s1 += prow[i + 1]; s2 += prow[i + 2]; s3 += prow[i + 3]; }
    const float cval = (s0 + s1) + (s2 + s3);
    if (lane < 16 || vB) st_slab(sl.mfcc + (size_t)(row + (lane >> 4)) * a.n16 + kg + (lane & 15), cval, wc.pol_slab);
    __syncwarp();
  }
}


// ---- sparse Slaney mel by SEGMENTS (afx_tables.h): (q, lane) walks the bins of its own segment once, accumulating the
// rising side U (band + 1's weights) and the falling side D (its own band's weights); the band value is D + the U of
// the lane that owns the band below.  loadp(k, p) fetches the NF power values of bin k.  d[f][q] = 10 log10 of band
// mel_band_of(q, lane): q0 = lane, q1 = 63 - lane, q2 = 64 + lane, q3 = 127 - lane.
template <int NF, typename MC, typename LoadP>
__device__ __forceinline__ void mel_db(const ClipArgs& a, const Smem& sm, LoadP loadp, float (&d)[NF][4], int lane) {
  const float2* mw = reinterpret_cast<const float2*>(sm.melw);
  float U[4][NF], D[4][NF];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
#pragma unroll
    for (int f = 0; f < NF; ++f) { U[q][f] = 0.f; D[q][f] = 0.f; }
    const int k0 = __ldg(a.tab.mel_k0 + 32 * q + lane);
    const float2* wq = mw + a.tab.mel_woff[q] + lane;
    auto body = [&](int i) {
      const float2 w = wq[32 * i];
      float p[NF];
      loadp(k0 + i, p);
#pragma unroll
      for (int f = 0; f < NF; ++f) { U[q][f] = fmaf(w.x, p[f], U[q][f]); D[q][f] = fmaf(w.y, p[f], D[q][f]); }
    };
    if constexpr (MC::kStatic) {
      constexpr int kCnt[4] = {mel_cnt_static<MC>(0), mel_cnt_static<MC>(1), mel_cnt_static<MC>(2), mel_cnt_static<MC>(3)};
#pragma unroll
      for (int i = 0; i < kCnt[q]; ++i) body(i);
    } else {
      const int cnt = a.tab.mel_cnt[q];
#pragma unroll 2
      for (int i = 0; i < cnt; ++i) body(i);
    }
  }
  float u00[NF];   // segment 0: the bins below the first mel point rise into band 0
#pragma unroll
  for (int f = 0; f < NF; ++f) u00[f] = 0.f;
  {
    auto body0 = [&](int i) {
      const float w = mw[a.tab.mel_woff0 + i].x;
      float p[NF];
      loadp(a.tab.mel_k00 + i, p);
#pragma unroll
      for (int f = 0; f < NF; ++f) u00[f] = fmaf(w, p[f], u00[f]);
    };
    if constexpr (MC::kStatic) {
#pragma unroll
      for (int i = 0; i < MC::cz; ++i) body0(i);
    } else {
      for (int i = 0; i < a.tab.mel_cnt0; ++i) body0(i);
    }
  }
#pragma unroll
  for (int f = 0; f < NF; ++f) {
    float up0 = __shfl_up_sync(0xffffffffu, U[0][f], 1);     // band lane - 1
    float up1 = __shfl_down_sync(0xffffffffu, U[1][f], 1);   // band 62 - lane = (q1, lane + 1)
    float up2 = __shfl_up_sync(0xffffffffu, U[2][f], 1);     // band 63 + lane = (q2, lane - 1)
    float up3 = __shfl_down_sync(0xffffffffu, U[3][f], 1);   // band 126 - lane = (q3, lane + 1)
    if (lane == 0) { up0 = u00[f]; up2 = U[1][f]; }          // band 63 is (q1, lane 0)
    if (lane == 31) { up1 = U[0][f]; up3 = U[2][f]; }        // bands 31 and 95 are (q0, 31) and (q2, 31)
    d[f][0] = to_db(D[0][f] + up0); d[f][1] = to_db(D[1][f] + up1);
    d[f][2] = to_db(D[2][f] + up2); d[f][3] = to_db(D[3][f] + up3);
  }
}

// ---- tail shared by the three front-ends: per-frame extrema of the unclamped log-mel rows, RMS, hop-block sums,
// clamp at floor_db, DCT.  d[f][q] = unclamped log-mel of band mel_band_of(q, lane), eg = lane-partial hop-group energies.
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

template <int NFFT, int HOP>
__device__ __forceinline__ void finish_step(const ClipArgs& a, const Smem& sm, const Slab& sl, const WarpCtx& wc,
                                            float (&d)[Geo<NFFT, HOP>::FPW][4], float (&eg)[Geo<NFFT, HOP>::NG], int nv, int row,
                                            bool own_sums, int usum_idx, float floor_db, int lane, int& bad) {
  using G = Geo<NFFT, HOP>;
  constexpr int GPF = NFFT / HOP;   // hop groups per frame
  constexpr int FPW = G::FPW;
  constexpr int M = 2 * FPW;        // values per reduction
  constexpr int LPV = 32 / M;       // lanes that end up holding one value
  {  // sums of squares: [frame energies | owned hop blocks (groups G0 + f)]
    float v[M];
#pragma unroll
    for (int f = 0; f < FPW; ++f) {
      float e = 0.f;
#pragma unroll
      for (int g = 0; g < GPF; ++g) e += eg[f + g];
      v[f] = e; v[FPW + f] = eg[G::G0 + f];
    }
    warp_reduce_scatter<M>(v, lane, [](float x, float y) { return x + y; });
    // a NaN / Inf sample poisons the sums of squares (librosa.util.valid_audio); the frame energies cover every group
    if (__any_sync(0xffffffffu, !isfinite(v[0]))) bad = 1;
    const int j = lane / LPV;
    if ((lane & (LPV - 1)) == 0) {
      if (j < FPW) { if (j < nv) __stcg(sl.rms + row + j, sqrtf(v[0] * (1.0f / NFFT))); }
      else if (own_sums) __stcg(sl.blk + usum_idx + (j - FPW), v[0]);
    }
  }
  {  // extrema of the unclamped log-mel rows: [row maxima | negated row minima]
    float v[M];
#pragma unroll
    for (int f = 0; f < FPW; ++f) {
      v[f] = fmaxf(fmaxf(d[f][0], d[f][1]), fmaxf(d[f][2], d[f][3]));
      v[FPW + f] = -fminf(fminf(d[f][0], d[f][1]), fminf(d[f][2], d[f][3]));
    }
    warp_reduce_scatter<M>(v, lane, [](float x, float y) { return fmaxf(x, y); });
    const int j = lane / LPV;
    if ((lane & (LPV - 1)) == 0) {
      if (j < FPW) { if (j < nv) __stcg(sl.rowmax + row + j, v[0]); }
      else if (j - FPW < nv) __stcg(sl.rowmin + row + (j - FPW), -v[0]);
    }
#pragma unroll
    for (int f = 0; f < FPW; ++f)
#pragma unroll
      for (int q = 0; q < 4; ++q) d[f][q] = fmaxf(d[f][q], floor_db);
  }
  if (G::FPW == 1) {
    dct_pair(a, sm, sl, wc, d[0], d[0], row, false, lane);
  } else {
#pragma unroll
    for (int f = 0; f < G::FPW; f += 2)
      if (f < nv) dct_pair(a, sm, sl, wc, d[f], d[f + 1 < G::FPW ? f + 1 : f], row + f, f + 1 < nv, lane);
  }
}

// =====================================================================================================
// n_fft = 1024: frames fA, fA + 1 as the real / imaginary lanes of one complex 1024-point FFT
template <int NFFT, int HOP> struct WarpStep;

template <int HOP>
struct WarpStep<1024, HOP> {
  using G = Geo<1024, HOP>;
  template <bool FAST>
  static __device__ __forceinline__ void load(float (&r)[32], float (&im)[32], float (&eg)[G::NG], const float* rp0, int gA,
                                              const Region& rg, int L0, float coef, const Smem& sm, int lane) {
    constexpr int SH = HOP / 32;            // rows by which frame B trails frame A
#pragma unroll
    for (int n1 = 0; n1 < 32 + SH; ++n1) {
      const float y = load_y<FAST>(rp0, 32 * n1 + lane, gA, rg, L0, coef);
      eg[n1 / SH] = fmaf(y, y, eg[n1 / SH]);
      if (n1 < 32) r[bitrev32(n1)] = sm.win[32 * n1 + lane] * y;
      if (n1 >= SH) im[bitrev32(n1 - SH)] = sm.win[32 * (n1 - SH) + lane] * y;
    }
  }
  template <typename MC>
  static __device__ __forceinline__ void run(const ClipArgs& a, const Smem& sm, const Slab& sl, WarpCtx& wc, const Region& rg,
                                             int gA, int L0, int nv, int row, bool own_sums, int usum_idx, float floor_db,
                                             unsigned next_bytes, const float* next_src, float* next_dst, int& bad) {
    const int lane = threadIdx.x & 31;
    float r[32], im[32];
    float eg[G::NG];
#pragma unroll
    for (int g = 0; g < G::NG; ++g) eg[g] = 0.f;
#ifdef AFX_FORCE_FAST   // development builds that count the instructions of the interior step (tools/dev)
    const bool fast = true;
#else
    const bool fast = (gA >= 1) && (gA >= rg.lo) && (gA + G::SPAN <= rg.hi);   // warp-uniform
#endif
    if (fast) load<true>(r, im, eg, wc.raw + 4, gA, rg, L0, a.coef, sm, lane);
    else load<false>(r, im, eg, wc.raw + 4, gA, rg, L0, a.coef, sm, lane);
    prefetch_next(wc, next_bytes, next_src, next_dst, lane);
    float* plane = wc.plane;
    c64 z[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) z[k] = pk2(r[k], im[k]);
    fft1024_core(z, sm, plane, lane);   // z[k2] = Z[lane + 32 k2]
    // split the two real spectra, power -> P2[k] = (|A_k|^2, |B_k|^2), k = 0..512.  With the partner Q = Z[1024 - k]:
    // X = Z + Q = (sr, si), Y = Z - Q = (dr, di); |A|^2 = sr^2 + di^2, |B|^2 = si^2 + dr^2, i.e. X^2 + swap(Y^2) on pairs
    c64* P2 = reinterpret_cast<c64*>(plane);
    const int src = (32 - lane) & 31;
#pragma unroll
    for (int k2 = 0; k2 <= 16; ++k2) {
      const int pk = (k2 < 16) ? 31 - k2 : 15;
      float pr = __shfl_sync(0xffffffffu, lo2(z[pk]), src);
      float pi = __shfl_sync(0xffffffffu, hi2(z[pk]), src);
      if (lane == 0) { pr = lo2(z[(32 - k2) & 31]); pi = hi2(z[(32 - k2) & 31]); }
      const c64 q = pk2(pr, pi);
      const c64 x = add2(z[k2], q), y = sub2(z[k2], q);
      const c64 pw = fma2(x, x, swap2(mul2(y, y)));
      if (k2 < 16 || lane == 0) P2[lane + 32 * k2] = pw;
    }
    __syncwarp();
    // sparse mel by segments, 10 log10
    float d[2][4];
    mel_db2<1, MC>(a, sm, [P2](int k, c64 (&p)[1]) { p[0] = P2[k]; }, d, lane);
    __syncwarp();   // P2 consumed: the plane is free for the DCT reduction
    finish_step<1024, HOP>(a, sm, sl, wc, d, eg, nv, row, own_sums, usum_idx, floor_db, lane, bad);
  }
};

// =====================================================================================================
// n_fft = 512: frames fA .. fA + 3 as two complex 512-point FFTs (signal s: re = frame 2s, im = frame 2s + 1)
template <int HOP>
struct WarpStep<512, HOP> {
  using G = Geo<512, HOP>;
  static constexpr int PSTRIDE = 260;   // float2 elements between the two signals' power spectra (257 used)
  template <bool FAST>
  static __device__ __forceinline__ void load(float (&r)[32], float (&im)[32], float (&eg)[G::NG], const float* rp0, int gA,
                                              const Region& rg, int L0, float coef, const Smem& sm, int lane) {
    constexpr int SH = HOP / 32;            // rows between consecutive frames
    constexpr int NR = G::SPAN / 32;        // rows of 32 samples in the step
#pragma unroll
    for (int row = 0; row < NR; ++row) {
      const float y = load_y<FAST>(rp0, 32 * row + lane, gA, rg, L0, coef);
      eg[row / SH] = fmaf(y, y, eg[row / SH]);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n1 = row - j * SH;
        if (n1 >= 0 && n1 < 16) {
          const float v = sm.win[32 * n1 + lane] * y;
          if (j & 1) im[16 * (j >> 1) + bitrev16(n1)] = v; else r[16 * (j >> 1) + bitrev16(n1)] = v;
        }
      }
    }
  }
  template <typename MC>
  static __device__ __forceinline__ void run(const ClipArgs& a, const Smem& sm, const Slab& sl, WarpCtx& wc, const Region& rg,
                                             int gA, int L0, int nv, int row, bool own_sums, int usum_idx, float floor_db,
                                             unsigned next_bytes, const float* next_src, float* next_dst, int& bad) {
    const int lane = threadIdx.x & 31;
    float r[32], im[32];
    float eg[G::NG];
#pragma unroll
    for (int g = 0; g < G::NG; ++g) eg[g] = 0.f;
#ifdef AFX_FORCE_FAST   // development builds that count the instructions of the interior step (tools/dev)
    const bool fast = true;
#else
    const bool fast = (gA >= 1) && (gA >= rg.lo) && (gA + G::SPAN <= rg.hi);   // warp-uniform
#endif
    if (fast) load<true>(r, im, eg, wc.raw + 4, gA, rg, L0, a.coef, sm, lane);
    else load<false>(r, im, eg, wc.raw + 4, gA, rg, L0, a.coef, sm, lane);
    prefetch_next(wc, next_bytes, next_src, next_dst, lane);
    float* plane = wc.plane;
    c64 z[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) z[k] = pk2(r[k], im[k]);
    fft512x2_core(z, sm, plane, lane);   // lane 16 s + k1: z[k2] = Z_s[k1 + 16 k2]
    // split each signal's two real spectra, power -> P2[s][k] = (|X_{2s}[k]|^2, |X_{2s+1}[k]|^2), k = 0..256
    c64* P2 = reinterpret_cast<c64*>(plane);
    const int k1 = lane & 15;
    const int src = (lane & 16) | ((16 - k1) & 15);
    c64* Ps = P2 + (lane >> 4) * PSTRIDE + k1;
#pragma unroll
    for (int k2 = 0; k2 <= 16; ++k2) {
      const int pk = (k2 < 16) ? 31 - k2 : 15;
      float pr = __shfl_sync(0xffffffffu, lo2(z[pk]), src);
      float pi = __shfl_sync(0xffffffffu, hi2(z[pk]), src);
      if (k1 == 0) { pr = lo2(z[(32 - k2) & 31]); pi = hi2(z[(32 - k2) & 31]); }
      const c64 q = pk2(pr, pi);
      const c64 x = add2(z[k2], q), y = sub2(z[k2], q);
      const c64 pw = fma2(x, x, swap2(mul2(y, y)));
      if (k2 < 16 || k1 == 0) Ps[16 * k2] = pw;
    }
    __syncwarp();
    float d[4][4];
    mel_db2<2, MC>(a, sm, [P2](int k, c64 (&p)[2]) { p[0] = P2[k]; p[1] = P2[PSTRIDE + k]; }, d, lane);
    __syncwarp();
    finish_step<512, HOP>(a, sm, sl, wc, d, eg, nv, row, own_sums, usum_idx, floor_db, lane, bad);
  }
};

// =====================================================================================================
// n_fft = 2048: one frame per step; z[m] = y[2m] + i y[2m+1] through the complex 1024-point FFT, then
// X[k] = E[k] + W_2048^k O[k], X[1024-k] = conj(E[k] - W_2048^k O[k]) with E / O the even / odd sample spectra
template <int HOP>
struct WarpStep<2048, HOP> {
  using G = Geo<2048, HOP>;
  template <bool FAST>
  static __device__ __forceinline__ void load(float (&r)[32], float (&im)[32], float (&eg)[G::NG], const float* rp0, int gA,
                                              const Region& rg, int L0, float coef, const Smem& sm, int lane) {
    const float2* win2 = reinterpret_cast<const float2*>(sm.win);
#pragma unroll
    for (int n1 = 0; n1 < 32; ++n1) {
      const int t = 64 * n1 + 2 * lane;
      float y0, y1;
      if (FAST) {
        const float2 x01 = *reinterpret_cast<const float2*>(rp0 + t);
        y0 = preemph_op(x01.x, rp0[t - 1], coef);
        y1 = preemph_op(x01.y, x01.x, coef);
      } else {
        y0 = load_y<false>(rp0, t, gA, rg, L0, coef);
        y1 = load_y<false>(rp0, t + 1, gA, rg, L0, coef);
      }
      const float2 w = win2[32 * n1 + lane];
      eg[(64 * n1) / HOP] = fmaf(y0, y0, fmaf(y1, y1, eg[(64 * n1) / HOP]));
      r[bitrev32(n1)] = w.x * y0;
      im[bitrev32(n1)] = w.y * y1;
    }
  }
  template <typename MC>
  static __device__ __forceinline__ void run(const ClipArgs& a, const Smem& sm, const Slab& sl, WarpCtx& wc, const Region& rg,
                                             int gA, int L0, int nv, int row, bool own_sums, int usum_idx, float floor_db,
                                             unsigned next_bytes, const float* next_src, float* next_dst, int& bad) {
    const int lane = threadIdx.x & 31;
    float r[32], im[32];
    float eg[G::NG];
#pragma unroll
    for (int g = 0; g < G::NG; ++g) eg[g] = 0.f;
#ifdef AFX_FORCE_FAST   // development builds that count the instructions of the interior step (tools/dev)
    const bool fast = true;
#else
    const bool fast = (gA >= 1) && (gA >= rg.lo) && (gA + G::SPAN <= rg.hi);   // warp-uniform
#endif
    if (fast) load<true>(r, im, eg, wc.raw + 4, gA, rg, L0, a.coef, sm, lane);
    else load<false>(r, im, eg, wc.raw + 4, gA, rg, L0, a.coef, sm, lane);
    prefetch_next(wc, next_bytes, next_src, next_dst, lane);
    float* plane = wc.plane;
    c64 z[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) z[k] = pk2(r[k], im[k]);
    fft1024_core(z, sm, plane, lane);   // z[k2] = Z[lane + 32 k2]
#pragma unroll
    for (int k = 0; k < 32; ++k) { r[k] = lo2(z[k]); im[k] = hi2(z[k]); }
    float* P = plane;                        // [1025] power spectrum
    const int src = (32 - lane) & 31;
#pragma unroll
    for (int k2 = 0; k2 <= 16; ++k2) {
      const int pk = (k2 < 16) ? 31 - k2 : 15;
      float pr = __shfl_sync(0xffffffffu, r[pk], src);
      float pi = __shfl_sync(0xffffffffu, im[pk], src);
      if (lane == 0) { pr = r[(32 - k2) & 31]; pi = im[(32 - k2) & 31]; }
      const int k = lane + 32 * k2;
      const float er = r[k2] + pr, ei = im[k2] - pi;        // E = Z_k + conj(Z_{1024-k})   (the 1/2 sits in the window)
      const float orr = im[k2] + pi, oi = pr - r[k2];       // O = -i (Z_k - conj(Z_{1024-k}))
      const float2 w = sm.ptw[k2 < 16 ? k : 512];           // W_2048^k = (cos, -sin)
      const float tr = orr * w.x - oi * w.y;
      const float ti = fmaf(orr, w.y, oi * w.x);
      const float ar = er + tr, ai = ei + ti, br = er - tr, bi = ei - ti;
      if (k2 < 16 || lane == 0) { P[k] = fmaf(ar, ar, ai * ai); P[1024 - k] = fmaf(br, br, bi * bi); }
    }
    __syncwarp();
    float d[1][4];
    mel_db<1, MC>(a, sm, [P](int k, float (&p)[1]) { p[0] = P[k]; }, d, lane);
    __syncwarp();
    finish_step<2048, HOP>(a, sm, sl, wc, d, eg, nv, row, own_sums, usum_idx, floor_db, lane, bad);
  }
};

}  // namespace afx
