// Warp steps of the streaming kernel (afx_stream.cuh), second generation.  Same arithmetic as afx_warp.cuh
// (register FFT on packed FP32 pairs, segment mel, mirror DCT) with a leaner instruction stream per step:
//
//   * the raw tile and the scratch plane of a warp alias: the 32 x 32 transposes of the FFT run on 64-bit
//     (re, im) pairs over the union of both (one STS.64 + one LDS.64 per value instead of two 32-bit round trips),
//     and the TMA prefetch of the next tile is issued once the transposes have released the tile region;
//   * pre-emphasis and the hop-block energies run on packed row pairs (FFMA2); a step reduces only the sums of
//     squares of the hop blocks it owns -- per-frame RMS is assembled from the block sums when the clip is finished;
//   * the DCT partials of both frames travel through shared memory as packed pairs and are summed with 128-bit loads;
//   * no per-step vote: a non-finite sample poisons its hop-block sum, which the trim decision inspects anyway.
//
// Replaces the framing / STFT / mel / dB / DCT chain of librosa.feature.mfcc and librosa.feature.rms reached from
// audio_feature_extraction_toolkit/core/feature_extractor.py:127-134,164-168 (reference).
#pragma once
#include "afx_warp.cuh"

namespace afx {

constexpr int kPlaneBFloats = 1152;           // region B of a warp buffer: power spectrum / DCT reduction (4608 B)
constexpr int kPlaneTBytes = 32 * 33 * 8;     // the 64-bit transpose plane spans the front of the buffer

// bytes of one warp's buffer: [raw tile | ... | region B], the transpose plane aliases the front
template <int NFFT, int HOP>
struct WarpBuf {
  static constexpr int kTile = Geo<NFFT, HOP>::RAWW * 4;
  static constexpr int kRaw = kTile + kPlaneBFloats * 4 > kPlaneTBytes ? kTile + kPlaneBFloats * 4 : kPlaneTBytes;
  static constexpr int kBytes = (kRaw + 15) & ~15;
};

// slab rows of one clip (per slot of the ring, afx_stream.cuh)
struct Slab2 {
  float* mfcc;   // [a_max][n16] DCT rows (speculative: no top_db clamp; rewritten by the redo pass where it bites)
  float* ext;    // [a_max][2] per-frame (max, min) of the unclamped log-mel row
  float* blk;    // [a_max + 4] sums of squares of the hop-sized blocks of the pre-emphasised clip
};

struct WarpCtx2 {
  float* tile;       // raw-sample tile (TMA destination); position p holds sample gA - 4 + p
  float* planeB;     // region B
  uint64_t* mbar;    // this warp's tile barrier
  unsigned parity;
  uint64_t pol_in, pol_slab;
};

// 32 x 32 transpose of packed complex values through the warp buffer: (lane a, reg k) -> (lane k, reg a), written in
// bit-reversed register order (the input order of the next 32-point DIT).  Row pitch 33 pairs: the column reads of a
// half warp touch 16 distinct even banks.
__device__ __forceinline__ void transpose32_bitrev64(c64 (&z)[32], float* buf, int lane) {
  c64* P = reinterpret_cast<c64*>(buf);
#pragma unroll
  for (int k = 0; k < 32; ++k) P[k * 33 + lane] = z[k];
  __syncwarp();
#pragma unroll
  for (int n = 0; n < 32; ++n) z[bitrev32(n)] = P[lane * 33 + n];
  __syncwarp();
}

__device__ __forceinline__ c64 shfl_xor2(c64 v, int m) {
  return pk2(__shfl_xor_sync(0xffffffffu, lo2(v), m), __shfl_xor_sync(0xffffffffu, hi2(v), m));
}

// ---- DCT-II (ortho) of the clamped log-mel rows of two frames, 16 coefficients at a time (mirror pairs as in
// afx_warp.cuh).  Lane partials are packed (frame A, frame B) pairs: Pc[k][lane]; lane j then sums one half of row
// j % 16 with 128-bit loads (two pairs each), halves meet by one shuffle, lanes < 16 hold frame A's coefficient.
__device__ __forceinline__ void dct_pair2(const ClipArgs& a, const Smem& sm, float* mfcc_rows, float* planeB,
                                          const float (&da)[4], const float (&db)[4], int row, bool vB, int lane, uint64_t pol) {
  c64* Pc = reinterpret_cast<c64*>(planeB);   // [16][34]
  const int stride = 2 * a.n16 + 4;
  const c64 d0 = pk2(da[0], db[0]), d1 = pk2(da[1], db[1]), d2 = pk2(da[2], db[2]), d3 = pk2(da[3], db[3]);
  const c64 e0 = add2(d0, d3), o0 = sub2(d0, d3), e1 = add2(d1, d2), o1 = sub2(d1, d2);
  const int r = lane & 15, h = lane >> 4;
  for (int kg = 0; kg < a.n16; kg += 16) {
    const float* c0row = sm.dct + lane * stride + kg;   // band lane
    const float* c1row = c0row + a.n16;                 // band 63 - lane
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) {
      const float4 c0 = *reinterpret_cast<const float4*>(c0row + 4 * k4);
      const float4 c1 = *reinterpret_cast<const float4*>(c1row + 4 * k4);
      Pc[(4 * k4 + 0) * 34 + lane] = fma2(e0, bc2(c0.x), mul2(e1, bc2(c1.x)));
      Pc[(4 * k4 + 1) * 34 + lane] = fma2(o0, bc2(c0.y), mul2(o1, bc2(c1.y)));
      Pc[(4 * k4 + 2) * 34 + lane] = fma2(e0, bc2(c0.z), mul2(e1, bc2(c1.z)));
      Pc[(4 * k4 + 3) * 34 + lane] = fma2(o0, bc2(c0.w), mul2(o1, bc2(c1.w)));
    }
    __syncwarp();
    const float4* src = reinterpret_cast<const float4*>(Pc + r * 34 + 16 * h);
    c64 s = 0ull, t = 0ull;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 v = src[i];
      s = add2(s, pk2(v.x, v.y)); t = add2(t, pk2(v.z, v.w));
    }
    s = add2(s, t);
    s = add2(s, shfl_xor2(s, 16));
    const float cval = h ? hi2(s) : lo2(s);
    if (h == 0 || vB) st_slab(mfcc_rows + (size_t)(row + h) * a.n16 + kg + r, cval, pol);
    __syncwarp();
  }
}

// ---- tail shared by the three front-ends: per-frame extrema of the unclamped log-mel rows, owned hop-block sums,
// clamp at floor_db (redo pass only), DCT.  d[f][q] = unclamped log-mel of band mel_band_of(q, lane); eb[f] = lane
// partial of the sum of squares of hop block (usum_idx + f).
template <int NFFT, int HOP>
__device__ __forceinline__ void finish_step2(const ClipArgs& a, const Smem& sm, const Slab2& sl, const WarpCtx2& wc,
                                             float (&d)[Geo<NFFT, HOP>::FPW][4], float (&eb)[Geo<NFFT, HOP>::FPW], int nv, int row,
                                             bool own_sums, int usum_idx, float floor_db, int lane) {
  constexpr int FPW = Geo<NFFT, HOP>::FPW;
  {  // owned hop-block sums
    warp_reduce_scatter<FPW>(eb, lane, [](float x, float y) { return x + y; });
    constexpr int LPV = 32 / FPW;
    if (own_sums && (lane & (LPV - 1)) == 0) __stcg(sl.blk + usum_idx + lane / LPV, eb[0]);
  }
  {  // extrema of the unclamped rows: [row maxima | negated row minima]
    constexpr int M = 2 * FPW, LPV = 32 / M;
    float v[M];
#pragma unroll
    for (int f = 0; f < FPW; ++f) {
      v[f] = fmaxf(fmaxf(d[f][0], d[f][1]), fmaxf(d[f][2], d[f][3]));
      v[FPW + f] = -fminf(fminf(d[f][0], d[f][1]), fminf(d[f][2], d[f][3]));
    }
    warp_reduce_scatter<M>(v, lane, [](float x, float y) { return fmaxf(x, y); });
    const int j = lane / LPV, jf = j & (FPW - 1), is_min = j >= FPW;
    if ((lane & (LPV - 1)) == 0 && jf < nv) __stcg(sl.ext + 2 * (row + jf) + is_min, is_min ? -v[0] : v[0]);
  }
  if (floor_db > -1.0e38f) {   // warp-uniform: only the redo pass clamps
#pragma unroll
    for (int f = 0; f < FPW; ++f)
#pragma unroll
      for (int q = 0; q < 4; ++q) d[f][q] = fmaxf(d[f][q], floor_db);
  }
  if (FPW == 1) {
    dct_pair2(a, sm, sl.mfcc, wc.planeB, d[0], d[0], row, false, lane, wc.pol_slab);
  } else {
#pragma unroll
    for (int f = 0; f < FPW; f += 2)
      if (f < nv) dct_pair2(a, sm, sl.mfcc, wc.planeB, d[f], d[f + 1 < FPW ? f + 1 : f], row + f, f + 1 < nv, lane, wc.pol_slab);
  }
}

// issue this warp's next tile: the transposes are done with the tile region (all lanes passed the __syncwarp)
__device__ __forceinline__ void prefetch_tile(WarpCtx2& wc, unsigned bytes, const float* src, float* dst, int lane) {
  if (lane == 0 && bytes) {
    fence_proxy_async();
    mbar_expect_tx(wc.mbar, bytes);
    tma_load_1d_hint(dst, src, bytes, wc.mbar, wc.pol_in);
  }
}

// pre-emphasised rows 2p, 2p + 1 of the tile (32 samples each) as one packed pair; FAST: no bounds checks
template <bool FAST>
__device__ __forceinline__ c64 load_rows2(const float* rp0, int p, int gA, const Region& rg, int L0, float coef, c64 ncoef2, int lane) {
  if (FAST) {
    const int t = 64 * p + lane;
    const c64 x = pk2(rp0[t], rp0[t + 32]);
    const c64 m = pk2(rp0[t - 1], rp0[t + 31]);
    return fma2(m, ncoef2, x);   // fl(x - c * x[-1]) per half, as preemph_op
  }
  return pk2(load_y<false>(rp0, 64 * p + lane, gA, rg, L0, coef), load_y<false>(rp0, 64 * p + 32 + lane, gA, rg, L0, coef));
}

template <int NFFT, int HOP> struct Step2;

// =====================================================================================================
// n_fft = 1024: frames fA, fA + 1 as the real / imaginary lanes of one complex 1024-point FFT
template <int HOP>
struct Step2<1024, HOP> {
  using G = Geo<1024, HOP>;
  template <bool FAST>
  static __device__ __forceinline__ void load(c64 (&z)[32], c64 (&E)[2], const float* rp0, int gA, const Region& rg, int L0,
                                              float coef, const Smem& sm, int lane) {
    constexpr int SH = HOP / 32;            // rows by which frame B trails frame A
    constexpr int NR = 32 + SH;             // rows of the step
    constexpr int R0 = G::G0 * SH;          // first row of the owned hop blocks
    const c64 nc2 = bc2(-coef);
    float y[NR];
#pragma unroll
    for (int p = 0; p < NR / 2; ++p) {
      const c64 Y = load_rows2<FAST>(rp0, p, gA, rg, L0, coef, nc2, lane);
      y[2 * p] = lo2(Y); y[2 * p + 1] = hi2(Y);
      if (2 * p >= R0 && 2 * p < R0 + 2 * SH) E[(2 * p - R0) / SH] = fma2(Y, Y, E[(2 * p - R0) / SH]);
    }
#pragma unroll
    for (int n1 = 0; n1 < 32; ++n1) {
      const float w = sm.win[32 * n1 + lane];
      z[bitrev32(n1)] = pk2(w * y[n1], w * y[n1 + SH]);
    }
  }
  template <typename MC>
  static __device__ __forceinline__ void run(const ClipArgs& a, const Smem& sm, const Slab2& sl, WarpCtx2& wc, const Region& rg,
                                             int gA, int L0, int nv, int row, bool own_sums, int usum_idx, float floor_db,
                                             unsigned next_bytes, const float* next_src, float* next_dst) {
    const int lane = threadIdx.x & 31;
    c64 z[32];
    c64 E[2] = {0ull, 0ull};
#ifdef AFX_FORCE_FAST   // development builds that count the instructions of the interior step (tools/dev)
    const bool fast = true;
#else
    const bool fast = (gA >= 1) && (gA >= rg.lo) && (gA + G::SPAN <= rg.hi);   // warp-uniform
#endif
    if (fast) load<true>(z, E, wc.tile + 4, gA, rg, L0, a.coef, sm, lane);
    else load<false>(z, E, wc.tile + 4, gA, rg, L0, a.coef, sm, lane);
    fft32_dit2(z);
#pragma unroll
    for (int k1 = 1; k1 < 32; ++k1) {   // twiddle by W_1024^(k1 * lane)
      const float2 w = sm.tw[k1 * 32 + lane];
      z[k1] = cmul2(z[k1], w.x, w.y);
    }
    transpose32_bitrev64(z, wc.tile, lane);
    prefetch_tile(wc, next_bytes, next_src, next_dst, lane);
    fft32_dit2(z);   // z[k2] = Z[lane + 32 k2]
    // split the two real spectra, power -> P2[k] = (|A_k|^2, |B_k|^2), k = 0..512 (afx_warp.cuh)
    c64* P2 = reinterpret_cast<c64*>(wc.planeB);
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
    float d[2][4];
    mel_db2<1, MC>(a, sm, [P2](int k, c64 (&p)[1]) { p[0] = P2[k]; }, d, lane);
    __syncwarp();   // P2 consumed: region B is free for the DCT reduction
    float eb[2] = {lo2(E[0]) + hi2(E[0]), lo2(E[1]) + hi2(E[1])};
    finish_step2<1024, HOP>(a, sm, sl, wc, d, eb, nv, row, own_sums, usum_idx, floor_db, lane);
  }
};

// =====================================================================================================
// n_fft = 512: frames fA .. fA + 3 as two complex 512-point FFTs (signal s: re = frame 2s, im = frame 2s + 1)
template <int HOP>
struct Step2<512, HOP> {
  using G = Geo<512, HOP>;
  static constexpr int PSTRIDE = 260;   // pairs between the two signals' power spectra (257 used)
  template <bool FAST>
  static __device__ __forceinline__ void load(c64 (&z)[32], c64 (&E)[4], const float* rp0, int gA, const Region& rg, int L0,
                                              float coef, const Smem& sm, int lane) {
    constexpr int SH = HOP / 32;            // rows between consecutive frames
    constexpr int NR = G::SPAN / 32;        // rows of the step (even)
    constexpr int R0 = G::G0 * SH;
    const c64 nc2 = bc2(-coef);
    float y[NR];
#pragma unroll
    for (int p = 0; p < NR / 2; ++p) {
      const c64 Y = load_rows2<FAST>(rp0, p, gA, rg, L0, coef, nc2, lane);
      y[2 * p] = lo2(Y); y[2 * p + 1] = hi2(Y);
      if (SH >= 2) {
        if (2 * p >= R0 && 2 * p < R0 + 4 * SH) E[(2 * p - R0) / SH] = fma2(Y, Y, E[(2 * p - R0) / SH]);
      }
    }
#pragma unroll
    for (int n1 = 0; n1 < 16; ++n1) {
      const float w = sm.win[32 * n1 + lane];
      z[bitrev16(n1)] = pk2(w * y[n1], w * y[n1 + SH]);
      z[16 + bitrev16(n1)] = pk2(w * y[n1 + 2 * SH], w * y[n1 + 3 * SH]);
    }
  }
  template <typename MC>
  static __device__ __forceinline__ void run(const ClipArgs& a, const Smem& sm, const Slab2& sl, WarpCtx2& wc, const Region& rg,
                                             int gA, int L0, int nv, int row, bool own_sums, int usum_idx, float floor_db,
                                             unsigned next_bytes, const float* next_src, float* next_dst) {
    static_assert(HOP >= 64, "hop-block sums are accumulated per row pair");
    const int lane = threadIdx.x & 31;
    c64 z[32];
    c64 E[4] = {0ull, 0ull, 0ull, 0ull};
#ifdef AFX_FORCE_FAST   // development builds that count the instructions of the interior step (tools/dev)
    const bool fast = true;
#else
    const bool fast = (gA >= 1) && (gA >= rg.lo) && (gA + G::SPAN <= rg.hi);   // warp-uniform
#endif
    if (fast) load<true>(z, E, wc.tile + 4, gA, rg, L0, a.coef, sm, lane);
    else load<false>(z, E, wc.tile + 4, gA, rg, L0, a.coef, sm, lane);
    fft16x2_dit2(z);
#pragma unroll
    for (int j = 0; j < 32; ++j) {      // twiddle by W_512^(k1 * lane), k1 = j & 15
      if ((j & 15) == 0) continue;
      const float2 w = sm.tw[(j & 15) * 32 + lane];
      z[j] = cmul2(z[j], w.x, w.y);
    }
    transpose32_bitrev64(z, wc.tile, lane);
    prefetch_tile(wc, next_bytes, next_src, next_dst, lane);
    fft32_dit2(z);   // lane 16 s + k1: z[k2] = Z_s[k1 + 16 k2]
    c64* P2 = reinterpret_cast<c64*>(wc.planeB);
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
    float eb[4];
#pragma unroll
    for (int f = 0; f < 4; ++f) eb[f] = lo2(E[f]) + hi2(E[f]);
    finish_step2<512, HOP>(a, sm, sl, wc, d, eb, nv, row, own_sums, usum_idx, floor_db, lane);
  }
};

// =====================================================================================================
// n_fft = 2048: one frame per step; z[m] = y[2m] + i y[2m+1] through the complex 1024-point FFT, then
// X[k] = E[k] + W_2048^k O[k], X[1024-k] = conj(E[k] - W_2048^k O[k]) with E / O the even / odd sample spectra
template <int HOP>
struct Step2<2048, HOP> {
  using G = Geo<2048, HOP>;
  template <bool FAST>
  static __device__ __forceinline__ void load(c64 (&z)[32], c64& E, const float* rp0, int gA, const Region& rg, int L0,
                                              float coef, const Smem& sm, int lane) {
    const c64* win2 = reinterpret_cast<const c64*>(sm.win);
    constexpr int R0 = G::G0 * (HOP / 64), R1 = R0 + HOP / 64;   // rows (64 samples each) of the owned hop block
#pragma unroll
    for (int n1 = 0; n1 < 32; ++n1) {
      const int t = 64 * n1 + 2 * lane;
      c64 Y;
      if (FAST) {
        const float2 x01 = *reinterpret_cast<const float2*>(rp0 + t);
        Y = pk2(preemph_op(x01.x, rp0[t - 1], coef), preemph_op(x01.y, x01.x, coef));
      } else {
        Y = pk2(load_y<false>(rp0, t, gA, rg, L0, coef), load_y<false>(rp0, t + 1, gA, rg, L0, coef));
      }
      if (n1 >= R0 && n1 < R1) E = fma2(Y, Y, E);
      z[bitrev32(n1)] = mul2(win2[32 * n1 + lane], Y);
    }
  }
  template <typename MC>
  static __device__ __forceinline__ void run(const ClipArgs& a, const Smem& sm, const Slab2& sl, WarpCtx2& wc, const Region& rg,
                                             int gA, int L0, int nv, int row, bool own_sums, int usum_idx, float floor_db,
                                             unsigned next_bytes, const float* next_src, float* next_dst) {
    static_assert(HOP >= 64, "hop-block sums are accumulated per 64-sample row");
    const int lane = threadIdx.x & 31;
    c64 z[32];
    c64 E = 0ull;
#ifdef AFX_FORCE_FAST   // development builds that count the instructions of the interior step (tools/dev)
    const bool fast = true;
#else
    const bool fast = (gA >= 1) && (gA >= rg.lo) && (gA + G::SPAN <= rg.hi);   // warp-uniform
#endif
    if (fast) load<true>(z, E, wc.tile + 4, gA, rg, L0, a.coef, sm, lane);
    else load<false>(z, E, wc.tile + 4, gA, rg, L0, a.coef, sm, lane);
    fft32_dit2(z);
#pragma unroll
    for (int k1 = 1; k1 < 32; ++k1) {
      const float2 w = sm.tw[k1 * 32 + lane];
      z[k1] = cmul2(z[k1], w.x, w.y);
    }
    transpose32_bitrev64(z, wc.tile, lane);
    prefetch_tile(wc, next_bytes, next_src, next_dst, lane);
    fft32_dit2(z);   // z[k2] = Z[lane + 32 k2]
    float* P = wc.planeB;                    // [1025] power spectrum
    const int src = (32 - lane) & 31;
#pragma unroll
    for (int k2 = 0; k2 <= 16; ++k2) {
      const int pk = (k2 < 16) ? 31 - k2 : 15;
      float pr = __shfl_sync(0xffffffffu, lo2(z[pk]), src);
      float pi = __shfl_sync(0xffffffffu, hi2(z[pk]), src);
      if (lane == 0) { pr = lo2(z[(32 - k2) & 31]); pi = hi2(z[(32 - k2) & 31]); }
      const float zr = lo2(z[k2]), zi = hi2(z[k2]);
      const int k = lane + 32 * k2;
      const float er = zr + pr, ei = zi - pi;               // E = Z_k + conj(Z_{1024-k})   (the 1/2 sits in the window)
      const float orr = zi + pi, oi = pr - zr;              // O = -i (Z_k - conj(Z_{1024-k}))
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
    float eb[1] = {lo2(E) + hi2(E)};
    finish_step2<2048, HOP>(a, sm, sl, wc, d, eb, nv, row, own_sums, usum_idx, floor_db, lane);
  }
};

}  // namespace afx
