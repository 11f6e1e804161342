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

// Reduce 16 packed (frame A, frame B) values per lane over the warp by shuffles alone: at each of four steps a lane keeps
// one half of its values and hands the other half to its partner (15 exchanges), the last step splits the pair itself.
// On return lane j holds the warp total of value j >> 1, frame A on even lanes and frame B on odd lanes.  The kernel is
// bound by the shared-memory data pipe: 31 SHFL (~0.6 data-pipe cycles each next to LDS traffic, tools/microbench/
// mio_bench.cu) replace the 64 wavefronts of a transposed reduction through shared memory.
__device__ __forceinline__ float warp_reduce_scatter_pairs16(c64 (&v)[16], int lane) {
  int dist = 16;
#pragma unroll
  for (int half = 8; half >= 1; half >>= 1, dist >>= 1) {
    const bool hi = (lane & dist) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      const c64 send = hi ? v[i] : v[i + half];
      const c64 keep = hi ? v[i + half] : v[i];
      v[i] = add2(keep, shfl_xor2(send, dist));
    }
  }
  const bool odd = (lane & 1) != 0;
  const float send = odd ? lo2(v[0]) : hi2(v[0]);
  const float keep = odd ? hi2(v[0]) : lo2(v[0]);
  return keep + __shfl_xor_sync(0xffffffffu, send, 1);
}

// ---- DCT-II (ortho) of the clamped log-mel rows of two frames, 16 coefficients at a time (mirror pairs as in
// afx_warp.cuh): lane partials as packed (frame A, frame B) pairs, reduced by warp_reduce_scatter_pairs16.
__device__ __forceinline__ void dct_pair2(const ClipArgs& a, const Smem& sm, float* mfcc_rows, float* planeB,
                                          const float (&da)[4], const float (&db)[4], int row, bool vB, int lane, uint64_t pol) {
  (void)planeB;
  const int stride = 2 * a.n16 + 4;
  const c64 d0 = pk2(da[0], db[0]), d1 = pk2(da[1], db[1]), d2 = pk2(da[2], db[2]), d3 = pk2(da[3], db[3]);
  const c64 e0 = add2(d0, d3), o0 = sub2(d0, d3), e1 = add2(d1, d2), o1 = sub2(d1, d2);
  for (int kg = 0; kg < a.n16; kg += 16) {
    const float* c0row = sm.dct + lane * stride + kg;   // band lane
    const float* c1row = c0row + a.n16;                 // band 63 - lane
    c64 p[16];
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) {
      const float4 c0 = *reinterpret_cast<const float4*>(c0row + 4 * k4);
      const float4 c1 = *reinterpret_cast<const float4*>(c1row + 4 * k4);
      p[4 * k4 + 0] = fma2(e0, bc2(c0.x), mul2(e1, bc2(c1.x)));
      p[4 * k4 + 1] = fma2(o0, bc2(c0.y), mul2(o1, bc2(c1.y)));
      p[4 * k4 + 2] = fma2(e0, bc2(c0.z), mul2(e1, bc2(c1.z)));
      p[4 * k4 + 3] = fma2(o0, bc2(c0.w), mul2(o1, bc2(c1.w)));
    }
    const float cval = warp_reduce_scatter_pairs16(p, lane);
    const int h = lane & 1;
    if (h == 0 || vB) st_slab(mfcc_rows + (size_t)(row + h) * a.n16 + kg + (lane >> 1), cval, pol);
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

// Inter-stage twiddles z[k1] *= W^(k1 * lane) for k1 < ROWS (registers k1 + ROWS, ... repeat the pattern).  Only the
// rows k1 = 1..7 and 8, 16, 24 of the table are read (10 instead of 31 shared-memory loads); W^((8 a + b) lane) is the
// product of two of them: one more rounding (6e-8) per twiddle, 21 complex products per step on a pipe with headroom.
template <int ROWS>
__device__ __forceinline__ void twiddle_rows(c64 (&z)[32], const float2* tw, int lane) {
  float2 B[8], A[4];
#pragma unroll
  for (int b = 1; b < 8; ++b) B[b] = tw[b * 32 + lane];
#pragma unroll
  for (int a = 1; a < ROWS / 8; ++a) A[a] = tw[8 * a * 32 + lane];
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const int k1 = j % ROWS, a = k1 >> 3, b = k1 & 7;
    if (k1 == 0) continue;
    float2 w;
    if (a == 0) w = B[b];
    else if (b == 0) w = A[a];
    else { w.x = fmaf(A[a].x, B[b].x, -A[a].y * B[b].y); w.y = fmaf(A[a].x, B[b].y, A[a].y * B[b].x); }
    z[j] = cmul2(z[j], w.x, w.y);
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

// ---- periodic Hamming window without a table: 0.5 w[n] = 0.27 - 0.23 cos(2 pi n / N) with n = 32 n1 + lane, i.e. the
// cosine of a sum of a compile-time angle (n1) and a per-lane angle whose (cos, -sin) sits in the twiddle table:
// two FFMA per value instead of one shared-memory wavefront (the kernel is bound by the shared-memory data pipe)
__device__ __forceinline__ constexpr float kcos32(int k) {
  constexpr float t[32] = {1.0f, 0.98078528040323043f, 0.92387953251128674f, 0.83146961230254524f, 0.70710678118654752f,
    0.55557023301960218f, 0.38268343236508978f, 0.19509032201612825f, 0.0f, -0.19509032201612825f, -0.38268343236508978f,
    -0.55557023301960218f, -0.70710678118654752f, -0.83146961230254524f, -0.92387953251128674f, -0.98078528040323043f, -1.0f,
    -0.98078528040323043f, -0.92387953251128674f, -0.83146961230254524f, -0.70710678118654752f, -0.55557023301960218f,
    -0.38268343236508978f, -0.19509032201612825f, 0.0f, 0.19509032201612825f, 0.38268343236508978f, 0.55557023301960218f,
    0.70710678118654752f, 0.83146961230254524f, 0.92387953251128674f, 0.98078528040323043f};
  return t[k & 31];
}
__device__ __forceinline__ constexpr float ksin32(int k) { return kcos32(k - 8); }
// half window at angle (2 pi k32 / 32 + lane angle); (cl, msl) = (cos, -sin) of the lane angle
__device__ __forceinline__ float half_hamming(int k32, float cl, float msl) {
  // cos(A + b) = cos A cos b - sin A sin b;  -0.23 cos(A + b) = (-0.23 cos A) cl + (-0.23 sin A) msl
  return fmaf(-0.23f * kcos32(k32), cl, fmaf(-0.23f * ksin32(k32), msl, 0.27f));
}

// predicated shared-memory loads: a lane whose predicate is off issues no access (and so takes part in no bank
// conflict) and gets zero
// (the predicate i < n is formed inside the asm block: one SETP, no bool -> int -> predicate round trip)
__device__ __forceinline__ c64 lds64_if(const c64* p, int i, int n) {
  c64 v;
  asm volatile("{\n\t.reg .pred q;\n\tsetp.lt.s32 q, %2, %3;\n\tmov.b64 %0, 0;\n\t@q ld.shared.b64 %0, [%1];\n\t}" : "=l"(v) : "r"(smem_u32(p)), "r"(i), "r"(n));
  return v;
}
__device__ __forceinline__ float lds32_if(const float* p, int i, int n) {
  float v;
  asm volatile("{\n\t.reg .pred q;\n\tsetp.lt.s32 q, %2, %3;\n\tmov.f32 %0, 0f00000000;\n\t@q ld.shared.f32 %0, [%1];\n\t}" : "=f"(v) : "r"(smem_u32(p)), "r"(i), "r"(n));
  return v;
}

// per-lane mel tables in shared memory (afx_tables.h: MelSparse::lin / kl / w0)
struct MelLane {
  const float4* lin;   // [128] (U0, dU, D0, dD)
  const int* kl;       // [128] first bin | bins << 16
  const float* w0;     // segment 0
};

// ---- sparse Slaney mel by segments with weights evaluated on the fly: inside its segment (q, lane) needs
// U(k0 + i) = U0 + i dU and D(k0 + i) = D0 + i dD (one FFMA each, the scalar broadcasts into the packed FFMA2 for free),
// and a lane whose segment has ended issues no load (lds64_if), which thins out the bank conflicts of the gather
// (the padded walk would otherwise fetch other lanes' bins only to weight them with zero).
// VT = c64: NP packed frame pairs per bin; d[2 p + h][q] as mel_db2.
template <int NP, typename MC, typename LoadP>
__device__ __forceinline__ void mel_db3(const ClipArgs& a, const MelLane& ml, LoadP loadp, float (&d)[2 * NP][4], int lane) {
  c64 U[4][NP], D[4][NP];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
#pragma unroll
    for (int f = 0; f < NP; ++f) { U[q][f] = 0ull; D[q][f] = 0ull; }
    const float4 c = ml.lin[32 * q + lane];
    const int kl = ml.kl[32 * q + lane];
    const int k0 = kl & 0xffff, len = kl >> 16;
    auto body = [&](int i) {
      const float u = fmaf((float)i, c.y, c.x), dd = fmaf((float)i, c.w, c.z);
      c64 p[NP];
      loadp(k0 + i, i, len, p);
#pragma unroll
      for (int f = 0; f < NP; ++f) { U[q][f] = fma2(p[f], bc2(u), U[q][f]); D[q][f] = fma2(p[f], bc2(dd), D[q][f]); }
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
      const float w = ml.w0[i];
      c64 p[NP];
      loadp(a.tab.mel_k00 + i, 0, 1, p);
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
// scalar variant (one frame per step, n_fft = 2048): loadp(k) returns the power of bin k
template <typename MC, typename LoadP>
__device__ __forceinline__ void mel_db3s(const ClipArgs& a, const MelLane& ml, LoadP loadp, float (&d)[1][4], int lane) {
  float U[4], D[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    U[q] = 0.f; D[q] = 0.f;
    const float4 c = ml.lin[32 * q + lane];
    const int kl = ml.kl[32 * q + lane];
    const int k0 = kl & 0xffff, len = kl >> 16;
    auto body = [&](int i) {
      const float p = loadp(k0 + i, i, len);
      U[q] = fmaf(fmaf((float)i, c.y, c.x), p, U[q]); D[q] = fmaf(fmaf((float)i, c.w, c.z), p, D[q]);
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
  float u00 = 0.f;
  if constexpr (MC::kStatic) {
#pragma unroll
    for (int i = 0; i < MC::cz; ++i) u00 = fmaf(ml.w0[i], loadp(a.tab.mel_k00 + i, 0, 1), u00);
  } else {
    for (int i = 0; i < a.tab.mel_cnt0; ++i) u00 = fmaf(ml.w0[i], loadp(a.tab.mel_k00 + i, 0, 1), u00);
  }
  float up0 = __shfl_up_sync(0xffffffffu, U[0], 1);
  float up1 = __shfl_down_sync(0xffffffffu, U[1], 1);
  float up2 = __shfl_up_sync(0xffffffffu, U[2], 1);
  float up3 = __shfl_down_sync(0xffffffffu, U[3], 1);
  if (lane == 0) { up0 = u00; up2 = U[1]; }
  if (lane == 31) { up1 = U[0]; up3 = U[2]; }
  d[0][0] = to_db(D[0] + up0); d[0][1] = to_db(D[1] + up1);
  d[0][2] = to_db(D[2] + up2); d[0][3] = to_db(D[3] + up3);
}

template <int NFFT, int HOP> struct Step2;

// =====================================================================================================
// n_fft = 1024: frames fA, fA + 1 as the real / imaginary lanes of one complex 1024-point FFT
template <int HOP>
struct Step2<1024, HOP> {
  using G = Geo<1024, HOP>;
  template <bool FAST>
  static __device__ __forceinline__ void load(c64 (&z)[32], c64 (&E)[2], const float* rp0, int gA, const Region& rg, int L0,
                                              float coef, int hop, const Smem& sm, int lane) {
    const c64 nc2 = bc2(-coef);
    const float2 wl = sm.tw[32 + lane];   // W_1024^lane = (cos, -sin) of the lane's share of the window angle
    if constexpr (HOP == 0) {
      // run-time hop: the two frames are read separately; E[f] = sum of squares of frame f (its RMS, not a hop block)
#pragma unroll
      for (int p = 0; p < 16; ++p) {
        const c64 YA = load_rows2<FAST>(rp0, p, gA, rg, L0, coef, nc2, lane);
        const c64 YB = load_rows2<FAST>(rp0 + hop, p, gA + hop, rg, L0, coef, nc2, lane);
        E[0] = fma2(YA, YA, E[0]); E[1] = fma2(YB, YB, E[1]);
        const float w0 = half_hamming(2 * p, wl.x, wl.y), w1 = half_hamming(2 * p + 1, wl.x, wl.y);
        z[bitrev32(2 * p)] = pk2(w0 * lo2(YA), w0 * lo2(YB));
        z[bitrev32(2 * p + 1)] = pk2(w1 * hi2(YA), w1 * hi2(YB));
      }
    } else {
      constexpr int SH = HOP / 32;            // rows by which frame B trails frame A
      constexpr int NR = 32 + SH;             // rows of the step
      constexpr int R0 = G::G0 * SH;          // first row of the owned hop blocks
      float y[NR];
#pragma unroll
      for (int p = 0; p < NR / 2; ++p) {
        const c64 Y = load_rows2<FAST>(rp0, p, gA, rg, L0, coef, nc2, lane);
        y[2 * p] = lo2(Y); y[2 * p + 1] = hi2(Y);
        if (2 * p >= R0 && 2 * p < R0 + 2 * SH) E[(2 * p - R0) / SH] = fma2(Y, Y, E[(2 * p - R0) / SH]);
      }
#pragma unroll
      for (int n1 = 0; n1 < 32; ++n1) {
        const float w = half_hamming(n1, wl.x, wl.y);
        z[bitrev32(n1)] = pk2(w * y[n1], w * y[n1 + SH]);
      }
    }
  }
  template <typename MC>
  static __device__ __forceinline__ void run(const ClipArgs& a, const Smem& sm, const MelLane& ml, const Slab2& sl, WarpCtx2& wc, const Region& rg,
                                             int gA, int L0, int nv, int row, bool own_sums, int usum_idx, float floor_db,
                                             unsigned next_bytes, const float* next_src, float* next_dst) {
    const int lane = threadIdx.x & 31;
    c64 z[32];
    c64 E[2] = {0ull, 0ull};
    const int hop = HOP ? HOP : a.hop;
    const int span = (G::FPW - 1) * hop + G::SPAN - (G::FPW - 1) * G::HOPMAX;
    const float* rp0 = wc.tile + 4 + (HOP ? 0 : (gA & 3));   // sample gA + t sits at rp0[t]
#ifdef AFX_FORCE_FAST   // development builds that count the instructions of the interior step (tools/dev)
    const bool fast = true;
#else
    const bool fast = (gA >= 1) && (gA >= rg.lo) && (gA + span <= rg.hi);   // warp-uniform
#endif
    if (fast) load<true>(z, E, rp0, gA, rg, L0, a.coef, hop, sm, lane);
    else load<false>(z, E, rp0, gA, rg, L0, a.coef, hop, sm, lane);
    fft32_dit2(z);
    twiddle_rows<32>(z, sm.tw, lane);   // z[k1] *= W_1024^(k1 * lane)
    transpose32_bitrev64(z, wc.tile, lane);
    prefetch_tile(wc, next_bytes, next_src, next_dst, lane);
    fft32_dit2(z);   // z[k2] = Z[lane + 32 k2]
    // split the two real spectra, power -> P2[k] = (|A_k|^2, |B_k|^2), k = 0..512 
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
    mel_db3<1, MC>(a, ml, [P2](int k, int i, int n, c64 (&p)[1]) { p[0] = lds64_if(P2 + k, i, n); }, d, lane);
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
                                              float coef, int hop, const Smem& sm, int lane) {
    const c64 nc2 = bc2(-coef);
    const float2 wl = sm.tw[32 + lane];   // W_512^lane
    if constexpr (HOP == 0) {
#pragma unroll
      for (int p = 0; p < 8; ++p) {
        c64 Y[4];
#pragma unroll
        for (int f = 0; f < 4; ++f) { Y[f] = load_rows2<FAST>(rp0 + f * hop, p, gA + f * hop, rg, L0, coef, nc2, lane); E[f] = fma2(Y[f], Y[f], E[f]); }
        const float w0 = half_hamming(4 * p, wl.x, wl.y), w1 = half_hamming(4 * p + 2, wl.x, wl.y);
        z[bitrev16(2 * p)] = pk2(w0 * lo2(Y[0]), w0 * lo2(Y[1]));
        z[bitrev16(2 * p + 1)] = pk2(w1 * hi2(Y[0]), w1 * hi2(Y[1]));
        z[16 + bitrev16(2 * p)] = pk2(w0 * lo2(Y[2]), w0 * lo2(Y[3]));
        z[16 + bitrev16(2 * p + 1)] = pk2(w1 * hi2(Y[2]), w1 * hi2(Y[3]));
      }
    } else {
      constexpr int SH = HOP / 32;            // rows between consecutive frames
      constexpr int NR = G::SPAN / 32;        // rows of the step (even)
      constexpr int R0 = G::G0 * SH;
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
        const float w = half_hamming(2 * n1, wl.x, wl.y);
        z[bitrev16(n1)] = pk2(w * y[n1], w * y[n1 + SH]);
        z[16 + bitrev16(n1)] = pk2(w * y[n1 + 2 * SH], w * y[n1 + 3 * SH]);
      }
    }
  }
  template <typename MC>
  static __device__ __forceinline__ void run(const ClipArgs& a, const Smem& sm, const MelLane& ml, const Slab2& sl, WarpCtx2& wc, const Region& rg,
                                             int gA, int L0, int nv, int row, bool own_sums, int usum_idx, float floor_db,
                                             unsigned next_bytes, const float* next_src, float* next_dst) {
    static_assert(HOP == 0 || HOP >= 64, "hop-block sums are accumulated per row pair");
    const int lane = threadIdx.x & 31;
    c64 z[32];
    c64 E[4] = {0ull, 0ull, 0ull, 0ull};
    const int hop = HOP ? HOP : a.hop;
    const int span = (G::FPW - 1) * hop + G::SPAN - (G::FPW - 1) * G::HOPMAX;
    const float* rp0 = wc.tile + 4 + (HOP ? 0 : (gA & 3));   // sample gA + t sits at rp0[t]
#ifdef AFX_FORCE_FAST   // development builds that count the instructions of the interior step (tools/dev)
    const bool fast = true;
#else
    const bool fast = (gA >= 1) && (gA >= rg.lo) && (gA + span <= rg.hi);   // warp-uniform
#endif
    if (fast) load<true>(z, E, rp0, gA, rg, L0, a.coef, hop, sm, lane);
    else load<false>(z, E, rp0, gA, rg, L0, a.coef, hop, sm, lane);
    fft16x2_dit2(z);
    twiddle_rows<16>(z, sm.tw, lane);   // z[16 s + k1] *= W_512^(k1 * lane)
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
    mel_db3<2, MC>(a, ml, [P2](int k, int i, int n, c64 (&p)[2]) { p[0] = lds64_if(P2 + k, i, n); p[1] = lds64_if(P2 + PSTRIDE + k, i, n); }, d, lane);
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
                                              float coef, int hop, const Smem& sm, int lane) {
    (void)hop;
    const float2 wa = sm.ptw[2 * lane], wb = sm.ptw[2 * lane + 1];   // W_2048^(2 lane), W_2048^(2 lane + 1)
    constexpr int R0 = HOP ? G::G0 * (HOP / 64) : 0, R1 = HOP ? R0 + HOP / 64 : 32;   // rows (64 samples each) of the owned hop block (run-time hop: the whole frame)
#pragma unroll
    for (int n1 = 0; n1 < 32; ++n1) {
      const int t = 64 * n1 + 2 * lane;
      c64 Y;
      if (FAST && HOP != 0) {
        const float2 x01 = *reinterpret_cast<const float2*>(rp0 + t);
        Y = pk2(preemph_op(x01.x, rp0[t - 1], coef), preemph_op(x01.y, x01.x, coef));
      } else if (FAST) {   // run-time hop: the tile offset may be odd, no 64-bit loads
        const float xa = rp0[t], xb = rp0[t + 1];
        Y = pk2(preemph_op(xa, rp0[t - 1], coef), preemph_op(xb, xa, coef));
      } else {
        Y = pk2(load_y<false>(rp0, t, gA, rg, L0, coef), load_y<false>(rp0, t + 1, gA, rg, L0, coef));
      }
      if (n1 >= R0 && n1 < R1) E = fma2(Y, Y, E);
      z[bitrev32(n1)] = mul2(pk2(half_hamming(n1, wa.x, wa.y), half_hamming(n1, wb.x, wb.y)), Y);
    }
  }
  template <typename MC>
  static __device__ __forceinline__ void run(const ClipArgs& a, const Smem& sm, const MelLane& ml, const Slab2& sl, WarpCtx2& wc, const Region& rg,
                                             int gA, int L0, int nv, int row, bool own_sums, int usum_idx, float floor_db,
                                             unsigned next_bytes, const float* next_src, float* next_dst) {
    static_assert(HOP == 0 || HOP >= 64, "hop-block sums are accumulated per 64-sample row");
    const int lane = threadIdx.x & 31;
    c64 z[32];
    c64 E = 0ull;
    const int hop = HOP ? HOP : a.hop;
    const int span = (G::FPW - 1) * hop + G::SPAN - (G::FPW - 1) * G::HOPMAX;
    const float* rp0 = wc.tile + 4 + (HOP ? 0 : (gA & 3));   // sample gA + t sits at rp0[t]
#ifdef AFX_FORCE_FAST   // development builds that count the instructions of the interior step (tools/dev)
    const bool fast = true;
#else
    const bool fast = (gA >= 1) && (gA >= rg.lo) && (gA + span <= rg.hi);   // warp-uniform
#endif
    if (fast) load<true>(z, E, rp0, gA, rg, L0, a.coef, hop, sm, lane);
    else load<false>(z, E, rp0, gA, rg, L0, a.coef, hop, sm, lane);
    fft32_dit2(z);
    twiddle_rows<32>(z, sm.tw, lane);
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
    mel_db3s<MC>(a, ml, [P](int k, int i, int n) { return lds32_if(P + k, i, n); }, d, lane);
    __syncwarp();
    float eb[1] = {lo2(E) + hi2(E)};
    finish_step2<2048, HOP>(a, sm, sl, wc, d, eb, nv, row, own_sums, usum_idx, floor_db, lane);
  }
};

}  // namespace afx
