// The fused single-pass hot path: ONE persistent kernel takes every clip from raw samples to its
// summary row.  Replaces, for a whole batch, preprocess_audio + extract_mfcc + extract_energy of
// audio_feature_extraction_toolkit/core/feature_extractor.py:58-74,116-151,153-179 (reference).
//
// One CTA (16 warps) owns one clip at a time (clips are handed out by an atomic counter):
//
//  S  streaming pass -- the clip is read from HBM exactly once, 32 frames per step, by 1-D TMA bulk
//     copies (cp.async.bulk + mbarrier) that overlap the FFTs of the previous step.  Per step:
//     pre-emphasis into shared memory; the sums of squares of the 512-sample blocks the step owns
//     (all librosa.effects.trim needs); two frames per warp as one complex 1024-point FFT in
//     registers -> power -> sparse Slaney mel -> 10 log10; log-mel rows, per-frame RMS and per-frame
//     row maxima go to the CTA's private scratch slab (441 KB per 10 s clip; reused clip after clip,
//     so it lives in L2, never in HBM).
//     The frame grid of the trimmed signal starts at trim_start, a multiple of 512; when hop divides
//     512 every trimmed frame centre is also a multiple of hop of the UNtrimmed signal, so the pass
//     runs speculatively on the absolute grid before the trim decision exists.
//  T  trim decision from the block sums (first / last 2048/512 RMS frame above max - top_db).
//  F  fix-up: only the first and last n_fft/(2 hop) (+1) trimmed frames see different zero padding
//     than their absolute-grid twins; they are recomputed (a few KB re-read, L2 hits).
//  C  power_to_db clamp at (clip max - 80 dB) -> DCT-II -> delta / delta-delta (SavGol 9) -> moments.
//  Z  statistics -> summary row, meta row.
//
// If hop does not divide 512 the host runs trim_kernel first and this kernel takes the trim as given
// (`pretrimmed`), skipping T and F.
#pragma once
#include "afx_common.cuh"
#include "afx_trim.cuh"
#include "afx_warp.cuh"

namespace afx {

constexpr int kMaxClipWarps = 16;                // CTA = 8 warps (two CTAs per SM) or 16 warps (one per SM), chosen by the host
constexpr int kMaxClipThreads = kMaxClipWarps * 32;
constexpr int kStatRows = kStatFrames + 2 * kStatHalo;   // frames per sub-chunk incl. halo
constexpr int kMaxRedo = 1024;                            // listed steps of the top_db redo pass
constexpr int kMoments = 6;                               // sum x, x^2, x^3, x^4, sum delta, sum delta2

// raw sample range a warp step starting at absolute sample gA needs; tile position p holds sample gA - 4 + p.
// Returns the bytes to copy (0: nothing of the step lies inside [lo, hi)).
__device__ __forceinline__ unsigned pair_copy_range(const Region& rg, int gA, int span, int L0, int& g_lo, int& s_lo) {
  const int lo = max(rg.lo, gA), hi = min(rg.hi, gA + span);
  if (hi <= lo) return 0u;
  g_lo = max(0, lo - 1) & ~3;
  const int g_hi4 = (min(hi, L0) + 3) & ~3;
  s_lo = g_lo - gA + 4;                                             // multiple of 4, >= 0
  return (unsigned)(g_hi4 - g_lo) * 4u;
}

// ---- streaming pass over frames [f_begin, f_end) of `rg`; slab row = row_shift + f.  Warp w takes the frame pairs
// f_begin + 2w, + 2*16, ...; no CTA-wide barrier inside.  own_sums: rg.origin == 0 and the hop-block sums are wanted.
// floor_db: power_to_db clamp applied before the DCT (-3e38 = none, the speculative pass).  redo: only pairs whose
// unclamped row minimum lies below floor_db are (re)computed -- the pass that makes the speculation exact.
template <int NFFT, int HOP, bool STOCK>
__device__ __forceinline__ void stream_frames(const ClipArgs& a, const Smem& sm, const Slab& sl, WarpCtx& wc,
                                              const float* __restrict__ clip_audio, int L0, Region rg, int f_begin, int f_end,
                                              int row_shift, bool own_sums, float floor_db, bool redo, int& bad,
                                              const int* redo_list = nullptr, int n_redo = 0) {
  using G = Geo<NFFT, HOP>;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nw = (int)(blockDim.x >> 5);
  const int stride = G::FPW * nw;   // frames per CTA step
  auto need = [&](int f) -> bool {   // warp-uniform
    if (!redo) return true;
    float m = 3.0e38f;
    for (int j = 0; j < G::FPW; ++j) if (f + j < f_end) m = fminf(m, __ldcg(sl.rowmin + row_shift + f + j));
    return m < floor_db;
  };
  // redo with a list: the steps to recompute were collected block-wide (redo_list[i] = first frame of a step)
  int li = warp;
  auto first_frame = [&]() -> int {
    if (redo_list) return li < n_redo ? redo_list[li] : f_end;
    int f = f_begin + G::FPW * warp;
    while (f < f_end && !need(f)) f += stride;
    return f;
  };
  auto next_frame = [&](int f) -> int {
    if (redo_list) { li += nw; return li < n_redo ? redo_list[li] : f_end; }
    f += stride;
    while (f < f_end && !need(f)) f += stride;
    return f;
  };
  int fA = first_frame();
  if (fA >= f_end) return;
  int gA = rg.origin + fA * HOP - NFFT / 2;
  unsigned cur_bytes;
  {
    int g_lo = 0, s_lo = 0;
    cur_bytes = pair_copy_range(rg, gA, G::SPAN, L0, g_lo, s_lo);
    if (lane == 0 && cur_bytes) {
      fence_proxy_async();
      mbar_expect_tx(wc.mbar, cur_bytes);
      tma_load_1d_hint(wc.raw + s_lo, clip_audio + g_lo, cur_bytes, wc.mbar, wc.pol_in);
    }
  }
  while (fA < f_end) {
    const int fN = next_frame(fA);
    const int gN = rg.origin + fN * HOP - NFFT / 2;
    unsigned next_bytes = 0;
    int g_lo = 0, s_lo = 0;
    if (fN < f_end) next_bytes = pair_copy_range(rg, gN, G::SPAN, L0, g_lo, s_lo);
    if (cur_bytes) { mbar_wait(wc.mbar, wc.parity); wc.parity ^= 1u; }
    WarpStep<NFFT, HOP>::template run<MelCnt<NFFT, STOCK>>(a, sm, sl, wc, rg, gA, L0, min(G::FPW, f_end - fA), row_shift + fA, own_sums, fA, floor_db,
                             next_bytes, clip_audio + g_lo, wc.raw + s_lo, bad);
    fA = fN; gA = gN; cur_bytes = next_bytes;
  }
}

// ---- block-wide helpers (all (int)blockDim.x threads call them; red = shared scratch of >= 64 doubles)
__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = red[0];
#pragma unroll
  for (int w = 1; w < (int)(blockDim.x >> 5); ++w) t = fmaxf(t, red[w]);
  return t;
}
__device__ __forceinline__ float block_min(float v, float* red) {
  v = warp_min(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = red[0];
#pragma unroll
  for (int w = 1; w < (int)(blockDim.x >> 5); ++w) t = fminf(t, red[w]);
  return t;
}
__device__ __forceinline__ double block_sum(double v, double* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0;
#pragma unroll
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
  return t;
}

// ---- T: librosa.effects.trim(top_db) from the block sums (core/feature_extractor.py:72)
//   r_i = sqrt(mean(y_pre[512 i - 1024 .. 512 i + 1024)^2)), zero padded, i = 0 .. L0/512
//   keep frames with 10 log10(max(1e-10, r_i^2)) - 10 log10(max(1e-10, max_i r_i^2)) > -top_db
//   start = 512 * first, end = min(L0, 512 * (last + 1))
__device__ __forceinline__ float trim_block(const float* usum, int b, int nb, int R, int nu) {
  // sum of squares of 512-sample block b from the hop-sized unit sums (fixed order); zero outside the clip
  if (b < 0 || b >= nb) return 0.f;
  float s = 0.f;
  for (int i = 0; i < R; ++i) { const int u = b * R + i; if (u < nu) s += __ldcg(usum + u); }
  return s;
}
__device__ __forceinline__ ClipInfo trim_decide(const float* usum, int nu, int L0, int hop, float top_db, int bad, float* red_f, int* red_i) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n_rms = 1 + L0 / kTrimHop;
  const int nb = (L0 + kTrimHop - 1) / kTrimHop;
  const int R = kTrimHop / hop;
  float mx = 0.f;
  for (int i = tid; i < n_rms; i += (int)blockDim.x) {
    // frame i covers blocks i-2 .. i+1
    const float s = (trim_block(usum, i - 2, nb, R, nu) + trim_block(usum, i - 1, nb, R, nu)) +
                    (trim_block(usum, i, nb, R, nu) + trim_block(usum, i + 1, nb, R, nu));
    mx = fmaxf(mx, sqrtf(s * (1.0f / kTrimFrame)));
  }
  mx = block_max(mx, red_f);
  const float ref_db = 10.0f * log10f(fmaxf(1e-10f, mx * mx));
  int first = 0x7fffffff, last = -1;
  for (int i = tid; i < n_rms; i += (int)blockDim.x) {
    const float s = (trim_block(usum, i - 2, nb, R, nu) + trim_block(usum, i - 1, nb, R, nu)) +
                    (trim_block(usum, i, nb, R, nu) + trim_block(usum, i + 1, nb, R, nu));
    const float r = sqrtf(s * (1.0f / kTrimFrame));
    float db = 10.0f * log10f(fmaxf(1e-10f, r * r));
    db -= ref_db;
    if (db > -top_db) { first = min(first, i); last = max(last, i); }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
    last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
  }
  __syncthreads();
  if (lane == 0) { red_i[warp] = first; red_i[kMaxClipWarps + warp] = last; }
  __syncthreads();
#pragma unroll
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { first = min(first, red_i[w]); last = max(last, red_i[kMaxClipWarps + w]); }
  ClipInfo ci;
  ci.start = 0; ci.end = 0; ci.len = 0; ci.n_frames = 0; ci.max_key = 0u; ci.pad0 = ci.pad1 = 0; ci.status = 0;
  if (L0 <= 0) ci.status = 1;                 // AFX_CLIP_EMPTY
  else if (bad) ci.status = 2;                // AFX_CLIP_NONFINITE
  else if (last < 0) ci.status = 1;           // every frame silent -> empty signal
  else {
    ci.start = first * kTrimHop;
    ci.end = min(L0, (last + 1) * kTrimHop);
    ci.len = ci.end - ci.start;
    const int T = 1 + ci.len / hop;
    if (ci.len <= 0) ci.status = 1;
    else if (T < kDeltaWidth) ci.status = 3;  // AFX_CLIP_TOO_SHORT
    else ci.n_frames = T;
  }
  return ci;
}

// ---- C: delta / delta2 (SavGol 9) -> moment sums for trimmed frames [0, T) from the slab's MFCC rows a0 + t.
// sM [256][n16+1] aliases the streaming buffers; acc [6][64] doubles.
__device__ __forceinline__ void mfcc_phase(const ClipArgs& a, const Slab& sl, int a0, int T, long long out_row0,
                                           float* sM, double* acc, uint64_t pol_slab) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int mstride = a.n16 + 1;
  for (int i = tid; i < kMoments * kMaxMfcc; i += (int)blockDim.x) acc[i] = 0.0;
  for (int f_lo = 0; f_lo < T; f_lo += kStatFrames) {
    const int f_hi = min(f_lo + kStatFrames, T);             // exclusive
    const int e_lo = min(max(f_lo, kStatHalo), T - 1 - kStatHalo) - kStatHalo;
    const int e_hi = min(max(f_hi - 1, kStatHalo), T - 1 - kStatHalo) + kStatHalo;   // inclusive
    const int ne = e_hi - e_lo + 1;                           // <= kStatRows
    __syncthreads();   // previous sub-chunk's readers are done with sM
    {
      const float* src = sl.mfcc + (size_t)(a0 + e_lo) * a.n16;
      for (int i = tid; i < ne * a.n16; i += (int)blockDim.x) {
        const int e = i / a.n16, c = i - e * a.n16;
        sM[e * mstride + c] = ld_slab(src + i, pol_slab);
      }
    }
    __syncthreads();
    // deltas + moments: warp w handles coefficients w, w+16, ...; lanes stride over the sub-chunk's frames
    for (int c = warp; c < a.n_mfcc; c += (int)(blockDim.x >> 5)) {
      double s1 = 0, s2 = 0, s3 = 0, s4 = 0, sd1 = 0, sd2 = 0;
      for (int f = f_lo + lane; f < f_hi; f += 32) {
        const int te = min(max(f, kStatHalo), T - 1 - kStatHalo) - e_lo;   // centre row of the 9-tap window
        const float* col = sM + c;
        // savgol(9, polyorder=deriv=1): sum k x[t+k] / 60 ; (9, 2, 2): sum c_k x[t+k] / 462  (double, cast to f32)
        const double xm4 = col[(te - 4) * mstride], xm3 = col[(te - 3) * mstride], xm2 = col[(te - 2) * mstride],
                     xm1 = col[(te - 1) * mstride], x0 = col[te * mstride], xp1 = col[(te + 1) * mstride],
                     xp2 = col[(te + 2) * mstride], xp3 = col[(te + 3) * mstride], xp4 = col[(te + 4) * mstride];
        const double d1 = (4.0 * (xp4 - xm4) + 3.0 * (xp3 - xm3) + 2.0 * (xp2 - xm2) + (xp1 - xm1)) / 60.0;
        const double d2 = (28.0 * (xp4 + xm4) + 7.0 * (xp3 + xm3) - 8.0 * (xp2 + xm2) - 17.0 * (xp1 + xm1) - 20.0 * x0) / 462.0;
        const float d1f = (float)d1, d2f = (float)d2;
        const float xf = col[(f - e_lo) * mstride];
        const double x = xf;
        s1 += x; s2 += x * x; s3 += x * x * x; s4 += (x * x) * (x * x);
        sd1 += (double)d1f; sd2 += (double)d2f;
        if (a.frames_out) {
          float* fr = a.frames_out + (size_t)(out_row0 + f) * a.frame_width;
          fr[c] = xf; fr[a.n_mfcc + c] = d1f; fr[2 * a.n_mfcc + c] = d2f;
        }
      }
      s1 = warp_sum(s1); s2 = warp_sum(s2); s3 = warp_sum(s3); s4 = warp_sum(s4);
      sd1 = warp_sum(sd1); sd2 = warp_sum(sd2);
      if (lane == 0) {   // coefficient c belongs to this warp alone: no race on acc[.][c]
        acc[0 * kMaxMfcc + c] += s1; acc[1 * kMaxMfcc + c] += s2; acc[2 * kMaxMfcc + c] += s3;
        acc[3 * kMaxMfcc + c] += s4; acc[4 * kMaxMfcc + c] += sd1; acc[5 * kMaxMfcc + c] += sd2;
      }
    }
  }
  __syncthreads();
}

// ---- the kernel
// STOCK: the mel tables carry the stock run lengths of this frame length (MelCnt<NFFT, true>), checked by the host
template <int NFFT, int HOP, bool STOCK = false>
__global__ void __launch_bounds__(kMaxClipThreads, 1) clip_kernel(ClipArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // layout: [16 mbarriers 128B][twiddle 8KB][window NFFT*4][acc 6*64 doubles][red 64 doubles]
  //         [sDct 32*(2*n16+4) floats][mel weights][raw tiles 16 x (HOP+NFFT+8) floats][planes 16 x 4224B]
  //         (phase C's sM aliases the raw tiles)
  constexpr int RAWW = Geo<NFFT, HOP>::RAWW;
  Smem sm;
  sm.mbar = reinterpret_cast<uint64_t*>(smem_raw);
  sm.tw = reinterpret_cast<float2*>(smem_raw + 128);
  constexpr int PTW_BYTES = NFFT == 2048 ? 4112 : 0;   // W_2048^k, k = 0..512 (+ pad)
  sm.ptw = reinterpret_cast<float2*>(smem_raw + 128 + 8192);
  sm.win = reinterpret_cast<float*>(smem_raw + 128 + 8192 + PTW_BYTES);
  double* acc = reinterpret_cast<double*>(sm.win + NFFT);
  double* red = acc + kMoments * kMaxMfcc;
  float* sDct = reinterpret_cast<float*>(red + 64);
  const int dstride = 2 * a.n16 + 4;   // per lane: DCT rows of band `lane`, then of band 63 - lane (+ pad: conflict-free float4 reads)
  float* sMel = sDct + 32 * dstride;
  sm.dct = sDct; sm.melw = sMel;
  sm.raw = sMel + a.mel_floats;
  sm.planes = sm.raw + (blockDim.x >> 5) * RAWW;
  float* sM = sm.raw;
  __shared__ int s_clip, s_bad, s_nredo;
  __shared__ int s_redo[kMaxRedo];

  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 1024; i += (int)blockDim.x) sm.tw[i] = a.tab.twiddle[i];
  for (int i = tid; i < NFFT; i += (int)blockDim.x) sm.win[i] = a.tab.window_half[i];
  if (NFFT == 2048) for (int i = tid; i <= 512; i += (int)blockDim.x) sm.ptw[i] = a.tab.post_twiddle[i];
  for (int i = tid; i < 64 * a.n16; i += (int)blockDim.x) {
    const int l = i / (2 * a.n16), rem = i - l * 2 * a.n16, hsel = rem / a.n16, k = rem - hsel * a.n16;
    sDct[l * dstride + rem] = a.dct_t[(hsel ? 63 - l : l) * a.n16 + k];
  }
  for (int i = tid; i < a.mel_floats; i += (int)blockDim.x) sMel[i] = a.tab.mel_w[i];
  if (tid < kMaxClipWarps) mbar_init(sm.mbar + tid, 1);
  if (tid == 0) fence_mbar_init();
  __syncthreads();

  // the slab's row pointers live in shared memory: one 64-bit load where a step needs one, instead of the pointer
  // arithmetic from the kernel parameters the register-starved step would otherwise redo every time
  __shared__ Slab s_slab;
  if (tid == 0) s_slab = carve_slab(a, blockIdx.x);
  __syncthreads();
  const Slab& sl = s_slab;
  WarpCtx wc;
  wc.raw = sm.raw + warp * RAWW;
  wc.plane = sm.planes + warp * kPlaneFloats;
  wc.mbar = sm.mbar + warp;
  wc.parity = 0;
  wc.pol_in = l2_policy_evict_first();
  wc.pol_slab = l2_policy_evict_last();
  const int hop = HOP;
  const int n = a.n_mfcc;

  for (;;) {
    __syncthreads();   // everyone is done with the previous clip (s_clip, aliased buffers)
    if (tid == 0) { s_clip = atomicAdd(a.counter, 1); s_bad = 0; }
    __syncthreads();
    const int clip = s_clip;
    if (clip >= a.n_clips) break;
    const int L0 = a.lengths[clip];
    const float* clip_audio = a.audio + a.offsets[clip];
    const long long out_row0 = a.frame_off[clip];
    const bool want_frames = (a.mask & 0x6u) != 0;   // MFCC or ENERGY

    ClipInfo ci;
    ci.start = 0; ci.end = 0; ci.len = 0; ci.n_frames = 0; ci.max_key = 0u; ci.pad0 = ci.pad1 = 0; ci.status = 1;
    int a0 = 0;          // slab row of trimmed frame 0
    int head_end = 0;
    int bad = 0;
    int n_redo = 0;
    // pass 0: S (speculative, absolute frame grid) or the whole pretrimmed clip; pass 1 / 2: head / tail fix-ups (F);
    // pass 3: R, redo with the top_db clamp the frames whose log-mel minimum lies below (clip max - top_db)
    float floor_db = -3.0e38f;
    for (int pass = 0; pass < 4; ++pass) {
      Region rg{0, 0, L0};
      rg.raw = a.raw;
      int fb = 0, fe = 0, shift = 0;
      bool own = false;
      if (pass == 0) {
        if (a.pretrimmed) {
          ci = a.info[clip];
          rg = Region{ci.start, ci.start, ci.end};
          rg.raw = a.raw;
          fe = (ci.status == 0 && want_frames) ? ci.n_frames : 0;
        } else {
          fe = L0 > 0 ? 1 + L0 / hop : 0;
          own = true;
        }
      } else {
        if (ci.status != 0 || !want_frames) break;
        rg = Region{ci.start, ci.start, ci.end};
        rg.raw = a.raw;
        shift = a0;                // 0 when pretrimmed
        if (pass == 1) {           // frames t with t*hop < n_fft/2 reach in front of the trimmed signal
          if (!a.pretrimmed && ci.start > 0) { head_end = min(ci.n_frames, (NFFT / 2 + hop - 1) / hop); fe = head_end; }
        } else if (pass == 2) {
          if (!a.pretrimmed && ci.end < L0) {       // frames t with t*hop + n_fft/2 > len reach behind it
            const int t_tail = (ci.len >= NFFT / 2) ? (ci.len - NFFT / 2) / hop + 1 : 0;
            fb = max(t_tail, head_end); fe = ci.n_frames;
          }
        } else {
          if (!(a.mask & 0x2u)) break;
          __syncthreads();         // every row's extrema are in the slab
          float mx = -3.0e38f, mn = 3.0e38f;
          for (int t = tid; t < ci.n_frames; t += (int)blockDim.x) {
            mx = fmaxf(mx, __ldcg(sl.rowmax + a0 + t)); mn = fminf(mn, __ldcg(sl.rowmin + a0 + t));
          }
          mx = block_max(mx, reinterpret_cast<float*>(red));
          mn = block_min(mn, reinterpret_cast<float*>(red));
          floor_db = mx - a.mfcc_top_db;   // power_to_db(top_db): clamp at the clip-global maximum - top_db
          if (mn >= floor_db) break;       // nothing is clamped: the speculative DCT rows are exact
          fe = ci.n_frames;
          // collect the steps whose unclamped row minimum lies below the floor (any order)
          if (tid == 0) s_nredo = 0;
          __syncthreads();
          constexpr int FPW = Geo<NFFT, HOP>::FPW;
          for (int f = FPW * tid; f < fe; f += FPW * (int)blockDim.x) {
            float m = 3.0e38f;
            for (int j = 0; j < FPW; ++j) if (f + j < fe) m = fminf(m, __ldcg(sl.rowmin + a0 + f + j));
            if (m < floor_db) { const int pos = atomicAdd(&s_nredo, 1); if (pos < kMaxRedo) s_redo[pos] = f; }
          }
          __syncthreads();
          n_redo = s_nredo;
        }
      }
      const bool listed = pass == 3 && n_redo <= kMaxRedo;   // else (very long clips) every warp filters its own steps
      stream_frames<NFFT, HOP, STOCK>(a, sm, sl, wc, clip_audio, L0, rg, fb, fe, shift, own, floor_db, pass == 3, bad,
                               listed ? s_redo : nullptr, n_redo);
      if (pass == 0 && !a.pretrimmed) {
        if (bad) s_bad = 1;
        __syncthreads();
        // ---- T
        const int nu = Geo<NFFT, HOP>::FPW * ((fe + Geo<NFFT, HOP>::FPW - 1) / Geo<NFFT, HOP>::FPW);   // hop-block sums written by the pass
        ci = trim_decide(sl.blk, nu, L0, hop, a.trim_top_db, s_bad, reinterpret_cast<float*>(red), reinterpret_cast<int*>(red + 32));
        a0 = ci.start / hop;
      }
    }
    __syncthreads();
    if (tid == 0) {
      a.info[clip] = ci;
      int* m = a.meta + (size_t)clip * 4;
      m[0] = ci.status; m[1] = ci.start; m[2] = ci.end; m[3] = ci.n_frames;
    }
    float* out = a.summaries + (size_t)clip * a.summary_width;
    for (int i = tid; i < a.summary_width; i += (int)blockDim.x) out[i] = 0.f;
    const int T = ci.n_frames;
    if (T <= 0) continue;   // status != OK: summary row stays zero (uniform branch)
    const double invT = 1.0 / (double)T;

    if (a.mask & 0x2u) {
      // ---- C
      mfcc_phase(a, sl, a0, T, out_row0, sM, acc, wc.pol_slab);
      if (tid < n) {
        const double s1 = acc[tid], s2 = acc[kMaxMfcc + tid], s3 = acc[2 * kMaxMfcc + tid], s4 = acc[3 * kMaxMfcc + tid];
        const double mean = s1 * invT;
        const double ex2 = s2 * invT, ex3 = s3 * invT, ex4 = s4 * invT;
        const double m2 = fmax(ex2 - mean * mean, 0.0);
        const double m3 = ex3 - 3.0 * mean * ex2 + 2.0 * mean * mean * mean;
        const double m4 = ex4 - 4.0 * mean * ex3 + 6.0 * mean * mean * ex2 - 3.0 * mean * mean * mean * mean;
        out[4 + tid] = (float)mean;
        out[4 + n + tid] = (float)sqrt(m2);
        out[4 + 2 * n + tid] = (float)(acc[4 * kMaxMfcc + tid] * invT);
        out[4 + 3 * n + tid] = (float)(acc[5 * kMaxMfcc + tid] * invT);
        out[4 + 4 * n + 3 + tid] = (m2 > 0.0) ? (float)(m3 / (m2 * sqrt(m2))) : 0.f;
        out[4 + 4 * n + 3 + n + tid] = (m2 > 0.0) ? (float)(m4 / (m2 * m2) - 3.0) : 0.f;
      }
    }
    if (a.mask & 0x4u) {   // RMS energy statistics (core/feature_extractor.py:171-173)
      double s1 = 0, s2 = 0, s3 = 0, s4 = 0;
      float mn = 3.0e38f, mx = -3.0e38f;
      for (int t = tid; t < T; t += (int)blockDim.x) {
        const float rf = __ldcg(sl.rms + a0 + t);
        const double r = rf;
        s1 += r; s2 += r * r; s3 += r * r * r; s4 += (r * r) * (r * r);
        mn = fminf(mn, rf); mx = fmaxf(mx, rf);
        if (a.frames_out) a.frames_out[(size_t)(out_row0 + t) * a.frame_width + 3 * n] = rf;
      }
      s1 = block_sum(s1, red); s2 = block_sum(s2, red); s3 = block_sum(s3, red); s4 = block_sum(s4, red);
      mn = block_min(mn, reinterpret_cast<float*>(red)); mx = block_max(mx, reinterpret_cast<float*>(red));
      if (tid == 0) {
        const double mean = s1 * invT;
        const double ex2 = s2 * invT, ex3 = s3 * invT, ex4 = s4 * invT;
        const double m2 = fmax(ex2 - mean * mean, 0.0);
        const double m3 = ex3 - 3.0 * mean * ex2 + 2.0 * mean * mean * mean;
        const double m4 = ex4 - 4.0 * mean * ex3 + 6.0 * mean * mean * ex2 - 3.0 * mean * mean * mean * mean;
        float* e = out + 4 + 4 * n;
        e[0] = (float)mean; e[1] = (float)sqrt(m2); e[2] = mx - mn;
        e[3 + 2 * n] = (m2 > 0.0) ? (float)(m3 / (m2 * sqrt(m2))) : 0.f;
        e[3 + 2 * n + 1] = (m2 > 0.0) ? (float)(m4 / (m2 * m2) - 3.0) : 0.f;
      }
    }
  }
}

}  // namespace afx
