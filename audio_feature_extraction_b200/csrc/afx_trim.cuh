// Phase A: pre-emphasis + librosa.effects.trim(top_db) decision, one CTA per clip.
//
// Replaces librosa.effects.preemphasis + librosa.effects.trim at
// audio_feature_extraction_toolkit/core/feature_extractor.py:69,72 (reference), i.e.
//   y_pre[0] = y[0] + (2 y[0] - y[1]);  y_pre[n] = y[n] + fl(-c * y[n-1])          (lfilter, float32)
//   r_i = sqrt(mean(y_pre[512 i - 1024 .. 512 i + 1024)^2)), zero padded, i = 0 .. L0/512
//   keep frames with 10 log10(max(1e-10, r_i^2)) - 10 log10(max(1e-10, max_i r_i^2)) > -top_db
//   start = 512 * first, end = min(L0, 512 * (last + 1))
// The clip is streamed from HBM once with 128-bit loads; the only state kept is one sum of squares
// per 512-sample block (shared memory), from which every 2048-sample RMS frame is four adds.
#pragma once
#include "afx_common.cuh"

namespace afx {

constexpr int kTrimThreads = 512;

struct TrimArgs {
  const float* audio;
  const long long* offsets;   // device, per clip
  const int* lengths;         // device, per clip
  ClipInfo* info;
  int clip_begin;
  int hop;                    // feature hop (for n_frames)
  float coef;                 // pre-emphasis
  float top_db;
  int raw;                    // input already pre-emphasised: no first-sample rule
  int min_frames;             // kDeltaWidth when the MFCC group is asked for (librosa.feature.delta needs 9 frames), else 1
};

__device__ __forceinline__ float preemph_sample(float x, float xm1, float coef) {
  // scipy.signal.lfilter with b=[1,-c], a=[1] in float32: fl(x + fl(-c * xm1)), no contraction
  return __fadd_rn(x, __fmul_rn(-coef, xm1));
}

__global__ void __launch_bounds__(kTrimThreads) trim_kernel(TrimArgs a) {
  extern __shared__ float blk[];  // [nb + 4], blk[j + 2] = sum of squares of block j; two zero blocks either side
  __shared__ float red_f[kTrimThreads / 32];
  __shared__ int red_i[2][kTrimThreads / 32];
  __shared__ int s_bad;

  const int clip = a.clip_begin + blockIdx.x;
  const int L0 = a.lengths[clip];
  const float* __restrict__ x = a.audio + a.offsets[clip];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nb = (L0 + kTrimHop - 1) / kTrimHop;

  if (tid == 0) s_bad = 0;
  for (int j = tid; j < nb + 4; j += kTrimThreads) blk[j] = 0.f;
  __syncthreads();

  float x0 = 0.f, x1 = 0.f;
  if (L0 > 0) x0 = x[0];
  if (L0 > 1) x1 = x[1];
  const float y_first = a.raw ? x0 : x0 + (2.f * x0 - x1);  // zi = 2*y[0]-y[1] handed to lfilter as state

  int bad = 0;
  for (int j = warp; j < nb; j += kTrimThreads / 32) {
    float acc = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int e = j * kTrimHop + q * 128 + lane * 4;
      if (e < L0) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(x + e));  // may over-read <= 3 padded floats
        const float xm1 = (e > 0) ? __ldg(x + e - 1) : 0.f;
        float y0 = (e == 0) ? y_first : preemph_sample(v.x, xm1, a.coef);
        float y1 = preemph_sample(v.y, v.x, a.coef);
        float y2 = preemph_sample(v.z, v.y, a.coef);
        float y3 = preemph_sample(v.w, v.z, a.coef);
        const int rem = L0 - e;  // >= 1
        bad |= !isfinite(v.x);
        if (rem > 1) bad |= !isfinite(v.y); else y1 = 0.f;
        if (rem > 2) bad |= !isfinite(v.z); else y2 = 0.f;
        if (rem > 3) bad |= !isfinite(v.w); else y3 = 0.f;
        acc = fmaf(y0, y0, acc);
        acc = fmaf(y1, y1, acc);
        acc = fmaf(y2, y2, acc);
        acc = fmaf(y3, y3, acc);
      }
    }
    acc = warp_sum(acc);
    if (lane == 0) blk[j + 2] = acc;
  }
  if (bad) s_bad = 1;
  __syncthreads();

  // RMS frames and their maximum
  const int n_rms = 1 + L0 / kTrimHop;
  float mx = 0.f;
  for (int i = tid; i < n_rms; i += kTrimThreads) {
    // frame i covers blocks i-2 .. i+1  -> blk[i] .. blk[i+3]; blocks >= nb are zero
    const float s = (blk[i] + blk[i + 1]) + (blk[i + 2] + blk[i + 3]);  // i + 3 <= nb + 3 always
    mx = fmaxf(mx, sqrtf(s * (1.0f / kTrimFrame)));
  }
  mx = warp_max(mx);
  if (lane == 0) red_f[warp] = mx;
  __syncthreads();
  mx = red_f[0];
#pragma unroll
  for (int w = 1; w < kTrimThreads / 32; ++w) mx = fmaxf(mx, red_f[w]);

  const float ref_db = 10.0f * log10f(fmaxf(1e-10f, mx * mx));
  int first = 0x7fffffff, last = -1;
  for (int i = tid; i < n_rms; i += kTrimThreads) {
    const float s = (blk[i] + blk[i + 1]) + (blk[i + 2] + blk[i + 3]);  // i + 3 <= nb + 3 always
    const float r = sqrtf(s * (1.0f / kTrimFrame));
    float db = 10.0f * log10f(fmaxf(1e-10f, r * r));
    db -= ref_db;
    if (db > -a.top_db) { first = min(first, i); last = max(last, i); }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
    last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
  }
  if (lane == 0) { red_i[0][warp] = first; red_i[1][warp] = last; }
  __syncthreads();
  if (tid == 0) {
    for (int w = 1; w < kTrimThreads / 32; ++w) { first = min(first, red_i[0][w]); last = max(last, red_i[1][w]); }
    ClipInfo ci;
    ci.start = 0; ci.end = 0; ci.len = 0; ci.n_frames = 0; ci.max_key = 0u; ci.pad0 = ci.pad1 = 0;
    ci.status = 0;
    if (L0 <= 0) {
      ci.status = 1;                     // AFX_CLIP_EMPTY
    } else if (s_bad) {
      ci.status = 2;                     // AFX_CLIP_NONFINITE
    } else if (last < 0) {
      ci.status = 1;                     // all frames silent -> empty signal
    } else {
      ci.start = first * kTrimHop;
      ci.end = min(L0, (last + 1) * kTrimHop);
      ci.len = ci.end - ci.start;
      const int T = 1 + ci.len / a.hop;
      if (ci.len <= 0) ci.status = 1;
      else if (T < a.min_frames) ci.status = 3;  // AFX_CLIP_TOO_SHORT
      else ci.n_frames = T;
    }
    a.info[clip] = ci;
  }
}

}  // namespace afx
