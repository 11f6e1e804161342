// Phase C: top_db clamp -> DCT-II (ortho) -> delta / delta-delta -> per-chunk moment partials.
// Phase D: per-clip statistics -> summary row / meta row.
//
// Replaces the second half of librosa.feature.mfcc (power_to_db's `max(D, D.max() - 80)` and
// scipy.fftpack.dct(norm='ortho')[:n_mfcc]), the two librosa.feature.delta calls and the numpy
// reductions at audio_feature_extraction_toolkit/core/feature_extractor.py:127-150,171-178,97-113.
#pragma once
#include "afx_common.cuh"

namespace afx {

constexpr int kStatThreads = 256;
constexpr int kStatRows = kStatFrames + 2 * kStatHalo;   // 64 frames incl. halo
constexpr int kDStride = kNMels + 1;                      // padded log-mel row in shared memory
constexpr int kMomentsPerCoef = 6;                        // sum x, x^2, x^3, x^4, sum delta, sum delta2

struct StatArgs {
  const ClipInfo* info;
  const int* chunk_off;          // device prefix [n_clips+1] of ceil(T_ub / kStatFrames)
  const long long* frame_off;    // device prefix [n_clips+1] of T_ub
  int clip_begin, clip_end;
  long long frame_base;
  const float* logmel;           // [wave frames][128]
  const float* dct_t;            // [128][n16] transposed DCT rows, n16 = n_mfcc rounded up to 4, zero padded
  double* partials;              // [wave chunks][6][n_mfcc]
  float* frames_out;             // nullable: global [sum T_ub][FW]
  int n_mfcc, n16, frame_width;
  float top_db;
};

__global__ void __launch_bounds__(kStatThreads) mfcc_stats_kernel(StatArgs a) {
  extern __shared__ __align__(16) float sm[];
  float* sD = sm;                                  // [64][129]
  float* sDct = sD + kStatRows * kDStride;         // [128][n16]
  float* sM = sDct + kNMels * a.n16;               // [64][n16 + 1]
  const int mstride = a.n16 + 1;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // decode (clip, chunk)
  const int base = a.chunk_off[a.clip_begin];
  const int task = blockIdx.x;
  int lo = a.clip_begin, hi = a.clip_end;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (a.chunk_off[mid] - base <= task) lo = mid; else hi = mid;
  }
  const int clip = lo;
  const int chunk = task - (a.chunk_off[clip] - base);
  const ClipInfo ci = a.info[clip];
  const int T = ci.n_frames;
  const int f_lo = chunk * kStatFrames;
  double* part = a.partials + (size_t)task * kMomentsPerCoef * a.n_mfcc;
  if (f_lo >= T) return;   // chunk beyond the trimmed clip (T_ub >= T); finalize never reads it
  const int f_hi = min(f_lo + kStatFrames, T);             // exclusive
  const int e_lo = min(max(f_lo, kStatHalo), T - 1 - kStatHalo) - kStatHalo;
  const int e_hi = min(max(f_hi - 1, kStatHalo), T - 1 - kStatHalo) + kStatHalo;   // inclusive
  const int ne = e_hi - e_lo + 1;                           // <= 64
  const float floor_db = key_to_float(ci.max_key) - a.top_db;
  const long long row0 = a.frame_off[clip] - a.frame_base;

  for (int i = tid; i < kNMels * a.n16; i += kStatThreads) sDct[i] = a.dct_t[i];
  // log-mel tile with the power_to_db clamp applied
  for (int e = warp; e < ne; e += kStatThreads / 32) {
    const float* src = a.logmel + (row0 + e_lo + e) * kNMels;
#pragma unroll
    for (int q = 0; q < 4; ++q) sD[e * kDStride + q * 32 + lane] = fmaxf(src[q * 32 + lane], floor_db);
  }
  __syncthreads();

  // DCT: thread (frame e = tid & 63, coefficient group g = tid >> 6) computes 4 coefficients at a time
  {
    const int e = tid & 63;
    for (int c0 = (tid >> 6) * 4; c0 < a.n16; c0 += 16) {
      float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
      if (e < ne) {
        const float* drow = sD + e * kDStride;
#pragma unroll 8
        for (int m = 0; m < kNMels; ++m) {
          const float d = drow[m];
          const float4 w = *reinterpret_cast<const float4*>(sDct + m * a.n16 + c0);
          acc0 = fmaf(d, w.x, acc0);
          acc1 = fmaf(d, w.y, acc1);
          acc2 = fmaf(d, w.z, acc2);
          acc3 = fmaf(d, w.w, acc3);
        }
      }
      float* mrow = sM + e * mstride + c0;
      mrow[0] = acc0; mrow[1] = acc1; mrow[2] = acc2; mrow[3] = acc3;
    }
  }
  __syncthreads();

  // deltas + moments: warp w handles coefficients w, w+8, ...; lanes stride over the chunk's frames
  for (int c = warp; c < a.n_mfcc; c += kStatThreads / 32) {
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
        float* fr = a.frames_out + (size_t)(a.frame_off[clip] + f) * a.frame_width;
        fr[c] = xf; fr[a.n_mfcc + c] = d1f; fr[2 * a.n_mfcc + c] = d2f;
      }
    }
    s1 = warp_sum(s1); s2 = warp_sum(s2); s3 = warp_sum(s3); s4 = warp_sum(s4);
    sd1 = warp_sum(sd1); sd2 = warp_sum(sd2);
    if (lane == 0) {
      part[0 * a.n_mfcc + c] = s1; part[1 * a.n_mfcc + c] = s2; part[2 * a.n_mfcc + c] = s3;
      part[3 * a.n_mfcc + c] = s4; part[4 * a.n_mfcc + c] = sd1; part[5 * a.n_mfcc + c] = sd2;
    }
  }
}

// ------------------------------------------------------------------------------------------------
constexpr int kFinalThreads = 128;

struct FinalArgs {
  const ClipInfo* info;
  const int* chunk_off;          // stat chunks prefix
  const long long* frame_off;
  int clip_begin;
  long long frame_base;
  const double* partials;        // wave-relative
  const float* rms;              // wave-relative
  const float* f0;               // wave-relative, nullable (NaN = unvoiced)
  const float* voiced_prob;      // wave-relative, nullable
  float* summaries;              // [n_clips][SS]
  int* meta;                     // [n_clips][4]
  float* frames_out;             // nullable
  int n_mfcc, summary_width, frame_width;
  unsigned mask;                 // AFX_FEAT_*
};

__device__ __forceinline__ double block_sum(double v, double* sh) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0;
  for (int w = 0; w < kFinalThreads / 32; ++w) t += sh[w];
  return t;
}

__global__ void __launch_bounds__(kFinalThreads) finalize_kernel(FinalArgs a) {
  __shared__ double sh[kFinalThreads / 32];
  __shared__ float shf[2][kFinalThreads / 32];
  const int clip = a.clip_begin + blockIdx.x;
  const int tid = threadIdx.x;
  const ClipInfo ci = a.info[clip];
  float* out = a.summaries + (size_t)clip * a.summary_width;
  const int n = a.n_mfcc;
  if (tid == 0) {
    int* m = a.meta + (size_t)clip * 4;
    m[0] = ci.status; m[1] = ci.start; m[2] = ci.end; m[3] = ci.n_frames;
  }
  for (int i = tid; i < a.summary_width; i += kFinalThreads) out[i] = 0.f;
  const int T = ci.n_frames;
  if (T <= 0) return;   // status != OK: summary row stays zero
  __syncthreads();
  const double invT = 1.0 / (double)T;
  const long long row0 = a.frame_off[clip] - a.frame_base;

  if (a.mask & 0x2u) {   // MFCC statistics: combine the chunk partials in chunk order (deterministic)
    const int cb = a.chunk_off[clip] - a.chunk_off[a.clip_begin];
    const int nchunks = (T + kStatFrames - 1) / kStatFrames;
    if (tid < n) {
      double s1 = 0, s2 = 0, s3 = 0, s4 = 0, sd1 = 0, sd2 = 0;
      for (int k = 0; k < nchunks; ++k) {
        const double* p = a.partials + (size_t)(cb + k) * kMomentsPerCoef * n;
        s1 += p[tid]; s2 += p[n + tid]; s3 += p[2 * n + tid]; s4 += p[3 * n + tid];
        sd1 += p[4 * n + tid]; sd2 += p[5 * n + tid];
      }
      const double mean = s1 * invT;
      const double m2 = fmax(s2 * invT - mean * mean, 0.0);
      const double ex3 = s3 * invT, ex4 = s4 * invT, ex2 = s2 * invT;
      const double m3 = ex3 - 3.0 * mean * ex2 + 2.0 * mean * mean * mean;
      const double m4 = ex4 - 4.0 * mean * ex3 + 6.0 * mean * mean * ex2 - 3.0 * mean * mean * mean * mean;
      out[4 + tid] = (float)mean;
      out[4 + n + tid] = (float)sqrt(m2);
      out[4 + 2 * n + tid] = (float)(sd1 * invT);
      out[4 + 3 * n + tid] = (float)(sd2 * invT);
      out[4 + 4 * n + 3 + tid] = (m2 > 0.0) ? (float)(m3 / (m2 * sqrt(m2))) : 0.f;
      out[4 + 4 * n + 3 + n + tid] = (m2 > 0.0) ? (float)(m4 / (m2 * m2) - 3.0) : 0.f;
    }
  }
  if (a.mask & 0x4u) {   // RMS energy statistics
    double s1 = 0, s2 = 0, s3 = 0, s4 = 0;
    float mn = 3.0e38f, mx = -3.0e38f;
    for (int t = tid; t < T; t += kFinalThreads) {
      const float rf = a.rms[row0 + t];
      const double r = rf;
      s1 += r; s2 += r * r; s3 += r * r * r; s4 += (r * r) * (r * r);
      mn = fminf(mn, rf); mx = fmaxf(mx, rf);
      if (a.frames_out) a.frames_out[(size_t)(a.frame_off[clip] + t) * a.frame_width + 3 * n] = rf;
    }
    s1 = block_sum(s1, sh); s2 = block_sum(s2, sh); s3 = block_sum(s3, sh); s4 = block_sum(s4, sh);
    mn = warp_min(mn); mx = warp_max(mx);
    __syncthreads();
    if ((tid & 31) == 0) { shf[0][tid >> 5] = mn; shf[1][tid >> 5] = mx; }
    __syncthreads();
    if (tid == 0) {
      for (int w = 1; w < kFinalThreads / 32; ++w) { mn = fminf(mn, shf[0][w]); mx = fmaxf(mx, shf[1][w]); }
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
  if (a.mask & 0x1u) {   // F0 statistics over voiced frames (core/feature_extractor.py:97-113)
    double s1 = 0, s2 = 0, cnt = 0;
    if (a.f0) {
      for (int t = tid; t < T; t += kFinalThreads) {
        const float f = a.f0[row0 + t];
        if (f == f) { const double d = f; s1 += d; s2 += d * d; cnt += 1.0; }
        if (a.frames_out) {
          float* fr = a.frames_out + (size_t)(a.frame_off[clip] + t) * a.frame_width + 3 * n;
          fr[1] = f;
          fr[2] = a.voiced_prob ? a.voiced_prob[row0 + t] : 0.f;
        }
      }
    }
    s1 = block_sum(s1, sh); s2 = block_sum(s2, sh); cnt = block_sum(cnt, sh);
    if (tid == 0) {
      if (cnt > 0.0) {
        const double mean = s1 / cnt;
        const double var = fmax(s2 / cnt - mean * mean, 0.0);
        const double missing = ((double)T - cnt) * invT;
        out[0] = (float)mean; out[1] = (float)sqrt(var); out[2] = (float)missing; out[3] = (float)(1.0 - missing);
      } else {
        out[0] = 0.f; out[1] = 0.f; out[2] = 1.f; out[3] = 0.f;
      }
    }
  }
}

}  // namespace afx
