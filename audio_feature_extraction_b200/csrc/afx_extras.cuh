// Framewise extras for the .npz feature set of the reference's experiment pipeline
// (04_feature_extraction_experiment/feature_extraction.py:191-292, stored :340-353): zero-crossing rate per frame
// (librosa.feature.zero_crossing_rate: edge padding, |x| <= 1e-10 counts as +0, first entry of a frame never a
// crossing) and the SavGol(9) deltas of the one-row features zcr, rms and nan_to_num(f0).
// Runs after afx_extract on the same batch: it reads the trim decision (meta), the frame rows (rms, f0) and
// recomputes the pre-emphasised samples on the fly.
#pragma once
#include "afx_common.cuh"

namespace afx {

constexpr int kExtrasThreads = 256;
constexpr int kExtrasWidth = 4;   // zcr, delta(zcr), delta(rms), delta(nan_to_num(f0))

struct ExtrasArgs {
  const float* audio;
  const long long* offsets;     // device [n_clips]
  const int* lengths;           // device [n_clips]
  const long long* frame_off;   // device [n_clips + 1]
  const int* meta;              // device [n_clips][4]: status, trim_start, trim_end, n_frames
  const float* frames;          // [rows][frame_width]
  float* extras;                // [rows][kExtrasWidth]
  int frame_width, n_mfcc, n_fft, hop;
  float coef;
};

__global__ void __launch_bounds__(kExtrasThreads) frame_extras_kernel(ExtrasArgs a) {
  const int clip = blockIdx.x;
  const int* m = a.meta + 4 * clip;
  const int T = m[3];
  if (m[0] != 0 || T <= 0) return;
  const int start = m[1], L = m[2] - m[1], L0 = a.lengths[clip];
  const float* __restrict__ x = a.audio + a.offsets[clip];
  const long long row0 = a.frame_off[clip];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // sign bit of the pre-emphasised, thresholded sample n of the trimmed signal (index clamped: edge padding)
  auto neg = [&](int n) -> bool {
    n = min(max(n, 0), L - 1);
    const int g = start + n;
    const float xv = __ldg(x + g);
    float y;
    if (g == 0) y = xv + (2.f * xv - (L0 > 1 ? __ldg(x + 1) : 0.f));
    else y = __fadd_rn(xv, __fmul_rn(-a.coef, __ldg(x + g - 1)));
    if (fabsf(y) <= 1e-10f) y = 0.f;
    return signbit(y);
  };
  for (int t = warp; t < T; t += kExtrasThreads / 32) {
    const int s = t * a.hop - a.n_fft / 2;
    int cnt = 0;
    for (int i = 1 + lane; i < a.n_fft; i += 32) cnt += (neg(s + i) != neg(s + i - 1)) ? 1 : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) a.extras[(size_t)(row0 + t) * kExtrasWidth] = (float)cnt / (float)a.n_fft;
  }
  __threadfence_block();
  __syncthreads();
  // librosa.feature.delta = savgol(9, polyorder 1, deriv 1, mode 'interp'): slope of the least-squares line through the
  // 9 frames around t, the window held inside [0, T) at the edges
  const int c_rms = 3 * a.n_mfcc, c_f0 = 3 * a.n_mfcc + 1;
  for (int t = tid; t < T; t += kExtrasThreads) {
    const int te = min(max(t, 4), T - 5);
    double dz = 0.0, dr = 0.0, df = 0.0;
#pragma unroll
    for (int k = 1; k <= 4; ++k) {
      const size_t rp = (size_t)(row0 + te + k), rm = (size_t)(row0 + te - k);
      dz += k * ((double)a.extras[rp * kExtrasWidth] - (double)a.extras[rm * kExtrasWidth]);
      dr += k * ((double)a.frames[rp * a.frame_width + c_rms] - (double)a.frames[rm * a.frame_width + c_rms]);
      float fp = a.frames[rp * a.frame_width + c_f0], fm = a.frames[rm * a.frame_width + c_f0];
      if (isnan(fp)) fp = 0.f;
      if (isnan(fm)) fm = 0.f;
      df += k * ((double)fp - (double)fm);
    }
    float* e = a.extras + (size_t)(row0 + t) * kExtrasWidth;
    e[1] = (float)(dz / 60.0); e[2] = (float)(dr / 60.0); e[3] = (float)(df / 60.0);
  }
}

}  // namespace afx
