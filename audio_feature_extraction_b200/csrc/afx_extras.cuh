// Framewise extras for the .npz feature set of the reference's experiment pipeline
// (04_feature_extraction_experiment/feature_extraction.py:191-292, stored :340-353): zero-crossing rate per frame
// (librosa.feature.zero_crossing_rate: edge padding, |x| <= 1e-10 counts as +0, first entry of a frame never a
// crossing), the SavGol(9) deltas of the one-row features zcr, rms and nan_to_num(f0), and the short-time energy
// sum(frame^2) over the UNPADDED frames of librosa.util.frame (README.md:33; 04_feature_extraction_experiment/
// 'audio_feature_extraction 2'/audio_feature_extraction/feature_extractor.py:234-238): unpadded frame t is the centred
// frame t + n_fft / (2 hop), so its energy is n_fft * rms^2 of that row.
// Runs after afx_extract on the same batch: it reads the trim decision (meta), the frame rows (rms, f0) and
// recomputes the pre-emphasised samples on the fly.
#pragma once
#include "afx_common.cuh"

namespace afx {

constexpr int kExtrasThreads = 256;
constexpr int kExtrasMaxBlocks = 2560;    // hop blocks whose crossing counts fit the shared-memory tables (2 ints each): 20 KB,
                                          // 8 CTAs per SM; 29 s at hop 256 / 22.05 kHz, longer clips take the per-frame count
constexpr int kExtrasWidth = 5;   // zcr, delta(zcr), delta(rms), delta(nan_to_num(f0)), short-time energy

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
  // Zero crossings of frame t are the crossings of the sample pairs (i - 1, i), i = s + 1 .. s + n_fft - 1 (s = t hop -
  // n_fft / 2; edge padding adds none).  Counted once per hop block - cin[b]: pairs inside block b, cb[b]: the pair
  // across its lower border - a frame is the sum of its n_fft / hop blocks and of the borders between them: every sample
  // is read once (one ballot per 32 samples) instead of once per frame it belongs to.
  extern __shared__ int ex_cnt[];
  const int nblk = (L + a.hop - 1) / a.hop;
  if (nblk <= kExtrasMaxBlocks && (a.hop & 31) == 0) {
    int* cin = ex_cnt;
    int* cb = ex_cnt + nblk + 1;
    for (int b = tid; b <= nblk; b += kExtrasThreads) { cin[b] = 0; cb[b] = 0; }
    __syncthreads();
    const int nwords = (L + 31) >> 5;
    constexpr int NW = kExtrasThreads / 32, UN = 4;             // 4 words per warp in flight: the loads overlap
    for (int w0 = warp; w0 < nwords; w0 += NW * UN) {
      unsigned bits[UN], valid[UN]; bool below[UN];
#pragma unroll
      for (int u = 0; u < UN; ++u) {
        const int w = w0 + u * NW;
        const int i = 32 * w + lane;
        const bool ok = w < nwords && i < L;
        bits[u] = __ballot_sync(0xffffffffu, ok && neg(i));
        valid[u] = __ballot_sync(0xffffffffu, ok);
        below[u] = (w > 0 && w < nwords) ? neg(32 * w - 1) : false;
      }
#pragma unroll
      for (int u = 0; u < UN; ++u) {
        const int w = w0 + u * NW;
        if (w >= nwords) break;
        unsigned prev = bits[u] << 1;                             // sign of sample i - 1 at bit position i
        prev |= w > 0 ? (below[u] ? 1u : 0u) : (bits[u] & 1u);    // sample 0 has no pair below it
        const unsigned cross = (bits[u] ^ prev) & valid[u];
        if (lane == 0) {
          const int b = (32 * w) / a.hop;
          const bool first = (32 * w) % a.hop == 0;               // bit 0 of a block's first word is its border pair
          const int inner = __popc(first ? (cross & ~1u) : cross);
          if (inner) atomicAdd(&cin[b], inner);
          if (first && (cross & 1u)) cb[b] = 1;
        }
      }
    }
    __syncthreads();
    const int K = a.n_fft / a.hop;
    for (int t = tid; t < T; t += kExtrasThreads) {
      const int b0 = t - K / 2;
      int cnt = 0;
      for (int k = 0; k < K; ++k) {
        const int b = b0 + k;
        if (b >= 0 && b < nblk) cnt += cin[b] + (k > 0 ? cb[b] : 0);
      }
      a.extras[(size_t)(row0 + t) * kExtrasWidth] = (float)cnt / (float)a.n_fft;
    }
  } else {
  for (int t = warp; t < T; t += kExtrasThreads / 32) {
    const int s = t * a.hop - a.n_fft / 2;
    int cnt = 0;
    for (int i = 1 + lane; i < a.n_fft; i += 32) cnt += (neg(s + i) != neg(s + i - 1)) ? 1 : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) a.extras[(size_t)(row0 + t) * kExtrasWidth] = (float)cnt / (float)a.n_fft;
  }
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
    float ste = 0.f;
    if (L >= a.n_fft && t <= (L - a.n_fft) / a.hop) {
      const float r = a.frames[(size_t)(row0 + t + a.n_fft / (2 * a.hop)) * a.frame_width + c_rms];
      ste = (float)a.n_fft * r * r;
    }
    e[4] = ste;
  }
}

}  // namespace afx

// =====================================================================================================
// F0 post-processing the reference advertises (README.md:12-15: median smoothing, range constraint, outlier handling;
// concrete recipe 04_feature_extraction_experiment/feature_extractor.py:283-301): on the VOICED frames of the pYIN
// track, scipy.signal.medfilt(kernel 5, zero padded) -> scipy.signal.savgol_filter(11, 3, mode='interp') when at least
// 11 voiced frames -> clip to [f0_min, f0_max] -> mean, std (= the recipe's global RMSE), local RMSE
// sqrt(mean(diff^2)) and stability std / (mean + 1e-8).  One CTA per clip; the voiced track sits in shared memory.
namespace afx {

constexpr int kF0PostThreads = 256;

struct F0PostArgs {
  const long long* frame_off;   // device [n_clips + 1]
  const int* meta;              // [n_clips][4]
  const float* frames;          // [rows][frame_width]; column 3 n_mfcc + 1 = f0 (NaN when unvoiced)
  float* track;                 // [rows] smoothed f0 at voiced frames (other rows untouched)
  float* stats;                 // [n_clips][4] mean, std, local_rmse, stability
  int frame_width, n_mfcc, max_frames;
  double fmin, fmax;
  double hat[6][11];            // rows 0..4: edge fit (cubic through the first 11 points, evaluated at 0..4); row 5: interior
};

__device__ __forceinline__ void cswap(double& x, double& y) { const double lo = fmin(x, y), hi = fmax(x, y); x = lo; y = hi; }

__global__ void __launch_bounds__(kF0PostThreads) f0_post_kernel(F0PostArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* A = reinterpret_cast<double*>(smem_raw);          // [max_frames]
  double* B = A + a.max_frames;                              // [max_frames]
  int* P = reinterpret_cast<int*>(B + a.max_frames);         // [max_frames] frame index of each voiced entry
  __shared__ int s_wcnt[kF0PostThreads / 32];
  __shared__ int s_base;
  __shared__ double s_red[3][kF0PostThreads / 32];
  const int clip = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int* m = a.meta + 4 * clip;
  const int T = m[3];
  float* st = a.stats + 4 * clip;
  if (m[0] != 0 || T <= 0) { if (tid < 4) st[tid] = 0.f; return; }
  const long long row0 = a.frame_off[clip];
  const int c_f0 = 3 * a.n_mfcc + 1;
  // ---- compact the voiced frames, order kept
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int t0 = 0; t0 < T; t0 += kF0PostThreads) {
    const int t = t0 + tid;
    float f = __int_as_float(0x7fc00000);
    if (t < T) f = a.frames[(size_t)(row0 + t) * a.frame_width + c_f0];
    const bool v = !isnan(f);
    const unsigned bal = __ballot_sync(0xffffffffu, v);
    if (lane == 0) s_wcnt[warp] = __popc(bal);
    __syncthreads();
    int pos = s_base + __popc(bal & ((1u << lane) - 1u));
    for (int w = 0; w < warp; ++w) pos += s_wcnt[w];
    if (v) { A[pos] = (double)f; P[pos] = t; }
    __syncthreads();
    if (tid == 0) { int s = 0; for (int w = 0; w < kF0PostThreads / 32; ++w) s += s_wcnt[w]; s_base += s; }
    __syncthreads();
  }
  const int nv = s_base;
  if (nv == 0) {
    if (tid == 0) { st[0] = 0.f; st[1] = st[2] = st[3] = __int_as_float(0x7f800000); }   // the recipe's np.inf branch
    return;
  }
  // ---- medfilt(kernel min(5, len)), zero padded: 5 from 5 voiced frames on, 3 for exactly 3, identity for 1; scipy
  // rejects the even kernels of 2 and 4 voiced frames (the recipe's except branch), which pass through here
  double* cur = A; double* nxt = B;
  if (nv >= 5) {
    for (int i = tid; i < nv; i += kF0PostThreads) {
      double v0 = i >= 2 ? cur[i - 2] : 0.0, v1 = i >= 1 ? cur[i - 1] : 0.0, v2 = cur[i];
      double v3 = i + 1 < nv ? cur[i + 1] : 0.0, v4 = i + 2 < nv ? cur[i + 2] : 0.0;
      cswap(v0, v1); cswap(v3, v4); cswap(v0, v3); cswap(v1, v4); cswap(v1, v2); cswap(v2, v3); cswap(v1, v2);   // median in v2
      nxt[i] = v2;
    }
    __syncthreads();
    double* tmp = cur; cur = nxt; nxt = tmp;
  } else if (nv == 3) {
    if (tid < 3) {
      const double lo = tid >= 1 ? cur[tid - 1] : 0.0, mid = cur[tid], hi = tid + 1 < nv ? cur[tid + 1] : 0.0;
      nxt[tid] = fmax(fmin(lo, mid), fmin(fmax(lo, mid), hi));
    }
    __syncthreads();
    double* tmp = cur; cur = nxt; nxt = tmp;
  }
  // ---- savgol_filter(11, 3, mode='interp')
  if (nv >= 11) {
    for (int i = tid; i < nv; i += kF0PostThreads) {
      double acc = 0.0;
      if (i < 5) { for (int j = 0; j < 11; ++j) acc += a.hat[i][j] * cur[j]; }
      else if (i >= nv - 5) { const int r = nv - 1 - i; for (int j = 0; j < 11; ++j) acc += a.hat[r][j] * cur[nv - 1 - j]; }
      else { for (int j = 0; j < 11; ++j) acc += a.hat[5][j] * cur[i - 5 + j]; }
      nxt[i] = acc;
    }
    __syncthreads();
    double* tmp = cur; cur = nxt; nxt = tmp;
  }
  // ---- range constraint, output track, statistics
  double s1 = 0.0, s2 = 0.0, sd = 0.0;
  for (int i = tid; i < nv; i += kF0PostThreads) {
    const double v = fmin(fmax(cur[i], a.fmin), a.fmax);
    nxt[i] = v;
    if (a.track) a.track[row0 + P[i]] = (float)v;
    s1 += v;
  }
  __syncthreads();   // nxt holds the clipped track
  s1 = warp_sum(s1);
  if (lane == 0) s_red[0][warp] = s1;
  __syncthreads();
  double mean = 0.0;
  for (int w = 0; w < kF0PostThreads / 32; ++w) mean += s_red[0][w];
  mean /= (double)nv;
  for (int i = tid; i < nv; i += kF0PostThreads) {
    const double d = nxt[i] - mean;
    s2 += d * d;
    if (i + 1 < nv) { const double e = nxt[i + 1] - nxt[i]; sd += e * e; }
  }
  s2 = warp_sum(s2); sd = warp_sum(sd);
  if (lane == 0) { s_red[1][warp] = s2; s_red[2][warp] = sd; }
  __syncthreads();
  if (tid == 0) {
    double v2 = 0.0, d2 = 0.0;
    for (int w = 0; w < kF0PostThreads / 32; ++w) { v2 += s_red[1][w]; d2 += s_red[2][w]; }
    const double sdev = sqrt(v2 / (double)nv);
    st[0] = (float)mean; st[1] = (float)sdev;
    st[2] = nv >= 2 ? (float)sqrt(d2 / (double)(nv - 1)) : __int_as_float(0x7fc00000);   // np.mean of an empty diff is NaN
    st[3] = (float)(sdev / (mean + 1e-8));
  }
}

}  // namespace afx
