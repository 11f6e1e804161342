// Sample-rate conversion on the device for files whose native rate differs from `sr`: the resampling half of
// librosa.load(path, sr=self.sr) (audio_feature_extraction_toolkit/core/feature_extractor.py:52, reference).
//
// librosa's default res_type 'soxr_hq' is the soxr C library, which is not in /root/reference (and not in this image),
// so its exact coefficients cannot be restated.  This is librosa's OTHER built-in rational resampler,
// res_type='polyphase' = scipy.signal.resample_poly(y, sr / g, native / g) with scipy's defaults: a Kaiser(beta 5)
// windowed-sinc low-pass of 2 * 10 * max(up, down) + 1 taps at cutoff 1 / max(up, down), DC gain `up`, output n centred
// on input position n * down / up, length ceil(n_in * up / down), zero extension at both ends, float32 arithmetic.
// scipy itself is the oracle for it (tests/test_gpu_parity.py).
//
// The filter is stored by phase: taps[p][j] = h_padded[p + up * j], so output m = (q * up + p) / down reads
// x[q], x[q - 1], ... x[q - J + 1] against one contiguous row of J coefficients.
#pragma once
#include <cmath>
#include <cstdint>
#include <numeric>
#include <vector>

namespace afx {

constexpr int kResampleThreads = 256;
constexpr int kResamplePerThread = 4;
constexpr int kResampleTilesPerCta = 8;

struct ResampleArgs {
  const float* in;
  float* out;
  const long long* in_off;    // device [n]
  const long long* out_off;   // device [n]
  const int* in_len;          // device [n]
  const int* out_len;         // device [n]
  const float* taps;          // device [up][J]
  int up, down, J, n_pre_remove;
  int clip0;
};

// One thread per output sample; products and sums rounded to float32 one by one, oldest input sample first, which is
// the order of scipy's upfirdn inner loop.  The phase table sits in shared memory (a row per phase, odd row length:
// the lanes' rows fall on different banks; from global memory every lane's row would be its own cache line and a warp
// load 32 wavefronts); a CTA loads it once for kResampleTilesPerCta consecutive tiles of 1024 outputs of clip blockIdx.y
// (CTAs in launch order walk one clip after the other: neighbours in time are neighbours in memory).
template <bool SMEM_TAPS>
__global__ void __launch_bounds__(kResampleThreads) resample_kernel(ResampleArgs a, int n_clips) {
  extern __shared__ float rs_taps[];
  const int Jp = a.J | 1;                                   // padded row length in shared memory
  if (SMEM_TAPS) {
    for (int e = threadIdx.x; e < a.up * a.J; e += kResampleThreads) {
      const int p = e / a.J, j = e - p * a.J;
      rs_taps[p * Jp + j] = a.taps[e];
    }
    __syncthreads();
  }
  const int tile = kResampleThreads * kResamplePerThread;
  {
    const int clip = a.clip0 + blockIdx.y;
    if (clip >= n_clips) return;
    const int n_in = a.in_len[clip], n_out = a.out_len[clip];
    if ((long long)blockIdx.x * kResampleTilesPerCta * tile >= n_out) return;
    const float* __restrict__ x = a.in + a.in_off[clip];
    float* __restrict__ y = a.out + a.out_off[clip];
    for (int tt = 0; tt < kResampleTilesPerCta; ++tt) {
      const long long base = ((long long)blockIdx.x * kResampleTilesPerCta + tt) * tile;
      if (base >= n_out) break;
#pragma unroll
      for (int u = 0; u < kResamplePerThread; ++u) {
        const long long n = base + u * kResampleThreads + threadIdx.x;
        if (n >= n_out) continue;
        const long long t = (n + a.n_pre_remove) * (long long)a.down;
        const long long q = t / a.up;
        const int p = (int)(t - q * a.up);
        const float* __restrict__ w = SMEM_TAPS ? rs_taps + p * Jp : a.taps + (size_t)p * a.J;
        float acc = 0.f;
        int j_hi = a.J - 1, j_lo = 0;
        if (q - j_hi < 0) j_hi = (int)q;                         // x index >= 0
        if (q >= n_in) j_lo = (int)(q - (n_in - 1));             // x index < n_in
        const float* xq = x + q;
        for (int j = j_hi; j >= j_lo; --j) acc = __fadd_rn(acc, __fmul_rn(__ldg(xq - j), w[j]));
        y[n] = acc;
      }
    }
  }
}

// ---- host: scipy.signal.firwin(2 * half + 1, 1 / max_rate, window=('kaiser', 5.0)) * up, in phase order
struct ResampleFilter {
  int up = 1, down = 1, J = 0, n_pre_remove = 0;
  std::vector<float> taps;   // [up][J]
};

inline double bessel_i0(double x) {
  const double q = 0.25 * x * x;
  double term = 1.0, sum = 1.0;
  for (int k = 1; k < 200; ++k) {
    term *= q / ((double)k * (double)k);
    sum += term;
    if (term < 1e-18 * sum) break;
  }
  return sum;
}

inline long long resample_out_len(long long n_in, int up, int down) {
  const long long n = n_in * up;
  return n / down + (n % down != 0);
}

inline ResampleFilter design_resample_filter(int target_sr, int native_sr) {
  ResampleFilter f;
  const int g = std::gcd(target_sr, native_sr);
  f.up = target_sr / g; f.down = native_sr / g;
  const int max_rate = f.up > f.down ? f.up : f.down;
  const double fc = 1.0 / (double)max_rate;
  const int half = 10 * max_rate, numtaps = 2 * half + 1;
  const double pi = 3.141592653589793238462643383279502884;
  std::vector<double> h(numtaps);
  const double alpha = 0.5 * (numtaps - 1), beta = 5.0, i0b = bessel_i0(beta);
  double s = 0.0;
  for (int n = 0; n < numtaps; ++n) {
    const double m = (double)n - alpha;
    const double xm = fc * m;
    const double yv = pi * (xm == 0.0 ? 1.0e-20 : xm);          // numpy.sinc
    const double r = (m / alpha);
    const double win = bessel_i0(beta * std::sqrt(std::fmax(0.0, 1.0 - r * r))) / i0b;
    h[n] = fc * (std::sin(yv) / yv) * win;
    s += h[n];
  }
  const int n_pre_pad = f.down - half % f.down;
  f.n_pre_remove = (half + n_pre_pad) / f.down;
  const int padded = n_pre_pad + numtaps;
  f.J = (padded + f.up - 1) / f.up;
  f.taps.assign((size_t)f.up * f.J, 0.f);
  for (int n = 0; n < numtaps; ++n) {
    const float hv = (float)(h[n] / s) * (float)f.up;           // cast to the data type first, then `h *= up`
    const int idx = n_pre_pad + n;
    f.taps[(size_t)(idx % f.up) * f.J + idx / f.up] = hv;
  }
  return f;
}

}  // namespace afx
