// Host-side construction of the read-only tables (double precision, cast once to float32).
// Mirrors what librosa builds inside librosa.feature.mfcc / librosa.filters.mel / scipy.fftpack.dct
// (reached from audio_feature_extraction_toolkit/core/feature_extractor.py:127-134).
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace afx {
namespace host {

inline double hz_to_mel(double f) {   // Slaney scale (htk=False)
  const double f_sp = 200.0 / 3.0, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp;
  const double logstep = std::log(6.4) / 27.0;
  return f >= min_log_hz ? min_log_mel + std::log(f / min_log_hz) / logstep : f / f_sp;
}
inline double mel_to_hz(double m) {
  const double f_sp = 200.0 / 3.0, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp;
  const double logstep = std::log(6.4) / 27.0;
  return m >= min_log_mel ? min_log_hz * std::exp(logstep * (m - min_log_mel)) : f_sp * m;
}

// 0.5 * scipy.signal.get_window('hamming', n, fftbins=True)
inline std::vector<float> window_half(int n_fft) {
  std::vector<float> w(n_fft);
  for (int k = 0; k < n_fft; ++k) w[k] = (float)(0.5 * (0.54 - 0.46 * std::cos(2.0 * M_PI * k / n_fft)));
  return w;
}

// T[row][col] = exp(-2 pi i row col / n) stored (cos, -sin), row, col in [0, 32)
inline std::vector<float> twiddle32x32(int n) {
  std::vector<float> t(2 * 32 * 32);
  for (int r = 0; r < 32; ++r)
    for (int c = 0; c < 32; ++c) {
      const double ang = 2.0 * M_PI * (double)(r * c) / n;
      t[2 * (r * 32 + c)] = (float)std::cos(ang);
      t[2 * (r * 32 + c) + 1] = (float)(-std::sin(ang));
    }
  return t;
}

// W_n^k = exp(-2 pi i k / n) stored (cos, -sin), k in [0, count)
inline std::vector<float> post_twiddle(int n, int count) {
  std::vector<float> t(2 * (size_t)count);
  for (int k = 0; k < count; ++k) {
    const double ang = 2.0 * M_PI * (double)k / n;
    t[2 * k] = (float)std::cos(ang);
    t[2 * k + 1] = (float)(-std::sin(ang));
  }
  return t;
}

// Band owned by (quarter q, lane) of the fused kernel: the DCT-II mirror pairs (b, 127 - b) share a lane
inline int mel_band_of(int q, int lane) { return q == 0 ? lane : q == 1 ? 63 - lane : q == 2 ? 64 + lane : 127 - lane; }

// Sparse Slaney filterbank organised by the 129 SEGMENTS between consecutive mel points instead of by band: every
// FFT bin lies in exactly one segment s (f_s < f_k <= f_{s+1}), where it feeds the rising side of band s (weight U)
// and the falling side of band s - 1 (weight D), so each bin is touched once.  The weights are the float32 entries of
// librosa.filters.mel unchanged.  (q, lane) owns band b = mel_band_of(q, lane) and segment b + 1; segment 0 (below the
// first mel point, rising side of band 0 only) is kept apart.
struct MelSparse {
  std::vector<float> dense;   // [128][nbins] float32 weights exactly as librosa.filters.mel builds them
  std::vector<float> w;       // float2 (U, D) per entry: padded [quarter][i][lane], then segment 0's [cnt0] entries
  std::vector<int> k0;        // [128] first FFT bin of the (padded) run of (q, lane)'s segment, index 32 q + lane
  int cnt[4];                 // padded run length per quarter
  int woff[4];                // float2 offset of each quarter in w
  int k00, cnt0, woff0;       // segment 0: first bin, run length, float2 offset
  int nnz;
  // streaming kernel (afx_step.cuh): inside a segment both weights are linear in the bin index, so (q, lane) carries
  // lin[4 (32 q + lane) ..] = (U0, dU, D0, dD) with U(k0 + i) = U0 + i dU, D(k0 + i) = D0 + i dD (double arithmetic,
  // rounded once), and kl[32 q + lane] = first bin | (bins in the segment << 16); w0[i] = segment 0's rising weights
  std::vector<float> lin;
  std::vector<int> kl;
  std::vector<float> w0;
  double lin_max_err = 0;     // largest |linear weight - librosa float32 weight| over all in-segment entries
};

inline MelSparse mel_sparse(int sr, int n_fft, int n_mels = 128) {
  MelSparse ms;
  const int nbins = 1 + n_fft / 2;
  const double fmin = 0.0, fmax = sr / 2.0;
  // np.fft.rfftfreq(n, d=1/sr): k * (1 / (n * d))
  const double d = 1.0 / sr, val = 1.0 / (n_fft * d);
  std::vector<double> fftfreqs(nbins);
  for (int k = 0; k < nbins; ++k) fftfreqs[k] = k * val;
  // np.linspace(min_mel, max_mel, n_mels + 2)
  const double min_mel = hz_to_mel(fmin), max_mel = hz_to_mel(fmax);
  const int np_ = n_mels + 2;
  std::vector<double> mel_f(np_);
  const double step = (max_mel - min_mel) / (np_ - 1);
  for (int i = 0; i < np_; ++i) mel_f[i] = mel_to_hz(i == np_ - 1 ? max_mel : min_mel + i * step);
  ms.dense.assign((size_t)n_mels * nbins, 0.f);
  ms.nnz = 0;
  for (int i = 0; i < n_mels; ++i) {
    const double fd0 = mel_f[i + 1] - mel_f[i], fd1 = mel_f[i + 2] - mel_f[i + 1];
    const double enorm = 2.0 / (mel_f[i + 2] - mel_f[i]);
    for (int k = 0; k < nbins; ++k) {
      const double lower = -(mel_f[i] - fftfreqs[k]) / fd0;
      const double upper = (mel_f[i + 2] - fftfreqs[k]) / fd1;
      const float tri = (float)std::fmax(0.0, std::fmin(lower, upper));
      const float wv = (float)((double)tri * enorm);     // weights *= enorm (float32 array, float64 factor)
      ms.dense[(size_t)i * nbins + k] = wv;
      if (wv != 0.f) ++ms.nnz;
    }
  }
  // segment of each bin: the s with band s - 1 falling / band s rising there.  A non-zero weight of band m at bin k
  // means f_m < f_k < f_{m+2}, i.e. s in {m, m + 1}; classify by the dense matrix itself so rounding cannot disagree.
  std::vector<int> seg(nbins, -1), sfirst(n_mels + 1, nbins), slast(n_mels + 1, -1);
  for (int k = 0; k < nbins; ++k) {
    int s = 0;
    while (s < n_mels && fftfreqs[k] > mel_f[s + 1]) ++s;   // f_s < f_k <= f_{s+1} (s = 0 also takes f_k <= f_0)
    seg[k] = s;
    sfirst[s] = std::min(sfirst[s], k); slast[s] = std::max(slast[s], k);
  }
  auto wU = [&](int s, int k) -> float { return (s < n_mels && seg[k] == s) ? ms.dense[(size_t)s * nbins + k] : 0.f; };
  auto wD = [&](int s, int k) -> float { return (s >= 1 && seg[k] == s) ? ms.dense[(size_t)(s - 1) * nbins + k] : 0.f; };
  ms.k0.assign(n_mels, 0);
  int off = 0;   // float2 units
  for (int q = 0; q < 4; ++q) {
    int cnt = 1;
    for (int l = 0; l < 32; ++l) { const int s = mel_band_of(q, l) + 1; if (slast[s] >= sfirst[s]) cnt = std::max(cnt, slast[s] - sfirst[s] + 1); }
    ms.cnt[q] = cnt;
    ms.woff[q] = off;
    ms.w.resize(2 * (size_t)(off + cnt * 32), 0.f);
    for (int l = 0; l < 32; ++l) {
      const int s = mel_band_of(q, l) + 1;
      int k0 = (slast[s] >= sfirst[s]) ? sfirst[s] : 0;
      k0 = std::max(0, std::min(k0, nbins - cnt));
      ms.k0[32 * q + l] = k0;
      for (int i = 0; i < cnt; ++i) {
        const int k = k0 + i;
        ms.w[2 * (size_t)(off + i * 32 + l)] = k < nbins ? wU(s, k) : 0.f;
        ms.w[2 * (size_t)(off + i * 32 + l) + 1] = k < nbins ? wD(s, k) : 0.f;
      }
    }
    off += cnt * 32;
  }
  ms.cnt0 = (slast[0] >= sfirst[0]) ? slast[0] - sfirst[0] + 1 : 0;
  ms.k00 = ms.cnt0 ? sfirst[0] : 0;
  ms.woff0 = off;
  ms.w.resize(2 * (size_t)(off + ((ms.cnt0 + 1) & ~1)), 0.f);   // keeps the float count a multiple of 4
  for (int i = 0; i < ms.cnt0; ++i) ms.w[2 * (size_t)(off + i)] = wU(0, ms.k00 + i);
  // linear form of the in-segment weights for the streaming kernel
  ms.lin.assign(4 * 128, 0.f);
  ms.kl.assign(128, 0);
  for (int q = 0; q < 4; ++q)
    for (int l = 0; l < 32; ++l) {
      const int s = mel_band_of(q, l) + 1;                 // segment between mel_f[s] and mel_f[s + 1]
      const bool has = slast[s] >= sfirst[s];
      const int k0 = has ? sfirst[s] : 0, len = has ? slast[s] - sfirst[s] + 1 : 0;
      const double width = mel_f[s + 1] - mel_f[s];
      // rising side of band s (absent for s = n_mels), falling side of band s - 1
      const double en_u = s < n_mels ? 2.0 / (mel_f[s + 2] - mel_f[s]) : 0.0;
      const double en_d = 2.0 / (mel_f[s + 1] - mel_f[s - 1]);
      const double U0 = (fftfreqs[k0] - mel_f[s]) / width * en_u, dU = val / width * en_u;
      const double D0 = (mel_f[s + 1] - fftfreqs[k0]) / width * en_d, dD = -val / width * en_d;
      float* o = &ms.lin[4 * (size_t)(32 * q + l)];
      o[0] = (float)U0; o[1] = (float)dU; o[2] = (float)D0; o[3] = (float)dD;
      ms.kl[32 * q + l] = k0 | (len << 16);
      for (int i = 0; i < len; ++i) {
        const float u = fmaf((float)i, o[1], o[0]), d = fmaf((float)i, o[3], o[2]);
        ms.lin_max_err = std::fmax(ms.lin_max_err, std::fabs((double)u - (double)wU(s, k0 + i)));
        ms.lin_max_err = std::fmax(ms.lin_max_err, std::fabs((double)d - (double)wD(s, k0 + i)));
      }
    }
  ms.w0.assign(4, 0.f);
  for (int i = 0; i < ms.cnt0 && i < 4; ++i) ms.w0[i] = wU(0, ms.k00 + i);
  // every non-zero weight must be covered exactly once
  {
    long covered = 0;
    for (int k = 0; k < nbins; ++k) { const int s = seg[k]; if (wU(s, k) != 0.f) ++covered; if (wD(s, k) != 0.f) ++covered; }
    if (covered != ms.nnz) ms.nnz = -1;   // flagged to the caller (afx_create fails)
  }
  return ms;
}

// scipy.fft.dct(type=2, norm='ortho') rows: dct[c][m] = s_c cos(pi c (2m+1) / (2 M))
inline std::vector<float> dct_rows(int n_out, int n_mels = 128) {
  std::vector<float> dct((size_t)n_out * n_mels);
  for (int c = 0; c < n_out; ++c) {
    const double s = c == 0 ? std::sqrt(1.0 / n_mels) : std::sqrt(2.0 / n_mels);
    for (int m = 0; m < n_mels; ++m) dct[(size_t)c * n_mels + m] = (float)(s * std::cos(M_PI * c * (2 * m + 1) / (2.0 * n_mels)));
  }
  return dct;
}

}  // namespace host
}  // namespace afx
