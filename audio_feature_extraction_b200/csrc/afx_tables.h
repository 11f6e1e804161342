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

struct MelSparse {
  std::vector<float> dense;   // [128][nbins] float32 weights exactly as librosa.filters.mel builds them
  std::vector<float> w;       // padded [quarter][i][lane]
  std::vector<int> k0;        // [128]
  int cnt[4];
  int woff[4];
  int nnz;
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
  std::vector<int> first(n_mels, nbins), last(n_mels, -1);
  for (int i = 0; i < n_mels; ++i) {
    const double fd0 = mel_f[i + 1] - mel_f[i], fd1 = mel_f[i + 2] - mel_f[i + 1];
    const double enorm = 2.0 / (mel_f[i + 2] - mel_f[i]);
    for (int k = 0; k < nbins; ++k) {
      const double lower = -(mel_f[i] - fftfreqs[k]) / fd0;
      const double upper = (mel_f[i + 2] - fftfreqs[k]) / fd1;
      const float tri = (float)std::fmax(0.0, std::fmin(lower, upper));
      const float wv = (float)((double)tri * enorm);     // weights *= enorm (float32 array, float64 factor)
      ms.dense[(size_t)i * nbins + k] = wv;
      if (wv != 0.f) { ++ms.nnz; first[i] = std::min(first[i], k); last[i] = std::max(last[i], k); }
    }
  }
  ms.k0.assign(n_mels, 0);
  int off = 0;
  for (int q = 0; q < 4; ++q) {
    int cnt = 1;
    for (int l = 0; l < 32; ++l) {
      const int m = 32 * q + l;
      if (last[m] >= first[m]) cnt = std::max(cnt, last[m] - first[m] + 1);
    }
    ms.cnt[q] = cnt;
    ms.woff[q] = off;
    ms.w.resize(off + cnt * 32, 0.f);
    for (int l = 0; l < 32; ++l) {
      const int m = 32 * q + l;
      int k0 = (last[m] >= first[m]) ? first[m] : 0;
      k0 = std::min(k0, nbins - cnt);
      if (k0 < 0) k0 = 0;
      ms.k0[m] = k0;
      for (int i = 0; i < cnt; ++i) {
        const int k = k0 + i;
        ms.w[off + i * 32 + l] = (k < nbins) ? ms.dense[(size_t)m * nbins + k] : 0.f;
      }
    }
    off += cnt * 32;
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
