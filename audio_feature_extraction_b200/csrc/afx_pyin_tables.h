// Host-side construction of the pYIN constants (double precision), following librosa.pyin's defaults as reached
// from audio_feature_extraction_toolkit/core/feature_extractor.py:87-94 (reference):
//   n_thresholds=100, beta_parameters=(2,18), boltzmann_parameter=2, resolution=0.1,
//   max_transition_rate=35.92, switch_prob=0.01, no_trough_prob=0.01.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace afx {
namespace host {

constexpr int kPyinThresholds = 100;
constexpr int kPyinMaxBins = 768;      // pitch bins supported by the Viterbi kernel
constexpr int kPyinMaxHalfWidth = 63;  // (transition_width - 1) / 2

struct PyinConsts {
  int win_length, min_period, max_period, n_lags;   // n_lags = max_period - min_period + 1
  int n_bins, half_width;                            // transition_width = 2 * half_width + 1
  std::vector<double> thresholds;   // [101] np.linspace(0, 1, 101)
  std::vector<double> beta;         // [100] diff(beta.cdf(thresholds, 2, 18))
  std::vector<double> beta_cum;     // [101] beta_cum[n] = sum(beta[:n])
  std::vector<double> beta_suf;     // [102] beta_suf[j] = sum(beta[j - 1:]) for j = 1 .. 100, 0 at 101 (index 0 = total)
  std::vector<double> boltz_fact;   // [n] (1 - exp(-2)) / (1 - exp(-2 n)), n = 0 unused
  std::vector<double> boltz_exp;    // [k] exp(-2 k)
  std::vector<double> ltrans_same;  // [2 hw + 1] log(p_stay   * tri[d] / rowsum_interior + tiny)
  std::vector<double> ltrans_sw;    // [2 hw + 1] log(p_switch * tri[d] / rowsum_interior + tiny)
  std::vector<double> edge_corr;    // [n_bins] log(rowsum_interior / rowsum_k): 0 away from the edges
  std::vector<double> freqs;        // [n_bins] fmin * 2^(b / 120)
  double log_tiny, log_init_unvoiced;
};

// numpy's pairwise summation of a contiguous double array (numpy/core/src/umath/loops_utils.h.src)
inline double np_pairwise_sum(const double* a, long n) {
  if (n < 8) {
    double res = 0.;
    for (long i = 0; i < n; ++i) res += a[i];
    return res;
  } else if (n <= 128) {
    double r[8];
    for (int k = 0; k < 8; ++k) r[k] = a[k];
    long i;
    for (i = 8; i < n - (n % 8); i += 8)
      for (int k = 0; k < 8; ++k) r[k] += a[i + k];
    double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < n; ++i) res += a[i];
    return res;
  } else {
    long n2 = n / 2;
    n2 -= n2 % 8;
    return np_pairwise_sum(a, n2) + np_pairwise_sum(a + n2, n - n2);
  }
}

// python round(): half to even
inline long py_round(double x) { return std::lrint(x); }   // default rounding mode is to-nearest-even

inline bool pyin_consts(int sr, int frame_length, int hop_length, double fmin, double fmax, PyinConsts& c) {
  const double tiny = 2.2250738585072014e-308;
  c.win_length = frame_length / 2;
  c.min_period = (int)std::floor((double)sr / fmax);
  c.max_period = std::min((int)std::ceil((double)sr / fmin), frame_length - c.win_length - 1);
  c.n_lags = c.max_period - c.min_period + 1;
  const int bins_per_semitone = (int)std::ceil(1.0 / 0.1);
  c.n_bins = (int)std::floor(12 * bins_per_semitone * std::log2(fmax / fmin)) + 1;
  const long max_semitones = py_round(35.92 * 12 * hop_length / sr);
  const int width = (int)max_semitones * bins_per_semitone + 1;
  c.half_width = (width - 1) / 2;
  if (c.min_period < 1 || c.n_lags < 3 || c.n_bins < 2 || c.n_bins > kPyinMaxBins || c.half_width > kPyinMaxHalfWidth ||
      width > c.n_bins)
    return false;
  // thresholds = np.linspace(0, 1, 101): arange * step + start, last element = stop
  c.thresholds.resize(kPyinThresholds + 1);
  const double step = 1.0 / kPyinThresholds;
  for (int j = 0; j <= kPyinThresholds; ++j) c.thresholds[j] = j * step + 0.0;
  c.thresholds[kPyinThresholds] = 1.0;
  // Beta(2, 18) CDF in closed form: I_x(2, 18) = 1 - (1-x)^19 - 19 x (1-x)^18
  std::vector<double> cdf(kPyinThresholds + 1);
  for (int j = 0; j <= kPyinThresholds; ++j) {
    const double x = c.thresholds[j], q = 1.0 - x;
    cdf[j] = 1.0 - std::pow(q, 19) - 19.0 * x * std::pow(q, 18);
  }
  cdf[0] = 0.0; cdf[kPyinThresholds] = 1.0;
  c.beta.resize(kPyinThresholds);
  c.beta_cum.assign(kPyinThresholds + 1, 0.0);
  for (int j = 0; j < kPyinThresholds; ++j) { c.beta[j] = cdf[j + 1] - cdf[j]; c.beta_cum[j + 1] = c.beta_cum[j] + c.beta[j]; }
  c.beta_suf.assign(kPyinThresholds + 2, 0.0);   // summed from the small tail upwards
  for (int j = kPyinThresholds; j >= 1; --j) c.beta_suf[j] = c.beta_suf[j + 1] + c.beta[j - 1];
  c.beta_suf[0] = c.beta_suf[1];
  // scipy.stats.boltzmann.pmf(k, 2, N) = (1 - exp(-2)) / (1 - exp(-2 N)) * exp(-2 k)
  const int max_troughs = c.n_lags / 2 + 2;
  c.boltz_fact.assign(max_troughs + 1, 0.0);
  c.boltz_exp.assign(max_troughs + 1, 0.0);
  for (int n = 1; n <= max_troughs; ++n) c.boltz_fact[n] = (1.0 - std::exp(-2.0)) / (1.0 - std::exp(-2.0 * n));
  for (int k = 0; k <= max_troughs; ++k) c.boltz_exp[k] = std::exp(-2.0 * k);
  // transition_local(n_bins, width, window='triangle', wrap=False), row-normalised, kron with transition_loop(2, 0.99)
  std::vector<double> win(width);   // scipy.signal.windows.triang(odd M): 2 n / (M + 1), n = 1 .. (M+1)/2, mirrored
  for (int n = 1; n <= (width + 1) / 2; ++n) { win[n - 1] = 2.0 * n / (width + 1.0); win[width - n] = win[n - 1]; }
  const int nb = c.n_bins, hw = c.half_width;
  std::vector<double> row(nb), rowsum(nb);
  for (int i = 0; i < nb; ++i) {
    std::fill(row.begin(), row.end(), 0.0);
    for (int d = -hw; d <= hw; ++d) { const int j = i + d; if (j >= 0 && j < nb) row[j] = win[d + hw]; }
    rowsum[i] = np_pairwise_sum(row.data(), nb);
  }
  const int mid = nb / 2;   // an interior row (width <= n_bins is checked above; rows hw .. nb-1-hw share one sum pattern only
                            // up to the pairwise grouping, so the tables take the true middle row and edge_corr absorbs the rest)
  const double p_stay = 1.0 - 0.01;
  const double p_switch = (1.0 - p_stay) / (2 - 1);
  c.ltrans_same.resize(width); c.ltrans_sw.resize(width);
  for (int d = 0; d < width; ++d) {
    const double local = win[d] / rowsum[mid];
    c.ltrans_same[d] = std::log(p_stay * local + tiny);
    c.ltrans_sw[d] = std::log(p_switch * local + tiny);
  }
  c.edge_corr.resize(nb);
  for (int i = 0; i < nb; ++i) c.edge_corr[i] = (rowsum[i] == rowsum[mid]) ? 0.0 : std::log(rowsum[mid] / rowsum[i]);
  c.freqs.resize(nb);
  for (int b = 0; b < nb; ++b) c.freqs[b] = fmin * std::pow(2.0, (double)b / (12.0 * bins_per_semitone));
  c.log_tiny = std::log(tiny);
  c.log_init_unvoiced = std::log(1.0 / nb + tiny);
  return true;
}

}  // namespace host
}  // namespace afx
