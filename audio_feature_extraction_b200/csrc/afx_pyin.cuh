// pYIN front end on the GPU: YIN difference function, cumulative-mean normalisation, parabolic refinement,
// trough probabilities (Beta(2,18) threshold prior x Boltzmann(2) trough prior) -> sparse observation rows.
//
// Replaces librosa.pyin's _cumulative_mean_normalized_difference / _parabolic_interpolation / __pyin_helper as reached
// from audio_feature_extraction_toolkit/core/feature_extractor.py:87-94 (reference), on the trimmed pre-emphasised clip.
//
// Work unit: (clip, chunk of F consecutive frames), one 16-warp CTA.
//   phase 0  the raw samples the chunk spans are read with coalesced loads, pre-emphasised into shared memory
//            (zero outside the trimmed signal = librosa's centre padding), and an exclusive prefix sum of y^2
//            (double) is built for the window energies.
//   phase 1  autocorrelation by hop-sized blocks: A[b][tau] = sum_{n in block b} y[n] y[n + tau]; one warp per block,
//            each lane owns 4 consecutive lags per group of 128, the block is walked 4 samples at a time with
//            128-bit shared-memory loads (one new load per 16 FMAs per lane).  A frame's window of n_fft/2 samples
//            is exactly n_fft/(2 hop) consecutive blocks, so every product is formed once although frames overlap.
//   phase 2  one warp per frame: d(tau) = E(0) + E(tau) - 2 acf(tau) in double, cumulative mean by a warp scan,
//            troughs, parabolic shifts, the 100-threshold sweep with ballot/popc ranks, pitch-bin candidates.
#pragma once
#include "afx_common.cuh"

namespace afx {

constexpr int kYinWarps = 16;
constexpr int kYinThreads = kYinWarps * 32;
constexpr int kYinBlocks = 16;          // hop blocks per chunk
constexpr int kYinMaxGroups = 11;       // trough groups of 32 (<= 352 troughs per frame: n_lags <= 700)

struct PyinTables {
  const double* thresholds;   // [101]
  const double* beta;         // [100]
  const double* beta_cum;     // [101]
  const double* boltz_fact;   // [cap + 1]
  const double* boltz_exp;    // [cap + 1]
};

struct YinArgs {
  const float* audio;
  const long long* offsets;
  const int* lengths;
  const ClipInfo* info;
  const int* chunk_off;        // device prefix [n_clips + 1] of ceil(T_ub / F)
  const long long* frame_off;  // device prefix [n_clips + 1] of T_ub
  int clip_begin, clip_end;
  long long frame_base;        // frame_off[clip_begin]
  int n_fft, hop, win, min_period, max_period, n_bins;
  int frames_per_chunk;        // F = 16 - (n_fft / (2 hop) - 1)
  int lag_pad;                 // lags 0 .. max_period padded to a multiple of 128
  int tile;                    // floats in the y tile: 16 hop + lag_pad + 8
  int cap;                     // candidate slots per frame record
  int rec_bytes;               // bytes per frame record
  float coef;
  double sr, fmin;
  PyinTables tab;
  unsigned char* obs;          // [wave frames][rec_bytes]
};

// frame record: int n_cand @0; double log_unvoiced @8; double voiced_prob @16; double logp[cap] @24; uint16 bin[cap]
__device__ __forceinline__ double* rec_logp(unsigned char* r) { return reinterpret_cast<double*>(r + 24); }
__device__ __forceinline__ unsigned short* rec_bins(unsigned char* r, int cap) { return reinterpret_cast<unsigned short*>(r + 24 + 8 * (size_t)cap); }

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
__device__ __forceinline__ double shfl_up_d(double v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }


template <int C>   // lags per lane in phase 2: 32 * C >= max_period
__global__ void __launch_bounds__(kYinThreads, 1) yin_obs_kernel(YinArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // phase 0/1 layout: [ys float tile][A float 16 x lag_pad][S double tile + 1]; phase 2b aliases it with trough lists
  float* ys = reinterpret_cast<float*>(smem_raw);
  float* A = ys + a.tile;
  double* S = reinterpret_cast<double*>(A + kYinBlocks * a.lag_pad);
  __shared__ double s_warp[kYinWarps];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // ---- decode (clip, chunk)
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
  const int F = a.frames_per_chunk;
  const int t0 = chunk * F;
  if (T <= 0 || t0 >= T) return;
  const int H = a.hop, W = a.win, RB = a.n_fft / (2 * H);   // RB blocks per frame window
  const float* __restrict__ x = a.audio + a.offsets[clip];
  const int L0 = a.lengths[clip];
  // tile index i holds y[n0 + 1 + i] (trimmed coordinates), n0 = (t0 - RB) * H: block k = tile [k H, (k+1) H)
  const int n0 = (t0 - RB) * H;

  // ---- phase 0: pre-emphasised tile, zero outside the trimmed signal
  for (int i = tid; i < a.tile; i += kYinThreads) {
    const int n = n0 + 1 + i;
    float y = 0.f;
    if (n >= 0 && n < ci.len) {
      const int g = ci.start + n;
      const float xv = __ldg(x + g);
      if (g == 0) {
        const float x1 = (L0 > 1) ? __ldg(x + 1) : 0.f;
        y = xv + (2.f * xv - x1);
      } else {
        y = __fadd_rn(xv, __fmul_rn(-a.coef, __ldg(x + g - 1)));
      }
    }
    ys[i] = y;
  }
  __syncthreads();
  {  // exclusive prefix sum of y^2 in double: S[i] = sum_{m < i} ys[m]^2, i = 0 .. tile
    const int per = (a.tile + kYinThreads - 1) / kYinThreads;
    const int b0 = min(tid * per, a.tile), b1 = min(b0 + per, a.tile);
    double loc = 0.0;
    for (int i = b0; i < b1; ++i) { const double v = ys[i]; loc += v * v; }
    double inc = loc;   // inclusive scan over the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double up = shfl_up_d(inc, o); if (lane >= o) inc += up; }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    double woff = 0.0;
    for (int w2 = 0; w2 < warp; ++w2) woff += s_warp[w2];
    double run = woff + inc - loc;
    for (int i = b0; i < b1; ++i) { S[i] = run; const double v = ys[i]; run += v * v; }
    if (tid == kYinThreads - 1) S[a.tile] = run;
  }
  // ---- phase 1: block autocorrelations A[k][tau], tau = 0 .. lag_pad - 1, warp k <-> block k
  {
    const int P = a.lag_pad >> 7;   // groups of 128 lags
    float* Ak = A + warp * a.lag_pad;
    const float4* yb = reinterpret_cast<const float4*>(ys + warp * H);        // block samples, 16-byte aligned
    for (int p = 0; p < P; ++p) {
      const int tau0 = 128 * p + 4 * lane;
      const float4* yl = reinterpret_cast<const float4*>(ys + warp * H + tau0);
      float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
      float4 v0 = yl[0];
      const int steps = H >> 2;
#pragma unroll 4
      for (int m = 0; m < steps; ++m) {
        const float4 s4 = yb[m];
        const float4 v1 = yl[m + 1];   // <= tile index 16 H + lag_pad - 1 (tile = 16 H + lag_pad + 8)
        // acc[c] += sum_r s[r] * y[i + r + tau0 + c]
        acc0 = fmaf(s4.x, v0.x, acc0); acc1 = fmaf(s4.x, v0.y, acc1); acc2 = fmaf(s4.x, v0.z, acc2); acc3 = fmaf(s4.x, v0.w, acc3);
        acc0 = fmaf(s4.y, v0.y, acc0); acc1 = fmaf(s4.y, v0.z, acc1); acc2 = fmaf(s4.y, v0.w, acc2); acc3 = fmaf(s4.y, v1.x, acc3);
        acc0 = fmaf(s4.z, v0.z, acc0); acc1 = fmaf(s4.z, v0.w, acc1); acc2 = fmaf(s4.z, v1.x, acc2); acc3 = fmaf(s4.z, v1.y, acc3);
        acc0 = fmaf(s4.w, v0.w, acc0); acc1 = fmaf(s4.w, v1.x, acc1); acc2 = fmaf(s4.w, v1.y, acc2); acc3 = fmaf(s4.w, v1.z, acc3);
        v0 = v1;
      }
      *reinterpret_cast<float4*>(Ak + tau0) = make_float4(acc0, acc1, acc2, acc3);
    }
  }
  __syncthreads();

  // ---- phase 2a: warp f <-> frame t0 + f: d(tau), cumulative-mean normalisation; lane owns tau = 1 + lane*C + e
  const int f = warp;
  const int t = t0 + f;
  const bool frame_ok = f < F && t < T;   // warp-uniform
  const double tiny = 2.2250738585072014e-308;
  double yv[C];        // CMND value at the lane's lags
  bool vld[C];         // lag inside [min_period, max_period]
  double y_prev = 0.0, y_next = 0.0;   // neighbours across the lane boundary
  if (frame_ok) {
    const int i0 = f * H;                        // tile index of x[1]
    const double E0r = S[i0 + W] - S[i0];
    const double E0 = fabs(E0r) < 1e-6 ? 0.0 : E0r;
    double dl[C];
    double loc = 0.0;
#pragma unroll
    for (int e = 0; e < C; ++e) {
      const int tau = 1 + lane * C + e;
      double d = 0.0;
      if (tau <= a.max_period) {
        double acf = 0.0;
        for (int k = 0; k < RB; ++k) acf += (double)A[(f + k) * a.lag_pad + tau];
        if (fabs(acf) < 1e-6) acf = 0.0;
        double Et = S[i0 + tau + W] - S[i0 + tau];
        if (fabs(Et) < 1e-6) Et = 0.0;
        d = E0 + Et - 2.0 * acf;
      }
      dl[e] = d;
      loc += d;
    }
    double inc = loc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double up = shfl_up_d(inc, o); if (lane >= o) inc += up; }
    double run = inc - loc;   // sum of d over all smaller lags
#pragma unroll
    for (int e = 0; e < C; ++e) {
      const int tau = 1 + lane * C + e;
      run += dl[e];
      const double cm = run / (double)tau;
      vld[e] = tau >= a.min_period && tau <= a.max_period;
      yv[e] = vld[e] ? dl[e] / (cm + tiny) : 0.0;
    }
    y_prev = shfl_up_d(yv[C - 1], 1);
    y_next = __shfl_down_sync(0xffffffffu, yv[0], 1);
  }
  __syncthreads();   // A / S consumed by every warp: the region is reused for the trough lists

  // ---- phase 2b: troughs, parabolic shifts, threshold sweep, pitch-bin candidates
  if (!frame_ok) return;
  double* lh = reinterpret_cast<double*>(smem_raw) + (size_t)warp * 2 * a.cap;   // trough heights
  double* lp = lh + a.cap;                                                       // refined periods
  const int n_lags = a.max_period - a.min_period + 1;
  unsigned tmask = 0u;      // bit e: lane's lag e is a trough
  double shift[C];
#pragma unroll
  for (int e = 0; e < C; ++e) {
    const int tau = 1 + lane * C + e;
    const int i = tau - a.min_period;          // index into the yin frame
    const double yc = yv[e];
    const double ym = e > 0 ? yv[e - 1] : y_prev;
    const double yp = e < C - 1 ? yv[e + 1] : y_next;
    bool tr = false;
    shift[e] = 0.0;
    if (vld[e]) {
      if (i == 0) tr = yc < yp;                                  // is_trough[0] = y[0] < y[1]
      else if (i == n_lags - 1) tr = yc < ym;                    // localmin's right edge
      else {
        tr = yc < ym && yc <= yp;
        const double pa = yp + ym - 2.0 * yc;
        const double pb = (yp - ym) / 2.0;
        if (fabs(pb) < fabs(pa)) shift[e] = -pb / pa;            // _parabolic_interpolation (0 at both ends)
      }
    }
    if (tr) tmask |= 1u << e;
  }
  int cnt = __popc(tmask);
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int up = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += up; }
  const int n_tr = __shfl_sync(0xffffffffu, incl, 31);
  {
    int pos = incl - cnt;
#pragma unroll
    for (int e = 0; e < C; ++e)
      if (tmask & (1u << e)) { lh[pos] = yv[e]; lp[pos] = (double)(1 + lane * C + e) + shift[e]; ++pos; }
  }
  __syncwarp();
  const long long slot = a.frame_off[clip] - a.frame_base + t;
  unsigned char* rec = a.obs + (size_t)slot * a.rec_bytes;
  if (n_tr == 0) {
    if (lane == 0) {
      *reinterpret_cast<int*>(rec) = 0;
      *reinterpret_cast<double*>(rec + 8) = log(1.0 / (double)a.n_bins + tiny);
      *reinterpret_cast<double*>(rec + 16) = 0.0;
    }
    return;
  }
  // lane-per-trough in groups of 32; jmin = first threshold index j (1..100) with h < thresholds[j], 101 if none
  double prob[kYinMaxGroups];
  int jmin[kYinMaxGroups];
  int jlo = 101;
#pragma unroll
  for (int g = 0; g < kYinMaxGroups; ++g) {
    prob[g] = 0.0; jmin[g] = 102;
    const int ai = 32 * g + lane;
    if (ai < n_tr) {
      const double h = lh[ai];
      int j = min(100, max(1, (int)(h * 100.0)));
      while (j > 1 && h < a.tab.thresholds[j - 1]) --j;
      while (j <= 100 && !(h < a.tab.thresholds[j])) ++j;
      jmin[g] = j;
      jlo = min(jlo, j);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) jlo = min(jlo, __shfl_xor_sync(0xffffffffu, jlo, o));
  const unsigned lt_mask = (1u << lane) - 1u;
  for (int j = jlo; j <= 100; ++j) {
    unsigned bal[kYinMaxGroups];
    int n_act = 0;
#pragma unroll
    for (int g = 0; g < kYinMaxGroups; ++g) {
      bal[g] = (32 * g < n_tr) ? __ballot_sync(0xffffffffu, jmin[g] <= j) : 0u;
      n_act += __popc(bal[g]);
    }
    const double cj = a.tab.beta[j - 1] * a.tab.boltz_fact[n_act];
    int carry = 0;
#pragma unroll
    for (int g = 0; g < kYinMaxGroups; ++g) {
      if (jmin[g] <= j) prob[g] = fma(cj, a.tab.boltz_exp[carry + __popc(bal[g] & lt_mask)], prob[g]);
      carry += __popc(bal[g]);
    }
  }
  {  // global minimum (first index): + no_trough_prob * sum(beta[:jmin - 1])
    double bh = 1.0e300; int bi = 0x7fffffff;
#pragma unroll
    for (int g = 0; g < kYinMaxGroups; ++g) {
      const int ai = 32 * g + lane;
      if (ai < n_tr) { const double h = lh[ai]; if (h < bh) { bh = h; bi = ai; } }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double oh = __shfl_xor_sync(0xffffffffu, bh, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (oh < bh || (oh == bh && oi < bi)) { bh = oh; bi = oi; }
    }
#pragma unroll
    for (int g = 0; g < kYinMaxGroups; ++g)
      if (32 * g + lane == bi) prob[g] += 0.01 * a.tab.beta_cum[min(jmin[g], 101) - 1];
  }
  __syncwarp();
  // pitch bins; a later trough (longer period) overwrites an earlier one in the same bin
  int* lbin = reinterpret_cast<int*>(lh);   // heights are dead: reuse as bin list (-1 = not a candidate)
  int bin_r[kYinMaxGroups];
#pragma unroll
  for (int g = 0; g < kYinMaxGroups; ++g) {
    const int ai = 32 * g + lane;
    bin_r[g] = -1;
    if (ai < n_tr && prob[g] > 0.0) {
      const double f0 = a.sr / lp[ai];
      double bf = rint(120.0 * log2(f0 / a.fmin));
      bf = fmin(fmax(bf, 0.0), (double)a.n_bins);
      const int b = (int)bf;
      if (b < a.n_bins) bin_r[g] = b;
    }
  }
  __syncwarp();
#pragma unroll
  for (int g = 0; g < kYinMaxGroups; ++g) { const int ai = 32 * g + lane; if (ai < n_tr) lbin[ai] = bin_r[g]; }
  __syncwarp();
  double vp = 0.0;
  int keep_cnt = 0;
  bool keep[kYinMaxGroups];
#pragma unroll
  for (int g = 0; g < kYinMaxGroups; ++g) {
    const int ai = 32 * g + lane;
    keep[g] = false;
    if (ai < n_tr && bin_r[g] >= 0) {
      keep[g] = true;
      for (int q = ai + 1; q < n_tr; ++q) { const int bq = lbin[q]; if (bq >= 0) { keep[g] = bq != bin_r[g]; break; } }
      if (keep[g]) { vp += prob[g]; ++keep_cnt; }
    }
  }
  vp = warp_sum(vp);
  vp = fmin(fmax(vp, 0.0), 1.0);
  int kinc = keep_cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int up = __shfl_up_sync(0xffffffffu, kinc, o); if (lane >= o) kinc += up; }
  const int n_keep = __shfl_sync(0xffffffffu, kinc, 31);
  {
    int pos = kinc - keep_cnt;
    double* out_lp = rec_logp(rec);
    unsigned short* out_b = rec_bins(rec, a.cap);
#pragma unroll
    for (int g = 0; g < kYinMaxGroups; ++g)
      if (keep[g]) { out_lp[pos] = log(prob[g] + tiny); out_b[pos] = (unsigned short)bin_r[g]; ++pos; }
  }
  if (lane == 0) {
    *reinterpret_cast<int*>(rec) = n_keep;
    *reinterpret_cast<double*>(rec + 8) = log((1.0 - vp) / (double)a.n_bins + tiny);
    *reinterpret_cast<double*>(rec + 16) = vp;
  }
}


// =====================================================================================================
// pYIN HMM decode (librosa.sequence.viterbi on 2 * n_bins states) + F0 statistics, one CTA per clip.
//   V_t[j] = logobs_t[j] + max_k (V_{t-1}[k] + logA[k][j]), first index on ties, float64 log domain.
// Every out-of-band entry of logA is exactly log(tiny) and every in-band entry is larger, so the dense maximum equals
// max(band maximum in both voicing blocks, G + log(tiny)) with G = max_k V_{t-1}[k] (first index k*).
// The band maximum is a (max,+) product with the 2*hw+1 triangle log-weights; each thread carries two neighbouring
// targets so every predecessor value is loaded once for two targets.
constexpr int kVitThreads = 640;

struct VitArgs {
  const ClipInfo* info;
  const long long* frame_off;   // device prefix [n_clips + 1] of T_ub (global rows of frames_out)
  int clip_begin;
  long long frame_base;         // frame_off[clip_begin]
  const unsigned char* obs;     // [wave frames][rec_bytes]
  int cap, rec_bytes;
  int n_bins, half_width;
  const double* lt_same;        // [2 hw + 1]
  const double* edge_corr;      // [n_bins]
  const double* freqs;          // [n_bins]
  double delta_switch;          // log-weight difference of a voicing switch
  double log_tiny, log_init_unvoiced;
  unsigned short* ptr;          // [wave frames][2 n_bins] back-pointers
  short* states;                // [wave frames] decoded state per frame
  float* summaries; int summary_width;
  float* frames_out; int frame_width, n_mfcc;
};

__global__ void __launch_bounds__(kVitThreads, 1) viterbi_kernel(VitArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int nb = a.n_bins, hw = a.half_width, ns = 2 * nb;
  double* Fp = reinterpret_cast<double*>(smem_raw);   // [2][nb] V_{t-1} + edge_corr
  double* hs = Fp + ns;                                // [2][nb] band maxima
  double* lobs = hs + ns;                              // [2][nb] dense log-observation rows (double buffered)
  double* corr = lobs + ns;                            // [nb]
  double* lt = corr + nb;                              // [2 hw + 1]
  unsigned short* harg = reinterpret_cast<unsigned short*>(lt + 2 * hw + 2);   // [2][nb]
  __shared__ double s_pv[kVitThreads / 32];
  __shared__ int s_pi[kVitThreads / 32];
  __shared__ double s_red[3][kVitThreads / 32];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int clip = a.clip_begin + blockIdx.x;
  const ClipInfo ci = a.info[clip];
  const int T = ci.n_frames;
  float* out = a.summaries + (size_t)clip * a.summary_width;
  if (T <= 0) return;   // status != OK: the fused kernel already zeroed the row
  const long long slot0 = a.frame_off[clip] - a.frame_base;
  const unsigned char* rec0 = a.obs + (size_t)slot0 * a.rec_bytes;
  unsigned short* ptr0 = a.ptr + (size_t)slot0 * ns;

  for (int i = tid; i < nb; i += kVitThreads) corr[i] = a.edge_corr[i];
  for (int i = tid; i < 2 * hw + 1; i += kVitThreads) lt[i] = a.lt_same[i];
  for (int i = tid; i < ns; i += kVitThreads) lobs[i] = a.log_tiny;
  __syncthreads();

  // block argmax (first index) of per-thread (v, i) pairs -> every thread gets (G, k*)
  auto block_argmax = [&](double v, int i, double& G, int& ks) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ov = __shfl_xor_sync(0xffffffffu, v, o);
      const int oi = __shfl_xor_sync(0xffffffffu, i, o);
      if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
    }
    if (lane == 0) { s_pv[warp] = v; s_pi[warp] = i; }
    __syncthreads();
    G = s_pv[0]; ks = s_pi[0];
    for (int w2 = 1; w2 < kVitThreads / 32; ++w2) {
      const double ov = s_pv[w2]; const int oi = s_pi[w2];
      if (ov > G || (ov == G && oi < ks)) { G = ov; ks = oi; }
    }
  };

  // ---- t = 0: V_0 = logobs_0 + log(p_init + tiny)
  double G; int kstar;
  {
    const int n0c = *reinterpret_cast<const int*>(rec0);
    const double lu0 = *reinterpret_cast<const double*>(rec0 + 8);
    if (tid < n0c) lobs[rec_bins(const_cast<unsigned char*>(rec0), a.cap)[tid]] = rec_logp(const_cast<unsigned char*>(rec0))[tid];
    __syncthreads();
    double bv = -1.0e308; int bi = 0x7fffffff;
    if (tid < nb) {
      const double vv = lobs[tid] + a.log_tiny;
      const double vu = lu0 + a.log_init_unvoiced;
      Fp[tid] = vv + corr[tid];
      Fp[nb + tid] = vu + corr[tid];
      bv = vv; bi = tid;
      if (vu > bv) { bv = vu; bi = nb + tid; }
    }
    block_argmax(bv, bi, G, kstar);
  }

  for (int t = 1; t < T; ++t) {
    const unsigned char* rec = rec0 + (size_t)t * a.rec_bytes;
    const unsigned char* recp = rec - a.rec_bytes;
    double* lob_cur = lobs + (t & 1) * nb;
    double* lob_old = lobs + ((t - 1) & 1) * nb;
    const int nc = *reinterpret_cast<const int*>(rec);
    const double lu = *reinterpret_cast<const double*>(rec + 8);
    // phase 1: band maxima h_s[j] = max_k (F_s[k] + lt[j - k + hw]); thread <-> (s, two neighbouring targets)
    {
      const int half = (nb + 1) >> 1;
      if (tid < 2 * half) {
        const int s = tid >= half ? 1 : 0;
        const int j0 = 2 * (tid - s * half);
        const double* Fs = Fp + s * nb;
        double m0 = -1.0e308, m1 = -1.0e308;
        int a0 = 0, a1 = 0;
        const int k_lo = max(0, j0 - hw), k_hi = min(nb - 1, j0 + 1 + hw);
        int idx = j0 - k_lo + hw;               // table index of target j0 at k = k_lo (<= 2 hw)
        double w1 = (idx + 1 <= 2 * hw) ? lt[idx + 1] : -1.0e308;   // weight of target j0 + 1
        for (int k = k_lo; k <= k_hi; ++k, --idx) {
          const double fk = Fs[k];
          const double w0 = (idx >= 0) ? lt[idx] : -1.0e308;         // weight of target j0
          const double c0 = fk + w0, c1 = fk + w1;
          if (idx >= 0 && c0 > m0) { m0 = c0; a0 = k; }
          if (idx + 1 <= 2 * hw && c1 > m1) { m1 = c1; a1 = k; }
          w1 = w0;
        }
        hs[s * nb + j0] = m0; harg[s * nb + j0] = (unsigned short)a0;
        if (j0 + 1 < nb) { hs[s * nb + j0 + 1] = m1; harg[s * nb + j0 + 1] = (unsigned short)a1; }
      }
      // this frame's sparse observations into the dense row; the previous frame's entries back to log(tiny)
      const int ncp = *reinterpret_cast<const int*>(recp);
      if (tid < ncp) lob_old[rec_bins(const_cast<unsigned char*>(recp), a.cap)[tid]] = a.log_tiny;
      if (tid < nc) lob_cur[rec_bins(const_cast<unsigned char*>(rec), a.cap)[tid]] = rec_logp(const_cast<unsigned char*>(rec))[tid];
    }
    __syncthreads();
    // phase 2: combine the voicing blocks and the global fallback, add the observation, store the back-pointers
    double bv = -1.0e308; int bi = 0x7fffffff;
    const double fb = G + a.log_tiny;   // best out-of-band predecessor: k* through a log(tiny) transition
    if (tid < nb) {
      const int j = tid;
      const double h0 = hs[j], h1 = hs[nb + j];
      const int k0 = harg[j], k1 = nb + harg[nb + j];
      // voiced target j: predecessors in index order are block 0 (same voicing), then block 1 (switch)
      double mv = h0; int pv = k0;
      { const double c = h1 + a.delta_switch; if (c > mv) { mv = c; pv = k1; } }
      if (fb > mv || (fb == mv && kstar < pv)) { mv = fb; pv = kstar; }
      // unvoiced target nb + j
      double mu = h0 + a.delta_switch; int pu = k0;
      if (h1 > mu) { mu = h1; pu = k1; }
      if (fb > mu || (fb == mu && kstar < pu)) { mu = fb; pu = kstar; }
      const double vv = lob_cur[j] + mv;
      const double vu = lu + mu;
      unsigned short* pt = ptr0 + (size_t)t * ns;
      pt[j] = (unsigned short)pv; pt[nb + j] = (unsigned short)pu;
      bv = vv; bi = j;
      if (vu > bv) { bv = vu; bi = nb + j; }
      // F for the next step (phase 1 readers are done: the barrier above)
      Fp[j] = vv + corr[j];
      Fp[nb + j] = vu + corr[j];
    }
    block_argmax(bv, bi, G, kstar);   // contains a barrier: Fp complete before the next phase 1
  }

  // ---- backtrace (thread 0), then statistics over voiced frames (core/feature_extractor.py:97-113)
  short* st = a.states + slot0;
  if (tid == 0) {
    int s = kstar;
    for (int t = T - 1; t >= 1; --t) { st[t] = (short)s; s = ptr0[(size_t)t * ns + s]; }
    st[0] = (short)s;
  }
  __syncthreads();
  double s1 = 0, s2 = 0, cntv = 0;
  for (int t = tid; t < T; t += kVitThreads) {
    const int s = st[t];
    const bool voiced = s < nb;
    const double f = voiced ? a.freqs[s] : 0.0;
    if (voiced) { s1 += f; s2 += f * f; cntv += 1.0; }
    if (a.frames_out) {
      float* fr = a.frames_out + (size_t)(a.frame_off[clip] + t) * a.frame_width + 3 * a.n_mfcc;
      fr[1] = voiced ? (float)f : __int_as_float(0x7fc00000);
      fr[2] = (float)*reinterpret_cast<const double*>(rec0 + (size_t)t * a.rec_bytes + 16);
    }
  }
  s1 = warp_sum(s1); s2 = warp_sum(s2); cntv = warp_sum(cntv);
  if (lane == 0) { s_red[0][warp] = s1; s_red[1][warp] = s2; s_red[2][warp] = cntv; }
  __syncthreads();
  if (tid == 0) {
    s1 = s2 = cntv = 0;
    for (int w2 = 0; w2 < kVitThreads / 32; ++w2) { s1 += s_red[0][w2]; s2 += s_red[1][w2]; cntv += s_red[2][w2]; }
    if (cntv > 0.0) {
      const double mean = s1 / cntv;
      const double var = fmax(s2 / cntv - mean * mean, 0.0);
      const double missing = ((double)T - cntv) / (double)T;
      out[0] = (float)mean; out[1] = (float)sqrt(var); out[2] = (float)missing; out[3] = (float)(1.0 - missing);
    } else {
      out[0] = 0.f; out[1] = 0.f; out[2] = 1.f; out[3] = 0.f;
    }
  }
}

}  // namespace afx
