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
  const double* beta_suf;     // [102] beta_suf[j] = sum(beta[j - 1:]), j = 1 .. 101 (0 at 101)
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
  int blk_len, blk_stride, rb; // autocorrelation blocks: block k = tile [k stride, k stride + len), a frame's window is rb
                               // consecutive blocks.  Block mode (hop | n_fft/2, hop % 4 == 0): len = stride = hop,
                               // rb = n_fft / (2 hop) -- every product formed once.  Frame mode (any other hop):
                               // len = n_fft / 2, stride = hop, rb = 1 -- one block per frame
  int frames_per_chunk;        // F = 16 - (rb - 1)
  int lag_pad;                 // lags 0 .. max_period padded to a multiple of 128
  int tile;                    // floats in the y tile: 15 stride + len + lag_pad + 12
  int cap;                     // candidate slots per frame record
  int rec_bytes;               // bytes per frame record
  float coef;
  int raw;                     // input already pre-emphasised: no first-sample rule
  double sr, fmin;
  PyinTables tab;
  unsigned char* obs;          // [wave frames][rec_bytes]
};

// frame record: int n_cand @0; double log_unvoiced @8; double voiced_prob @16; double logp[cap] @24; uint16 bin[cap]
__device__ __forceinline__ double* rec_logp(unsigned char* r) { return reinterpret_cast<double*>(r + 24); }
__device__ __forceinline__ unsigned short* rec_bins(unsigned char* r, int cap) { return reinterpret_cast<unsigned short*>(r + 24 + 8 * (size_t)cap); }

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
__device__ __forceinline__ double shfl_up_d(double v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }



// ---- phase 2b of yin_obs_kernel for one frame (one warp): global minimum, compaction, threshold sweep, pitch bins, record.
// MAXG = compile-time bound of the trough / candidate groups of 32: the kernel instantiates it for 4 groups (<= 128 troughs:
// nearly every frame) and for kYinMaxGroups; the group loops are fully unrolled (their arrays live in registers), so the
// common instantiation is a third of the general one's code -- the kernel stalled on instruction fetch (no_instruction 0.9
// per issue with 16 warps on data-dependent paths through 146 KB of SASS).
template <int MAXG>
__device__ __forceinline__ void yin_phase2b(const YinArgs& a, unsigned char* smem_raw, double* lh, double* lp, int n_tr,
                                            unsigned char* rec, int warp, int lane) {
  const double tiny = 2.2250738585072014e-308;
  // ---- global minimum trough (first index): it gets no_trough_prob * sum(beta[:jmin - 1]) on top of its sweep sum
  double bh = 1.0e300; int bi = 0x7fffffff;
#pragma unroll
  for (int g = 0; g < MAXG; ++g) {
    if (32 * g >= n_tr) break;
    const int ai = 32 * g + lane;
    if (ai < n_tr) { const double h = lh[ai]; if (h < bh) { bh = h; bi = ai; } }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double oh = __shfl_xor_sync(0xffffffffu, bh, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (oh < bh || (oh == bh && oi < bi)) { bh = oh; bi = oi; }
  }
  // ---- compaction (order kept): only troughs below some threshold (jmin <= 100; jmin = first j in 1..100 with
  // h < thresholds[j]) and the global minimum can get a non-zero probability.  packed entry = jmin << 16 | trough index;
  // am = 128-bit set of the activation points {jmin}
  unsigned* cpk = reinterpret_cast<unsigned*>(smem_raw + (size_t)kYinWarps * 2 * a.cap * 8) + (size_t)warp * a.cap;
  const unsigned lt_mask = (1u << lane) - 1u;
  unsigned am0 = 0u, am1 = 0u, am2 = 0u, am3 = 0u;
  int n_c = 0;
#pragma unroll
  for (int g = 0; g < MAXG; ++g) {
    if (32 * g >= n_tr) break;
    const int ai = 32 * g + lane;
    int j = 102;
    bool act = false;
    if (ai < n_tr) {
      const double h = lh[ai];
      j = min(100, max(1, (int)(h * 100.0)));
      while (j > 1 && h < __ldg(a.tab.thresholds + j - 1)) --j;
      while (j <= 100 && !(h < __ldg(a.tab.thresholds + j))) ++j;
      act = j <= 100 || ai == bi;
    }
    const unsigned bal = __ballot_sync(0xffffffffu, act);
    if (act) {
      cpk[n_c + __popc(bal & lt_mask)] = ((unsigned)j << 16) | (unsigned)ai;
      if (j <= 100) {
        const unsigned bit = 1u << (j & 31);
        const int w = j >> 5;
        am0 |= w == 0 ? bit : 0u; am1 |= w == 1 ? bit : 0u; am2 |= w == 2 ? bit : 0u; am3 |= w == 3 ? bit : 0u;
      }
    }
    n_c += __popc(bal);
  }
  am0 = __reduce_or_sync(0xffffffffu, am0); am1 = __reduce_or_sync(0xffffffffu, am1);
  am2 = __reduce_or_sync(0xffffffffu, am2); am3 = __reduce_or_sync(0xffffffffu, am3);
  __syncwarp();
  // the activation points in ascending order, popped from the 128-bit set (warp-uniform); 101 when exhausted
  unsigned cw = am0; int cwi = 0;
  auto pop_act = [&]() -> int {
    while (cw == 0u && cwi < 3) { ++cwi; cw = cwi == 1 ? am1 : cwi == 2 ? am2 : am3; }
    if (cw == 0u) return 101;
    const int j = 32 * cwi + __ffs(cw) - 1;
    cw &= cw - 1u;
    return j;
  };
  // ---- threshold sweep over the compact list, lane-per-candidate in groups of 32.  Between two activation points the
  // active set (hence every rank and the Boltzmann normalisation) is constant, so the thresholds of a segment [j, jn)
  // contribute (sum of their beta weights) * boltzmann.pmf(rank; 2, n_active) in one step.
  int cjm[MAXG], cai[MAXG];
  double prob[MAXG];
#pragma unroll
  for (int g = 0; g < MAXG; ++g) {
    cjm[g] = 102; cai[g] = -1; prob[g] = 0.0;
    const int c = 32 * g + lane;
    if (c < n_c) { const unsigned pk = cpk[c]; cjm[g] = (int)(pk >> 16); cai[g] = (int)(pk & 0xffffu); }
  }
  for (int j = pop_act(); j <= 100;) {
    const int jn = pop_act();
    unsigned bal[MAXG];
    int n_act = 0;
#pragma unroll
    for (int g = 0; g < MAXG; ++g) {
      if (32 * g >= n_c) break;
      bal[g] = __ballot_sync(0xffffffffu, cjm[g] <= j);
      n_act += __popc(bal[g]);
    }
    const double cj = (__ldg(a.tab.beta_suf + j) - __ldg(a.tab.beta_suf + jn)) * __ldg(a.tab.boltz_fact + n_act);
    int carry = 0;
#pragma unroll
    for (int g = 0; g < MAXG; ++g) {
      if (32 * g >= n_c) break;
      if (cjm[g] <= j) prob[g] = fma(cj, __ldg(a.tab.boltz_exp + carry + __popc(bal[g] & lt_mask)), prob[g]);
      carry += __popc(bal[g]);
    }
    j = jn;
  }
#pragma unroll
  for (int g = 0; g < MAXG; ++g)
    if (cai[g] == bi) prob[g] += 0.01 * __ldg(a.tab.beta_cum + min(cjm[g], 101) - 1);
  // ---- pitch bins; a later candidate (longer period) overwrites an earlier one in the same bin
  int* lbin = reinterpret_cast<int*>(lh);   // heights are dead: reuse as bin list (-1 = not a candidate)
  int bin_r[MAXG];
  __syncwarp();
#pragma unroll
  for (int g = 0; g < MAXG; ++g) {
    bin_r[g] = -1;
    if (32 * g >= n_c) continue;
    const int c = 32 * g + lane;
    if (c < n_c && prob[g] > 0.0) {
      const double f0 = a.sr / lp[cai[g]];
      double bf = rint(120.0 * log2(f0 / a.fmin));
      bf = fmin(fmax(bf, 0.0), (double)a.n_bins);
      const int b = (int)bf;
      if (b < a.n_bins) bin_r[g] = b;
    }
    if (c < n_c) lbin[c] = bin_r[g];
  }
  __syncwarp();
  double vp = 0.0;
  int keep_cnt = 0;
  bool keep[MAXG];
#pragma unroll
  for (int g = 0; g < MAXG; ++g) {
    const int c = 32 * g + lane;
    keep[g] = false;
    if (c < n_c && bin_r[g] >= 0) {
      keep[g] = true;
      for (int q = c + 1; q < n_c; ++q) { const int bq = lbin[q]; if (bq >= 0) { keep[g] = bq != bin_r[g]; break; } }
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
    for (int g = 0; g < MAXG; ++g)
      if (keep[g]) { out_lp[pos] = log(prob[g] + tiny); out_b[pos] = (unsigned short)bin_r[g]; ++pos; }
  }
  if (lane == 0) {
    *reinterpret_cast<int*>(rec) = n_keep;
    *reinterpret_cast<double*>(rec + 8) = log((1.0 - vp) / (double)a.n_bins + tiny);
    *reinterpret_cast<double*>(rec + 16) = vp;
  }
}

// One pass of the block autocorrelation: lane owns the LPL consecutive lags tau_base + LPL lane + c and slides over the
// block four samples at a time; acc[c] += s[r] * y[i + r + tau + c] in sample order (the same sum whatever LPL is).
// yb = tile at the (16-byte aligned) start of the walk; sh / e: first / one-past-last walk index that belongs to the block.
template <int LPL, bool MASKED>
__device__ __forceinline__ void acf_pass(const float* yb_f, float* Ak, int tau_base, int lane, int steps, int sh, int e) {
  const int tau0 = tau_base + LPL * lane;
  const float4* yb = reinterpret_cast<const float4*>(yb_f);
  const float4* yl = reinterpret_cast<const float4*>(yb_f + tau0);
  float acc[LPL], v[LPL + 4];
#pragma unroll
  for (int c = 0; c < LPL; ++c) acc[c] = 0.f;
#pragma unroll
  for (int q = 0; q < LPL / 4; ++q) { const float4 t = yl[q]; v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w; }
#pragma unroll(LPL / 4 + 1)
  for (int m = 0; m < steps; ++m) {
    float4 s4 = yb[m];
    if (MASKED) {
      const int i = 4 * m;
      if (i < sh || i + 3 >= e) {   // first / last step
        if (i < sh || i >= e) s4.x = 0.f;
        if (i + 1 < sh || i + 1 >= e) s4.y = 0.f;
        if (i + 2 < sh || i + 2 >= e) s4.z = 0.f;
        if (i + 3 < sh || i + 3 >= e) s4.w = 0.f;
      }
    }
    const float4 nv = yl[m + LPL / 4];   // <= tile index of the walk start + 4 steps + lag_pad + 3 < tile
    v[LPL] = nv.x; v[LPL + 1] = nv.y; v[LPL + 2] = nv.z; v[LPL + 3] = nv.w;
    const float sr[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < LPL; ++c) acc[c] = fmaf(sr[r], v[r + c], acc[c]);
#pragma unroll
    for (int c = 0; c < LPL; ++c) v[c] = v[c + 4];
  }
#pragma unroll
  for (int q = 0; q < LPL / 4; ++q)
    *reinterpret_cast<float4*>(Ak + tau0 + 4 * q) = make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
}


template <int C, bool MASKED>   // C: lags per lane in phase 2, 32 * C >= max_period; MASKED: blocks at any alignment / length (frame mode)
// (C >= 22, i.e. max_period > 512 -- 44.1 kHz / 2048: the tile, the block sums and the prefix sums take 157 KB, one CTA per
// SM fits anyway, so the kernel may use up to 128 registers instead of spilling the C float64 values a lane carries)
__global__ void __launch_bounds__(kYinThreads, C >= 22 ? 1 : 2) yin_obs_kernel(YinArgs a) {
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
  {  // clips of similar length: the clip is near task / (chunks per clip); bracket it there before bisecting
    const int g = a.clip_begin + (int)(((long long)task * (a.clip_end - a.clip_begin)) / max(1, a.chunk_off[a.clip_end] - base));
    const int g0 = max(a.clip_begin, g - 1), g1 = min(a.clip_end, g + 2);
    if (a.chunk_off[g0] - base <= task) lo = g0;
    if (g1 < a.clip_end && a.chunk_off[g1] - base > task && g1 > lo) hi = g1;
  }
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
  const int BS = a.blk_stride, W = a.win, RB = a.rb;   // RB blocks per frame window
  const float* __restrict__ x = a.audio + a.offsets[clip];
  const int L0 = a.lengths[clip];
  // tile index i holds y[n0 + 1 + i] (trimmed coordinates), n0 = t0 hop - W: frame t0 + f owns tile [f BS, f BS + W)
  const int n0 = t0 * a.hop - W;

  // ---- phase 0: pre-emphasised tile, zero outside the trimmed signal.  Eight samples per thread and round: the 16 global
  // loads of a round are issued back to back (one exposed L2 latency per round instead of one per sample; the per-sample
  // loop spent 7 % of the kernel's stall samples on its second load)
  for (int i0 = 0; i0 < a.tile; i0 += 8 * kYinThreads) {
    float xv[8], xp[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int i = i0 + tid + k * kYinThreads;
      const int n = n0 + 1 + i;
      const bool in = i < a.tile && n >= 0 && n < ci.len;
      const int g = ci.start + n;
      xv[k] = in ? __ldg(x + g) : 0.f;
      xp[k] = (in && g > 0) ? __ldg(x + g - 1) : 0.f;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int i = i0 + tid + k * kYinThreads;
      const int n = n0 + 1 + i;
      float y = 0.f;
      if (n >= 0 && n < ci.len) {
        if (ci.start + n == 0) {
          const float x1 = (L0 > 1) ? __ldg(x + 1) : 0.f;
          y = a.raw ? xv[k] : xv[k] + (2.f * xv[k] - x1);
        } else {
          y = __fadd_rn(xv[k], __fmul_rn(-a.coef, xp[k]));
        }
      }
      if (i < a.tile) ys[i] = y;
    }
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
    float* Ak = A + warp * a.lag_pad;
    // the block starts at tile index b; the walk starts at the 16-byte boundary below it and the samples before the
    // block / past its end enter with weight zero (fmaf(0, y, acc) = acc: the sums are those of the unshifted walk;
    // block mode: b is a multiple of 4 and so is the length, nothing is masked)
    const int b = warp * BS, b4 = MASKED ? b & ~3 : b, sh = MASKED ? b & 3 : 0, e = sh + a.blk_len;
    const int steps = MASKED ? (e + 3) >> 2 : a.blk_len >> 2;
    // lags per lane and pass: 12 (48-byte lane pitch: conflict-free 128-bit loads, 48 FMAs per 2 loads) when the padded
    // lag count is a multiple of 384 (22.05 kHz / 1024: one pass; 44.1 kHz / 2048: two), else 8 or 4
    if (a.lag_pad % 384 == 0) { for (int t0l = 0; t0l < a.lag_pad; t0l += 384) acf_pass<12, MASKED>(ys + b4, Ak, t0l, lane, steps, sh, e); }
    else if (a.lag_pad % 256 == 0) { for (int t0l = 0; t0l < a.lag_pad; t0l += 256) acf_pass<8, MASKED>(ys + b4, Ak, t0l, lane, steps, sh, e); }
    else { for (int t0l = 0; t0l < a.lag_pad; t0l += 128) acf_pass<4, MASKED>(ys + b4, Ak, t0l, lane, steps, sh, e); }
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
    const int i0 = f * BS;                       // tile index of x[1]
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
  if (n_tr <= 128) yin_phase2b<4>(a, smem_raw, lh, lp, n_tr, rec, warp, lane);
  else yin_phase2b<kYinMaxGroups>(a, smem_raw, lh, lp, n_tr, rec, warp, lane);
}


// =====================================================================================================
// pYIN HMM decode (librosa.sequence.viterbi on 2 * n_bins states) + F0 statistics, one 4- or 5-warp CTA per clip, several
// clips resident per SM (the time recursion is sequential; parallelism comes from co-resident clips).
//   V_t[j] = logobs_t[j] + max_k (V_{t-1}[k] + logA[k][j]), first index on ties, float64 log domain.
// Every out-of-band entry of logA is exactly log(tiny) and every in-band entry is larger, so the dense maximum equals
// max(band maximum in both voicing blocks, G + log(tiny)) with G = max_k V_{t-1}[k] (first index k*).
// The band maximum h[j] = max_{|k-j|<=hw} (F[k] + lt[j-k+hw]) is a (max,+) product with the log of a triangle, a
// concave kernel: M[j][k] = F[k] + lt[j-k+hw] is inverse-Monge, so the (leftmost) arg-maxima are non-decreasing in j.
//   phase A  brute force over the band for every kVitStride-th target (anchors), thread per anchor, four
//            independent compare chains;
//   phase B  every other target searches only [argmax(anchor below), argmax(anchor above)] (a handful of entries);
//   phase C  voicing blocks + global fallback + observation, back-pointers, next F, block arg-max.
// Declared inexactness: monotonicity holds for exact sums; two band entries whose float64 sums differ by rounding only
// could in principle be resolved differently from the dense search (never observed; a full-band fallback covers
// empty search ranges).
constexpr int kVitThreads = 160;
constexpr int kVitWarps = kVitThreads / 32;
constexpr int kVitStride = 8;
constexpr int kVitMinCtas = 5;

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
  int wk[64];                   // filter path: (floor(ltr[d] * 2^14) << 6) | (63 - d), d = 0 .. 2 hw
};

// shared-memory geometry, host and device: F rows are padded by `pad` (multiple of 8 >= hw) on the left and 4 lq + 8
// on the right (pads hold -1e308) and skewed by one slot per 8 (anchors sit 8 apart: conflict-free 64-bit loads)
struct VitGeo {
  int pad, lq, fs, nb_even, ks;
  __host__ __device__ VitGeo(int nb, int hw, bool filter_path) {
    pad = (hw + 7) & ~7;
    lq = (2 * hw + 1 + 3) / 4;
    const int n = pad + nb + 4 * lq + 8;
    fs = filter_path ? (n + 1) & ~1 : (n + (n >> 3) + 8) & ~1;   // only the generic path skews its rows
    nb_even = (nb + 1) & ~1;
    ks = (2 * hw + nb + 16 + 3) & ~3;   // int32 key row of the filter path: index hw + k, k = -hw .. nb + hw + 12
  }
  __host__ __device__ size_t bytes(int nb) const {
    return (size_t)(2 * fs + 2 * nb_even /*hs*/ + 2 * nb_even /*lobs*/ + nb_even /*corr*/ + 4 * lq + 8 /*ltr*/) * 8 +
           (size_t)2 * ks * 4 /*keys*/ + (size_t)2 * nb_even * 2 /*harg*/;
  }
};

// HW = 0: generic band search (anchors + monotone ranges).  HW > 0 (compile-time half width, 2 HW + 1 <= 63): the band
// maximum is found by an exact two-stage search -- an int32 fixed-point pre-filter (keys floor((F - G) 2^14) << 6 with
// the band offset in the low bits, one VIADDMNMX per (target, predecessor), maxima kept per chunk of kVitChunk
// predecessors), then the float64 scan of only those chunks whose filtered maximum lies within 3 quantisation steps of
// the best one (any chunk that could hold the true float64 maximum: the filter's error is < 2 steps per entry).
constexpr int kVitChunk = 8;
// NW: warps per clip.  5 (thread <-> 4 consecutive targets in the filter) or 4 (thread <-> 2 targets: 19 rounds of 32 tasks
// split 5 / 5 / 5 / 4 over the warps, and every scheduler of the SM hosts exactly one warp of every resident CTA)
template <int HW, int NW = kVitWarps>
__global__ void __launch_bounds__(NW * 32, kVitMinCtas) viterbi_kernel(VitArgs a) {
  constexpr int kThreads = NW * 32, kWarps = NW;
  constexpr int TPT = NW == 4 ? 2 : 4;   // targets per thread in the filter stage
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int nb = a.n_bins, hw = a.half_width, ns = 2 * nb;
  const VitGeo geo(nb, hw, HW > 0);
  const int FS = geo.fs, NBE = geo.nb_even, PAD = geo.pad, LQ = geo.lq;
  double* Fp = reinterpret_cast<double*>(smem_raw);   // [2][FS] V_{t-1} + edge_corr, padded + skewed
  double* hs = Fp + 2 * FS;                            // [2][NBE] band maxima
  double* lobs = hs + 2 * NBE;                         // [2][NBE] dense log-observation rows (double buffered)
  double* corr = lobs + 2 * NBE;                       // [NBE]
  double* ltr = corr + NBE;                            // [4 LQ + 8] ltr[d] = lt[2 hw - d]: weight of predecessor j - hw + d
  int* Kq = reinterpret_cast<int*>(ltr + 4 * LQ + 8);  // [2][KS] filter keys, index hw + k
  const int KS = geo.ks;
  unsigned short* harg = reinterpret_cast<unsigned short*>(Kq + 2 * KS);   // [2][NBE]
  __shared__ double s_pv[kWarps];
  __shared__ int s_pi[kWarps];
  __shared__ double s_red[3][kWarps];
  __shared__ unsigned short s_queue[HW > 0 ? kWarps : 1][HW > 0 ? 128 : 1];   // pass-2 work items of a warp's round

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int clip = a.clip_begin + blockIdx.x;
  const ClipInfo ci = a.info[clip];
  const int T = ci.n_frames;
  float* out = a.summaries + (size_t)clip * a.summary_width;
  if (T <= 0) return;   // status != OK: the fused kernel already zeroed the row
  const long long slot0 = a.frame_off[clip] - a.frame_base;
  const unsigned char* rec0 = a.obs + (size_t)slot0 * a.rec_bytes;
  unsigned short* ptr0 = a.ptr + (size_t)slot0 * ns;
  auto fidx = [PAD](int k) { const int kp = k + PAD; return HW > 0 ? kp : kp + (kp >> 3); };   // the filter path needs no skew

  for (int i = tid; i < 2 * FS; i += kThreads) Fp[i] = -1.0e308;
  for (int i = tid; i < nb; i += kThreads) corr[i] = a.edge_corr[i];
  for (int i = tid; i < 4 * LQ + 8; i += kThreads) ltr[i] = (i <= 2 * hw) ? a.lt_same[2 * hw - i] : -1.0e308;
  if (HW > 0) for (int i = tid; i < 2 * KS; i += kThreads) Kq[i] = -(1 << 30);   // pads never win
  for (int i = tid; i < 2 * NBE; i += kThreads) lobs[i] = a.log_tiny;
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
#pragma unroll
    for (int w2 = 1; w2 < kWarps; ++w2) {
      const double ov = s_pv[w2]; const int oi = s_pi[w2];
      if (ov > G || (ov == G && oi < ks)) { G = ov; ks = oi; }
    }
  };

  // ---- t = 0: V_0 = logobs_0 + log(p_init + tiny)
  double G, Gkey = 0.0; int kstar;   // Gkey: reference of the filter keys
  {
    const int n0c = *reinterpret_cast<const int*>(rec0);
    const double lu0 = *reinterpret_cast<const double*>(rec0 + 8);
    for (int i = tid; i < n0c; i += kThreads)
      lobs[rec_bins(const_cast<unsigned char*>(rec0), a.cap)[i]] = rec_logp(const_cast<unsigned char*>(rec0))[i];
    __syncthreads();
    double bv = -1.0e308; int bi = 0x7fffffff;
    for (int j = tid; j < nb; j += kThreads) {
      const double vv = lobs[j] + a.log_tiny;
      const double vu = lu0 + a.log_init_unvoiced;
      Fp[fidx(j)] = vv + corr[j];
      Fp[FS + fidx(j)] = vu + corr[j];
      if (vv > bv) { bv = vv; bi = j; }          // j ascending within a thread: strict > keeps the first index
    }
    double bu = -1.0e308; int biu = 0x7fffffff;   // unvoiced states come after every voiced state
    for (int j = tid; j < nb; j += kThreads) { const double vu = lu0 + a.log_init_unvoiced; if (vu > bu) { bu = vu; biu = nb + j; } }
    if (bu > bv) { bv = bu; bi = biu; }
    block_argmax(bv, bi, G, kstar);
    if (HW > 0) {   // filter keys of F_0 relative to G_0 (later steps: relative to the G of the step before, in phase C)
      for (int j = tid; j < nb; j += kThreads) {
        Kq[HW + j] = max(__double2int_rd((Fp[fidx(j)] - G) * 16384.0), -33554431) * 64;
        Kq[KS + HW + j] = max(__double2int_rd((Fp[FS + fidx(j)] - G) * 16384.0), -33554431) * 64;
      }
      Gkey = G;
      __syncthreads();   // the keys are complete before the first stage 1 reads its neighbours' entries
    }
  }

  const int nA = (nb - 1 + kVitStride - 1) / kVitStride + 1;   // anchors per voicing block: 0, S, 2S, ..., nb - 1
  for (int t = 1; t < T; ++t) {
    const unsigned char* rec = rec0 + (size_t)t * a.rec_bytes;
    const unsigned char* recp = rec - a.rec_bytes;
    double* lob_cur = lobs + (t & 1) * NBE;
    double* lob_old = lobs + ((t - 1) & 1) * NBE;
    const int nc = *reinterpret_cast<const int*>(rec);
    const double lu = *reinterpret_cast<const double*>(rec + 8);
    if constexpr (HW > 0) {
      // ---- filter path.  Stage 1 (int32 pre-filter): thread <-> 4 consecutive targets of one voicing block, so that
      // each 128-bit key load feeds 16 (target, predecessor) pairs.  Stage 2 (float64 scan of the surviving chunks):
      // lane <-> ONE target, consecutive lanes on consecutive targets -- with the stage-1 assignment (lanes 4 targets =
      // 32 bytes apart) every 64-bit load of F ran into 4-way bank conflicts, which made this kernel shared-memory
      // bound (ncu r02h: 1 217 conflict wavefronts of ~3 400 per time step).  The chunk masks of a thread's 4 targets
      // travel packed in one register: one shuffle per round of 32 targets.
      constexpr int W = 2 * HW + 1, NCH = (W + kVitChunk - 1) / kVitChunk, NI4 = (W + 3 + 3) / 4;
      static_assert(NCH <= 8, "chunk masks are packed 8 bits per target");
      constexpr int NLD = (W + TPT - 1 + TPT - 1) / TPT;   // key loads of TPT keys each: predecessors j0 - HW + i, i < W + TPT - 1
      const int NT = (nb + TPT - 1) / TPT;                 // tasks per voicing block
      // A target whose filtered band maximum lies a safe margin below fb = G + log(tiny) (the global fallback) cannot
      // win in phase C: its chunks are not scanned (true h < (m >> 6) + 2 steps; one more step covers the float64
      // roundings of fb).  Keys are relative to Gkey, the maximum of the step before the previous one.
      int fbq = -2147483647;
      { const double fbx = (G + a.log_tiny - Gkey) * 16384.0; if (fbx > -2.0e9) fbq = __double2int_rd(fbx); }
      for (int t0 = 32 * warp; t0 < 2 * NT; t0 += kThreads) {   // warp-uniform: 32 tasks = 32 TPT targets per round
        const int task = t0 + lane;
        unsigned packed = 0u;
        if (task < 2 * NT) {
          const int s = task >= NT ? 1 : 0;
          const int j0 = TPT * (task - s * NT);
          const int* kp = Kq + s * KS + j0;   // keys of predecessors j0 - HW + i, i = 0, 1, ... (16 / 8-byte aligned)
          int acc[TPT][NCH];
#pragma unroll
          for (int r = 0; r < TPT; ++r)
#pragma unroll
            for (int c = 0; c < NCH; ++c) acc[r][c] = -2147483647;
#pragma unroll
          for (int il = 0; il < NLD; ++il) {
            int keys[TPT];
            if constexpr (TPT == 4) { const int4 kv = reinterpret_cast<const int4*>(kp)[il]; keys[0] = kv.x; keys[1] = kv.y; keys[2] = kv.z; keys[3] = kv.w; }
            else { const int2 kv = reinterpret_cast<const int2*>(kp)[il]; keys[0] = kv.x; keys[1] = kv.y; }
#pragma unroll
            for (int u = 0; u < TPT; ++u)
#pragma unroll
              for (int r = 0; r < TPT; ++r) {
                const int d = TPT * il + u - r;
                if (d >= 0 && d < W) acc[r][d / kVitChunk] = __viaddmax_s32(keys[u], a.wk[d], acc[r][d / kVitChunk]);
              }
          }
#pragma unroll
          for (int r = 0; r < TPT; ++r) {
            if (j0 + r >= nb) break;
            int m = acc[r][0];
#pragma unroll
            for (int c = 1; c < NCH; ++c) m = max(m, acc[r][c]);
            const int thr = ((m >> 6) - 3) * 64;      // every chunk that could hold the true float64 maximum
            unsigned cmask = 0u;
#pragma unroll
            for (int c = 0; c < NCH; ++c) cmask |= (acc[r][c] >= thr) ? (1u << c) : 0u;
            if ((m >> 6) + 3 <= fbq) cmask = 0u;
            packed |= cmask << (8 * r);
          }
        }
        // chunk scan of one target, ascending entries, strict >: the first index wins float64 ties
        auto scan_chunk = [&](const double* Fj, int c, double& best, int& arg) {
          const double* fc = Fj + c * kVitChunk;
          const double* wc = ltr + c * kVitChunk;   // ltr is -1e308 beyond d = 2 HW: those entries never win
#pragma unroll
          for (int dd = 0; dd < kVitChunk; ++dd) {
            const double cand = fc[dd] + wc[dd];
            if (cand > best) { best = cand; arg = c * kVitChunk + dd; }
          }
        };
        // pass 1: every target scans its first surviving chunk (one chunk for ~90 % of the targets: adjacent predecessors
        // tie exactly on the log-triangle ramps F is made of, and one tie in eight straddles a chunk boundary)
        unsigned rest_all = 0u;   // remaining chunk masks of this lane's four pass-1 targets, 8 bits each
#pragma unroll
        for (int q = 0; q < TPT; ++q) {
          const int src = (32 / TPT) * q + lane / TPT, r = lane % TPT;
          unsigned cmask = (__shfl_sync(0xffffffffu, packed, src) >> (8 * r)) & 0xffu;   // 0: skipped or no such target
          const int task2 = t0 + src;
          const int s = task2 >= NT ? 1 : 0;
          const int j = TPT * (task2 - s * NT) + r;
          if (task2 < 2 * NT && j < nb) {
            double best = -1.0e308; int arg = HW;
            if (cmask) {
              const int c = __ffs(cmask) - 1;
              cmask &= cmask - 1;
              scan_chunk(Fp + s * FS + PAD - HW + j, c, best, arg);
            }
            hs[s * NBE + j] = best;
            harg[s * NBE + j] = (unsigned short)(j - HW + arg);
            rest_all |= cmask << (8 * q);
          }
        }
        // pass 2: the few targets with more chunks, compacted over the warp (in a lane-per-target loop all 32 lanes
        // wait for them: 1.4 lock-step scans per target instead of 1.1)
        if (__any_sync(0xffffffffu, rest_all != 0u)) {
          unsigned short* qu = s_queue[warp];
          const unsigned lt_mask = (1u << lane) - 1u;
          int L = 0;
#pragma unroll
          for (int q = 0; q < TPT; ++q) {
            const unsigned rq = (rest_all >> (8 * q)) & 0xffu;
            const unsigned bal = __ballot_sync(0xffffffffu, rq != 0u);
            if (rq) qu[L + __popc(bal & lt_mask)] = (unsigned short)((q << 13) | (lane << 8) | rq);
            L += __popc(bal);
          }
          __syncwarp();   // queue and the pass-1 results of this warp are visible
          for (int i = lane; i < L; i += 32) {
            const unsigned it = qu[i];
            unsigned rq = it & 0xffu;
            const int ln = (it >> 8) & 31, q = it >> 13;
            const int task2 = t0 + (32 / TPT) * q + ln / TPT;
            const int s = task2 >= NT ? 1 : 0;
            const int j = TPT * (task2 - s * NT) + ln % TPT;
            double best = hs[s * NBE + j];
            int arg = (int)harg[s * NBE + j] - (j - HW);
            while (rq) {
              const int c = __ffs(rq) - 1;
              rq &= rq - 1;
              scan_chunk(Fp + s * FS + PAD - HW + j, c, best, arg);
            }
            hs[s * NBE + j] = best;
            harg[s * NBE + j] = (unsigned short)(j - HW + arg);
          }
        }
      }
    } else {
    // ---- phase A: anchors, full band
    for (int ax = tid; ax < 2 * nA; ax += kThreads) {
      const int s = ax >= nA ? 1 : 0;
      const int j = min((ax - s * nA) * kVitStride, nb - 1);
      const double* Fs = Fp + s * FS;
      const int kp0 = j - hw + PAD;                 // padded index of predecessor d = 0
      double b0 = -1.0e308, b1 = -1.0e308, b2 = -1.0e308, b3 = -1.0e308;
      int a0 = 0, a1 = LQ, a2 = 2 * LQ, a3 = 3 * LQ;
      for (int i = 0; i < LQ; ++i) {
        const int k0 = kp0 + i, k1 = k0 + LQ, k2 = k1 + LQ, k3 = k2 + LQ;
        const double c0 = Fs[k0 + (k0 >> 3)] + ltr[i];
        const double c1 = Fs[k1 + (k1 >> 3)] + ltr[LQ + i];
        const double c2 = Fs[k2 + (k2 >> 3)] + ltr[2 * LQ + i];
        const double c3 = Fs[k3 + (k3 >> 3)] + ltr[3 * LQ + i];
        if (c0 > b0) { b0 = c0; a0 = i; }
        if (c1 > b1) { b1 = c1; a1 = LQ + i; }
        if (c2 > b2) { b2 = c2; a2 = 2 * LQ + i; }
        if (c3 > b3) { b3 = c3; a3 = 3 * LQ + i; }
      }
      if (b1 > b0) { b0 = b1; a0 = a1; }
      if (b2 > b0) { b0 = b2; a0 = a2; }
      if (b3 > b0) { b0 = b3; a0 = a3; }
      hs[s * NBE + j] = b0;
      harg[s * NBE + j] = (unsigned short)(j - hw + a0);
    }
    }
    // this frame's sparse observations into the dense row; the previous frame's entries back to log(tiny)
    {
      const int ncp = *reinterpret_cast<const int*>(recp);
      for (int i = tid; i < ncp; i += kThreads) lob_old[rec_bins(const_cast<unsigned char*>(recp), a.cap)[i]] = a.log_tiny;
      for (int i = tid; i < nc; i += kThreads) lob_cur[rec_bins(const_cast<unsigned char*>(rec), a.cap)[i]] = rec_logp(const_cast<unsigned char*>(rec))[i];
    }
    __syncthreads();
    if constexpr (HW == 0) {
    // ---- phase B: the other targets, restricted to [argmax of the anchor below, argmax of the anchor above]
    for (int idx = tid; idx < ns; idx += kThreads) {
      const int s = idx >= nb ? 1 : 0;
      const int j = idx - s * nb;
      if ((j & (kVitStride - 1)) == 0 || j == nb - 1) continue;
      const int ja = j & ~(kVitStride - 1), jb = min(ja + kVitStride, nb - 1);
      int lo = max(j - hw, (int)harg[s * NBE + ja]), hi = min(j + hw, (int)harg[s * NBE + jb]);
      if (lo > hi) { lo = max(j - hw, 0); hi = min(j + hw, nb - 1); }
      const double* Fs = Fp + s * FS;
      double best = -1.0e308; int arg = lo;
      for (int k = lo; k <= hi; ++k) {
        const int kp = k + PAD;
        const double c = Fs[kp + (kp >> 3)] + ltr[k - j + hw];
        if (c > best) { best = c; arg = k; }
      }
      hs[s * NBE + j] = best;
      harg[s * NBE + j] = (unsigned short)arg;
    }
    __syncthreads();
    }
    // ---- phase C: combine the voicing blocks and the global fallback, add the observation, store the back-pointers
    double bv = -1.0e308, bu = -1.0e308; int bi = 0x7fffffff, biu = 0x7fffffff;
    const double fb = G + a.log_tiny;   // best out-of-band predecessor: k* through a log(tiny) transition
    unsigned short* pt = ptr0 + (size_t)t * ns;
    for (int j = tid; j < nb; j += kThreads) {
      const double h0 = hs[j], h1 = hs[NBE + j];
      const int k0 = harg[j], k1 = nb + harg[NBE + j];
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
      pt[j] = (unsigned short)pv; pt[nb + j] = (unsigned short)pu;
      if (vv > bv) { bv = vv; bi = j; }
      if (vu > bu) { bu = vu; biu = nb + j; }
      // F for the next step (phase A / B readers are done: the barrier above)
      const int fi = fidx(j);
      const double fv = vv + corr[j], fu = vu + corr[j];
      Fp[fi] = fv;
      Fp[FS + fi] = fu;
      if (HW > 0) {   // filter keys relative to the previous step's maximum (G still holds it here)
        Kq[HW + j] = max(__double2int_rd((fv - G) * 16384.0), -33554431) * 64;
        Kq[KS + HW + j] = max(__double2int_rd((fu - G) * 16384.0), -33554431) * 64;
      }
    }
    if (bu > bv) { bv = bu; bi = biu; }   // every unvoiced index is larger than every voiced one: strict >
    Gkey = G;                         // the keys just written are relative to this step's G
    block_argmax(bv, bi, G, kstar);   // contains a barrier: Fp complete before the next phase A
  }

  // ---- backtrace (warp 0; the row 8 frames ahead is pulled into L2 around the current state), then statistics over
  // voiced frames (core/feature_extractor.py:97-113)
  short* st = a.states + slot0;
  if (warp == 0) {
    int s = kstar;
    for (int t = T - 1; t >= 1; --t) {
      const int tp = t - 8;
      if (tp >= 1 && lane < 4) {
        const int jj = s >= nb ? s - nb : s;
        const int col = (lane >> 1) * nb + min(nb - 1, max(0, jj - 32 + 64 * (lane & 1)));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr0 + (size_t)tp * ns + col));
      }
      if (lane == 0) st[t] = (short)s;
      s = ptr0[(size_t)t * ns + s];
    }
    if (lane == 0) st[0] = (short)s;
  }
  __syncthreads();
  double s1 = 0, s2 = 0, cntv = 0;
  for (int t = tid; t < T; t += kThreads) {
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
    for (int w2 = 0; w2 < kWarps; ++w2) { s1 += s_red[0][w2]; s2 += s_red[1][w2]; cntv += s_red[2][w2]; }
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
