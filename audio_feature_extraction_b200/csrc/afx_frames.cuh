// Phase B: pre-emphasis -> framing -> Hamming -> 1024-point real FFT -> power -> sparse Slaney mel
// -> 10 log10, plus the per-frame RMS, fused in one persistent kernel.
//
// Replaces, for the trimmed signal, the framing/STFT/mel/dB half of librosa.feature.mfcc and all of
// librosa.feature.rms at audio_feature_extraction_toolkit/core/feature_extractor.py:127-134,164-168.
//
// Work unit ("chunk") = 16 consecutive frames of one clip.  Per chunk a CTA
//   1. receives the raw samples the 16 frames span ((16-1)*hop + n_fft + halo) by ONE 1-D TMA bulk
//      copy (cp.async.bulk + mbarrier) -- the 4x-overlapping frame matrix is never materialised,
//   2. pre-emphasises them once into a second shared buffer (zero outside the trimmed signal, so the
//      centre padding costs nothing), then immediately re-arms the TMA for its next chunk so the copy
//      overlaps the FFTs,
//   3. each of the 8 warps transforms TWO frames as one complex 1024-point FFT held in registers
//      (32 complex values per lane): 32-point in-register FFT, twiddle, transpose through shared
//      memory, second 32-point in-register FFT; the two real spectra are separated with one
//      partner exchange by warp shuffle and turned into power,
//   4. applies the sparse triangular mel filters (each lane owns 4 bands), takes 10 log10 and writes
//      128 floats per frame to the L2-resident log-mel scratch; the clip-global maximum needed by
//      power_to_db(top_db=80) is folded into ClipInfo::max_key with one atomicMax per warp.
#pragma once
#include "afx_common.cuh"
#include "fft_gen.cuh"

namespace afx {

constexpr int kTilePad = 4;       // floats in front of the first frame sample inside the tile
constexpr int kPlaneFloats = 32 * 33;

struct FramesArgs {
  const float* audio;
  const long long* offsets;    // device [n_clips]
  const int* lengths;          // device [n_clips]
  const ClipInfo* info_ro;
  ClipInfo* info;
  const int* chunk_off;        // device prefix [n_clips + 1] of ceil(T_ub / 16)
  const long long* frame_off;  // device prefix [n_clips + 1] of T_ub
  int clip_begin, clip_end;
  int n_tasks;                 // chunk_off[clip_end] - chunk_off[clip_begin]
  long long frame_base;        // frame_off[clip_begin]
  float* logmel;               // [wave frames][128]
  float* rms;                  // [wave frames]
  Tables tab;
  int hop;
  int tile_floats;             // (16-1)*hop + n_fft + 8, multiple of 4
  float coef;
};

struct ChunkTask {
  int clip;
  int f0;          // first frame of the chunk
  int start, len, T;
  int L0;
  long long goff;  // clip offset in audio
  long long frow;  // scratch row of frame 0 of this clip
};

__device__ __forceinline__ bool decode_chunk(const FramesArgs& a, int task, ChunkTask& t) {
  // binary search: largest clip with chunk_off[clip] - base <= task
  const int base = a.chunk_off[a.clip_begin];
  int lo = a.clip_begin, hi = a.clip_end;  // invariant: chunk_off[lo]-base <= task < chunk_off[hi]-base
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (a.chunk_off[mid] - base <= task) lo = mid; else hi = mid;
  }
  t.clip = lo;
  t.f0 = (task - (a.chunk_off[lo] - base)) * kFramesPerChunk;
  const ClipInfo ci = a.info_ro[lo];
  t.start = ci.start; t.len = ci.len; t.T = ci.n_frames;
  t.L0 = a.lengths[lo];
  t.goff = a.offsets[lo];
  t.frow = a.frame_off[lo] - a.frame_base;
  return t.f0 < t.T;
}

// global sample range [g_lo, g_hi4) the chunk needs and where it lands in the tile
__device__ __forceinline__ void chunk_copy_range(const ChunkTask& t, int hop, int n_fft, int& g_lo, int& g_hi4, int& s_lo) {
  const int tbase = t.f0 * hop - n_fft / 2;                       // trimmed index of tile position kTilePad
  const int span = (kFramesPerChunk - 1) * hop + n_fft;
  const int lo_idx = max(tbase, 0);
  const int hi_idx = min(tbase + span, t.len);
  g_lo = max(0, t.start + lo_idx - 1) & ~3;
  const int g_hi = min(t.L0, t.start + hi_idx);
  g_hi4 = (g_hi + 3) & ~3;
  s_lo = g_lo - t.start - tbase + kTilePad;                        // multiple of 4, >= 0
}

template <int NFFT>
__global__ void __launch_bounds__(kFramesWarps * 32, 2) frames_kernel(FramesArgs a);

template <>
__global__ void __launch_bounds__(kFramesWarps * 32, 2) frames_kernel<1024>(FramesArgs a) {
  constexpr int NFFT = 1024;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // layout: [mbar 16B][twiddle 8KB][window 4KB][raw tile][y tile][planes 8 x 4224B]
  uint64_t* mbar = reinterpret_cast<uint64_t*>(smem_raw);
  float2* s_tw = reinterpret_cast<float2*>(smem_raw + 16);
  float* s_win = reinterpret_cast<float*>(smem_raw + 16 + 8192);
  float* s_rawt = s_win + NFFT;
  float* s_y = s_rawt + a.tile_floats;
  float* s_planes = s_y + a.tile_floats;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float* plane = s_planes + warp * kPlaneFloats;

  for (int i = tid; i < 1024; i += blockDim.x) s_tw[i] = a.tab.twiddle[i];
  for (int i = tid; i < NFFT; i += blockDim.x) s_win[i] = a.tab.window_half[i];
  if (tid == 0) { mbar_init(mbar, 1); fence_mbar_init(); }
  __syncthreads();

  const int hop = a.hop;
  const int n_tile = (kFramesPerChunk - 1) * hop + NFFT + kTilePad;   // positions the frames read

  // prologue: arm the first tile
  int task = blockIdx.x;
  ChunkTask cur;
  bool cur_valid = false;
  if (task < a.n_tasks) cur_valid = decode_chunk(a, task, cur);
  if (tid == 0 && task < a.n_tasks && cur_valid) {
    int g_lo, g_hi4, s_lo;
    chunk_copy_range(cur, hop, NFFT, g_lo, g_hi4, s_lo);
    const unsigned bytes = (unsigned)(g_hi4 - g_lo) * 4u;
    mbar_expect_tx(mbar, bytes);
    tma_load_1d(s_rawt + s_lo, a.audio + cur.goff + g_lo, bytes, mbar);
  }
  unsigned parity = 0;

  while (task < a.n_tasks) {
    const int next_task = task + gridDim.x;
    ChunkTask nxt;
    bool nxt_valid = false;
    if (next_task < a.n_tasks) nxt_valid = decode_chunk(a, next_task, nxt);

    if (cur_valid) {
      // ---- 1. wait for the tile, pre-emphasise into s_y (zero outside the trimmed signal)
      mbar_wait(mbar, parity);
      parity ^= 1u;
      {
        const int tbase = cur.f0 * hop - NFFT / 2;
        for (int s = tid; s < n_tile; s += blockDim.x) {
          const int idx = tbase + s - kTilePad;
          float y = 0.f;
          if (idx >= 0 && idx < cur.len) {
            const float xv = s_rawt[s];
            const int g = cur.start + idx;
            if (g == 0) {
              const float x1 = (cur.L0 > 1) ? s_rawt[s + 1] : 0.f;
              y = xv + (2.f * xv - x1);
            } else {
              y = __fadd_rn(xv, __fmul_rn(-a.coef, s_rawt[s - 1]));
            }
          }
          s_y[s] = y;
        }
      }
    }
    __syncthreads();   // s_y complete, raw tile free
    // ---- 2. re-arm the TMA for the next chunk (overlaps the FFTs below)
    if (tid == 0 && nxt_valid) {
      int g_lo, g_hi4, s_lo;
      chunk_copy_range(nxt, hop, NFFT, g_lo, g_hi4, s_lo);
      const unsigned bytes = (unsigned)(g_hi4 - g_lo) * 4u;
      fence_proxy_async();
      mbar_expect_tx(mbar, bytes);
      tma_load_1d(s_rawt + s_lo, a.audio + nxt.goff + g_lo, bytes, mbar);
    }

    const int tA = cur.f0 + 2 * warp;
    if (cur_valid && tA < cur.T) {
      const bool vB = (tA + 1) < cur.T;
      const float* yA = s_y + kTilePad + (2 * warp) * hop;
      const float* yB = yA + hop;
      float r[32], im[32];
      float eA = 0.f, eB = 0.f;
      // ---- 3a. load two frames (re = frame A, im = frame B), window, energy
#pragma unroll
      for (int n1 = 0; n1 < 32; ++n1) {
        const float va = yA[32 * n1 + lane];
        const float vb = yB[32 * n1 + lane];
        const float w = s_win[32 * n1 + lane];
        eA = fmaf(va, va, eA);
        eB = fmaf(vb, vb, eB);
        r[bitrev32(n1)] = w * va;
        im[bitrev32(n1)] = w * vb;
      }
      fft32_dit(r, im);
      // ---- 3b. twiddle by W_1024^(k1 * lane)
#pragma unroll
      for (int k1 = 1; k1 < 32; ++k1) {
        const float2 w = s_tw[k1 * 32 + lane];
        const float tr = r[k1] * w.x - im[k1] * w.y;
        const float ti = fmaf(r[k1], w.y, im[k1] * w.x);
        r[k1] = tr; im[k1] = ti;
      }
      // ---- 3c. transpose (lane n2, reg k1) -> (lane k1, reg n2) through the warp's plane, re then im
#pragma unroll
      for (int k1 = 0; k1 < 32; ++k1) plane[k1 * 33 + lane] = r[k1];
      __syncwarp();
#pragma unroll
      for (int n2 = 0; n2 < 32; ++n2) r[bitrev32(n2)] = plane[lane * 33 + n2];
      __syncwarp();
#pragma unroll
      for (int k1 = 0; k1 < 32; ++k1) plane[k1 * 33 + lane] = im[k1];
      __syncwarp();
#pragma unroll
      for (int n2 = 0; n2 < 32; ++n2) im[bitrev32(n2)] = plane[lane * 33 + n2];
      __syncwarp();
      fft32_dit(r, im);   // now (r[k2], im[k2]) = Z[lane + 32 k2]
      // ---- 3d. split the two real spectra, power, -> P2[k] = (|A_k|^2, |B_k|^2), k = 0..512
      float2* P2 = reinterpret_cast<float2*>(plane);
      const int src = (32 - lane) & 31;
#pragma unroll
      for (int k2 = 0; k2 <= 16; ++k2) {
        const int pk = (k2 < 16) ? 31 - k2 : 15;
        float pr = __shfl_sync(0xffffffffu, r[pk], src);
        float pi = __shfl_sync(0xffffffffu, im[pk], src);
        if (lane == 0) { pr = r[(32 - k2) & 31]; pi = im[(32 - k2) & 31]; }
        const float sr = r[k2] + pr, di = im[k2] - pi;
        const float si = im[k2] + pi, dr = r[k2] - pr;
        const float pa = fmaf(sr, sr, di * di);
        const float pb = fmaf(si, si, dr * dr);
        if (k2 < 16 || lane == 0) P2[lane + 32 * k2] = make_float2(pa, pb);
      }
      __syncwarp();
      // ---- 4. sparse mel (4 bands per lane), log, store, running max
      const long long rowA = cur.frow + tA;
      float mxv = -3.0e38f;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int band = 32 * q + lane;
        const int k0 = __ldg(a.tab.mel_k0 + band);
        const float* wq = a.tab.mel_w + a.tab.mel_woff[q] + lane;
        const int cnt = a.tab.mel_cnt[q];
        float ma = 0.f, mb = 0.f;
        for (int i = 0; i < cnt; ++i) {
          const float w = __ldg(wq + 32 * i);
          const float2 p = P2[k0 + i];
          ma = fmaf(w, p.x, ma);
          mb = fmaf(w, p.y, mb);
        }
        const float da = 3.0102999566398120f * __log2f(fmaxf(ma, 1e-10f));
        const float db = 3.0102999566398120f * __log2f(fmaxf(mb, 1e-10f));
        a.logmel[rowA * kNMels + band] = da;
        mxv = fmaxf(mxv, da);
        if (vB) { a.logmel[(rowA + 1) * kNMels + band] = db; mxv = fmaxf(mxv, db); }
      }
      mxv = warp_max(mxv);
      eA = warp_sum(eA);
      eB = warp_sum(eB);
      if (lane == 0) {
        atomicMax(&a.info[cur.clip].max_key, float_to_key(mxv));
        a.rms[rowA] = sqrtf(eA * (1.0f / NFFT));
        if (vB) a.rms[rowA + 1] = sqrtf(eB * (1.0f / NFFT));
      }
    }
    __syncthreads();   // planes / s_y reusable
    task = next_task;
    cur = nxt;
    cur_valid = nxt_valid;
  }
}

}  // namespace afx
