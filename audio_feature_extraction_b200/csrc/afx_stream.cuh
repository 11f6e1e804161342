// The fused single-pass hot path, second generation: ONE persistent kernel in which every WARP is autonomous.
// Replaces, for a whole batch, preprocess_audio + extract_mfcc + extract_energy of
// audio_feature_extraction_toolkit/core/feature_extractor.py:58-74,116-151,153-179 (reference).
//
// Work is cut into ITEMS (a run of consecutive warp steps of one clip, afx_step.cuh) listed in a device-built table;
// warps claim items from one atomic counter.  There is no CTA-wide barrier after the table set-up:
//
//  S  a warp streams its item: raw samples by 1-D TMA bulk copies into its private tile (the next tile -- of this item
//     or of the item the warp has already claimed -- is in flight while the current one is transformed), MFCC rows,
//     per-frame log-mel extrema and hop-block sums of squares go to the clip's slab (a slot of an L2-resident ring).
//     The frame grid of the trimmed signal starts at trim_start, a multiple of 512; when hop divides 512 every trimmed
//     frame centre is also a hop multiple of the UNtrimmed signal, so the pass runs speculatively on the absolute grid
//     before the trim decision exists (else trim_kernel runs first and the items cover the trimmed frames: `pretrimmed`).
//  The warp that completes the LAST item of a clip (per-slot arrival counter, release/acquire through L2) finishes it
//  alone, while every other warp keeps streaming:
//  T  trim decision from the hop-block sums;  F  head / tail fix-up steps;  R  top_db clamp: steps whose log-mel
//  minimum lies below (clip max - 80 dB) are redone with the clamp;  C  delta / delta-delta + moment sums;
//  Z  RMS per frame from the block sums, statistics, summary + meta rows.  Then the slot is handed to clip + n_slots.
#pragma once
#include "afx_common.cuh"
#include "afx_trim.cuh"
#include "afx_step.cuh"

namespace afx {

constexpr int kMoments = 6;                               // sum x, x^2, x^3, x^4, sum delta, sum delta2

// raw sample range a warp step starting at absolute sample gA needs; tile position p holds sample gA - 4 + p.
// Returns the bytes to copy (0: nothing of the step lies inside [lo, hi)).
__device__ __forceinline__ unsigned pair_copy_range(const Region& rg, int gA, int span, int L0, int& g_lo, int& s_lo) {
  const int lo = max(rg.lo, gA), hi = min(rg.hi, gA + span);
  if (hi <= lo) return 0u;
  g_lo = max(0, lo - 1) & ~3;
  const int g_hi4 = (min(hi, L0) + 3) & ~3;
  s_lo = g_lo - (gA & ~3) + 4;                                      // multiple of 4, >= 0 (gA itself is one unless the hop is a run-time value)
  return (unsigned)(g_hi4 - g_lo) * 4u;
}

// ---- T: librosa.effects.trim(top_db) from the block sums (core/feature_extractor.py:72)
//   r_i = sqrt(mean(y_pre[512 i - 1024 .. 512 i + 1024)^2)), zero padded, i = 0 .. L0/512
//   keep frames with 10 log10(max(1e-10, r_i^2)) - 10 log10(max(1e-10, max_i r_i^2)) > -top_db
//   start = 512 * first, end = min(L0, 512 * (last + 1))
__device__ __forceinline__ float trim_block(const float* usum, int b, int nb, int R, int nu) {
  // sum of squares of 512-sample block b from the hop-sized unit sums (fixed order); zero outside the clip
  if (b < 0 || b >= nb) return 0.f;
  float s = 0.f;
  for (int i = 0; i < R; ++i) { const int u = b * R + i; if (u < nu) s += __ldcg(usum + u); }
  return s;
}
constexpr int kStreamMaxWarps = 16;
constexpr int kMelW0Max = 64;          // floats reserved for the rising weights of mel segment 0 (bins below the first mel point)
constexpr int kStreamWarpSlots = 16;   // per-CTA mbarrier / scratch slots

struct __align__(16) WorkItem {       // 32 bytes, built on the device (item_build_kernel)
  long long audio_off;                // sample offset of the clip in the packed buffer
  int L0;                             // clip length
  int clip;
  int f_begin;                        // first frame of the item (absolute grid; trimmed grid when pretrimmed)
  int lo, hi;                         // sample range of the signal ([0, L0) or the trim range when pretrimmed)
  unsigned nitems_cnt;                // items of this clip << 8 | frames in this item
};

struct __align__(16) WarpScratch {    // per-warp shared-memory state that need not live in registers across a step
  WorkItem item[2];                   // current / next table entry (TMA destinations)
  Slab2 sl;                           // slab rows of the current clip
  int pad[2];
};

struct StreamArgs {
  ClipArgs c;
  const WorkItem* items;
  int n_items;
  int n_slots;
  int* slot_owner;                    // [n_slots] clip that may use the slot (initially the slot index)
  int* slot_done;                     // [n_slots] items of the owner that have been completed
  int mel_w0_floats;                  // floats of Tables::mel_w0 (multiple of 4)
  int l2_mode;                        // bit 0: input tiles evict_normal instead of evict_first; bit 1: slab rows evict_normal instead of evict_last
};

// one thread per clip: the clip's items (item_off: host-built prefix; frames_per_item: multiple of the step width)
__global__ void item_build_kernel(StreamArgs sa, const int* __restrict__ item_off, int frames_per_item) {
  const ClipArgs& a = sa.c;
  const int clip = blockIdx.x * blockDim.x + threadIdx.x;
  if (clip < sa.n_slots) { sa.slot_owner[clip] = clip; sa.slot_done[clip] = 0; }
  if (clip >= a.n_clips) return;
  const int L0 = a.lengths[clip];
  int T = L0 > 0 ? 1 + L0 / a.hop : 0, lo = 0, hi = L0;
  if (a.pretrimmed) {
    const ClipInfo ci = a.info[clip];
    T = (ci.status == 0 && (a.mask & 0x6u)) ? ci.n_frames : 0;
    lo = ci.start; hi = ci.end;
  }
  const int i0 = item_off[clip], ni = item_off[clip + 1] - i0;
  WorkItem it;
  it.audio_off = a.offsets[clip]; it.L0 = L0; it.clip = clip; it.lo = lo; it.hi = hi;
  for (int j = 0; j < ni; ++j) {
    const long long fb = (long long)j * frames_per_item;
    const int b = (int)(fb < T ? fb : T), e = (int)(fb + frames_per_item < T ? fb + frames_per_item : T);
    it.f_begin = b;
    it.nitems_cnt = ((unsigned)ni << 8) | (unsigned)(e - b);
    const_cast<WorkItem*>(sa.items)[i0 + j] = it;
  }
}

__device__ __forceinline__ Slab2 carve_slab2(const ClipArgs& a, int slot) {
  Slab2 s;
  char* p = a.slabs + (size_t)slot * a.slab_bytes;
  s.mfcc = reinterpret_cast<float*>(p);
  s.ext = s.mfcc + (size_t)a.a_max * a.n16;
  s.blk = s.ext + 2 * (size_t)a.a_max;
  return s;
}

// ---- T: librosa.effects.trim(top_db) from the block sums, one warp (same arithmetic as trim_kernel)
__device__ __forceinline__ ClipInfo trim_decide_warp(const float* usum, int nu, int L0, int hop, float top_db, int min_frames, int lane) {
  const int n_rms = 1 + L0 / kTrimHop;
  const int nb = (L0 + kTrimHop - 1) / kTrimHop;
  const int R = kTrimHop / hop;
  float mx = 0.f;
  int bad = 0;
  for (int i = lane; i < n_rms; i += 32) {
    const float s = (trim_block(usum, i - 2, nb, R, nu) + trim_block(usum, i - 1, nb, R, nu)) +
                    (trim_block(usum, i, nb, R, nu) + trim_block(usum, i + 1, nb, R, nu));
    bad |= !isfinite(s);     // a NaN / Inf sample poisons its block sum (librosa.util.valid_audio)
    mx = fmaxf(mx, sqrtf(s * (1.0f / kTrimFrame)));
  }
  mx = warp_max(mx);
  bad = __any_sync(0xffffffffu, bad);
  const float ref_db = 10.0f * log10f(fmaxf(1e-10f, mx * mx));
  int first = 0x7fffffff, last = -1;
  for (int i = lane; i < n_rms; i += 32) {
    const float s = (trim_block(usum, i - 2, nb, R, nu) + trim_block(usum, i - 1, nb, R, nu)) +
                    (trim_block(usum, i, nb, R, nu) + trim_block(usum, i + 1, nb, R, nu));
    const float r = sqrtf(s * (1.0f / kTrimFrame));
    float db = 10.0f * log10f(fmaxf(1e-10f, r * r));
    db -= ref_db;
    if (db > -top_db) { first = min(first, i); last = max(last, i); }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
    last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
  }
  ClipInfo ci;
  ci.start = 0; ci.end = 0; ci.len = 0; ci.n_frames = 0; ci.max_key = 0u; ci.pad0 = ci.pad1 = 0; ci.status = 0;
  if (L0 <= 0) ci.status = 1;                 // AFX_CLIP_EMPTY
  else if (bad) ci.status = 2;                // AFX_CLIP_NONFINITE
  else if (last < 0) ci.status = 1;           // every frame silent -> empty signal
  else {
    ci.start = first * kTrimHop;
    ci.end = min(L0, (last + 1) * kTrimHop);
    ci.len = ci.end - ci.start;
    const int T = 1 + ci.len / hop;
    if (ci.len <= 0) ci.status = 1;
    else if (T < min_frames) ci.status = 3;   // AFX_CLIP_TOO_SHORT: only librosa.feature.delta (MFCC group) needs 9 frames
    else ci.n_frames = T;
  }
  return ci;
}

// ---- C: delta / delta2 (SavGol 9) -> moments of coefficients [kg, kg + 16) for trimmed frames [0, T), one warp.
// Lane = (coefficient c = lane & 15, half h = lane >> 4): half h takes the frames f = h (mod 2), so a lane carries its
// six sums across the whole clip and the halves meet in one shuffle at the end (no reduction per sub-chunk).  sM
// [stat_rows][16] floats is the warp's own buffer (row pitch 16: the two halves read adjacent rows = disjoint banks).
// The 9-tap filters run in float32: scipy accumulates them in float64 and rounds to float32; evaluating in float32
// moves a delta by <= a few ulp of the largest tap product (~1e-5 for |mfcc| ~ 500), far inside the 1e-3 tolerance.
// Lanes c < 16 of half 0 return the six sums of coefficient kg + c (valid where kg + c < n_mfcc).
__device__ __forceinline__ void mfcc_phase_warp(const ClipArgs& a, const float* mfcc_rows, int a0, int T, long long out_row0,
                                                int kg, float* sM, int stat_frames, double (&acc)[kMoments], uint64_t pol, int lane) {
  const int c = lane & 15, h = lane >> 4;
  const bool on = kg + c < a.n_mfcc;
  double s1 = 0, s2 = 0, s3 = 0, s4 = 0;
  float sd1 = 0.f, sd2 = 0.f;   // running sums of the float32 deltas (a few hundred terms of magnitude <~ 10)
  double sd1d = 0, sd2d = 0;
  for (int f_lo = 0; f_lo < T; f_lo += stat_frames) {
    const int f_hi = min(f_lo + stat_frames, T);             // exclusive
    const int e_lo = min(max(f_lo, kStatHalo), T - 1 - kStatHalo) - kStatHalo;
    const int e_hi = min(max(f_hi - 1, kStatHalo), T - 1 - kStatHalo) + kStatHalo;   // inclusive
    const int ne = e_hi - e_lo + 1;                           // <= stat_frames + 8
    __syncwarp();
    {
      // rows of 16 coefficients (64 B): four lanes per row, 128-bit L2 loads, eight rows per warp instruction
      const float* src = mfcc_rows + (size_t)(a0 + e_lo) * a.n16 + kg;
#pragma unroll 4
      for (int i = lane; i < ne * 4; i += 32)
        *reinterpret_cast<float4*>(sM + 4 * i) = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(i >> 2) * a.n16) + (i & 3));
    }
    __syncwarp();
    const float* col = sM + c;
    for (int f = f_lo + ((f_lo ^ h) & 1); f < f_hi; f += 2) {
      const int te = min(max(f, kStatHalo), T - 1 - kStatHalo) - e_lo;   // centre row of the 9-tap window
      const float xm4 = col[(te - 4) * 16], xm3 = col[(te - 3) * 16], xm2 = col[(te - 2) * 16], xm1 = col[(te - 1) * 16],
                  x0 = col[te * 16], xp1 = col[(te + 1) * 16], xp2 = col[(te + 2) * 16], xp3 = col[(te + 3) * 16],
                  xp4 = col[(te + 4) * 16];
      // savgol(9, polyorder = deriv = 1): sum k x[t+k] / 60 ; (9, 2, 2): sum c_k x[t+k] / 462
      const float d1f = fmaf(4.0f / 60.0f, xp4 - xm4, fmaf(3.0f / 60.0f, xp3 - xm3, fmaf(2.0f / 60.0f, xp2 - xm2, (1.0f / 60.0f) * (xp1 - xm1))));
      const float d2f = fmaf(28.0f / 462.0f, xp4 + xm4, fmaf(7.0f / 462.0f, xp3 + xm3, fmaf(-8.0f / 462.0f, xp2 + xm2,
                        fmaf(-17.0f / 462.0f, xp1 + xm1, (-20.0f / 462.0f) * x0))));
      const float xf = col[(f - e_lo) * 16];
      const double x = xf, x2 = x * x;
      s1 += x; s2 += x2; s3 += x2 * x; s4 += x2 * x2;
      sd1 += d1f; sd2 += d2f;
      if (a.frames_out && on) {
        float* fr = a.frames_out + (size_t)(out_row0 + f) * a.frame_width;
        fr[kg + c] = xf; fr[a.n_mfcc + kg + c] = d1f; fr[2 * a.n_mfcc + kg + c] = d2f;
      }
    }
    sd1d += (double)sd1; sd2d += (double)sd2; sd1 = 0.f; sd2 = 0.f;   // fold per sub-chunk: keeps the float sums short
  }
  acc[0] = s1 + __shfl_xor_sync(0xffffffffu, s1, 16); acc[1] = s2 + __shfl_xor_sync(0xffffffffu, s2, 16);
  acc[2] = s3 + __shfl_xor_sync(0xffffffffu, s3, 16); acc[3] = s4 + __shfl_xor_sync(0xffffffffu, s4, 16);
  acc[4] = sd1d + __shfl_xor_sync(0xffffffffu, sd1d, 16); acc[5] = sd2d + __shfl_xor_sync(0xffffffffu, sd2d, 16);
  __syncwarp();
}

// ---- the kernel.  STOCK: the mel tables carry the stock run lengths of this frame length (MelCnt<NFFT, true>)
template <int NFFT, int HOP, bool STOCK = false, int MAXW = kStreamMaxWarps>
__global__ void __launch_bounds__(MAXW * 32, 1) stream_kernel(StreamArgs sa) {
  const ClipArgs& a = sa.c;
  using G = Geo<NFFT, HOP>;
  constexpr int FPW = G::FPW;
  constexpr int WB = WarpBuf<NFFT, HOP>::kBytes;
  const int hop = HOP ? HOP : a.hop;                    // HOP == 0: any hop <= n_fft / 2, trim done beforehand (pretrimmed)
  const int span = (FPW - 1) * hop + NFFT;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // layout: [16 tile mbarriers][16 item mbarriers][16 warp scratch blocks][twiddle 8 KB][W_2048 table]
  //         [DCT rows per lane][mel lane tables][per-warp buffers]
  Smem sm;
  sm.mbar = reinterpret_cast<uint64_t*>(smem_raw);
  constexpr int SCR = kStreamWarpSlots * (int)sizeof(WarpScratch);
  WarpScratch* s_scr = reinterpret_cast<WarpScratch*>(smem_raw + 512);
  sm.tw = reinterpret_cast<float2*>(smem_raw + 512 + SCR);
  constexpr int PTW_BYTES = NFFT == 2048 ? 4112 : 0;
  sm.ptw = reinterpret_cast<float2*>(smem_raw + 512 + SCR + 8192);
  // every table of fixed size first, the per-warp buffers next, the n16-dependent DCT rows last: all the addresses
  // the step uses are then compile-time offsets (+ warp * WB), none has to live in a register or be recomputed
  constexpr int FIXED = 512 + SCR + 8192 + PTW_BYTES;
  float4* sLin = reinterpret_cast<float4*>(smem_raw + FIXED);
  int* sKl = reinterpret_cast<int*>(smem_raw + FIXED + 2048);
  float* sW0 = reinterpret_cast<float*>(smem_raw + FIXED + 2048 + 512);
  unsigned char* wbufs = smem_raw + FIXED + 2048 + 512 + kMelW0Max * 4;
  float* sDct = reinterpret_cast<float*>(wbufs + (size_t)(blockDim.x >> 5) * WB);
  const int dstride = 2 * a.n16 + 4;
  sm.dct = sDct;
  MelLane ml;
  ml.lin = sLin; ml.kl = sKl; ml.w0 = sW0;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 1024; i += (int)blockDim.x) sm.tw[i] = a.tab.twiddle[i];
  if (NFFT == 2048) for (int i = tid; i <= 512; i += (int)blockDim.x) sm.ptw[i] = a.tab.post_twiddle[i];
  for (int i = tid; i < 64 * a.n16; i += (int)blockDim.x) {
    const int l = i / (2 * a.n16), rem = i - l * 2 * a.n16, hsel = rem / a.n16, k = rem - hsel * a.n16;
    sDct[l * dstride + rem] = a.dct_t[(hsel ? 63 - l : l) * a.n16 + k];
  }
  for (int i = tid; i < 128; i += (int)blockDim.x) { sLin[i] = a.tab.mel_lin[i]; sKl[i] = a.tab.mel_kl[i]; }
  for (int i = tid; i < sa.mel_w0_floats; i += (int)blockDim.x) sW0[i] = a.tab.mel_w0[i];
  if (tid < 2 * kStreamWarpSlots) mbar_init(sm.mbar + tid, 1);
  if (tid == 0) fence_mbar_init();
  __syncthreads();
  // ---- from here on every warp is on its own

  WarpCtx2 wc;
  wc.tile = reinterpret_cast<float*>(wbufs + (size_t)warp * WB);
  wc.planeB = reinterpret_cast<float*>(wbufs + (size_t)warp * WB + (WB - kPlaneBFloats * 4));
  wc.mbar = sm.mbar + warp;
  wc.parity = 0;
  wc.pol_in = (sa.l2_mode & 1) ? l2_policy_evict_normal() : l2_policy_evict_first();
  wc.pol_slab = (sa.l2_mode & 2) ? l2_policy_evict_normal() : l2_policy_evict_last();
  uint64_t* mb_item = sm.mbar + kStreamWarpSlots + warp;
  unsigned par_item = 0;
  WarpScratch& scr = s_scr[warp];
  WorkItem* my_items = scr.item;
  const int n = a.n_mfcc;

  auto claim = [&]() -> int {
    int v = 0;
    if (lane == 0) v = atomicAdd(a.counter, 1);
    return __shfl_sync(0xffffffffu, v, 0);
  };
  auto fetch_item = [&](int idx, int buf) {   // async copy of one table entry into this warp's slot `buf`
    if (lane == 0) {
      fence_proxy_async();
      mbar_expect_tx(mb_item, (unsigned)sizeof(WorkItem));
      tma_load_1d(my_items + buf, sa.items + idx, (unsigned)sizeof(WorkItem), mb_item);
    }
  };

  int cur = claim();
  if (cur >= sa.n_items) return;
  fetch_item(cur, 0);
  int buf = 0;
  bool item_in_flight = true;    // the entry of `cur` has been requested but not waited for
  bool tile_in_flight = false;   // the first tile of `cur` was prefetched by the previous item's last step
  unsigned tile_bytes = 0;

  for (;;) {   // ---- one item per iteration
    if (item_in_flight) { mbar_wait(mb_item, par_item); par_item ^= 1u; item_in_flight = false; }
    __syncwarp();
    const WorkItem& it = my_items[buf];
    const int clip = it.clip, L0 = it.L0;
    const float* clip_audio = a.audio + it.audio_off;
    const int slot = clip % sa.n_slots;
    if (lane == 0) scr.sl = carve_slab2(a, slot);
    __syncwarp();
    const Slab2& sl = scr.sl;
    const int nxt = claim();     // the item after this one; its entry is requested after the first step
    bool nxt_requested = false;
    {  // the slot belongs to `clip` once the finisher of clip - n_slots has released it
      if (lane == 0) {
        const volatile int* ow = sa.slot_owner + slot;
        while (*ow != clip) __nanosleep(256);
      }
      __syncwarp();
      __threadfence();
    }

    // job state: pass 0 = the item (S); 1 / 2 = head / tail fix-ups (F); 3 = redo with the clamp (R)
    int pass = 0;
    Region rg{it.lo, it.lo, it.hi};
    rg.raw = a.raw;
    int f_begin = it.f_begin, f_end = it.f_begin + (int)(it.nitems_cnt & 0xffu);
    int shift = 0;
    bool own = true;
    float floor_db = -3.0e38f;
    ClipInfo ci;
    ci.start = 0; ci.end = 0; ci.len = 0; ci.n_frames = 0; ci.max_key = 0u; ci.pad0 = ci.pad1 = 0; ci.status = 1;
    int a0 = 0, head_end = 0;
    const bool want_frames = (a.mask & 0x6u) != 0;
    bool fin = false;            // this warp completed the clip's last item and finishes the clip

    for (;;) {   // ---- one job per iteration: steps [f_begin, f_end) of rg, FPW frames each
      // step iterator: groups of 32 steps; in the redo pass only steps whose unclamped row minimum lies below the floor
      int g0 = f_begin;
      unsigned bits = 0u;
      auto load_group = [&]() {
        const int f = g0 + FPW * lane;
        bool v = f < f_end;
        if (pass == 3 && v) {
          float m = 3.0e38f;
          for (int j = 0; j < FPW; ++j) if (f + j < f_end) m = fminf(m, __ldcg(sl.ext + 2 * (shift + f + j) + 1));
          v = m < floor_db;
        }
        bits = __ballot_sync(0xffffffffu, v);
      };
      auto next_frame = [&]() -> int {   // pops the next step of the job; f_end when exhausted
        while (bits == 0u) {
          if (g0 + 32 * FPW >= f_end) return f_end;
          g0 += 32 * FPW;
          load_group();
        }
        const int j = __ffs(bits) - 1;
        bits &= bits - 1u;
        return g0 + FPW * j;
      };
      if (f_begin < f_end) load_group();
      int fA = f_begin < f_end ? next_frame() : f_end;
      unsigned cur_bytes = 0;
      if (fA < f_end) {
        const int gA0 = rg.origin + fA * hop - NFFT / 2;
        if (pass == 0 && tile_in_flight) {
          cur_bytes = tile_bytes;
        } else {
          int g_lo = 0, s_lo = 0;
          cur_bytes = pair_copy_range(rg, gA0, span, L0, g_lo, s_lo);
          if (lane == 0 && cur_bytes) {
            fence_proxy_async();
            mbar_expect_tx(wc.mbar, cur_bytes);
            tma_load_1d_hint(wc.tile + s_lo, clip_audio + g_lo, cur_bytes, wc.mbar, wc.pol_in);
          }
        }
      } else if (pass == 0 && tile_in_flight) {
        // cannot happen: a prefetched item has at least one step
      }
      tile_in_flight = false;
      while (fA < f_end) {
        const int gA = rg.origin + fA * hop - NFFT / 2;
        const int fN = next_frame();
        unsigned next_bytes = 0;
        const float* next_src = clip_audio;
        int s_lo = 0;
        if (fN < f_end) {
          int g_lo = 0;
          next_bytes = pair_copy_range(rg, rg.origin + fN * hop - NFFT / 2, span, L0, g_lo, s_lo);
          next_src = clip_audio + g_lo;
        } else if (pass == 0 && nxt_requested) {
          // last step of the item: the next item's first tile rides this step's prefetch slot
          if (item_in_flight) { mbar_wait(mb_item, par_item); par_item ^= 1u; item_in_flight = false; }
          __syncwarp();
          const WorkItem nx = my_items[buf ^ 1];
          if ((nx.nitems_cnt & 0xffu) != 0u) {
            Region rn{nx.lo, nx.lo, nx.hi};
            int g_lo = 0;
            next_bytes = pair_copy_range(rn, nx.lo + nx.f_begin * hop - NFFT / 2, span, nx.L0, g_lo, s_lo);
            next_src = a.audio + nx.audio_off + g_lo;
            tile_in_flight = next_bytes != 0u;
            tile_bytes = next_bytes;
          }
        }
        if (cur_bytes) { mbar_wait(wc.mbar, wc.parity); wc.parity ^= 1u; }
        Step2<NFFT, HOP>::template run<MelCnt<NFFT, STOCK>>(a, sm, ml, sl, wc, rg, gA, L0, min(FPW, f_end - fA), shift + fA, own, fA,
                                                            floor_db, next_bytes, next_src, wc.tile + s_lo);
        if (pass == 0 && !nxt_requested && nxt < sa.n_items) { fetch_item(nxt, buf ^ 1); nxt_requested = true; item_in_flight = true; }
        fA = fN; cur_bytes = next_bytes;
      }

      // ---- the job is done: what next?
      if (pass == 0) {
        __threadfence();
        int last = 0;
        if (lane == 0) last = atomicAdd(sa.slot_done + slot, 1) + 1 == (int)(it.nitems_cnt >> 8);
        last = __shfl_sync(0xffffffffu, last, 0);
        if (!last) break;
        fin = true;
        __threadfence();
        // this warp finishes the clip.  A tile prefetched for the next item would be overwritten by the fix-up
        // steps: let it land and drop it (the next item re-issues it)
        if (tile_in_flight) { mbar_wait(wc.mbar, wc.parity); wc.parity ^= 1u; tile_in_flight = false; }
        if (a.pretrimmed) {
          ci = a.info[clip];
          a0 = 0;
          pass = 2;   // no T, no F
        } else {
          const int fe0 = L0 > 0 ? 1 + L0 / hop : 0;
          const int nu = FPW * ((fe0 + FPW - 1) / FPW);   // hop-block sums written by pass 0
          ci = trim_decide_warp(sl.blk, nu, L0, hop, a.trim_top_db, (a.mask & 0x2u) ? kDeltaWidth : 1, lane);
          a0 = ci.start / hop;
        }
      }
      if (ci.status != 0 || !want_frames) break;
      rg = Region{ci.start, ci.start, ci.end};
      rg.raw = a.raw;
      shift = a0;
      own = false;
      f_begin = 0; f_end = 0;
      if (pass == 0) {          // -> head fix-up: frames t with t*hop < n_fft/2 reach in front of the trimmed signal
        pass = 1;
        if (ci.start > 0) { head_end = min(ci.n_frames, (NFFT / 2 + hop - 1) / hop); f_end = head_end; }
        continue;
      }
      if (pass == 1) {          // -> tail fix-up: frames t with t*hop + n_fft/2 > len reach behind it
        pass = 2;
        if (ci.end < L0) {
          const int t_tail = (ci.len >= NFFT / 2) ? (ci.len - NFFT / 2) / hop + 1 : 0;
          f_begin = max(t_tail, head_end); f_end = ci.n_frames;
        }
        continue;
      }
      if (pass == 2) {          // -> R: power_to_db(top_db) clamps at the clip-global maximum - top_db
        pass = 3;
        if (!(a.mask & 0x2u)) break;
        __threadfence();
        float mx = -3.0e38f, mn = 3.0e38f;
        for (int t = lane; t < ci.n_frames; t += 32) {
          const float2 e = __ldcg(reinterpret_cast<const float2*>(sl.ext) + a0 + t);
          mx = fmaxf(mx, e.x); mn = fminf(mn, e.y);
        }
        mx = warp_max(mx); mn = warp_min(mn);
        floor_db = mx - a.mfcc_top_db;
        if (mn >= floor_db) break;       // nothing is clamped: the speculative DCT rows are exact
        f_end = ci.n_frames;
        continue;
      }
      break;   // pass 3 done
    }

    if (fin) {
      // ---- this warp owns the finished clip: C, Z, then release the slot
      __threadfence();
      if (lane == 0) {
        a.info[clip] = ci;
        int* m = a.meta + (size_t)clip * 4;
        m[0] = ci.status; m[1] = ci.start; m[2] = ci.end; m[3] = ci.n_frames;
      }
      float* out = a.summaries + (size_t)clip * a.summary_width;
      for (int i = lane; i < a.summary_width; i += 32) out[i] = 0.f;
      __syncwarp();
      const int T = ci.n_frames;
      const long long out_row0 = a.frame_off[clip];
      if (T > 0) {
        const double invT = 1.0 / (double)T;
        if (a.mask & 0x2u) {
          for (int kg = 0; kg < n; kg += 16) {
            double acc[kMoments];
            mfcc_phase_warp(a, sl.mfcc, a0, T, out_row0, kg, wc.tile, ((WB / 4) / 16 - 2 * kStatHalo) & ~1, acc, wc.pol_slab, lane);
            const int c = kg + lane;
            if (lane < 16 && c < n) {
              const double s1 = acc[0], s2 = acc[1], s3 = acc[2], s4 = acc[3];
              const double mean = s1 * invT;
              const double ex2 = s2 * invT, ex3 = s3 * invT, ex4 = s4 * invT;
              const double m2 = fmax(ex2 - mean * mean, 0.0);
              const double m3 = ex3 - 3.0 * mean * ex2 + 2.0 * mean * mean * mean;
              const double m4 = ex4 - 4.0 * mean * ex3 + 6.0 * mean * mean * ex2 - 3.0 * mean * mean * mean * mean;
              out[4 + c] = (float)mean;
              out[4 + n + c] = (float)sqrt(m2);
              out[4 + 2 * n + c] = (float)(acc[4] * invT);
              out[4 + 3 * n + c] = (float)(acc[5] * invT);
              out[4 + 4 * n + 3 + c] = (m2 > 0.0) ? (float)(m3 / (m2 * sqrt(m2))) : 0.f;
              out[4 + 4 * n + 3 + n + c] = (m2 > 0.0) ? (float)(m4 / (m2 * m2) - 3.0) : 0.f;
            }
          }
        }
        if (a.mask & 0x4u) {   // RMS energy (core/feature_extractor.py:164-173): frame t covers hop blocks t - G0 .. t + G0 - 1
          constexpr int GPF = HOP ? NFFT / HOP : 1;               // run-time hop: blk[t] is the sum of squares of frame t itself
          const int ub = a.pretrimmed ? 0 : a0;                   // first block of the trimmed signal
          const int ue = HOP ? ub + (ci.len + hop - 1) / hop : T; // blocks at or past ue hold no trimmed sample
          double s1 = 0, s2 = 0, s3 = 0, s4 = 0;
          float mn = 3.0e38f, mx = -3.0e38f;
          for (int t = lane; t < T; t += 32) {
            const int u0 = ub + t - (HOP ? G::G0 : 0);
            float bs[GPF];
#pragma unroll
            for (int g = 0; g < GPF; ++g) { const int u = u0 + g; bs[g] = (u >= ub && u < ue) ? __ldcg(sl.blk + u) : 0.f; }
            float s = 0.f;
            if (GPF >= 4) {
#pragma unroll
              for (int g = 0; g + 3 < GPF; g += 4) s += (bs[g] + bs[g + 1]) + (bs[g + 2] + bs[g + 3]);
            } else {
#pragma unroll
              for (int g = 0; g < GPF; ++g) s += bs[g];
            }
            const float rf = sqrtf(s * (1.0f / NFFT));
            const double r = rf;
            s1 += r; s2 += r * r; s3 += r * r * r; s4 += (r * r) * (r * r);
            mn = fminf(mn, rf); mx = fmaxf(mx, rf);
            if (a.frames_out) a.frames_out[(size_t)(out_row0 + t) * a.frame_width + 3 * n] = rf;
          }
          s1 = warp_sum(s1); s2 = warp_sum(s2); s3 = warp_sum(s3); s4 = warp_sum(s4);
          mn = warp_min(mn); mx = warp_max(mx);
          if (lane == 0) {
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
      }
      // release the slot to clip + n_slots
      __syncwarp();
      if (lane == 0) {
        sa.slot_done[slot] = 0;
        __threadfence();
        *reinterpret_cast<volatile int*>(sa.slot_owner + slot) = clip + sa.n_slots;
      }
      __syncwarp();
    }

    // ---- next item
    if (nxt >= sa.n_items) break;
    if (!nxt_requested) { fetch_item(nxt, buf ^ 1); item_in_flight = true; }
    cur = nxt; buf ^= 1;
  }
}

}  // namespace afx
