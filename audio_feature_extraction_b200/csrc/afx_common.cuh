// Shared device-side types and helpers for the AudioFeatureExtractor hot path (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace afx {

constexpr int kNMels = 128;
constexpr int kTrimFrame = 2048;   // librosa.effects.trim defaults (core/feature_extractor.py:72)
constexpr int kTrimHop = 512;
constexpr int kDeltaWidth = 9;     // librosa.feature.delta default width (:137-138)
constexpr int kMaxMfcc = 64;

constexpr int kFramesPerChunk = 16;    // frames one CTA iteration of the frames kernel covers (8 warps x 2)
constexpr int kFramesWarps = 8;
constexpr int kStatFrames = 248;       // frames per sub-chunk of the delta/statistics phase (+8 halo = 256)
constexpr int kStatHalo = 4;

// per-clip record produced by the trim kernel and consumed by everything after it
struct ClipInfo {
  int start;        // trim start (samples, multiple of 512)
  int len;          // trimmed length L = end - start
  int n_frames;     // T = 1 + L / hop (0 when status != OK)
  int status;       // AFX_CLIP_*
  unsigned max_key; // order-preserving key of the clip-global max log-mel (power_to_db top_db clamp)
  int end;
  int pad0, pad1;
};

// read-only tables built once per handle (afx_tables.cuh)
struct Tables {
  const float*  window_half;   // [n_fft] 0.5 * periodic Hamming (0.5 folds the 1/4 of the two-real-FFT split)
  const float2* twiddle;       // [32][32] W_1024^(row*col) as (cos, -sin)
  const float2* post_twiddle;  // n_fft=2048 only: W_2048^k, k = 0..512
  const float*  mel_w;         // float2 (rising-side weight U, falling-side weight D) per entry, padded [quarter][i][lane]
  const int*    mel_k0;        // [128] first FFT bin of the (padded) run of (q, lane)'s mel segment
  int           mel_cnt[4];    // padded run length per quarter of 32 segments
  int           mel_woff[4];   // float2 offset of each quarter in mel_w
  int           mel_k00, mel_cnt0, mel_woff0;   // segment 0 (below the first mel point): first bin, run, float2 offset
  const float*  dct;           // [n_mfcc][128] ortho DCT-II rows
  // streaming kernel: linear form of the in-segment mel weights (afx_tables.h, MelSparse::lin / kl / w0)
  const float4* mel_lin;       // [128] (U0, dU, D0, dD) of (q, lane) at index 32 q + lane
  const int*    mel_kl;        // [128] first bin | bins << 16
  const float*  mel_w0;        // [mel_cnt0 rounded up to 4] rising weights of segment 0
};

__device__ __forceinline__ unsigned float_to_key(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_to_float(unsigned k) {
  unsigned u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(u);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- mbarrier + 1-D TMA bulk copy (global -> shared), sm_90+/sm_100a PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// bytes and both addresses must be multiples of 16
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gsrc, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- L2 eviction-priority hints: the input stream is read once (evict_first), the per-CTA scratch slab is
// rewritten clip after clip and must stay resident (evict_last)
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_normal() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void tma_load_1d_hint(void* smem_dst, const void* gsrc, unsigned bytes, uint64_t* bar, uint64_t policy) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ void st_slab(float* p, float v, uint64_t policy) {
  asm volatile("st.global.L2::cache_hint.f32 [%0], %1, %2;" ::"l"(p), "f"(v), "l"(policy) : "memory");
}
__device__ __forceinline__ float ld_slab(const float* p, uint64_t policy) {
  float v;
  asm volatile("ld.global.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(v) : "l"(p), "l"(policy) : "memory");
  return v;
}

}  // namespace afx
