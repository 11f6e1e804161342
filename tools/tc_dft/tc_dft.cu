// Prototype: the two 32-point DFT stages of a frame pair on the 5th-generation tensor cores
// (tcgen05.mma, accumulators in TMEM) with split-precision fp16 operands, against the register FFT the fused kernel
// uses (packed FP32 butterflies + one shared-memory transpose).  Both kernels do the same job per step:
//     z <- DFT_1024(z) / 32       (z = 1024 complex float32 = two real frames packed as re / im)
// from registers to registers, R times, so the measurement is the DFT core and not the I/O around it.
//
//   stage 1   Y[(p, n2), (k1, c)] = sum_{n1, c'} z_p[32 n1 + n2][c'] * W32r[(n1, c'), (k1, c)]     M = 128 rows = 4 pairs x 32
//   twiddle   Y'[p][k1, n2] = Y * W_1024^(n2 k1)                                                    (registers, from TMEM)
//   stage 2   Z[(p, k1), (k2, c)] = sum_{n2, c'} Y'_p[k1, n2][c'] * W32r[(n2, c'), (k2, c)]        K = 64, N = 64
// Each product runs as three fp16 MMAs, A_hi B_hi + A_lo B_hi + A_hi B_lo (x = hi + lo with hi = fp16(x)): 22 mantissa
// bits, the error floor a single-pass fp16 / TF32 product lacks.  Operands are staged in shared memory in the canonical
// no-swizzle K-major layout of the UMMA descriptor (8 x 16 B core matrices).
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o tc_dft tc_dft.cu
// Run:   ./tc_dft [pairs] [repeats]   -> one JSON line per kernel: ms, pairs/s, max error against a float64 DFT
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "../../audio_feature_extraction_b200/csrc/fft_gen.cuh"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

namespace {

constexpr int kPairs = 4;                 // frame pairs per CTA tile: M = 4 x 32 = 128 rows
constexpr int kLBO = 2048 + 16;           // bytes between K chunks (8 elements) of an operand tile; +16: conflict-free scatter
constexpr int kSBO = 128;                 // bytes between 8-row groups
constexpr int kATile = 8 * kLBO;          // one [128 x 64] fp16 operand tile
constexpr int kBTile = 8 * (1024 + 16);   // one [64 x 64] fp16 operand tile (LBO 1040)
constexpr int kBLBO = 1024 + 16;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// UMMA shared-memory descriptor, no swizzle, K-major: start >> 4 | LBO >> 4 << 16 | SBO >> 4 << 32 | version 1 << 46
__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
// instruction descriptor, kind::f16: D = F32 (bit 4), A = B = F16 (0), K-major both, N >> 3 at bit 17, M >> 4 at bit 24
constexpr uint32_t kIdesc = (1u << 4) | ((64u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      :: "r"(tmem_d), "l"(da), "l"(db), "r"(kIdesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}"
      :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 8 consecutive TMEM columns of this thread's lane
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr));
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// x = hi + lo, both fp16 (packed pairs: two values per call)
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
  const __half2 h = __floats2half2_rn(a, b);
  const float2 back = __half22float2(h);
  const __half2 l = __floats2half2_rn(a - back.x, b - back.y);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = *reinterpret_cast<const uint32_t*>(&l);
}

// ---- tensor-core kernel: one CTA = 128 threads = 4 frame pairs; thread (p, n2) owns row 32 p + n2
__global__ void __launch_bounds__(128, 2) tc_dft_kernel(const float2* __restrict__ zin, float2* __restrict__ zout,
                                                         const __half* __restrict__ w_hi, const __half* __restrict__ w_lo,
                                                         const float2* __restrict__ tw1024, int n_tiles, int repeats) {
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* sA_hi = smem;                       // [128 x 64] fp16, stage operand (reused by both stages)
  unsigned char* sA_lo = sA_hi + kATile;
  unsigned char* sB_hi = sA_lo + kATile;             // [64 x 64] W32 in real form, N x K, K-major
  unsigned char* sB_lo = sB_hi + kBTile;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sB_lo + kBTile);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // B tiles: element (n, k) at (k / 8) * LBO + (n / 8) * 128 + (n % 8) * 16 + (k % 8) * 2
  for (int i = tid; i < 64 * 64; i += 128) {
    const int n = i >> 6, k = i & 63;
    const int off = (k >> 3) * kBLBO + (n >> 3) * kSBO + (n & 7) * 16 + (k & 7) * 2;
    *reinterpret_cast<__half*>(sB_hi + off) = w_hi[i];
    *reinterpret_cast<__half*>(sB_lo + off) = w_lo[i];
  }
  if (tid == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(tmem_slot)), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;
  unsigned parity = 0;

  const int p = warp, n2 = lane, m = tid;
  // twiddles of this row: W_1024^(n2 k1), k1 = 0..31 -> registers would be 64; read from the L1-resident table instead
  const float2* twr = tw1024 + n2 * 32;
  const uint32_t a_hi = smem_u32(sA_hi), a_lo = smem_u32(sA_lo), b_hi = smem_u32(sB_hi), b_lo = smem_u32(sB_lo);

  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    float2 z[32];   // z[n1] = z_p[32 n1 + n2]
#pragma unroll
    for (int n1 = 0; n1 < 32; ++n1) z[n1] = zin[((size_t)tile * kPairs + p) * 1024 + 32 * n1 + n2];
    for (int rep = 0; rep < repeats; ++rep) {
#pragma unroll
      for (int stage = 0; stage < 2; ++stage) {
        // ---- operand A of this stage from the registers.  Stage 1: row m = (p, n2), column k = 2 n1 + c: the thread's own
        // row, 16 B at a time.  Stage 2: row (p, k1), column 2 n2 + c: a scatter of 4-byte (re, im) pairs down a column.
        if (stage == 0) {
#pragma unroll
          for (int kc = 0; kc < 8; ++kc) {
            uint32_t h[4], l[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) split2(z[4 * kc + j].x, z[4 * kc + j].y, h[j], l[j]);
            const int off = kc * kLBO + (m >> 3) * kSBO + (m & 7) * 16;
            *reinterpret_cast<uint4*>(sA_hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(sA_lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
          }
        } else {
#pragma unroll
          for (int k1 = 0; k1 < 32; ++k1) {
            uint32_t h, l;
            split2(z[k1].x, z[k1].y, h, l);
            const int row = 32 * p + k1;
            const int off = (n2 >> 2) * kLBO + (row >> 3) * kSBO + (row & 7) * 16 + (n2 & 3) * 4;
            *reinterpret_cast<uint32_t*>(sA_hi + off) = h;
            *reinterpret_cast<uint32_t*>(sA_lo + off) = l;
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the tensor core
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t d = tmem + stage * 64;
#pragma unroll
          for (int prod = 0; prod < 3; ++prod) {   // A_hi B_hi + A_lo B_hi + A_hi B_lo
            const uint32_t aa = prod == 1 ? a_lo : a_hi, bb = prod == 2 ? b_lo : b_hi;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)          // K = 64 = 4 x 16
              umma_f16(d, umma_desc(aa + 2 * ks * kLBO, kLBO, kSBO), umma_desc(bb + 2 * ks * kBLBO, kBLBO, kSBO), (prod | ks) != 0);
          }
          umma_commit(bar);
        }
        mbar_wait(bar, parity); parity ^= 1u;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // ---- D of this stage: this thread's row (lane 32 warp + lane), 64 columns = 32 complex values
        const uint32_t taddr = tmem + ((uint32_t)(32 * warp) << 16) + stage * 64;
        float v[64];
#pragma unroll
        for (int c8 = 0; c8 < 8; ++c8) tmem_ld8(taddr + 8 * c8, v + 8 * c8);
        tmem_ld_wait();
        if (stage == 0) {   // twiddle by W_1024^(n2 k1)
#pragma unroll
          for (int k1 = 0; k1 < 32; ++k1) {
            const float2 w = twr[k1];
            z[k1] = make_float2(v[2 * k1] * w.x - v[2 * k1 + 1] * w.y, fmaf(v[2 * k1], w.y, v[2 * k1 + 1] * w.x));
          }
        } else {            // Z[k1 + 32 k2] / 32: the next repeat reads it as z[n1 = k2] of row n2 = k1
#pragma unroll
          for (int k2 = 0; k2 < 32; ++k2) z[k2] = make_float2(v[2 * k2] * 0.03125f, v[2 * k2 + 1] * 0.03125f);
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      }
    }
#pragma unroll
    for (int k2 = 0; k2 < 32; ++k2) zout[((size_t)tile * kPairs + p) * 1024 + 32 * k2 + n2] = z[k2];
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(128u) : "memory");
}

// ---- reference: the register FFT of the fused kernel (afx_step.cuh), one warp per frame pair, 16 warps per CTA
__global__ void __launch_bounds__(512, 1) fft_dft_kernel(const float2* __restrict__ zin, float2* __restrict__ zout,
                                                          const float2* __restrict__ tw1024, int n_pairs, int repeats) {
  extern __shared__ __align__(16) unsigned char smem[];
  using afx::c64;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  c64* P = reinterpret_cast<c64*>(smem) + warp * (32 * 33);
  for (int pair = blockIdx.x * 16 + warp; pair < n_pairs; pair += gridDim.x * 16) {
    c64 z[32];
#pragma unroll
    for (int n1 = 0; n1 < 32; ++n1) { const float2 v = zin[(size_t)pair * 1024 + 32 * n1 + lane]; z[afx::bitrev32(n1)] = afx::pk2(v.x, v.y); }
    for (int rep = 0; rep < repeats; ++rep) {
      afx::fft32_dit2(z);
#pragma unroll
      for (int k1 = 1; k1 < 32; ++k1) { const float2 w = tw1024[lane * 32 + k1]; z[k1] = afx::cmul2(z[k1], w.x, w.y); }
#pragma unroll
      for (int k = 0; k < 32; ++k) P[k * 33 + lane] = z[k];
      __syncwarp();
#pragma unroll
      for (int n = 0; n < 32; ++n) z[afx::bitrev32(n)] = P[lane * 33 + n];
      __syncwarp();
      afx::fft32_dit2(z);   // z[k2] = Z[lane + 32 k2]
      if (rep + 1 < repeats) {   // next input: z[n1 = k2] of lane n2 = k1, bit-reversed register order, scaled by 1 / 32
        c64 t[32];
#pragma unroll
        for (int k2 = 0; k2 < 32; ++k2) t[afx::bitrev32(k2)] = afx::mul2(z[k2], afx::bc2(0.03125f));
#pragma unroll
        for (int k = 0; k < 32; ++k) z[k] = t[k];
      }
    }
#pragma unroll
    for (int k2 = 0; k2 < 32; ++k2) zout[(size_t)pair * 1024 + 32 * k2 + lane] = make_float2(afx::lo2(z[k2]) * 0.03125f, afx::hi2(z[k2]) * 0.03125f);
  }
}

}  // namespace

int main(int argc, char** argv) {
  const int n_pairs = argc > 1 ? atoi(argv[1]) : 148 * 2 * 4 * 16;   // 18 944 pairs
  const int repeats = argc > 2 ? atoi(argv[2]) : 64;
  const int n_tiles = n_pairs / kPairs;
  // inputs: a strong harmonic + components 60 .. 100 dB below it + noise at -110 dB: bins far below the peak must survive
  std::vector<float2> h_in((size_t)n_pairs * 1024);
  uint64_t rng = 88172645463325252ull;
  auto unif = [&]() { rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17; return (double)(rng >> 11) / 9007199254740992.0; };
  for (int p = 0; p < n_pairs; ++p) {
    const double f1 = 20 + 400 * unif(), f2 = 30 + 450 * unif(), f3 = 5 + 500 * unif();
    for (int n = 0; n < 1024; ++n) {
      const double w = 0.54 - 0.46 * cos(2 * M_PI * n / 1024);
      const double a = w * (0.9 * sin(2 * M_PI * f1 * n / 1024 + 0.3) + 1e-3 * sin(2 * M_PI * f2 * n / 1024) + 1e-5 * sin(2 * M_PI * f3 * n / 1024) + 3e-6 * (unif() - 0.5));
      const double b = w * (0.7 * sin(2 * M_PI * f2 * n / 1024 + 1.1) + 1e-4 * sin(2 * M_PI * f1 * n / 1024) + 3e-6 * (unif() - 0.5));
      h_in[(size_t)p * 1024 + n] = make_float2((float)a, (float)b);
    }
  }
  // W32 in real form as the B operand [N = (k, c)][K = (n, c')]: out (k, 0) = sum zr wr - zi wi ; out (k, 1) = sum zr wi + zi wr
  std::vector<__half> w_hi(64 * 64), w_lo(64 * 64);
  for (int k = 0; k < 32; ++k)
    for (int n = 0; n < 32; ++n) {
      const double ang = -2 * M_PI * (double)(k * n) / 32, wr = cos(ang), wi = sin(ang);
      const double e[2][2] = {{wr, -wi}, {wi, wr}};   // [c][c']
      for (int c = 0; c < 2; ++c)
        for (int cp = 0; cp < 2; ++cp) {
          const float v = (float)e[c][cp];
          const __half hi = __float2half_rn(v);
          w_hi[(2 * k + c) * 64 + 2 * n + cp] = hi;
          w_lo[(2 * k + c) * 64 + 2 * n + cp] = __float2half_rn(v - __half2float(hi));
        }
    }
  std::vector<float2> tw(32 * 32);   // tw[n2 * 32 + k1] = W_1024^(n2 k1)
  for (int a = 0; a < 32; ++a)
    for (int b = 0; b < 32; ++b) { const double ang = -2 * M_PI * (double)(a * b) / 1024; tw[a * 32 + b] = make_float2((float)cos(ang), (float)sin(ang)); }

  float2 *d_in, *d_out, *d_tw; __half *d_whi, *d_wlo;
  CK(cudaMalloc(&d_in, h_in.size() * sizeof(float2))); CK(cudaMalloc(&d_out, h_in.size() * sizeof(float2)));
  CK(cudaMalloc(&d_tw, tw.size() * sizeof(float2))); CK(cudaMalloc(&d_whi, w_hi.size() * 2)); CK(cudaMalloc(&d_wlo, w_lo.size() * 2));
  CK(cudaMemcpy(d_in, h_in.data(), h_in.size() * sizeof(float2), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_tw, tw.data(), tw.size() * sizeof(float2), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_whi, w_hi.data(), w_hi.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_wlo, w_lo.data(), w_lo.size() * 2, cudaMemcpyHostToDevice));

  // float64 reference of DFT / 32 for the first few pairs
  const int n_check = 8;
  std::vector<double> ref((size_t)n_check * 2048);
  for (int p = 0; p < n_check; ++p)
    for (int k = 0; k < 1024; ++k) {
      double sr = 0, si = 0;
      for (int n = 0; n < 1024; ++n) {
        const double ang = -2 * M_PI * (double)((k * n) & 1023) / 1024, c = cos(ang), s = sin(ang);
        const float2 v = h_in[(size_t)p * 1024 + n];
        sr += v.x * c - v.y * s; si += v.x * s + v.y * c;
      }
      ref[((size_t)p * 1024 + k) * 2] = sr / 32; ref[((size_t)p * 1024 + k) * 2 + 1] = si / 32;
    }
  auto check = [&](const char* name) {
    std::vector<float2> out((size_t)n_check * 1024);
    CK(cudaMemcpy(out.data(), d_out, out.size() * sizeof(float2), cudaMemcpyDeviceToHost));
    double worst_abs = 0, worst_rel_weak = 0, peak = 0;
    for (int p = 0; p < n_check; ++p)
      for (int k = 0; k < 1024; ++k) peak = fmax(peak, hypot(ref[((size_t)p * 1024 + k) * 2], ref[((size_t)p * 1024 + k) * 2 + 1]));
    for (int p = 0; p < n_check; ++p)
      for (int k = 0; k < 1024; ++k) {
        const double rr = ref[((size_t)p * 1024 + k) * 2], ri = ref[((size_t)p * 1024 + k) * 2 + 1];
        const double err = hypot(out[(size_t)p * 1024 + k].x - rr, out[(size_t)p * 1024 + k].y - ri), mag = hypot(rr, ri);
        worst_abs = fmax(worst_abs, err);
        if (mag > peak * 1e-4) worst_rel_weak = fmax(worst_rel_weak, err / mag);   // bins down to 80 dB below the peak
      }
    printf("{\"kernel\": \"%s\", \"check\": \"repeat 1 vs float64 DFT, %d pairs\", \"max_abs_err_rel_to_peak\": %.3e, \"error_floor_db\": %.1f, "
           "\"max_rel_err_bins_within_80dB\": %.3e}\n", name, n_check, worst_abs / peak, 20 * log10(worst_abs / peak), worst_rel_weak);
  };
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));

  // ---- tensor-core kernel
  const size_t tc_smem = 2 * kATile + 2 * kBTile + 64;
  CK(cudaFuncSetAttribute(tc_dft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_smem));
  tc_dft_kernel<<<2 * sms, 128, tc_smem>>>(d_in, d_out, d_whi, d_wlo, d_tw, n_tiles, 1);
  CK(cudaDeviceSynchronize());
  check("tc_dft_kernel (tcgen05, fp16 x3)");
  for (int w = 0; w < 2; ++w) tc_dft_kernel<<<2 * sms, 128, tc_smem>>>(d_in, d_out, d_whi, d_wlo, d_tw, n_tiles, repeats);
  CK(cudaEventRecord(e0));
  for (int i = 0; i < 5; ++i) tc_dft_kernel<<<2 * sms, 128, tc_smem>>>(d_in, d_out, d_whi, d_wlo, d_tw, n_tiles, repeats);
  CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
  float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= 5;
  printf("{\"kernel\": \"tc_dft_kernel (tcgen05, fp16 x3)\", \"pairs\": %d, \"repeats\": %d, \"ms\": %.4f, \"pair_dfts_per_s\": %.4e, \"ms_per_4.31M_pairs\": %.3f}\n",
         n_pairs, repeats, ms, (double)n_pairs * repeats / (ms * 1e-3), 4.31e6 / ((double)n_pairs * repeats / ms));

  // ---- register FFT
  const size_t fft_smem = 16 * 32 * 33 * 8;
  CK(cudaFuncSetAttribute(fft_dft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fft_smem));
  fft_dft_kernel<<<sms, 512, fft_smem>>>(d_in, d_out, d_tw, n_pairs, 1);
  CK(cudaDeviceSynchronize());
  check("fft_dft_kernel (FFMA2 register FFT)");
  for (int w = 0; w < 2; ++w) fft_dft_kernel<<<sms, 512, fft_smem>>>(d_in, d_out, d_tw, n_pairs, repeats);
  CK(cudaEventRecord(e0));
  for (int i = 0; i < 5; ++i) fft_dft_kernel<<<sms, 512, fft_smem>>>(d_in, d_out, d_tw, n_pairs, repeats);
  CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
  CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= 5;
  printf("{\"kernel\": \"fft_dft_kernel (FFMA2 register FFT)\", \"pairs\": %d, \"repeats\": %d, \"ms\": %.4f, \"pair_dfts_per_s\": %.4e, \"ms_per_4.31M_pairs\": %.3f}\n",
         n_pairs, repeats, ms, (double)n_pairs * repeats / (ms * 1e-3), 4.31e6 / ((double)n_pairs * repeats / ms));
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
