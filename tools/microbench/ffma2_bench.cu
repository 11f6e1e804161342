// Microbenchmark: issue rate of packed FP32 (FFMA2 / FADD2, plain and with swapped / negated halves) against scalar
// FFMA on sm_100a.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_bench ffma2_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void up(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

constexpr int ITERS = 4096;
template <int MODE>
__global__ void __launch_bounds__(256) bench(float* out, float seed) {
  const float s = seed + threadIdx.x * 1e-7f;
  if (MODE == 0) {          // 16 independent scalar FFMA chains
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = s + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], 0.999f, s);
    }
    float t = 0; for (int i = 0; i < 16; ++i) t += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
  } else if (MODE == 1) {   // 8 independent FFMA2 chains (same flops)
    u64 a[8]; const u64 m = pk(0.999f, 0.998f), c = pk(s, s + 1);
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = pk(s + i, s - i);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = fma2(a[i], m, c);
    }
    float t = 0; for (int i = 0; i < 8; ++i) { float l, h; up(a[i], l, h); t += l + h; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
  } else if (MODE == 2) {   // 16 independent FFMA2 chains (twice the flops of mode 0, same instruction count)
    u64 a[16]; const u64 m = pk(0.999f, 0.998f), c = pk(s, s + 1);
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = pk(s + i, s - i);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fma2(a[i], m, c);
    }
    float t = 0; for (int i = 0; i < 16; ++i) { float l, h; up(a[i], l, h); t += l + h; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
  } else if (MODE == 3) {   // 16 FADD2 chains with a swapped + half-negated operand (the -i butterfly)
    u64 a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = pk(s + i, s - i);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
      for (int i = 0; i < 16; ++i) { float l, h; up(a[(i + 1) & 15], l, h); a[i] = add2(a[i], pk(h, -l)); }
    }
    float t = 0; for (int i = 0; i < 16; ++i) { float l, h; up(a[i], l, h); t += l + h; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
  } else {                  // 16 scalar FADD chains
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = s + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = a[i] + a[(i + 1) & 15];
    }
    float t = 0; for (int i = 0; i < 16; ++i) t += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
  }
}
template <int MODE> void run(const char* name, double ops_per_thread_iter, float* d) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int blocks = 148 * 8, threads = 256;
  bench<MODE><<<blocks, threads>>>(d, 1.0f); cudaDeviceSynchronize();
  cudaEventRecord(e0); bench<MODE><<<blocks, threads>>>(d, 1.0f); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double inst = (double)blocks * threads / 32 * ITERS * 16 * (MODE == 1 ? 0.5 : 1.0);
  printf("%-34s %8.3f ms  %7.2f warp-inst/clk/SM (at 1.965 GHz)  %7.2f TFLOP/s\n", name, ms, inst / (ms * 1e-3 * 1.965e9 * 148),
         (double)blocks * threads * ITERS * ops_per_thread_iter / (ms * 1e-3) / 1e12);
}
int main() {
  float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  run<0>("FFMA x16", 32, d); run<1>("FFMA2 x8", 32, d); run<2>("FFMA2 x16", 64, d); run<3>("FADD2 x16 swapped/negated", 32, d); run<4>("FADD x16", 16, d);
  return 0;
}
