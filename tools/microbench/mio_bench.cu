// Microbenchmark: does SHFL share the shared-memory data pipe with LDS / STS on sm_100a, and what do 32 / 64 / 128-bit
// shared-memory accesses cost per SM?  One 512-thread CTA per SM, every warp runs the same unrolled loop.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mio_bench mio_bench.cu ; prints SM-cycles per warp-instruction.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float2 lds64(const void* p) {
  float2 v; asm volatile("ld.volatile.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"((unsigned)__cvta_generic_to_shared(p)));
  return v;
}
__device__ __forceinline__ float4 lds128(const void* p) {
  float4 v; asm volatile("ld.volatile.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"((unsigned)__cvta_generic_to_shared(p)));
  return v;
}
constexpr int ITERS = 2048;
constexpr int U = 8;

template <int MODE>
__global__ void __launch_bounds__(512, 1) bench(float* out, long long* cycles, int seed) {
  __shared__ __align__(16) float sm[512 * 8 + 64];
  const int tid = threadIdx.x, lane = tid & 31;
  for (int i = tid; i < 512 * 8 + 64; i += 512) sm[i] = (float)(i + seed);
  __syncthreads();
  float acc[U];
  float4 acc4[U];
#pragma unroll
  for (int u = 0; u < U; ++u) { acc[u] = (float)(tid + u); acc4[u] = make_float4(0, 0, 0, 0); }
  const float* p32 = sm + tid;                                   // conflict-free 32-bit
  const float2* p64 = reinterpret_cast<const float2*>(sm) + tid; // conflict-free 64-bit
  const float4* p128 = reinterpret_cast<const float4*>(sm) + tid;
  const long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
    if (MODE == 0) {          // LDS.32 only
#pragma unroll
      for (int u = 0; u < U; ++u) acc[u] += *((volatile const float*)(p32 + ((it + u) & 7) * 512));
    } else if (MODE == 1) {   // SHFL only
#pragma unroll
      for (int u = 0; u < U; ++u) acc[u] = __shfl_xor_sync(0xffffffffu, acc[u], 1 + u) + 1.0f;
    } else if (MODE == 2) {   // LDS.32 + SHFL, same counts as modes 0 and 1 together
#pragma unroll
      for (int u = 0; u < U; ++u) {
        acc[u] += *((volatile const float*)(p32 + ((it + u) & 7) * 512));
        acc4[u].x = __shfl_xor_sync(0xffffffffu, acc4[u].x, 1 + u) + 1.0f;
      }
    } else if (MODE == 3) {   // LDS.64
#pragma unroll
      for (int u = 0; u < U; ++u) { const float2 v = lds64(p64 + ((it + u) & 3) * 512); acc[u] += v.x + v.y; }
    } else if (MODE == 4) {   // LDS.128
#pragma unroll
      for (int u = 0; u < U; ++u) { const float4 v = lds128(p128 + ((it + u) & 1) * 512); acc[u] += v.x + v.y + v.z + v.w; }
    } else if (MODE == 5) {   // STS.32
#pragma unroll
      for (int u = 0; u < U; ++u) *((volatile float*)(sm + tid + ((it + u) & 7) * 512)) = acc[u];
    } else if (MODE == 6) {   // LDS.32 + FFMA2-free FMA work: 4 FFMA per LDS (does math hide under the LSU?)
#pragma unroll
      for (int u = 0; u < U; ++u) {
        acc[u] += *((volatile const float*)(p32 + ((it + u) & 7) * 512));
        acc4[u].x = fmaf(acc4[u].x, 0.999f, 1.0f); acc4[u].y = fmaf(acc4[u].y, 0.999f, 1.0f);
        acc4[u].z = fmaf(acc4[u].z, 0.999f, 1.0f); acc4[u].w = fmaf(acc4[u].w, 0.999f, 1.0f);
      }
    } else if (MODE == 7) {   // broadcast LDS.128: all lanes of a half warp read the same 16 bytes
#pragma unroll
      for (int u = 0; u < U; ++u) { const float4 v = lds128(reinterpret_cast<const float4*>(sm) + (lane >> 4) + 2 * ((it + u) & 31)); acc[u] += v.x + v.y + v.z + v.w; }
    }
  }
  const long long t1 = clock64();
  float t = 0;
#pragma unroll
  for (int u = 0; u < U; ++u) t += acc[u] + acc4[u].x + acc4[u].y + acc4[u].z + acc4[u].w;
  out[blockIdx.x * 512 + tid] = t;
  if (tid == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int per_iter, float* out, long long* cyc, int sms) {
  bench<MODE><<<sms, 512>>>(out, cyc, 1);
  bench<MODE><<<sms, 512>>>(out, cyc, 2);
  cudaDeviceSynchronize();
  long long h[256];
  cudaMemcpy(h, cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < sms; ++i) avg += (double)h[i]; avg /= sms;
  // 16 warps each issue ITERS * per_iter instructions of the measured kind
  printf("{\"mode\": \"%s\", \"sm_cycles_per_warp_instr\": %.3f}\n", name, avg / (16.0 * ITERS * per_iter));
}

int main() {
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  float* out; long long* cyc;
  cudaMalloc(&out, sizeof(float) * 512 * sms); cudaMalloc(&cyc, sizeof(long long) * sms);
  run<0>("LDS.32", U, out, cyc, sms);
  run<1>("SHFL", U, out, cyc, sms);
  run<2>("LDS.32+SHFL (per pair)", U, out, cyc, sms);
  run<3>("LDS.64", U, out, cyc, sms);
  run<4>("LDS.128", U, out, cyc, sms);
  run<5>("STS.32", U, out, cyc, sms);
  run<6>("LDS.32 + 4 FFMA (per group)", U, out, cyc, sms);
  run<7>("LDS.128 broadcast per half warp", U, out, cyc, sms);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
