// development aid: compiles ONE instantiation of the streaming kernel (seconds instead of minutes) for -Xptxas -v / cuobjdump
#include "../../audio_feature_extraction_b200/csrc/afx_stream.cuh"
#ifndef ONE_NFFT
#define ONE_NFFT 1024
#define ONE_HOP 256
#define ONE_STOCK true
#endif
template __global__ void afx::stream_kernel<ONE_NFFT, ONE_HOP, ONE_STOCK>(afx::StreamArgs);

// the warp step alone (static instruction count of one step = everything attributed to this kernel minus its short frame)
namespace afx {
template <int NFFT, int HOP, bool STOCK>
__global__ void __launch_bounds__(512, 1) step_only_kernel(ClipArgs a, Slab2 sl, Smem sm, MelLane ml, float* tile_base, int steps, int L0) {
  WarpCtx2 wc;
  wc.tile = tile_base + threadIdx.x / 32 * 4096; wc.planeB = wc.tile + 2048; wc.mbar = sm.mbar; wc.parity = 0;
  wc.pol_in = 0; wc.pol_slab = 0;
  Region rg{0, 0, L0};
  for (int s = 0; s < steps; ++s)
    Step2<NFFT, HOP>::template run<MelCnt<NFFT, STOCK>>(a, sm, ml, sl, wc, rg, s * 512, L0, 2, s * 2, true, s * 2, -3.0e38f, 5152u, a.audio + s, wc.tile);
}
template __global__ void step_only_kernel<ONE_NFFT, ONE_HOP, ONE_STOCK>(ClipArgs, Slab2, Smem, MelLane, float*, int, int);
}
