// Kernel argument block and per-CTA scratch slab of the fused kernel.
#pragma once
#include "afx_common.cuh"

namespace afx {

struct ClipArgs {
  const float* audio;
  const long long* offsets;     // device [n_clips] sample offsets
  const int* lengths;           // device [n_clips]
  const long long* frame_off;   // device [n_clips + 1] prefix of 1 + len/hop (rows of frames_out)
  ClipInfo* info;               // [n_clips]: input when pretrimmed, always written back
  int* counter;                 // work counter (zeroed before launch)
  int n_clips;
  char* slabs;                  // per-CTA scratch
  size_t slab_bytes;
  int a_max;                    // frame rows per slab (multiple of 4)
  Tables tab;
  const float* dct_t;           // [128][n16]
  int hop;
  int mel_floats;               // floats in Tables::mel_w (multiple of 4)
  float coef, trim_top_db, mfcc_top_db;
  int n_mfcc, n16, summary_width, frame_width;
  unsigned mask;                // AFX_FEAT_*
  float* summaries;             // [n_clips][SS]
  int* meta;                    // [n_clips][4]
  float* frames_out;            // nullable
  int pretrimmed;
  int raw;                      // clips are already pre-emphasised and trimmed (AFX_FEAT_PREPROCESSED)
};

struct Slab {
  float* mfcc;     // [a_max][n16] DCT rows (speculative: no top_db clamp; rewritten by the redo pass where it bites)
  float* rms;      // [a_max]
  float* rowmax;   // [a_max] per-frame max of the unclamped log-mel row
  float* rowmin;   // [a_max] per-frame min
  float* blk;      // [rows + 2] sums of squares of the hop-sized blocks [u*hop, (u+1)*hop) of the pre-emphasised clip
};

__device__ __forceinline__ Slab carve_slab(const ClipArgs& a, int cta) {
  Slab s;
  char* p = a.slabs + (size_t)cta * a.slab_bytes;
  s.mfcc = reinterpret_cast<float*>(p);
  s.rms = s.mfcc + (size_t)a.a_max * a.n16;
  s.rowmax = s.rms + a.a_max;
  s.rowmin = s.rowmax + a.a_max;
  s.blk = s.rowmin + a.a_max;
  return s;
}

}  // namespace afx
