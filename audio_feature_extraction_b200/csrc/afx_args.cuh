// Kernel argument block of the fused (streaming) kernel.
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
  char* slabs;                  // ring of per-clip scratch slabs (afx_stream.cuh)
  size_t slab_bytes;
  int a_max;                    // frame rows per slab (multiple of 4)
  Tables tab;
  const float* dct_t;           // [128][n16]
  int hop;
  float coef, trim_top_db, mfcc_top_db;
  int n_mfcc, n16, summary_width, frame_width;
  unsigned mask;                // AFX_FEAT_*
  float* summaries;             // [n_clips][SS]
  int* meta;                    // [n_clips][4]
  float* frames_out;            // nullable
  int pretrimmed;
  int raw;                      // clips are already pre-emphasised and trimmed (AFX_FEAT_PREPROCESSED)
};

}  // namespace afx
