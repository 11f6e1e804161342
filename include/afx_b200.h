/*
 * afx_b200.h -- C ABI of the B200-native AudioFeatureExtractor hot path (libafx_b200.so).
 *
 * The reference (chiy48308/audio_feature_extraction) has no FFI layer: its hot path is the pure
 * Python class `AudioFeatureExtractor` in
 *   audio_feature_extraction_toolkit/core/feature_extractor.py
 * which hands every number to librosa.  This header is the boundary a maintainer binds instead
 * (ctypes stub: INTEGRATION.md); each entry point names the reference lines it replaces.
 *
 * Conventions: plain C types only (no torch / CUDA types in signatures; a CUDA stream is passed
 * as void*), every function returns an afx_status_t, afx_last_error() gives the message of the
 * last failure on the calling thread.  There is NO CPU fallback: every compute entry point
 * requires an sm_100a device and fails with AFX_ERR_CUDA otherwise.
 */
#ifndef AFX_B200_H_
#define AFX_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AFX_ABI_VERSION 1

typedef enum {
  AFX_OK = 0,
  AFX_ERR_INVALID_ARG = 1,   /* null pointer, unsupported frame_length, misaligned clip offset ... */
  AFX_ERR_CUDA = 2,          /* CUDA runtime failure (message carries cudaGetErrorString) */
  AFX_ERR_WORKSPACE = 3,     /* caller workspace smaller than afx_workspace_bytes() */
  AFX_ERR_UNSUPPORTED = 4
} afx_status_t;

/* feature groups: the reference's extract_f0 / extract_mfcc / extract_energy
 * (core/feature_extractor.py:76-114, :116-151, :153-179) */
#define AFX_FEAT_F0      0x1u
#define AFX_FEAT_MFCC    0x2u
#define AFX_FEAT_ENERGY  0x4u
#define AFX_FEAT_ALL     0x7u
/* modifier: the clips are outputs of preprocess_audio (:58-74) -- already pre-emphasised and trimmed -- which is what
 * the reference's extract_f0(y) / extract_mfcc(y) / extract_energy(y) receive: no filter, no trim */
#define AFX_FEAT_PREPROCESSED 0x100u

/* per-clip status written to meta[clip][0]; mirrors the librosa exceptions that make the
 * reference's extract_features raise (:211-213) and batch_process skip (:233-235) */
#define AFX_CLIP_OK          0
#define AFX_CLIP_EMPTY       1   /* nothing left after trim (all frames <= -30 dB) or zero-length input */
#define AFX_CLIP_NONFINITE   2   /* NaN/Inf sample: librosa.util.valid_audio raises */
#define AFX_CLIP_TOO_SHORT   3   /* fewer than 9 frames: librosa.feature.delta(width=9) raises */

/* Constructor arguments of AudioFeatureExtractor.__init__ (core/feature_extractor.py:10-17) plus the
 * constants the reference hard-codes at its librosa call sites. Zero-initialise, then
 * afx_config_default(). */
typedef struct {
  int32_t sr;               /* 22050 */
  int32_t frame_length;     /* 1024: n_fft of mfcc (:131), frame_length of rms (:166) and pyin (:91); 512, 1024 or 2048 */
  int32_t hop_length;       /* 256 */
  int32_t n_mfcc;           /* 13 (1..64) */
  double  f0_min;           /* librosa.note_to_hz('C2') = 65.40639132514966 (:15) */
  double  f0_max;           /* librosa.note_to_hz('C7') = 2093.004522404789 (:16) */
  float   pre_emphasis;     /* 0.97 (:17, used :69) */
  float   trim_top_db;      /* 30 (:72) */
  int32_t n_mels;           /* 128 (librosa.feature.melspectrogram default); only 128 is supported */
  float   mfcc_top_db;      /* 80 (librosa.power_to_db default inside librosa.feature.mfcc) */
  int32_t device;           /* CUDA device ordinal */
  int32_t reserved[7];
} afx_config_t;

typedef struct afx_handle_s* afx_handle_t;

void         afx_config_default(afx_config_t* cfg);
int          afx_abi_version(void);
const char*  afx_last_error(void);

/* AudioFeatureExtractor.__init__ (core/feature_extractor.py:10-39): validates the configuration,
 * builds the device tables (Hamming window, FFT twiddles, sparse Slaney mel filterbank, DCT-II
 * matrix, pYIN beta weights / transition band) on cfg->device. */
int afx_create(const afx_config_t* cfg, afx_handle_t* out);
int afx_destroy(afx_handle_t h);

/* Row widths of the outputs below.
 *   summary row (float32, SS = afx_summary_width): in the key order of extract_features (:202-207)
 *     [0] f0_mean [1] f0_std [2] f0_missing_rate [3] f0_quality            (:110-113)
 *     [4 .. 4+n)   mfcc_mean   [4+n .. 4+2n) mfcc_std                      (:147-148)
 *     [4+2n .. 4+3n) mfcc_delta_mean  [4+3n .. 4+4n) mfcc_delta2_mean      (:149-150)
 *     [4+4n] energy_mean [4+4n+1] energy_std [4+4n+2] energy_range         (:176-178)
 *     extension statistics (README.md:25-28 of the reference; biased moments, Fisher kurtosis):
 *     [4+4n+3 .. +n) mfcc_skew  [.. +n) mfcc_kurtosis  then energy_skew, energy_kurtosis
 *   frame row (float32, FW = afx_frame_width): [0..n) mfcc [n..2n) delta [2n..3n) delta2
 *     [3n] rms [3n+1] f0 (Hz, NaN when unvoiced) [3n+2] voiced probability
 *   meta row (int32 x 4): status, trim_start, trim_end, n_frames                        (:72)
 */
int afx_summary_width(afx_handle_t h);
int afx_frame_width(afx_handle_t h);

/* Number of frame rows clip i can produce at most: 1 + length/hop_length (librosa centre padding). */
int64_t afx_max_frames(afx_handle_t h, int64_t length);

/* Bytes of device scratch afx_extract needs for this batch: a header, the work-item table of the streaming kernel and a
 * ring of per-clip slabs (sized by the longest clip, kept L2-resident).  afx_workspace_min_bytes returns the same. */
size_t afx_workspace_bytes(afx_handle_t h, int64_t n_clips, const int32_t* h_lengths);
size_t afx_workspace_min_bytes(afx_handle_t h, int64_t n_clips, const int32_t* h_lengths);

/*
 * The hot path: preprocess_audio + extract_f0 + extract_mfcc + extract_energy of
 * extract_features (core/feature_extractor.py:194-199) for a whole batch of clips resident in
 * device memory -- what batch_process (:215-237) loops over, minus file decoding.
 *
 *   d_audio      device, float32 mono PCM in [-1,1] already at cfg.sr; clip i occupies
 *                d_audio[h_offsets[i] .. h_offsets[i] + h_lengths[i]).  h_offsets[i] must be a
 *                multiple of 4 samples and the buffer must be readable up to the next multiple of 4
 *                past each clip (TMA bulk copies are 16-byte granular).
 *   h_offsets    HOST, n_clips int64 sample offsets;  h_lengths: HOST, n_clips int32 lengths
 *   feature_mask AFX_FEAT_* groups to compute (others are written as 0 in the summary row)
 *   d_summaries  device float32 [n_clips, SS]
 *   d_meta       device int32   [n_clips, 4]
 *   d_frames     device float32 [sum_i afx_max_frames(len_i), FW] or NULL; clip i's rows start at
 *                row sum_{j<i} afx_max_frames(len_j); only the first meta[i][3] rows are written
 *   d_workspace  device scratch of workspace_bytes (>= afx_workspace_bytes), 256-byte aligned
 *   stream       cudaStream_t as void* (NULL = legacy default stream).  The call only enqueues work, except that the
 *                pYIN group may synchronise `stream` once to grow the handle-owned F0 scratch.
 * A handle is NOT thread-safe and serves ONE stream at a time: its pinned header staging and F0 scratch are shared by
 * all calls on it; use one handle per stream / host thread.
 */
int afx_extract(afx_handle_t h, const float* d_audio, const int64_t* h_offsets, const int32_t* h_lengths,
                int64_t n_clips, uint32_t feature_mask, float* d_summaries, int32_t* d_meta,
                float* d_frames, void* d_workspace, size_t workspace_bytes, void* stream);

/*
 * afx_extract without allocation.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same computation and same results as afx_extract(..., d_frames = NULL, ...) for the AFX_FEAT_MFCC / AFX_FEAT_ENERGY
 * groups (AFX_FEAT_PREPROCESSED allowed; the F0 group is not served: its scratch is sized per frame), but the scratch
 * the fused pass needs (batch header, work-item table, ring of per-clip slabs) is a file-scope static device array of
 * AFX_STATIC_WORKSPACE_BYTES that every GPU which loads the library holds until exit, so the caller passes only inputs
 * and outputs.  The call performs no cudaMalloc / cudaFree / cudaMallocHost and creates no event or stream; it enqueues
 * one small host->device copy of the header, one memset of the work counter and the kernels on `stream`.
 * Limits: n_clips <= AFX_STATIC_MAX_CLIPS and afx_workspace_bytes(h, n_clips, h_lengths) - 1024 <= AFX_STATIC_WORKSPACE_BYTES
 * (at the defaults 22.05 kHz / 1024 / 256 / n_mfcc 13: about 100 clips of 10 s or 35 clips of 30 s); beyond either the
 * call fails with AFX_ERR_UNSUPPORTED / AFX_ERR_WORKSPACE and a message that names the sizes.
 * All calls on one device share the array, which keeps the last call's contents (the call clears what must start at
 * zero): calls whose GPU work overlaps -- two streams, two host threads, two handles on one device -- race.
 * afx_static_workspace_capacity returns AFX_STATIC_WORKSPACE_BYTES as built.
 */
#define AFX_STATIC_WORKSPACE_BYTES (8u << 20)
#define AFX_STATIC_MAX_CLIPS 256
int afx_extract_static(afx_handle_t h, const float* d_audio, const int64_t* h_offsets, const int32_t* h_lengths,
                       int64_t n_clips, uint32_t feature_mask, float* d_summaries, int32_t* d_meta, void* stream);
size_t afx_static_workspace_capacity(void);

/*
 * preprocess_audio (core/feature_extractor.py:58-74) on its own: librosa.effects.preemphasis (:69) followed by
 * librosa.effects.trim(top_db=30) (:72).  d_out has the packing of d_audio; clip i's result is
 * d_out[h_offsets[i] + meta[i][1] .. h_offsets[i] + meta[i][2]) (zeros elsewhere inside the clip).
 * d_summaries_scratch: device float32 [n_clips, SS] scratch (the fused kernel zeroes its rows).
 */
int afx_preprocess(afx_handle_t h, const float* d_audio, const int64_t* h_offsets, const int32_t* h_lengths, int64_t n_clips,
                   float* d_out, int32_t* d_meta, float* d_summaries_scratch, void* d_workspace, size_t workspace_bytes,
                   void* stream);

/*
 * Framewise extras of the .npz feature set the reference's experiment pipeline stores
 * (04_feature_extraction_experiment/feature_extraction.py:191-292, np.savez :340-353): per frame
 *   [0] zero-crossing rate (librosa.feature.zero_crossing_rate, :279-283)  [1] its delta (:286)
 *   [2] delta of the RMS energy (:265)   [3] delta of nan_to_num(f0) (:241-244)
 *   [4] short-time energy sum(frame^2) of the UNPADDED frame t (README.md:33; librosa.util.frame, '04_.../audio_feature_
 *       extraction 2/audio_feature_extraction/feature_extractor.py:234-238'): 1 + (L - n_fft) / hop values, then zeros
 * Call right after afx_extract(..., d_frames != NULL, ...) on the same stream with the SAME d_audio, d_meta, d_frames
 * and d_workspace (the batch header afx_extract left there is reused).  d_extras: float32 [rows of d_frames, 5].
 */
#define AFX_EXTRAS_WIDTH 5
int afx_frame_extras(afx_handle_t h, const float* d_audio, int64_t n_clips, const int32_t* d_meta, const float* d_frames,
                     float* d_extras, const void* d_workspace, void* stream);

/*
 * F0 post-processing the reference advertises (README.md:12-15; recipe 04_feature_extraction_experiment/
 * feature_extractor.py:283-301) on the voiced frames of the pYIN track in d_frames: scipy.signal.medfilt(5) ->
 * scipy.signal.savgol_filter(11, 3) (>= 11 voiced frames) -> clip to [f0_min, f0_max].  Call after
 * afx_extract(AFX_FEAT_F0, d_frames != NULL) with the same d_meta / d_frames / d_workspace.
 *   d_track  float32 [rows of d_frames] or NULL: smoothed F0 at the voiced rows (other rows untouched)
 *   d_stats  float32 [n_clips, 4]: mean, std (global RMSE), local RMSE sqrt(mean(diff^2)), stability std / (mean + 1e-8)
 */
int afx_f0_postprocess(afx_handle_t h, int64_t n_clips, const int32_t* h_lengths, const int32_t* d_meta, const float* d_frames,
                       float* d_track, float* d_stats, const void* d_workspace, void* stream);

/*
 * Sample-rate conversion of a batch to the handle's `sr`: the resampling half of librosa.load(path, sr=self.sr)
 * (core/feature_extractor.py:52) for files whose native rate differs.  librosa's default res_type 'soxr_hq' is the soxr
 * C library (absent from the reference tree); this is librosa's res_type='polyphase', i.e.
 * scipy.signal.resample_poly(y, sr / g, native_sr / g) with scipy's default Kaiser(5) design, float32, bit-exact.
 *   afx_resample_length   ceil(n_in * sr / native_sr): samples clip of n_in samples resamples to (-1 on bad arguments)
 *   d_in / d_out          device float32; clip i at d_in[h_in_offsets[i] .. + h_in_lengths[i]) -> d_out[h_out_offsets[i] ..)
 * Only enqueues work on `stream` (after waiting for a previous afx_resample on the same handle).
 */
int64_t afx_resample_length(afx_handle_t h, int32_t native_sr, int64_t n_in);
int afx_resample(afx_handle_t h, const float* d_in, const int64_t* h_in_offsets, const int32_t* h_in_lengths, int64_t n_clips,
                 int32_t native_sr, float* d_out, const int64_t* h_out_offsets, void* stream);
/* host only (no device needed): the phase-ordered filter afx_resample uses for native_sr -> sr, taps[up][taps_per_phase] =
 * h_padded[p + up * j] of scipy.signal.resample_poly's padded Kaiser design; returns up * taps_per_phase (-1 on bad
 * arguments) and copies the taps when taps_capacity is large enough */
int64_t afx_resample_filter(int32_t sr, int32_t native_sr, float* taps, int64_t taps_capacity, int32_t* up, int32_t* down,
                            int32_t* taps_per_phase, int32_t* n_pre_remove);

/* host only (no device needed): which instantiation of the fused (streaming) kernel afx_create picks for
 * (sr, frame_length, hop_length), the reference's constructor arguments (core/feature_extractor.py:10-17):
 *    1 = the stock variant (compile-time run lengths of the mel segment walk: 22.05 kHz / 1024 / 256, 16 kHz / 512 / 128,
 *        44.1 kHz / 2048 / 512),
 *    0 = the table-driven variant (hop = frame_length / 2, / 4, / 8 at any sample rate),
 *    2 = the run-time-hop variant (any other hop_length <= frame_length / 2; every feature group: the pYIN front end
 *        then forms one autocorrelation block per frame instead of sharing hop-sized blocks between frames),
 *   -1 = no kernel (frame_length other than 512 / 1024 / 2048, or hop_length > frame_length / 2).
 * mel_runs (nullable) receives the five run lengths of the mel tables built for (sr, frame_length): quarters 0..3 and
 * segment 0. */
int afx_clip_kernel_variant(int32_t sr, int32_t frame_length, int32_t hop_length, int32_t* mel_runs);

/* diagnostic aid: bytes of the batch header at the front of a workspace for n_clips clips (offsets, lengths, prefixes,
 * per-clip records, work counters); the work-item table and the slab ring follow it */
size_t afx_debug_header_bytes(int64_t n_clips);

/*
 * DTW stage downstream of the extractor, batched over pairs of feature sequences, float64:
 *   afx_dtw_frame_distance    FastDTWAligner.compute_distance_matrix, 05_dtw_alignment_experiment/dtw_alignment.py:532-552
 *                             (0.6 euclid - 0.25 (1 + cos) cos - 0.15 (1 + |corr|) corr; without the serial positional
 *                             penalties of _apply_local_constraints :559-668)
 *   afx_dtw_accumulated_cost  FastDTWAligner._compute_accumulated_cost :930-969 (bit-exact operation order); the DTW
 *                             distance of align_features :847 is the last element of each matrix
 * Pair p: x rows at d_x[h_x_off[p] ..) as [h_n[p], dim] row-major, y at d_y[h_y_off[p] ..) as [h_m[p], dim], matrices
 * at d_D / d_A[h_D_off[p] ..) as [h_n[p], h_m[p]] row-major (offsets in elements).  Only enqueue work on `stream`.
 */
int afx_dtw_frame_distance(afx_handle_t h, const double* d_x, const int64_t* h_x_off, const int32_t* h_n, const double* d_y,
                           const int64_t* h_y_off, const int32_t* h_m, int32_t dim, int64_t n_pairs, double* d_D,
                           const int64_t* h_D_off, void* stream);
int afx_dtw_accumulated_cost(afx_handle_t h, const double* d_D, const int64_t* h_D_off, const int32_t* h_n, const int32_t* h_m,
                             int64_t n_pairs, double* d_A, void* stream);

/*
 * Same computation from HOST memory (what a caller holding decoded files has): clips are staged
 * through pinned buffers, copied host->device in waves on a copy stream that overlaps the compute
 * stream, and the summaries / meta rows are copied back.  Blocks until the results are in
 * h_summaries / h_meta.  h_clips[i] points at clip i's float32 samples (any alignment).
 */
int afx_extract_host(afx_handle_t h, const float* const* h_clips, const int32_t* h_lengths, int64_t n_clips,
                     uint32_t feature_mask, float* h_summaries, int32_t* h_meta);

/* sample formats of afx_extract_host_packed */
#define AFX_SAMPLE_F32 0   /* float32 in [-1, 1] */
#define AFX_SAMPLE_S16 1   /* 16-bit PCM as stored in a WAV data chunk; converted on the device as
                              x / 32768, the libsndfile scaling librosa.load applies (:52) */

/*
 * Host path without the gather: the caller already holds the clips in ONE host buffer (ideally
 * pinned -- cudaHostRegister / torch pin_memory -- so the H2D copies run at PCIe speed and truly
 * overlap the kernels).  Clip i occupies h_audio[h_offsets[i] .. + h_lengths[i]) in units of
 * samples; offsets must be ascending, non-overlapping multiples of 8 samples.  h_audio_samples is
 * the buffer's size in samples.  Copies waves of clips on a copy stream, double buffered against
 * the compute stream; blocks until h_summaries [n_clips, SS] / h_meta [n_clips, 4] are filled.
 * This is the call batch_process (core/feature_extractor.py:215-237) makes once per directory.
 */
int afx_extract_host_packed(afx_handle_t h, const void* h_audio, int sample_format, int64_t h_audio_samples,
                            const int64_t* h_offsets, const int32_t* h_lengths, int64_t n_clips,
                            uint32_t feature_mask, float* h_summaries, int32_t* h_meta);

/*
 * Host ingest for batch_process (core/feature_extractor.py:215-237): the decode half of librosa.load (:52) for a list of
 * RIFF/WAVE files on `threads` host threads (<= 0: all cores).  PCM 8/16/24/32 and IEEE float32/64, mono mix by channel
 * mean, no resampling.  Two steps so the caller can size ONE packed (pinned) buffer for afx_extract_host_packed:
 *   afx_wav_probe_batch  headers only: sample rate, channels, frames per channel, format = (format tag << 8) | bits
 *   afx_wav_load_batch   decodes file i into h_out[h_offsets[i] .. + h_frames[i]) as float32 (AFX_SAMPLE_F32), or copies
 *                        its raw samples (AFX_SAMPLE_S16; 16-bit mono PCM files only -- converted on the device later)
 * status[i]: 0 ok, 1 cannot open, 2 not RIFF/WAVE, 3 fmt/data chunk missing, 4 unsupported encoding,
 *            5 length differs from h_frames[i], 6 not 16-bit mono PCM (S16 requested).  Pure host code (no GPU needed).
 */
int afx_wav_probe_batch(const char* const* paths, int64_t n, int threads, int32_t* sr, int32_t* channels, int64_t* frames,
                        int32_t* format, int32_t* status);
int afx_wav_load_batch(const char* const* paths, int64_t n, int threads, int sample_format, void* h_out,
                       const int64_t* h_offsets, const int64_t* h_frames, int32_t* status);

/* Per-phase device-time accounting, for bench.py's roofline line: when enabled, every kernel launch
 * of afx_extract is bracketed by two CUDA events on the launching stream.  afx_profile_read
 * synchronises on the recorded events and returns, per phase, the accumulated milliseconds and the
 * number of launches since the last reset. */
#define AFX_PHASE_TRIM     0   /* separate trim pass, only when hop_length does not divide 512 (trim_kernel) */
#define AFX_PHASE_FRAMES   1   /* the fused single-pass kernel: pre-emphasis, trim, frames -> FFT -> mel -> dB -> DCT,
                                  energies, delta / delta2, statistics -> summary rows  (item_build_kernel + stream_kernel) */
/* slots 2 and 3 are reserved (always zero) */
#define AFX_PHASE_PCM      4   /* PCM16 -> float32                         (pcm16_to_f32_kernel) */
#define AFX_PHASE_YIN      5   /* YIN difference / CMND / pYIN observations */
#define AFX_PHASE_VITERBI  6   /* pYIN HMM decode */
#define AFX_N_PHASES       8
int afx_profile_enable(afx_handle_t h, int on);
int afx_profile_read(afx_handle_t h, double* ms /*[AFX_N_PHASES]*/, int64_t* launches /*[AFX_N_PHASES]*/, int reset);

/* Kernels launched by this library since process start (bench.py reports it as gpu_launches). */
int64_t afx_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* AFX_B200_H_ */
