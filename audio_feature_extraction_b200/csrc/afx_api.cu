// C ABI (include/afx_b200.h) + host-side wave scheduler of the B200 AudioFeatureExtractor hot path.
//
// Reference boundary: audio_feature_extraction_toolkit/core/feature_extractor.py (pure Python over
// librosa).  Everything here is new; there is no CPU fallback -- every entry point that computes
// needs an sm_100a device.
#include "../../include/afx_b200.h"

#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <map>
#include <numeric>
#include <vector>

#include "afx_common.cuh"
#include "afx_stream.cuh"
#include "afx_extras.cuh"
#include "afx_pyin.cuh"
#include "afx_pyin_tables.h"
#include "afx_tables.h"
#include "afx_trim.cuh"
#include "afx_wav.h"
#include "afx_resample.cuh"
#include "afx_dtw.cuh"

namespace {

thread_local std::string g_err;
std::atomic<long long> g_launches{0};

int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define CUDA_TRY(expr)                                                                         \
  do {                                                                                         \
    cudaError_t e__ = (expr);                                                                  \
    if (e__ != cudaSuccess)                                                                    \
      return fail(AFX_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

template <typename T>
cudaError_t upload(const std::vector<T>& v, T** dptr) {
  cudaError_t e = cudaMalloc((void**)dptr, std::max<size_t>(v.size(), 1) * sizeof(T));
  if (e != cudaSuccess) return e;
  return cudaMemcpy(*dptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
}

}  // namespace

struct afx_handle_s {
  afx_config_t cfg;
  int n16;                 // n_mfcc rounded up to a multiple of 16 (the in-pass DCT works on groups of 16)
  // pYIN (F0 group)
  bool f0_ok = false;      // pYIN constants valid for this (sr, frame_length, hop_length, f0_min, f0_max)
  afx::host::PyinConsts pc;
  double* d_pyin = nullptr;            // all double tables, one allocation
  afx::PyinTables ptab;
  const double *d_lt_same = nullptr, *d_edge_corr = nullptr, *d_freqs = nullptr;
  int yin_c = 0, yin_cap = 0, yin_rec_bytes = 0, yin_lag_pad = 0, yin_tile = 0, yin_frames_per_chunk = 0, yin_blk_len = 0, yin_rb = 1;
  size_t yin_smem = 0, vit_smem = 0;
  void* d_f0_ws = nullptr; size_t d_f0_ws_bytes = 0;   // handle-owned scratch of the F0 path (observation rows, back-pointers)
  int mel_floats = 0;      // floats in the padded sparse mel weight table
  int sm_count;
  void (*stream_fn)(afx::StreamArgs) = nullptr;   // stream_kernel<frame_length, hop_length>: warp-autonomous streaming kernel
  int stream_warps = 16;   // warps per CTA of the streaming kernel (one CTA per SM)
  int item_steps = 0;      // warp steps per work item; 0 = chosen per batch (item_steps_for)
  int slab_ring_mb = 48;   // budget of the slab ring (AFX_SLAB_RING_MB)
  int l2_mode = 1;         // L2 eviction hints of the streaming kernel (AFX_L2_MODE, StreamArgs::l2_mode)
  void (*yin_fn)(afx::YinArgs) = nullptr;     // yin_obs_kernel<lags per lane, frame mode>
  void (*vit_fn)(afx::VitArgs) = nullptr;     // viterbi_kernel<half width or 0>
  int vit_threads = afx::kVitThreads;
  // device tables
  float* d_window = nullptr;
  float* d_twiddle = nullptr;
  float* d_ptw = nullptr;
  float* d_mel_w = nullptr;
  int* d_mel_k0 = nullptr;
  float* d_mel_lin = nullptr; int* d_mel_kl = nullptr; float* d_mel_w0 = nullptr;
  int mel_w0_floats = 4;
  float* d_dct_t = nullptr;
  afx::Tables tab;
  // pinned staging for the per-call header arrays
  void* h_stage = nullptr;
  size_t h_stage_bytes = 0;
  cudaEvent_t stage_ev = nullptr;
  // resources of the host-buffer paths (afx_extract_host / afx_extract_host_packed)
  void* d_ws = nullptr;       size_t d_ws_bytes = 0;
  float* d_audio[2] = {nullptr, nullptr}; size_t d_audio_floats = 0;
  int16_t* d_pcm[2] = {nullptr, nullptr}; size_t d_pcm_samples = 0;
  float* h_pin[2] = {nullptr, nullptr};   size_t h_pin_floats = 0;
  float* d_sum = nullptr;     size_t d_sum_bytes = 0;
  int* d_meta = nullptr;      size_t d_meta_bytes = 0;
  cudaStream_t copy_stream = nullptr, compute_stream = nullptr;
  cudaEvent_t copied[2] = {nullptr, nullptr}, computed[2] = {nullptr, nullptr};
  long long host_wave_samples = 96ll << 20;   // samples per host->device wave (AFX_HOST_WAVE_SAMPLES): several clips per
                                              // resident CTA of the fused kernel; the first waves are shorter (pipeline fill)
  // sample-rate conversion (afx_resample): one phase-ordered filter per native rate, header arrays of the last call
  struct ResampleDev { afx::ResampleFilter f; float* d_taps = nullptr; };
  std::map<int, ResampleDev> resamplers;
  void* d_rs_hdr = nullptr; size_t d_rs_hdr_bytes = 0;
  // per-phase device-time accounting (afx_profile_enable / afx_profile_read)
  bool profiling = false;
  struct Span { cudaEvent_t a, b; int phase; };
  std::vector<Span> spans;            // recorded, not yet folded into prof_ms
  std::vector<cudaEvent_t> ev_pool;   // recycled events
  double prof_ms[AFX_N_PHASES] = {0};
  long long prof_launches[AFX_N_PHASES] = {0};
};

namespace {

cudaEvent_t prof_event(afx_handle_s* h) {
  cudaEvent_t e = nullptr;
  if (!h->ev_pool.empty()) { e = h->ev_pool.back(); h->ev_pool.pop_back(); }
  else cudaEventCreate(&e);
  return e;
}
// RAII: brackets the launches of one phase with two events on `stream` when profiling is on
struct PhaseTimer {
  afx_handle_s* h; cudaStream_t s; int phase; cudaEvent_t a = nullptr;
  PhaseTimer(afx_handle_s* h_, cudaStream_t s_, int phase_) : h(h_), s(s_), phase(phase_) {
    ++h->prof_launches[phase];
    if (h->profiling) { a = prof_event(h); cudaEventRecord(a, s); }
  }
  ~PhaseTimer() {
    if (a) { cudaEvent_t b = prof_event(h); cudaEventRecord(b, s); h->spans.push_back({a, b, phase}); }
  }
};

// The batch header travels from the pinned staging buffer to the workspace through a kernel that reads host memory
// directly (unified addressing), not through a copy engine: on the host-buffer paths the host->device engine is busy
// with the NEXT wave's samples, and a header memcpy queued behind that 100+ MB transfer delayed this wave's kernels until
// the next wave had landed, which serialised compute with the copies.  Also zeroes the work counter.
__global__ void header_fetch_kernel(int* __restrict__ dst, const int* __restrict__ src_host, long long words, int* __restrict__ counter) {
  const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  for (long long i = i0; i < words; i += (long long)gridDim.x * blockDim.x) dst[i] = src_host[i];
  if (i0 < 64) counter[i0] = 0;
}

struct Layout {   // carve-up of the caller's workspace
  size_t off_offsets, off_frame_off, off_lengths, off_yin_chunk, off_item_off, off_info, off_counter, off_slots, header_bytes;
};
constexpr int kMaxSlots = 4096;   // slab ring of the streaming kernel

Layout header_layout(int64_t n) {
  Layout L;
  size_t o = 0;
  L.off_offsets = o;   o += align_up(sizeof(long long) * n, 16);
  L.off_frame_off = o; o += align_up(sizeof(long long) * (n + 1), 16);
  L.off_lengths = o;   o += align_up(sizeof(int) * n, 16);
  L.off_yin_chunk = o; o += align_up(sizeof(int) * (n + 1), 16);   // prefix of pYIN frame chunks per clip
  L.off_item_off = o;  o += align_up(sizeof(int) * (n + 1), 16);   // prefix of work items per clip (streaming kernel)
  L.off_info = o;      o += align_up(sizeof(afx::ClipInfo) * n, 16);   // produced on device; not staged from the host
  L.off_counter = o;   o += 256;
  L.off_slots = o;     o += 2 * sizeof(int) * kMaxSlots;              // slot owner / arrival counters (device-initialised)
  L.header_bytes = align_up(o, 256);
  return L;
}

inline long long frames_ub(const afx_handle_s* h, long long len) { return 1 + std::max<long long>(len, 0) / h->cfg.hop_length; }
inline long long ceil_div(long long a, long long b) { return (a + b - 1) / b; }

// per-CTA scratch slab of the fused kernel for clips of at most max_len samples: log-mel rows, RMS, row maxima, block sums
inline int slab_rows(const afx_handle_s* h, long long max_len) { return (int)((frames_ub(h, max_len) + 3) & ~3LL); }
size_t slab_bytes_for(const afx_handle_s* h, long long max_len) {
  const size_t rows = (size_t)slab_rows(h, max_len);
  return align_up((rows * ((size_t)h->n16 + 4) + 8) * sizeof(float), 256);   // MFCC rows, rms, row max / min, hop-block sums
}

// the speculative single pass needs every trimmed frame centre (start = k*512) on the absolute hop grid
inline bool runtime_hop(int n_fft, int hop) { return !(hop * 2 == n_fft || hop * 4 == n_fft || hop * 8 == n_fft); }
inline bool single_pass(const afx_handle_s* h) { return afx::kTrimHop % h->cfg.hop_length == 0 && !runtime_hop(h->cfg.frame_length, h->cfg.hop_length); }

// the streaming kernel is instantiated per (frame_length, hop_length) for hop = frame_length / 2, / 4, / 8, plus one
// run-time-hop variant per frame length.
// stock: the mel tables have MelCnt<n_fft, true>'s run lengths -> the variant with the unrolled segment walk
// (instantiated for the stock hop of each frame length: 1024/256, 512/128, 2048/512)
template <int NFFT>
bool mel_counts_are_stock(const afx::host::MelSparse& m) {
  using MC = afx::MelCnt<NFFT, true>;
  return m.cnt[0] == MC::c0 && m.cnt[1] == MC::c1 && m.cnt[2] == MC::c2 && m.cnt[3] == MC::c3 && m.cnt0 == MC::cz;
}
typedef void (*stream_fn_t)(afx::StreamArgs);
stream_fn_t stream_kernel_for(int n_fft, int hop, const afx::host::MelSparse& mel) {
  if (n_fft == 1024 && hop == 256 && mel_counts_are_stock<1024>(mel)) return afx::stream_kernel<1024, 256, true>;
  if (n_fft == 512 && hop == 128 && mel_counts_are_stock<512>(mel)) return afx::stream_kernel<512, 128, true>;
  if (n_fft == 2048 && hop == 512 && mel_counts_are_stock<2048>(mel)) return afx::stream_kernel<2048, 512, true>;
  if (n_fft == 1024) {
    if (hop == 512) return afx::stream_kernel<1024, 512>;
    if (hop == 256) return afx::stream_kernel<1024, 256>;
    if (hop == 128) return afx::stream_kernel<1024, 128>;
  } else if (n_fft == 512) {
    if (hop == 256) return afx::stream_kernel<512, 256>;
    if (hop == 128) return afx::stream_kernel<512, 128>;
    if (hop == 64) return afx::stream_kernel<512, 64>;
  } else if (n_fft == 2048) {
    if (hop == 1024) return afx::stream_kernel<2048, 1024>;
    if (hop == 512) return afx::stream_kernel<2048, 512>;
    if (hop == 256) return afx::stream_kernel<2048, 256>;
  }
  // any other hop <= n_fft / 2: the run-time-hop variant (frames of a step are read separately, the trim runs first)
  if (hop >= 1 && hop <= n_fft / 2) {
    if (n_fft == 1024) return afx::stream_kernel<1024, 0>;
    if (n_fft == 512) return afx::stream_kernel<512, 0>;
    if (n_fft == 2048) return afx::stream_kernel<2048, 0>;
  }
  return nullptr;
}


inline int frames_per_step(const afx_handle_s* h) { return h->cfg.frame_length == 2048 ? 1 : (h->cfg.frame_length == 1024 ? 2 : 4); }

// bytes of one warp's buffer of the streaming kernel (afx::WarpBuf<>::kBytes)
inline size_t stream_warp_bytes(const afx_handle_s* h) {
  const bool rt = runtime_hop(h->cfg.frame_length, h->cfg.hop_length);   // afx::Geo<n_fft, 0>: sized for hop = n_fft / 2
  const size_t tile = (size_t)((frames_per_step(h) - 1) * (rt ? h->cfg.frame_length / 2 : h->cfg.hop_length) + h->cfg.frame_length + (rt ? 12 : 8)) * 4;
  return align_up(std::max<size_t>(tile + afx::kPlaneBFloats * 4, afx::kPlaneTBytes), 16);
}
size_t stream_smem_bytes(const afx_handle_s* h, int warps) {
  const int n_fft = h->cfg.frame_length;
  return 512 + (size_t)afx::kStreamWarpSlots * sizeof(afx::WarpScratch) + 8192 + (n_fft == 2048 ? 4112 : 0) +
         (size_t)32 * (2 * h->n16 + 4) * 4 + (size_t)(128 * 5 + afx::kMelW0Max) * 4 + (size_t)warps * stream_warp_bytes(h);
}

// work items of the streaming kernel: runs of item_steps warp steps over the frames of the untrimmed clip (an upper
// bound of the trimmed frames when the trim runs first); every clip has at least one item, whose completion
// triggers the clip's finishing phases
// ring of per-clip slabs: L2-resident when it fits the budget, but never fewer than ~1024 slots (capped at 1 GB):
// every resident warp streams a different clip when clips are short, and a finishing clip keeps its slot for the
// duration of its single-warp finisher, so a starved ring stalls the streaming warps (measured on C3: 184 slots 30 ms,
// 1970 slots 11.6 ms per 8000 clips)
inline int slots_for(const afx_handle_s* h, long long n_clips, long long max_len) {
  const size_t sb = slab_bytes_for(h, max_len);
  const long long by_l2 = (long long)(((size_t)h->slab_ring_mb << 20) / sb);
  const long long floor_slots = std::min<long long>(1024, (long long)(((size_t)1024 << 20) / sb));
  return (int)std::max<long long>(2, std::min<long long>({n_clips, std::max<long long>({by_l2, floor_slots, 64LL}), (long long)kMaxSlots}));
}

// steps per item: large enough that the per-item costs (slot check, arrival atomic + fences) vanish, small enough that
// every resident warp still gets >= 8 items (a single 5 s clip must spread over the whole GPU: 2-step items) and that the
// clips being streamed at any moment (resident warps x item steps / steps per clip) fill at most half of the slab ring
inline int item_steps_for(const afx_handle_s* h, int64_t n_clips, const int32_t* h_lengths) {
  if (h->item_steps > 0) return h->item_steps;
  const int fpw = frames_per_step(h);
  long long steps = 0, max_len = 0;
  for (int64_t i = 0; i < n_clips; ++i) { steps += ceil_div(h_lengths[i] > 0 ? frames_ub(h, h_lengths[i]) : 0, fpw); max_len = std::max<long long>(max_len, h_lengths[i]); }
  const long long warps = (long long)h->sm_count * h->stream_warps;
  const long long by_balance = steps / (8LL * warps);
  const long long by_ring = n_clips > 0 ? (long long)slots_for(h, n_clips, max_len) * (steps / std::max<int64_t>(n_clips, 1)) / (2LL * warps) : 2;
  // (an item boundary costs about 0.6 warp steps -- claim, slot hand-over, arrival count: measured on C2, 10 000 clips, with
  // AFX_ITEM_STEPS = 6 / 16 / 32 / 64 / 96: 12.20 / 11.54 / 11.31 / 11.17 / 11.31 ms)
  // the slab rows of the clips in flight should stay L2-resident until their finisher reads them: at most 64 MB of slabs
  // in flight (C3, 240 KB per clip: 63-step items put 114 MB in flight and ran 5 % slower than 32-step items)
  const long long spc = steps / std::max<int64_t>(n_clips, 1);
  const long long by_l2 = n_clips > 0 ? (long long)(((size_t)64 << 20) / slab_bytes_for(h, max_len)) * spc / warps : 2;
  return (int)std::max<long long>(2, std::min<long long>({by_balance, by_ring, by_l2, 64LL, 255LL / fpw}));
}
inline long long items_of(const afx_handle_s* h, long long len, int item_steps) {
  const long long fpi = (long long)frames_per_step(h) * item_steps;
  return std::max<long long>(1, ceil_div(len > 0 ? frames_ub(h, len) : 0, fpi));
}

}  // namespace

extern "C" {

int afx_abi_version(void) { return AFX_ABI_VERSION; }
const char* afx_last_error(void) { return g_err.c_str(); }
int64_t afx_launch_count(void) { return g_launches.load(); }

void afx_config_default(afx_config_t* cfg) {
  if (!cfg) return;
  memset(cfg, 0, sizeof *cfg);
  cfg->sr = 22050;
  cfg->frame_length = 1024;
  cfg->hop_length = 256;
  cfg->n_mfcc = 13;
  cfg->f0_min = 65.40639132514966;     // librosa.note_to_hz('C2')
  cfg->f0_max = 2093.004522404789;     // librosa.note_to_hz('C7')
  cfg->pre_emphasis = 0.97f;
  cfg->trim_top_db = 30.f;
  cfg->n_mels = 128;
  cfg->mfcc_top_db = 80.f;
  cfg->device = 0;
}

int afx_create(const afx_config_t* cfg, afx_handle_t* out) {
  if (!cfg || !out) return fail(AFX_ERR_INVALID_ARG, "afx_create: null argument");
  if (cfg->frame_length != 512 && cfg->frame_length != 1024 && cfg->frame_length != 2048)
    return fail(AFX_ERR_UNSUPPORTED, "frame_length %d not supported (512, 1024 or 2048)", cfg->frame_length);
  if (cfg->hop_length < 1 || cfg->hop_length > cfg->frame_length / 2)
    return fail(AFX_ERR_INVALID_ARG, "hop_length %d out of range [1, frame_length/2]", cfg->hop_length);
  if (cfg->n_mfcc < 1 || cfg->n_mfcc > afx::kMaxMfcc) return fail(AFX_ERR_INVALID_ARG, "n_mfcc %d out of range [1,64]", cfg->n_mfcc);
  if (cfg->n_mels != afx::kNMels) return fail(AFX_ERR_UNSUPPORTED, "n_mels must be 128");
  if (cfg->sr < 1000) return fail(AFX_ERR_INVALID_ARG, "sr %d too small", cfg->sr);

  int ndev = 0;
  CUDA_TRY(cudaGetDeviceCount(&ndev));
  if (cfg->device < 0 || cfg->device >= ndev) return fail(AFX_ERR_CUDA, "CUDA device %d not present (%d devices)", cfg->device, ndev);
  CUDA_TRY(cudaSetDevice(cfg->device));
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major < 10) return fail(AFX_ERR_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", cfg->device, prop.major, prop.minor);

  afx_handle_s* h = new afx_handle_s();
  h->cfg = *cfg;
  h->n16 = (cfg->n_mfcc + 15) & ~15;
  h->sm_count = prop.multiProcessorCount;
  if (const char* e = getenv("AFX_HOST_WAVE_SAMPLES")) { const long long v = atoll(e); if (v >= (1 << 16)) h->host_wave_samples = v; }

  const auto win = afx::host::window_half(cfg->frame_length);
  const auto tw = afx::host::twiddle32x32(cfg->frame_length == 512 ? 512 : 1024);
  const auto ptw = afx::host::post_twiddle(2048, 513);
  const auto mel = afx::host::mel_sparse(cfg->sr, cfg->frame_length, cfg->n_mels);
  if (mel.nnz < 0) { delete h; return fail(AFX_ERR_UNSUPPORTED, "mel filterbank: segment decomposition does not cover every weight"); }
  const auto dct = afx::host::dct_rows(cfg->n_mfcc, cfg->n_mels);
  std::vector<float> dct_t((size_t)cfg->n_mels * h->n16, 0.f);
  for (int c = 0; c < cfg->n_mfcc; ++c)
    for (int m = 0; m < cfg->n_mels; ++m) dct_t[(size_t)m * h->n16 + c] = dct[(size_t)c * cfg->n_mels + m];
  CUDA_TRY(upload(win, &h->d_window));
  CUDA_TRY(upload(tw, &h->d_twiddle));
  CUDA_TRY(upload(ptw, &h->d_ptw));
  h->mel_floats = (int)mel.w.size();
  CUDA_TRY(upload(mel.w, &h->d_mel_w));
  CUDA_TRY(upload(mel.k0, &h->d_mel_k0));
  CUDA_TRY(upload(dct_t, &h->d_dct_t));
  if (mel.lin_max_err > 2e-6) { delete h; return fail(AFX_ERR_UNSUPPORTED, "mel filterbank: linear weight form is off by %.3g", mel.lin_max_err); }
  {
    std::vector<float> w0 = mel.w0;
    w0.resize((std::max<size_t>(w0.size(), (size_t)mel.cnt0) + 3) & ~(size_t)3, 0.f);
    for (int i = 0; i < mel.cnt0; ++i) w0[i] = mel.dense[(size_t)0 * (1 + cfg->frame_length / 2) + mel.k00 + i];
    h->mel_w0_floats = (int)w0.size();
    if (h->mel_w0_floats > afx::kMelW0Max) { delete h; return fail(AFX_ERR_UNSUPPORTED, "mel filterbank: %d bins below the first mel point (at most %d)", mel.cnt0, afx::kMelW0Max); }
    CUDA_TRY(upload(mel.lin, &h->d_mel_lin));
    CUDA_TRY(upload(mel.kl, &h->d_mel_kl));
    CUDA_TRY(upload(w0, &h->d_mel_w0));
  }
  h->tab.mel_lin = reinterpret_cast<const float4*>(h->d_mel_lin); h->tab.mel_kl = h->d_mel_kl; h->tab.mel_w0 = h->d_mel_w0;
  h->tab.window_half = h->d_window;
  h->tab.twiddle = reinterpret_cast<const float2*>(h->d_twiddle);
  h->tab.post_twiddle = reinterpret_cast<const float2*>(h->d_ptw);
  h->tab.mel_w = h->d_mel_w;
  h->tab.mel_k0 = h->d_mel_k0;
  for (int q = 0; q < 4; ++q) { h->tab.mel_cnt[q] = mel.cnt[q]; h->tab.mel_woff[q] = mel.woff[q]; }
  h->tab.mel_k00 = mel.k00; h->tab.mel_cnt0 = mel.cnt0; h->tab.mel_woff0 = mel.woff0;
  h->tab.dct = nullptr;

  CUDA_TRY(cudaFuncSetAttribute(afx::trim_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  {  // the warp-autonomous streaming kernel: one CTA per SM with as many warps as shared memory (and 128 registers) allow
    h->stream_fn = stream_kernel_for(cfg->frame_length, cfg->hop_length, mel);
    if (const char* e = getenv("AFX_L2_MODE")) h->l2_mode = atoi(e) & 3;
    if (const char* e = getenv("AFX_SLAB_RING_MB")) { const int v = atoi(e); if (v >= 1 && v <= 4096) h->slab_ring_mb = v; }
    if (const char* e = getenv("AFX_ITEM_STEPS")) { const int v = atoi(e); if (v >= 1 && v * frames_per_step(h) <= 255) h->item_steps = v; }
    int w = afx::kStreamMaxWarps;
    if (const char* e = getenv("AFX_STREAM_WARPS")) { const int v = atoi(e); if (v >= 1 && v <= afx::kStreamMaxWarps) w = v; }
    // (measured on C2: 12 warps 13.3 ms, 14 warps 11.7 ms, 16 warps 11.3 ms; variants compiled for 18 / 20 warps at 96
    // registers spill in the step and lose: 12.7 / 13.7 ms; 12 warps at 157 registers without spills: 12.0 ms)
    while (w > 1 && stream_smem_bytes(h, w) + 1024 > (size_t)prop.sharedMemPerBlockOptin) --w;
    if (stream_smem_bytes(h, w) + 1024 > (size_t)prop.sharedMemPerBlockOptin) { delete h; return fail(AFX_ERR_UNSUPPORTED, "the streaming kernel does not fit this device's shared memory"); }
    h->stream_warps = w;
    CUDA_TRY(cudaFuncSetAttribute(h->stream_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stream_smem_bytes(h, w)));
  }
  {  // pYIN constants and tables (F0 group); an unsupported f0 range only disables that group
    afx::host::PyinConsts& pc = h->pc;
    // yin_obs_kernel shares the autocorrelation products between overlapping frames when the analysis window
    // (frame_length / 2) is a whole number of hops (block mode); any other hop gets one block per frame (frame mode)
    h->f0_ok = afx::host::pyin_consts(cfg->sr, cfg->frame_length, cfg->hop_length, cfg->f0_min, cfg->f0_max, pc) &&
               pc.n_lags <= 32 * afx::kYinMaxGroups * 2 - 4 && pc.max_period <= 32 * 22;
    if (h->f0_ok) {
      std::vector<double> all;
      auto push = [&](const std::vector<double>& v) { const size_t o = all.size(); all.insert(all.end(), v.begin(), v.end()); while (all.size() & 1) all.push_back(0.0); return o; };
      const size_t o_th = push(pc.thresholds), o_b = push(pc.beta), o_bc = push(pc.beta_cum), o_bs = push(pc.beta_suf), o_bf = push(pc.boltz_fact),
                   o_be = push(pc.boltz_exp), o_lt = push(pc.ltrans_same), o_ec = push(pc.edge_corr), o_fr = push(pc.freqs);
      CUDA_TRY(upload(all, &h->d_pyin));
      h->ptab.thresholds = h->d_pyin + o_th; h->ptab.beta = h->d_pyin + o_b; h->ptab.beta_cum = h->d_pyin + o_bc;
      h->ptab.boltz_fact = h->d_pyin + o_bf; h->ptab.boltz_exp = h->d_pyin + o_be; h->ptab.beta_suf = h->d_pyin + o_bs;
      h->d_lt_same = h->d_pyin + o_lt; h->d_edge_corr = h->d_pyin + o_ec; h->d_freqs = h->d_pyin + o_fr;
      const int need_c = (pc.max_period + 31) / 32;
      h->yin_c = need_c <= 8 ? 8 : need_c <= 11 ? 11 : need_c <= 16 ? 16 : 22;
      h->yin_cap = ((pc.n_lags / 2 + 2) + 3) & ~3;
      h->yin_rec_bytes = (int)align_up(24 + 10 * (size_t)h->yin_cap, 16);
      h->yin_lag_pad = (pc.max_period + 1 + 127) & ~127;
      {
        const int W = cfg->frame_length / 2, hop = cfg->hop_length;
        const bool block_mode = W % hop == 0 && hop % 4 == 0 && W / hop <= 8;
        h->yin_blk_len = block_mode ? hop : W;
        h->yin_rb = block_mode ? W / hop : 1;
      }
      h->yin_tile = ((afx::kYinBlocks - 1) * cfg->hop_length + h->yin_blk_len + h->yin_lag_pad + 12 + 3) & ~3;   // A and S follow the tile: 16-byte rows
      h->yin_frames_per_chunk = afx::kYinBlocks - (h->yin_rb - 1);
      const size_t ph1 = (size_t)h->yin_tile * 4 + (size_t)afx::kYinBlocks * h->yin_lag_pad * 4 + ((size_t)h->yin_tile + 2) * 8;
      const size_t ph2 = (size_t)afx::kYinWarps * h->yin_cap * (2 * 8 + 4);   // trough heights + periods (double) + packed compact list
      h->yin_smem = align_up(std::max(ph1, ph2), 16);
      // Viterbi band search: the int32-filtered variant for the half widths of the stock geometries, else the generic one
      h->vit_fn = afx::viterbi_kernel<0>;
      const char* vsel = getenv("AFX_VITERBI");
      if (!(vsel && !strcmp(vsel, "generic"))) {
        // half width = 5 * round(35.92 * 12 * hop / sr): 25 at 22.05 kHz / 256 and 44.1 kHz / 512, 15 at 16 kHz / 128,
        // 20 at the 10 ms hops (16 kHz / 160, 22.05 kHz / 200, 44.1 kHz / 441), 10 / 30 at half / 1.5 x those
        // four warps per clip (thread <-> 2 targets in the filter): every scheduler of the SM hosts one warp of every
        // resident CTA, so no warp of a CTA reaches the barriers late for sharing its scheduler with a sibling: 36.1 -> 34.4 ms
        // per 2 960 clips against the five-warp shape (AFX_VIT_WARPS=5 selects it, half width 25 only: development A/B)
        const char* vw = getenv("AFX_VIT_WARPS");
        h->vit_threads = 128;
        switch (pc.half_width) {
          case 10: h->vit_fn = afx::viterbi_kernel<10, 4>; break;
          case 15: h->vit_fn = afx::viterbi_kernel<15, 4>; break;
          case 20: h->vit_fn = afx::viterbi_kernel<20, 4>; break;
          case 25: h->vit_fn = afx::viterbi_kernel<25, 4>; break;
          case 30: h->vit_fn = afx::viterbi_kernel<30, 4>; break;
          default: h->vit_threads = afx::kVitThreads; break;   // generic band search, five warps
        }
        if (vw && atoi(vw) == 5 && pc.half_width == 25) { h->vit_fn = afx::viterbi_kernel<25, 5>; h->vit_threads = afx::kVitThreads; }
      }
      h->vit_smem = align_up(afx::VitGeo(pc.n_bins, pc.half_width, h->vit_fn != afx::viterbi_kernel<0>).bytes(pc.n_bins), 16);
      if (h->yin_smem > (size_t)prop.sharedMemPerBlockOptin || h->vit_smem > (size_t)prop.sharedMemPerBlockOptin) h->f0_ok = false;
    }
    if (h->f0_ok) {
      const bool fm = h->yin_rb == 1 && h->yin_blk_len != cfg->hop_length;   // frame mode: blocks at any alignment
      h->yin_fn = h->yin_c == 8 ? (fm ? afx::yin_obs_kernel<8, true> : afx::yin_obs_kernel<8, false>)
                : h->yin_c == 11 ? (fm ? afx::yin_obs_kernel<11, true> : afx::yin_obs_kernel<11, false>)
                : h->yin_c == 16 ? (fm ? afx::yin_obs_kernel<16, true> : afx::yin_obs_kernel<16, false>)
                                 : (fm ? afx::yin_obs_kernel<22, true> : afx::yin_obs_kernel<22, false>);
      CUDA_TRY(cudaFuncSetAttribute(h->yin_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->yin_smem));
      CUDA_TRY(cudaFuncSetAttribute(h->vit_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->vit_smem));
    }
  }
  CUDA_TRY(cudaEventCreateWithFlags(&h->stage_ev, cudaEventDisableTiming));
  *out = h;
  return AFX_OK;
}

int afx_destroy(afx_handle_t h) {
  if (!h) return AFX_OK;
  cudaSetDevice(h->cfg.device);
  cudaDeviceSynchronize();
  cudaFree(h->d_window); cudaFree(h->d_twiddle); cudaFree(h->d_ptw); cudaFree(h->d_pyin); cudaFree(h->d_f0_ws); cudaFree(h->d_mel_w); cudaFree(h->d_mel_k0); cudaFree(h->d_dct_t);
  cudaFree(h->d_mel_lin); cudaFree(h->d_mel_kl); cudaFree(h->d_mel_w0);
  if (h->h_stage) cudaFreeHost(h->h_stage);
  if (h->stage_ev) cudaEventDestroy(h->stage_ev);
  cudaFree(h->d_ws); cudaFree(h->d_audio[0]); cudaFree(h->d_audio[1]); cudaFree(h->d_sum); cudaFree(h->d_meta);
  for (auto& sp : h->spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
  for (auto e : h->ev_pool) cudaEventDestroy(e);
  cudaFree(h->d_pcm[0]); cudaFree(h->d_pcm[1]);
  for (auto& kv : h->resamplers) cudaFree(kv.second.d_taps);
  cudaFree(h->d_rs_hdr);
  if (h->h_pin[0]) cudaFreeHost(h->h_pin[0]);
  if (h->h_pin[1]) cudaFreeHost(h->h_pin[1]);
  for (int i = 0; i < 2; ++i) { if (h->copied[i]) cudaEventDestroy(h->copied[i]); if (h->computed[i]) cudaEventDestroy(h->computed[i]); }
  if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
  if (h->compute_stream) cudaStreamDestroy(h->compute_stream);
  delete h;
  return AFX_OK;
}

int afx_summary_width(afx_handle_t h) { return h ? 4 + 4 * h->cfg.n_mfcc + 3 + 2 * h->cfg.n_mfcc + 2 : 0; }
int afx_frame_width(afx_handle_t h) { return h ? 3 * h->cfg.n_mfcc + 3 : 0; }
int64_t afx_max_frames(afx_handle_t h, int64_t length) { return h ? frames_ub(h, length) : 0; }

static size_t workspace_bytes_impl(afx_handle_t h, int64_t n_clips, const int32_t* h_lengths) {
  if (!h || n_clips < 0 || (n_clips > 0 && !h_lengths)) return 0;
  long long max_len = 0, items = 0;
  const int isteps = item_steps_for(h, n_clips, h_lengths);
  for (int64_t i = 0; i < n_clips; ++i) { max_len = std::max<long long>(max_len, h_lengths[i]); items += items_of(h, h_lengths[i], isteps); }
  const size_t stream = align_up((size_t)items * sizeof(afx::WorkItem), 256) + (size_t)slots_for(h, n_clips, max_len) * slab_bytes_for(h, max_len);
  return header_layout(n_clips).header_bytes + stream + 1024;
}
// header + work-item table + the ring of per-clip slabs; there is no smaller configuration, so both agree
size_t afx_workspace_bytes(afx_handle_t h, int64_t n_clips, const int32_t* h_lengths) { return workspace_bytes_impl(h, n_clips, h_lengths); }
size_t afx_workspace_min_bytes(afx_handle_t h, int64_t n_clips, const int32_t* h_lengths) { return workspace_bytes_impl(h, n_clips, h_lengths); }

int afx_extract(afx_handle_t h, const float* d_audio, const int64_t* h_offsets, const int32_t* h_lengths,
                int64_t n_clips, uint32_t feature_mask, float* d_summaries, int32_t* d_meta,
                float* d_frames, void* d_workspace, size_t workspace_bytes, void* stream_v) {
  if (!h) return fail(AFX_ERR_INVALID_ARG, "afx_extract: null handle");
  if (n_clips == 0) return AFX_OK;
  if (n_clips < 0 || n_clips > 0x7fffff00LL) return fail(AFX_ERR_INVALID_ARG, "afx_extract: bad n_clips");
  if (!d_audio || !h_offsets || !h_lengths || !d_summaries || !d_meta || !d_workspace)
    return fail(AFX_ERR_INVALID_ARG, "afx_extract: null pointer argument");
  if ((feature_mask & ~(AFX_FEAT_ALL | AFX_FEAT_PREPROCESSED)) != 0) return fail(AFX_ERR_INVALID_ARG, "afx_extract: unknown feature bits 0x%x", feature_mask);
  // AFX_FEAT_PREPROCESSED: the clips are outputs of preprocess_audio (already pre-emphasised and trimmed), the way the
  // reference's extract_f0 / extract_mfcc / extract_energy receive them: no filter, no first-sample rule, no trim
  const bool raw = (feature_mask & AFX_FEAT_PREPROCESSED) != 0;
  feature_mask &= AFX_FEAT_ALL;
  const float coef = raw ? 0.f : h->cfg.pre_emphasis;
  const float trim_top_db = raw ? 3.0e38f : h->cfg.trim_top_db;
  if ((feature_mask & AFX_FEAT_F0) && !h->f0_ok)
    return fail(AFX_ERR_UNSUPPORTED, "afx_extract: the F0 (pYIN) group does not support this f0_min / f0_max / frame geometry");
  if ((reinterpret_cast<uintptr_t>(d_audio) & 15u) != 0) return fail(AFX_ERR_INVALID_ARG, "afx_extract: d_audio must be 16-byte aligned");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  const int n = (int)n_clips;
  const int hop = h->cfg.hop_length;

  // ---- host-side header: offsets, lengths, output-row prefix (pinned staging, one async copy)
  const Layout L = header_layout(n);
  const size_t staged_bytes = L.off_info;
  if (h->h_stage_bytes < staged_bytes) {
    if (h->h_stage) { CUDA_TRY(cudaEventSynchronize(h->stage_ev)); CUDA_TRY(cudaFreeHost(h->h_stage)); h->h_stage = nullptr; }
    const size_t cap = std::max<size_t>(staged_bytes * 2, 1 << 16);
    CUDA_TRY(cudaMallocHost(&h->h_stage, cap));
    h->h_stage_bytes = cap;
  } else {
    CUDA_TRY(cudaEventSynchronize(h->stage_ev));   // previous call's upload has drained
  }
  char* st = static_cast<char*>(h->h_stage);
  long long* s_offsets = reinterpret_cast<long long*>(st + L.off_offsets);
  long long* s_frame_off = reinterpret_cast<long long*>(st + L.off_frame_off);
  int* s_lengths = reinterpret_cast<int*>(st + L.off_lengths);
  int* s_ychunk = reinterpret_cast<int*>(st + L.off_yin_chunk);
  int* s_item_off = reinterpret_cast<int*>(st + L.off_item_off);
  const int yF = std::max(1, h->yin_frames_per_chunk);
  long long fo = 0, max_len = 0, yc = 0, n_items = 0;
  const int isteps = item_steps_for(h, n_clips, h_lengths);
  for (int i = 0; i < n; ++i) {
    if (h_lengths[i] < 0) return fail(AFX_ERR_INVALID_ARG, "afx_extract: clip %d has negative length", i);
    if (h_offsets[i] < 0 || (h_offsets[i] & 3) != 0) return fail(AFX_ERR_INVALID_ARG, "afx_extract: clip %d offset %lld is not a multiple of 4 samples", i, (long long)h_offsets[i]);
    s_offsets[i] = h_offsets[i];
    s_lengths[i] = h_lengths[i];
    s_frame_off[i] = fo;
    s_ychunk[i] = (int)yc;
    s_item_off[i] = (int)n_items;
    n_items += items_of(h, h_lengths[i], isteps);
    if (n_items > 0x7fffff00LL) return fail(AFX_ERR_UNSUPPORTED, "afx_extract: batch too large");
    yc += ceil_div(frames_ub(h, h_lengths[i]), yF);
    if (yc > 0x7fffff00LL) return fail(AFX_ERR_UNSUPPORTED, "afx_extract: batch too large");
    fo += frames_ub(h, h_lengths[i]);
    max_len = std::max<long long>(max_len, h_lengths[i]);
  }
  s_frame_off[n] = fo;
  s_ychunk[n] = (int)yc;
  s_item_off[n] = (int)n_items;
  const bool one_pass = single_pass(h);
  const long long max_trim_blocks = (200 * 1024) / 4 - 8;
  if (!one_pass && ceil_div(max_len, afx::kTrimHop) > max_trim_blocks)
    return fail(AFX_ERR_UNSUPPORTED, "afx_extract: clips longer than %lld samples need hop_length dividing 512", max_trim_blocks * afx::kTrimHop);

  const size_t slab_bytes = slab_bytes_for(h, max_len);
  const int n_slots = slots_for(h, n, max_len);
  const size_t items_bytes = align_up((size_t)n_items * sizeof(afx::WorkItem), 256);
  if (workspace_bytes < L.header_bytes + items_bytes + (size_t)n_slots * slab_bytes)
    return fail(AFX_ERR_WORKSPACE, "afx_extract: workspace of %zu bytes is below afx_workspace_bytes", workspace_bytes);
  char* ws = static_cast<char*>(d_workspace);
  if ((reinterpret_cast<uintptr_t>(ws) & 255u) != 0) return fail(AFX_ERR_INVALID_ARG, "afx_extract: workspace must be 256-byte aligned");
  {
    const long long words = (long long)((staged_bytes + 3) / 4);
    const int blocks = (int)std::min<long long>((words + 255) / 256, 4LL * h->sm_count);
    header_fetch_kernel<<<std::max(blocks, 1), 256, 0, stream>>>(reinterpret_cast<int*>(ws), reinterpret_cast<const int*>(st), words,
                                                                 reinterpret_cast<int*>(ws + L.off_counter));
    ++g_launches;
  }
  CUDA_TRY(cudaEventRecord(h->stage_ev, stream));
  const long long* d_offsets = reinterpret_cast<const long long*>(ws + L.off_offsets);
  const long long* d_frame_off = reinterpret_cast<const long long*>(ws + L.off_frame_off);
  const int* d_lengths = reinterpret_cast<const int*>(ws + L.off_lengths);
  afx::ClipInfo* d_info = reinterpret_cast<afx::ClipInfo*>(ws + L.off_info);

  if (!one_pass) {  // hop does not divide 512: the trim decision needs its own pass over the batch first
    afx::TrimArgs ta;
    ta.audio = d_audio; ta.offsets = d_offsets; ta.lengths = d_lengths; ta.info = d_info;
    ta.clip_begin = 0; ta.hop = hop; ta.coef = coef; ta.top_db = trim_top_db; ta.raw = raw ? 1 : 0;
    ta.min_frames = (feature_mask & AFX_FEAT_MFCC) ? afx::kDeltaWidth : 1;
    const size_t smem = (size_t)(ceil_div(max_len, afx::kTrimHop) + 4) * sizeof(float);
    PhaseTimer pt(h, stream, AFX_PHASE_TRIM);
    afx::trim_kernel<<<n, afx::kTrimThreads, smem, stream>>>(ta);
    ++g_launches;
  }
  {
    afx::ClipArgs ca;
    ca.audio = d_audio; ca.offsets = d_offsets; ca.lengths = d_lengths; ca.frame_off = d_frame_off; ca.info = d_info;
    ca.counter = reinterpret_cast<int*>(ws + L.off_counter); ca.n_clips = n;
    ca.slabs = nullptr; ca.slab_bytes = slab_bytes; ca.a_max = slab_rows(h, max_len);
    ca.tab = h->tab; ca.dct_t = h->d_dct_t; ca.hop = hop;
    ca.coef = coef; ca.trim_top_db = trim_top_db; ca.mfcc_top_db = h->cfg.mfcc_top_db; ca.raw = raw ? 1 : 0;
    ca.n_mfcc = h->cfg.n_mfcc; ca.n16 = h->n16; ca.summary_width = afx_summary_width(h); ca.frame_width = afx_frame_width(h);
    ca.mask = feature_mask; ca.summaries = d_summaries; ca.meta = d_meta; ca.frames_out = d_frames;
    ca.pretrimmed = one_pass ? 0 : 1;
    PhaseTimer pt(h, stream, AFX_PHASE_FRAMES);
    {
      afx::StreamArgs sa;
      sa.c = ca;
      sa.c.slabs = ws + L.header_bytes + items_bytes;
      sa.items = reinterpret_cast<const afx::WorkItem*>(ws + L.header_bytes);
      sa.n_items = (int)n_items; sa.n_slots = n_slots; sa.mel_w0_floats = h->mel_w0_floats; sa.l2_mode = h->l2_mode;
      sa.slot_owner = reinterpret_cast<int*>(ws + L.off_slots); sa.slot_done = sa.slot_owner + kMaxSlots;
      const int threads = std::max(n, n_slots);
      afx::item_build_kernel<<<(threads + 255) / 256, 256, 0, stream>>>(sa, reinterpret_cast<const int*>(ws + L.off_item_off),
                                                                         frames_per_step(h) * isteps);
      const long long want = ceil_div(n_items, h->stream_warps);
      const int sgrid = (int)std::max<long long>(1, std::min<long long>(want, h->sm_count));
      h->stream_fn<<<sgrid, h->stream_warps * 32, stream_smem_bytes(h, h->stream_warps), stream>>>(sa);
      g_launches += 2;
    }
  }
  if (feature_mask & AFX_FEAT_F0) {
    // ---- pYIN: waves of clips sized by the handle-owned scratch (observation rows + back-pointers + states)
    const afx::host::PyinConsts& pc = h->pc;
    const size_t per_frame = (size_t)h->yin_rec_bytes + (size_t)4 * pc.n_bins + 2;
    size_t budget = (size_t)12288 << 20;   // observation rows + back-pointers: ~3.6 MB per 10 s clip; 12 GB = four full rounds of the Viterbi kernel
    if (const char* e = getenv("AFX_F0_SCRATCH_MB")) { const long long v = atoll(e); if (v >= 16) budget = (size_t)v << 20; }
    long long max_f = 1;
    for (int i = 0; i < n; ++i) max_f = std::max(max_f, s_frame_off[i + 1] - s_frame_off[i]);
    long long cap_frames = std::max<long long>(max_f, std::min<long long>(fo, (long long)(budget / per_frame)));
    const size_t need = (size_t)cap_frames * per_frame + 4096;
    if (h->d_f0_ws_bytes < need) {
      CUDA_TRY(cudaStreamSynchronize(stream));
      if (h->d_f0_ws) CUDA_TRY(cudaFree(h->d_f0_ws));
      h->d_f0_ws = nullptr; h->d_f0_ws_bytes = 0;
      CUDA_TRY(cudaMalloc(&h->d_f0_ws, need));
      h->d_f0_ws_bytes = need;
    }
    unsigned char* f0ws = static_cast<unsigned char*>(h->d_f0_ws);
    unsigned char* d_obs = f0ws;
    unsigned short* d_ptr = reinterpret_cast<unsigned short*>(f0ws + align_up((size_t)cap_frames * h->yin_rec_bytes, 256));
    short* d_states = reinterpret_cast<short*>(reinterpret_cast<unsigned char*>(d_ptr) + align_up((size_t)cap_frames * 4 * pc.n_bins, 256));
    const int* d_ychunk = reinterpret_cast<const int*>(ws + L.off_yin_chunk);
    int c0 = 0;
    while (c0 < n) {
      int c1 = c0; long long F = 0;
      while (c1 < n) {
        const long long f = s_frame_off[c1 + 1] - s_frame_off[c1];
        if (c1 > c0 && F + f > cap_frames) break;
        F += f; ++c1;
      }
      // a wave that is not the last one is cut to whole rounds of the Viterbi kernel (one CTA per clip, kVitMinCtas CTAs
      // per SM): a partial round at the end of every wave would leave SMs idle for a whole clip's decode
      {
        const int unit = h->sm_count * afx::kVitMinCtas;
        if (c1 < n && c1 - c0 > unit) c1 = c0 + (c1 - c0) / unit * unit;
      }
      {
        afx::YinArgs ya;
        ya.audio = d_audio; ya.offsets = d_offsets; ya.lengths = d_lengths; ya.info = d_info;
        ya.chunk_off = d_ychunk; ya.frame_off = d_frame_off; ya.clip_begin = c0; ya.clip_end = c1; ya.frame_base = s_frame_off[c0];
        ya.n_fft = h->cfg.frame_length; ya.hop = hop; ya.win = pc.win_length; ya.min_period = pc.min_period; ya.max_period = pc.max_period;
        ya.blk_len = h->yin_blk_len; ya.blk_stride = hop; ya.rb = h->yin_rb;
        ya.n_bins = pc.n_bins; ya.frames_per_chunk = h->yin_frames_per_chunk; ya.lag_pad = h->yin_lag_pad; ya.tile = h->yin_tile;
        ya.cap = h->yin_cap; ya.rec_bytes = h->yin_rec_bytes; ya.coef = coef; ya.raw = raw ? 1 : 0; ya.sr = (double)h->cfg.sr; ya.fmin = h->cfg.f0_min;
        ya.tab = h->ptab; ya.obs = d_obs;
        const int grid_y = s_ychunk[c1] - s_ychunk[c0];
        PhaseTimer pt(h, stream, AFX_PHASE_YIN);
        if (grid_y > 0) {
          h->yin_fn<<<grid_y, afx::kYinThreads, h->yin_smem, stream>>>(ya);
          ++g_launches;
        }
      }
      {
        afx::VitArgs va;
        va.info = d_info; va.frame_off = d_frame_off; va.clip_begin = c0; va.frame_base = s_frame_off[c0];
        va.obs = d_obs; va.cap = h->yin_cap; va.rec_bytes = h->yin_rec_bytes; va.n_bins = pc.n_bins; va.half_width = pc.half_width;
        va.lt_same = h->d_lt_same; va.edge_corr = h->d_edge_corr; va.freqs = h->d_freqs;
        va.delta_switch = pc.ltrans_sw[pc.half_width] - pc.ltrans_same[pc.half_width];
        va.log_tiny = pc.log_tiny; va.log_init_unvoiced = pc.log_init_unvoiced;
        va.ptr = d_ptr; va.states = d_states; va.summaries = d_summaries; va.summary_width = afx_summary_width(h);
        va.frames_out = d_frames; va.frame_width = afx_frame_width(h); va.n_mfcc = h->cfg.n_mfcc;
        PhaseTimer pt(h, stream, AFX_PHASE_VITERBI);
        for (int d = 0; d < 64; ++d) {
          const int w = 2 * pc.half_width + 1;
          va.wk[d] = d < w ? (int)std::floor(pc.ltrans_same[w - 1 - d] * 16384.0) * 64 + (63 - d) : -(1 << 30);
        }
        h->vit_fn<<<c1 - c0, h->vit_threads, h->vit_smem, stream>>>(va);
        ++g_launches;
      }
      c0 = c1;
    }
  }
  CUDA_TRY(cudaGetLastError());
  return AFX_OK;
}

// y[n] = x[n] - c x[n-1] (librosa's first-sample rule at n = 0) for the samples of the trimmed range, same packing as x
__global__ void preemph_kernel(const float* __restrict__ x, float* __restrict__ y, const long long* __restrict__ offsets,
                               const int* __restrict__ lengths, const int* __restrict__ meta, float coef) {
  const int clip = blockIdx.y;
  const int L0 = lengths[clip];
  const int start = meta[4 * clip + 1], end = meta[4 * clip + 2];
  const float* xc = x + offsets[clip];
  float* yc = y + offsets[clip];
  for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < L0; n += gridDim.x * blockDim.x) {
    float v = 0.f;
    if (n >= start && n < end) {
      const float xv = xc[n];
      if (n == 0) v = xv + (2.f * xv - (L0 > 1 ? xc[1] : 0.f));
      else v = __fadd_rn(xv, __fmul_rn(-coef, xc[n - 1]));   // scipy.signal.lfilter in float32
    }
    yc[n] = v;
  }
}

int afx_preprocess(afx_handle_t h, const float* d_audio, const int64_t* h_offsets, const int32_t* h_lengths, int64_t n_clips,
                   float* d_out, int32_t* d_meta, float* d_summaries_scratch, void* d_workspace, size_t workspace_bytes, void* stream_v) {
  if (!h || !d_out || !d_summaries_scratch) return fail(AFX_ERR_INVALID_ARG, "afx_preprocess: null argument");
  if (n_clips == 0) return AFX_OK;
  // the trim decision comes from the fused kernel run with no feature group (it still streams the clip once)
  const int rc = afx_extract(h, d_audio, h_offsets, h_lengths, n_clips, 0u, d_summaries_scratch, d_meta, nullptr, d_workspace,
                             workspace_bytes, stream_v);
  if (rc != AFX_OK) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  const Layout L = header_layout(n_clips);
  const char* ws = static_cast<const char*>(d_workspace);
  long long max_len = 1;
  for (int64_t i = 0; i < n_clips; ++i) max_len = std::max<long long>(max_len, h_lengths[i]);
  const dim3 grid((unsigned)std::min<long long>(ceil_div(max_len, 256), 1024), (unsigned)n_clips);
  if (n_clips > 65535) return fail(AFX_ERR_UNSUPPORTED, "afx_preprocess: at most 65535 clips per call");
  preemph_kernel<<<grid, 256, 0, stream>>>(d_audio, d_out, reinterpret_cast<const long long*>(ws + L.off_offsets),
                                           reinterpret_cast<const int*>(ws + L.off_lengths), d_meta, h->cfg.pre_emphasis);
  ++g_launches;
  CUDA_TRY(cudaGetLastError());
  return AFX_OK;
}

int afx_frame_extras(afx_handle_t h, const float* d_audio, int64_t n_clips, const int32_t* d_meta, const float* d_frames,
                     float* d_extras, const void* d_workspace, void* stream_v) {
  if (!h || !d_audio || !d_meta || !d_frames || !d_extras || !d_workspace) return fail(AFX_ERR_INVALID_ARG, "afx_frame_extras: null argument");
  if (n_clips <= 0) return AFX_OK;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  const Layout L = header_layout(n_clips);
  const char* ws = static_cast<const char*>(d_workspace);
  afx::ExtrasArgs ea;
  ea.audio = d_audio;
  ea.offsets = reinterpret_cast<const long long*>(ws + L.off_offsets);
  ea.lengths = reinterpret_cast<const int*>(ws + L.off_lengths);
  ea.frame_off = reinterpret_cast<const long long*>(ws + L.off_frame_off);
  ea.meta = d_meta; ea.frames = d_frames; ea.extras = d_extras;
  ea.frame_width = afx_frame_width(h); ea.n_mfcc = h->cfg.n_mfcc; ea.n_fft = h->cfg.frame_length; ea.hop = h->cfg.hop_length;
  ea.coef = h->cfg.pre_emphasis;
  const int ex_smem = 2 * (afx::kExtrasMaxBlocks + 1) * (int)sizeof(int);
  CUDA_TRY(cudaFuncSetAttribute(afx::frame_extras_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ex_smem));
  afx::frame_extras_kernel<<<(unsigned)n_clips, afx::kExtrasThreads, ex_smem, static_cast<cudaStream_t>(stream_v)>>>(ea);
  ++g_launches;
  CUDA_TRY(cudaGetLastError());
  return AFX_OK;
}

int afx_wav_probe_batch(const char* const* paths, int64_t n, int threads, int32_t* sr, int32_t* channels, int64_t* frames,
                        int32_t* format, int32_t* status) {
  if (n < 0 || (n > 0 && (!paths || !sr || !channels || !frames || !format || !status)))
    return fail(AFX_ERR_INVALID_ARG, "afx_wav_probe_batch: null argument");
  afx::wav::parallel_for(n, threads, [&](int64_t i) {
    afx::wav::Info info;
    try { status[i] = paths[i] ? afx::wav::probe(paths[i], info) : afx::wav::OPEN_FAILED; }
    catch (...) { status[i] = afx::wav::UNSUPPORTED; }   // e.g. std::bad_alloc: reported per file, never across the C boundary
    sr[i] = info.sr; channels[i] = info.channels; frames[i] = info.frames; format[i] = (info.tag << 8) | info.bits;
  });
  return AFX_OK;
}

int afx_wav_load_batch(const char* const* paths, int64_t n, int threads, int sample_format, void* h_out,
                       const int64_t* h_offsets, const int64_t* h_frames, int32_t* status) {
  if (n < 0 || (n > 0 && (!paths || !h_out || !h_offsets || !h_frames || !status)))
    return fail(AFX_ERR_INVALID_ARG, "afx_wav_load_batch: null argument");
  if (sample_format != AFX_SAMPLE_F32 && sample_format != AFX_SAMPLE_S16)
    return fail(AFX_ERR_INVALID_ARG, "afx_wav_load_batch: unknown sample format %d", sample_format);
  afx::wav::parallel_for(n, threads, [&](int64_t i) {
    if (!paths[i] || h_offsets[i] < 0 || h_frames[i] < 0) { status[i] = afx::wav::OPEN_FAILED; return; }
    try {
      if (sample_format == AFX_SAMPLE_S16) status[i] = afx::wav::load(paths[i], h_frames[i], nullptr, static_cast<int16_t*>(h_out) + h_offsets[i]);
      else status[i] = afx::wav::load(paths[i], h_frames[i], static_cast<float*>(h_out) + h_offsets[i], nullptr);
    } catch (...) { status[i] = afx::wav::UNSUPPORTED; }   // the reference's batch_process logs and skips a bad file (:233-235)
  });
  return AFX_OK;
}

int afx_f0_postprocess(afx_handle_t h, int64_t n_clips, const int32_t* h_lengths, const int32_t* d_meta, const float* d_frames,
                       float* d_track, float* d_stats, const void* d_workspace, void* stream_v) {
  if (!h || !h_lengths || !d_meta || !d_frames || !d_stats || !d_workspace) return fail(AFX_ERR_INVALID_ARG, "afx_f0_postprocess: null argument");
  if (n_clips <= 0) return AFX_OK;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  const Layout L = header_layout(n_clips);
  const char* ws = static_cast<const char*>(d_workspace);
  afx::F0PostArgs fa;
  fa.frame_off = reinterpret_cast<const long long*>(ws + L.off_frame_off);
  fa.meta = d_meta; fa.frames = d_frames; fa.track = d_track; fa.stats = d_stats;
  fa.frame_width = afx_frame_width(h); fa.n_mfcc = h->cfg.n_mfcc;
  long long max_f = 16;
  for (int64_t i = 0; i < n_clips; ++i) max_f = std::max<long long>(max_f, frames_ub(h, h_lengths[i]));
  fa.max_frames = (int)((max_f + 1) & ~1LL);
  fa.fmin = h->cfg.f0_min; fa.fmax = h->cfg.f0_max;
  // Savitzky-Golay(11, 3) as a least-squares hat matrix H = V (V^T V)^-1 V^T on x = -5 .. 5 (centred: the fit is shift
  // invariant): row 5 is the interior filter, rows 0 .. 4 scipy's mode='interp' edge fit evaluated at the first 5 points
  {
    double V[11][4], G[4][8];
    for (int i = 0; i < 11; ++i) { const double x = i - 5; V[i][0] = 1; V[i][1] = x; V[i][2] = x * x; V[i][3] = x * x * x; }
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { double acc = 0; for (int i = 0; i < 11; ++i) acc += V[i][r] * V[i][c]; G[r][c] = acc; G[r][4 + c] = r == c; }
    for (int p = 0; p < 4; ++p) {   // Gauss-Jordan with partial pivoting on [V^T V | I]
      int best = p; for (int r = p + 1; r < 4; ++r) if (std::fabs(G[r][p]) > std::fabs(G[best][p])) best = r;
      for (int c = 0; c < 8; ++c) std::swap(G[p][c], G[best][c]);
      const double piv = G[p][p]; for (int c = 0; c < 8; ++c) G[p][c] /= piv;
      for (int r = 0; r < 4; ++r) if (r != p) { const double f = G[r][p]; for (int c = 0; c < 8; ++c) G[r][c] -= f * G[p][c]; }
    }
    for (int i = 0; i < 6; ++i) for (int j = 0; j < 11; ++j) {
      double acc = 0;
      for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) acc += V[i][r] * G[r][4 + c] * V[j][c];
      fa.hat[i][j] = acc;
    }
  }
  const size_t smem = (size_t)fa.max_frames * (8 + 8 + 4);
  if (smem > 200 * 1024) return fail(AFX_ERR_UNSUPPORTED, "afx_f0_postprocess: clips of more than %d frames are not supported", (200 * 1024) / 20);
  CUDA_TRY(cudaFuncSetAttribute(afx::f0_post_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  afx::f0_post_kernel<<<(unsigned)n_clips, afx::kF0PostThreads, smem, static_cast<cudaStream_t>(stream_v)>>>(fa);
  ++g_launches;
  CUDA_TRY(cudaGetLastError());
  return AFX_OK;
}

static int upload_small_header(afx_handle_t h, const void* host, size_t bytes, cudaStream_t stream);

int afx_clip_kernel_variant(int32_t sr, int32_t frame_length, int32_t hop_length, int32_t* mel_runs) {
  if (sr < 1 || (frame_length != 512 && frame_length != 1024 && frame_length != 2048)) return -1;
  const auto mel = afx::host::mel_sparse(sr, frame_length, afx::kNMels);
  if (mel_runs) { for (int q = 0; q < 4; ++q) mel_runs[q] = mel.cnt[q]; mel_runs[4] = mel.cnt0; }
  const stream_fn_t fn = stream_kernel_for(frame_length, hop_length, mel);
  if (!fn) return -1;
  if (runtime_hop(frame_length, hop_length)) return 2;
  return fn == stream_kernel_for(frame_length, hop_length, afx::host::MelSparse{}) ? 0 : 1;
}

int64_t afx_resample_filter(int32_t sr, int32_t native_sr, float* taps, int64_t taps_capacity, int32_t* up, int32_t* down,
                            int32_t* taps_per_phase, int32_t* n_pre_remove) {
  if (sr < 1 || native_sr < 1 || sr > (1 << 22) || native_sr > (1 << 22)) { fail(AFX_ERR_INVALID_ARG, "afx_resample_filter: bad sample rates %d, %d", sr, native_sr); return -1; }
  const afx::ResampleFilter f = afx::design_resample_filter(sr, native_sr);
  if (up) *up = f.up;
  if (down) *down = f.down;
  if (taps_per_phase) *taps_per_phase = f.J;
  if (n_pre_remove) *n_pre_remove = f.n_pre_remove;
  const int64_t count = (int64_t)f.taps.size();
  if (taps && taps_capacity >= count) std::memcpy(taps, f.taps.data(), f.taps.size() * sizeof(float));
  return count;
}

int64_t afx_resample_length(afx_handle_t h, int32_t native_sr, int64_t n_in) {
  if (!h || native_sr < 1 || n_in < 0) return -1;
  const int g = std::gcd(h->cfg.sr, native_sr);
  return afx::resample_out_len(n_in, h->cfg.sr / g, native_sr / g);
}

int afx_resample(afx_handle_t h, const float* d_in, const int64_t* h_in_offsets, const int32_t* h_in_lengths, int64_t n_clips,
                 int32_t native_sr, float* d_out, const int64_t* h_out_offsets, void* stream_v) {
  if (!h || !d_in || !d_out || !h_in_offsets || !h_in_lengths || !h_out_offsets) return fail(AFX_ERR_INVALID_ARG, "afx_resample: null argument");
  if (native_sr < 1 || native_sr > (1 << 22)) return fail(AFX_ERR_INVALID_ARG, "afx_resample: bad native sample rate %d", native_sr);
  if (n_clips < 0 || n_clips > 0x7fffff00LL) return fail(AFX_ERR_INVALID_ARG, "afx_resample: bad n_clips");
  if (n_clips == 0) return AFX_OK;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  auto it = h->resamplers.find(native_sr);
  if (it == h->resamplers.end()) {
    afx_handle_s::ResampleDev rd;
    rd.f = afx::design_resample_filter(h->cfg.sr, native_sr);
    if ((size_t)rd.f.up * rd.f.J > (64u << 20)) return fail(AFX_ERR_UNSUPPORTED, "afx_resample: %d -> %d Hz needs a %d-phase filter", native_sr, h->cfg.sr, rd.f.up);
    CUDA_TRY(cudaMalloc(&rd.d_taps, rd.f.taps.size() * sizeof(float)));
    CUDA_TRY(cudaMemcpy(rd.d_taps, rd.f.taps.data(), rd.f.taps.size() * sizeof(float), cudaMemcpyHostToDevice));
    it = h->resamplers.emplace(native_sr, std::move(rd)).first;
  }
  const afx::ResampleFilter& f = it->second.f;
  // header: in_off | out_off (long long), in_len | out_len (int)
  const size_t n = (size_t)n_clips, bytes = n * (2 * sizeof(long long) + 2 * sizeof(int));
  std::vector<unsigned char> hdr(bytes);
  long long* ho = reinterpret_cast<long long*>(hdr.data());
  int* hl = reinterpret_cast<int*>(hdr.data() + 2 * n * sizeof(long long));
  long long max_out = 0;
  for (size_t i = 0; i < n; ++i) {
    if (h_in_lengths[i] < 0 || h_in_offsets[i] < 0 || h_out_offsets[i] < 0) return fail(AFX_ERR_INVALID_ARG, "afx_resample: clip %zu has a negative length or offset", i);
    const long long no = afx::resample_out_len(h_in_lengths[i], f.up, f.down);
    if (no > 0x7fffffffLL) return fail(AFX_ERR_UNSUPPORTED, "afx_resample: clip %zu is too long", i);
    ho[i] = h_in_offsets[i]; ho[n + i] = h_out_offsets[i]; hl[i] = h_in_lengths[i]; hl[n + i] = (int)no;
    max_out = std::max(max_out, no);
  }
  if (max_out == 0) return AFX_OK;
  if (int rc = upload_small_header(h, hdr.data(), bytes, stream)) return rc;
  afx::ResampleArgs ra;
  ra.in = d_in; ra.out = d_out;
  ra.in_off = reinterpret_cast<const long long*>(h->d_rs_hdr);
  ra.out_off = ra.in_off + n;
  ra.in_len = reinterpret_cast<const int*>(static_cast<const unsigned char*>(h->d_rs_hdr) + 2 * n * sizeof(long long));
  ra.out_len = ra.in_len + n;
  ra.taps = it->second.d_taps; ra.up = f.up; ra.down = f.down; ra.J = f.J; ra.n_pre_remove = f.n_pre_remove;
  const long long tile = afx::kResampleThreads * afx::kResamplePerThread;
  const long long tiles = (max_out + tile - 1) / tile;
  const unsigned gx = (unsigned)((tiles + afx::kResampleTilesPerCta - 1) / afx::kResampleTilesPerCta);
  const size_t smem = (size_t)f.up * (f.J | 1) * sizeof(float);
  const bool smem_taps = smem <= 200 * 1024;
  if (smem_taps) CUDA_TRY(cudaFuncSetAttribute(afx::resample_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  for (int64_t c0 = 0; c0 < n_clips; c0 += 65535) {
    ra.clip0 = (int)c0;
    const unsigned gy = (unsigned)std::min<int64_t>(65535, n_clips - c0);
    if (smem_taps) afx::resample_kernel<true><<<dim3(gx, gy), afx::kResampleThreads, smem, stream>>>(ra, (int)n_clips);
    else afx::resample_kernel<false><<<dim3(gx, gy), afx::kResampleThreads, 0, stream>>>(ra, (int)n_clips);
    ++g_launches;
  }
  CUDA_TRY(cudaGetLastError());
  return AFX_OK;
}

// header arrays of the small batched entry points (afx_resample, afx_dtw_*): one handle-owned device buffer
static int upload_small_header(afx_handle_t h, const void* host, size_t bytes, cudaStream_t stream) {
  if (h->d_rs_hdr_bytes < bytes) {
    CUDA_TRY(cudaFree(h->d_rs_hdr)); h->d_rs_hdr = nullptr; h->d_rs_hdr_bytes = 0;
    CUDA_TRY(cudaMalloc(&h->d_rs_hdr, bytes * 2));
    h->d_rs_hdr_bytes = bytes * 2;
  }
  CUDA_TRY(cudaStreamSynchronize(stream));   // a previous call on this stream may still read the header
  CUDA_TRY(cudaMemcpyAsync(h->d_rs_hdr, host, bytes, cudaMemcpyHostToDevice, stream));
  return AFX_OK;
}

int afx_dtw_frame_distance(afx_handle_t h, const double* d_x, const int64_t* h_x_off, const int32_t* h_n, const double* d_y,
                           const int64_t* h_y_off, const int32_t* h_m, int32_t dim, int64_t n_pairs, double* d_D,
                           const int64_t* h_D_off, void* stream_v) {
  if (!h || !d_x || !d_y || !d_D || !h_x_off || !h_y_off || !h_D_off || !h_n || !h_m) return fail(AFX_ERR_INVALID_ARG, "afx_dtw_frame_distance: null argument");
  if (dim < 1 || dim > 200) return fail(AFX_ERR_INVALID_ARG, "afx_dtw_frame_distance: feature dimension %d out of range [1, 200]", dim);
  if (n_pairs < 0 || n_pairs > 0x7fffff00LL) return fail(AFX_ERR_INVALID_ARG, "afx_dtw_frame_distance: bad n_pairs");
  if (n_pairs == 0) return AFX_OK;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  const size_t n = (size_t)n_pairs, bytes = n * (3 * sizeof(long long) + 2 * sizeof(int));
  std::vector<unsigned char> hdr(bytes);
  long long* ho = reinterpret_cast<long long*>(hdr.data());
  int* hl = reinterpret_cast<int*>(hdr.data() + 3 * n * sizeof(long long));
  long long max_tiles = 0;
  for (size_t i = 0; i < n; ++i) {
    if (h_n[i] < 0 || h_m[i] < 0 || h_x_off[i] < 0 || h_y_off[i] < 0 || h_D_off[i] < 0) return fail(AFX_ERR_INVALID_ARG, "afx_dtw_frame_distance: pair %zu has a negative size or offset", i);
    ho[i] = h_x_off[i]; ho[n + i] = h_y_off[i]; ho[2 * n + i] = h_D_off[i]; hl[i] = h_n[i]; hl[n + i] = h_m[i];
    const long long ti = (h_n[i] + afx::kDtwTile - 1) / afx::kDtwTile, tj = (h_m[i] + afx::kDtwTile - 1) / afx::kDtwTile;
    max_tiles = std::max(max_tiles, ti * tj);
  }
  if (max_tiles == 0) return AFX_OK;
  if (int rc = upload_small_header(h, hdr.data(), bytes, stream)) return rc;
  afx::DtwDistArgs da;
  da.x = d_x; da.y = d_y; da.D = d_D;
  da.x_off = reinterpret_cast<const long long*>(h->d_rs_hdr); da.y_off = da.x_off + n; da.D_off = da.y_off + n;
  da.n = reinterpret_cast<const int*>(static_cast<const unsigned char*>(h->d_rs_hdr) + 3 * n * sizeof(long long)); da.m = da.n + n;
  da.dim = dim;
  int ld = (dim + 3) & ~3; while ((ld & 7) != 4) ld += 4;   // DMMA fragment reads (8 rows x 4 columns) hit distinct banks
  da.ld = ld;
  const size_t smem = ((size_t)2 * afx::kDtwTile * ld + 2 * afx::kDtwTile * 4) * sizeof(double);
  CUDA_TRY(cudaFuncSetAttribute(afx::dtw_distance_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // a few thousand CTAs in flight are plenty: the tiles of a pair are strided over gridDim.x
  const unsigned gx = (unsigned)std::min<long long>(max_tiles, std::max<long long>(1, (long long)h->sm_count * 16 / std::min<int64_t>(n_pairs, 65535) + 1));
  for (int64_t c0 = 0; c0 < n_pairs; c0 += 65535) {
    da.pair0 = (int)c0;
    const unsigned gy = (unsigned)std::min<int64_t>(65535, n_pairs - c0);
    afx::dtw_distance_kernel<<<dim3(gx, gy), afx::kDtwDistThreads, smem, stream>>>(da);
    ++g_launches;
  }
  CUDA_TRY(cudaGetLastError());
  return AFX_OK;
}

int afx_dtw_accumulated_cost(afx_handle_t h, const double* d_D, const int64_t* h_D_off, const int32_t* h_n, const int32_t* h_m,
                             int64_t n_pairs, double* d_A, void* stream_v) {
  if (!h || !d_D || !d_A || !h_D_off || !h_n || !h_m) return fail(AFX_ERR_INVALID_ARG, "afx_dtw_accumulated_cost: null argument");
  if (n_pairs < 0 || n_pairs > 0x7fffff00LL) return fail(AFX_ERR_INVALID_ARG, "afx_dtw_accumulated_cost: bad n_pairs");
  if (n_pairs == 0) return AFX_OK;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  const size_t n = (size_t)n_pairs, bytes = n * (sizeof(long long) + 2 * sizeof(int));
  std::vector<unsigned char> hdr(bytes);
  long long* ho = reinterpret_cast<long long*>(hdr.data());
  int* hl = reinterpret_cast<int*>(hdr.data() + n * sizeof(long long));
  int max_n = 0;
  for (size_t i = 0; i < n; ++i) {
    if (h_n[i] < 0 || h_m[i] < 0 || h_D_off[i] < 0) return fail(AFX_ERR_INVALID_ARG, "afx_dtw_accumulated_cost: pair %zu has a negative size or offset", i);
    ho[i] = h_D_off[i]; hl[i] = h_n[i]; hl[n + i] = h_m[i];
    max_n = std::max(max_n, h_n[i]);
  }
  if (int rc = upload_small_header(h, hdr.data(), bytes, stream)) return rc;
  afx::DtwAccArgs aa;
  aa.D = d_D; aa.A = d_A;
  aa.D_off = reinterpret_cast<const long long*>(h->d_rs_hdr);
  aa.n = reinterpret_cast<const int*>(static_cast<const unsigned char*>(h->d_rs_hdr) + n * sizeof(long long)); aa.m = aa.n + n;
  (void)max_n;
  CUDA_TRY(cudaFuncSetAttribute(afx::dtw_accumulate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, afx::kDtwAccSmem));
  aa.pair0 = 0;
  const unsigned g = (unsigned)((n_pairs + afx::kDtwAccWarps - 1) / afx::kDtwAccWarps);
  afx::dtw_accumulate_kernel<<<g, afx::kDtwAccWarps * 32, afx::kDtwAccSmem, stream>>>(aa, (int)n_pairs);
  ++g_launches;
  CUDA_TRY(cudaGetLastError());
  return AFX_OK;
}

int afx_profile_enable(afx_handle_t h, int on) {
  if (!h) return fail(AFX_ERR_INVALID_ARG, "afx_profile_enable: null handle");
  h->profiling = on != 0;
  return AFX_OK;
}

int afx_profile_read(afx_handle_t h, double* ms, int64_t* launches, int reset) {
  if (!h) return fail(AFX_ERR_INVALID_ARG, "afx_profile_read: null handle");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  for (auto& sp : h->spans) {
    CUDA_TRY(cudaEventSynchronize(sp.b));
    float t = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&t, sp.a, sp.b));
    h->prof_ms[sp.phase] += t;
    h->ev_pool.push_back(sp.a); h->ev_pool.push_back(sp.b);
  }
  h->spans.clear();
  for (int p = 0; p < AFX_N_PHASES; ++p) {
    if (ms) ms[p] = h->prof_ms[p];
    if (launches) launches[p] = h->prof_launches[p];
    if (reset) { h->prof_ms[p] = 0; h->prof_launches[p] = 0; }
  }
  return AFX_OK;
}

// test/diagnostic aid: byte offset of the log-mel scratch inside a workspace for n_clips clips
size_t afx_debug_header_bytes(int64_t n_clips) { return header_layout(n_clips).header_bytes; }

}  // extern "C"

namespace {

__global__ void pcm16_to_f32_kernel(const int16_t* __restrict__ in, float* __restrict__ out, long long n8) {
  // libsndfile / librosa.load scaling of 16-bit PCM: x / 32768 (exact in float32); 8 samples per thread
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const int4 v = __ldg(reinterpret_cast<const int4*>(in) + i);
    const int w[4] = {v.x, v.y, v.z, v.w};
    float o[8];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      o[2 * k] = (float)(short)(w[k] & 0xffff) * (1.0f / 32768.0f);
      o[2 * k + 1] = (float)(short)(w[k] >> 16) * (1.0f / 32768.0f);
    }
    float4* dst = reinterpret_cast<float4*>(out) + 2 * i;
    dst[0] = make_float4(o[0], o[1], o[2], o[3]);
    dst[1] = make_float4(o[4], o[5], o[6], o[7]);
  }
}

int ensure_host_path(afx_handle_s* h, size_t wave_floats, bool need_pcm, bool need_pin, int64_t n_clips) {
  if (!h->copy_stream) {
    CUDA_TRY(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&h->compute_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      CUDA_TRY(cudaEventCreateWithFlags(&h->copied[i], cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&h->computed[i], cudaEventDisableTiming));
    }
  }
  if (h->d_audio_floats < wave_floats) {
    for (int i = 0; i < 2; ++i) { if (h->d_audio[i]) CUDA_TRY(cudaFree(h->d_audio[i])); h->d_audio[i] = nullptr; }
    for (int i = 0; i < 2; ++i) CUDA_TRY(cudaMalloc((void**)&h->d_audio[i], wave_floats * sizeof(float)));
    h->d_audio_floats = wave_floats;
  }
  if (need_pcm && h->d_pcm_samples < wave_floats) {
    for (int i = 0; i < 2; ++i) { if (h->d_pcm[i]) CUDA_TRY(cudaFree(h->d_pcm[i])); h->d_pcm[i] = nullptr; }
    for (int i = 0; i < 2; ++i) CUDA_TRY(cudaMalloc((void**)&h->d_pcm[i], wave_floats * sizeof(int16_t)));
    h->d_pcm_samples = wave_floats;
  }
  if (need_pin && h->h_pin_floats < wave_floats) {
    for (int i = 0; i < 2; ++i) { if (h->h_pin[i]) CUDA_TRY(cudaFreeHost(h->h_pin[i])); h->h_pin[i] = nullptr; }
    for (int i = 0; i < 2; ++i) CUDA_TRY(cudaMallocHost((void**)&h->h_pin[i], wave_floats * sizeof(float)));
    h->h_pin_floats = wave_floats;
  }
  const size_t sum_bytes = (size_t)n_clips * afx_summary_width(h) * sizeof(float);
  const size_t meta_bytes = (size_t)n_clips * 4 * sizeof(int);
  if (h->d_sum_bytes < sum_bytes) { if (h->d_sum) CUDA_TRY(cudaFree(h->d_sum)); h->d_sum = nullptr; CUDA_TRY(cudaMalloc((void**)&h->d_sum, sum_bytes)); h->d_sum_bytes = sum_bytes; }
  if (h->d_meta_bytes < meta_bytes) { if (h->d_meta) CUDA_TRY(cudaFree(h->d_meta)); h->d_meta = nullptr; CUDA_TRY(cudaMalloc((void**)&h->d_meta, meta_bytes)); h->d_meta_bytes = meta_bytes; }
  return AFX_OK;
}

struct HostWave { int64_t c0, c1; long long base, span; };   // clips [c0,c1); samples [base, base+span) of the packed buffer

// Shared driver of the two host entry points.  `fill(slot, wave)` makes wave's samples available for the
// H2D copy and returns the host source pointer (the caller's packed buffer, or a pinned staging slot).
template <typename Fill>
int run_host_waves(afx_handle_s* h, const std::vector<HostWave>& waves, const std::vector<long long>& rel_off,
                   const int32_t* h_lengths, int64_t n_clips, uint32_t feature_mask, int sample_format,
                   float* h_summaries, int32_t* h_meta, Fill fill) {
  const int SS = afx_summary_width(h);
  const size_t elt = sample_format == AFX_SAMPLE_S16 ? sizeof(int16_t) : sizeof(float);
  // workspace for the largest wave
  size_t ws_need = 0;
  for (const HostWave& w : waves) ws_need = std::max(ws_need, afx_workspace_bytes(h, w.c1 - w.c0, h_lengths + w.c0));
  if (h->d_ws_bytes < ws_need) { if (h->d_ws) CUDA_TRY(cudaFree(h->d_ws)); h->d_ws = nullptr; CUDA_TRY(cudaMalloc(&h->d_ws, ws_need)); h->d_ws_bytes = ws_need; }

  auto enqueue_copy = [&](size_t wi) -> int {
    const HostWave& w = waves[wi];
    const int slot = (int)(wi & 1);
    CUDA_TRY(cudaStreamWaitEvent(h->copy_stream, h->computed[slot], 0));   // slot's previous wave has been consumed
    const void* src = fill(slot, w);
    if (!src) return AFX_ERR_CUDA;
    void* dst = sample_format == AFX_SAMPLE_S16 ? (void*)h->d_pcm[slot] : (void*)h->d_audio[slot];
    CUDA_TRY(cudaMemcpyAsync(dst, src, (size_t)w.span * elt, cudaMemcpyHostToDevice, h->copy_stream));
    CUDA_TRY(cudaEventRecord(h->copied[slot], h->copy_stream));
    return AFX_OK;
  };

  if (int rc = enqueue_copy(0)) return rc;
  for (size_t wi = 0; wi < waves.size(); ++wi) {
    const HostWave& w = waves[wi];
    const int slot = (int)(wi & 1);
    if (wi + 1 < waves.size()) { if (int rc = enqueue_copy(wi + 1)) return rc; }
    CUDA_TRY(cudaStreamWaitEvent(h->compute_stream, h->copied[slot], 0));
    if (sample_format == AFX_SAMPLE_S16) {
      const long long n8 = (w.span + 7) / 8;
      const int grid = (int)std::min<long long>((n8 + 255) / 256, 8LL * h->sm_count);
      PhaseTimer pt(h, h->compute_stream, AFX_PHASE_PCM);
      pcm16_to_f32_kernel<<<grid, 256, 0, h->compute_stream>>>(h->d_pcm[slot], h->d_audio[slot], n8);
      ++g_launches;
    }
    const int rc = afx_extract(h, h->d_audio[slot], reinterpret_cast<const int64_t*>(rel_off.data() + w.c0), h_lengths + w.c0, w.c1 - w.c0,
                               feature_mask, h->d_sum + (size_t)w.c0 * SS, h->d_meta + (size_t)w.c0 * 4, nullptr,
                               h->d_ws, h->d_ws_bytes, h->compute_stream);
    if (rc != AFX_OK) { cudaDeviceSynchronize(); return rc; }
    CUDA_TRY(cudaEventRecord(h->computed[slot], h->compute_stream));
  }
  CUDA_TRY(cudaMemcpyAsync(h_summaries, h->d_sum, (size_t)n_clips * SS * sizeof(float), cudaMemcpyDeviceToHost, h->compute_stream));
  CUDA_TRY(cudaMemcpyAsync(h_meta, h->d_meta, (size_t)n_clips * 4 * sizeof(int), cudaMemcpyDeviceToHost, h->compute_stream));
  CUDA_TRY(cudaStreamSynchronize(h->compute_stream));
  CUDA_TRY(cudaStreamSynchronize(h->copy_stream));
  return AFX_OK;
}

// split clips [0,n) with (4-aligned) sample offsets `off` into waves of at most `cap` samples (a longer clip gets its own wave)
int plan_waves(const int64_t* off, const int32_t* len, int64_t n, long long cap_full, std::vector<HostWave>& waves,
               std::vector<long long>& rel_off, long long* max_span) {
  rel_off.resize(n);
  *max_span = 0;
  int64_t c0 = 0;
  while (c0 < n) {
    const long long base = off[c0];
    int64_t c1 = c0;
    long long end = base;
    // ramp: the first copy cannot overlap any compute, so it is kept short; sizes then double up to `cap_full`
    const long long cap = waves.size() == 0 ? cap_full / 8 : waves.size() == 1 ? cap_full / 4 : waves.size() == 2 ? cap_full / 2 : cap_full;
    while (c1 < n) {
      if (c1 > c0 && off[c1] < off[c1 - 1] + len[c1 - 1]) return fail(AFX_ERR_INVALID_ARG, "host path: clip offsets must be ascending and non-overlapping (clip %lld)", (long long)c1);
      const long long e = (off[c1] + len[c1] + 7) & ~7LL;   // 8-sample granularity (s16 conversion, TMA tail reads)
      if (c1 > c0 && e - base > cap) break;
      end = e; rel_off[c1] = off[c1] - base; ++c1;
    }
    waves.push_back(HostWave{c0, c1, base, end - base});
    *max_span = std::max(*max_span, end - base);
    c0 = c1;
  }
  return AFX_OK;
}

}  // namespace

extern "C" {

int afx_extract_host_packed(afx_handle_t h, const void* h_audio, int sample_format, int64_t h_audio_samples,
                            const int64_t* h_offsets, const int32_t* h_lengths, int64_t n_clips,
                            uint32_t feature_mask, float* h_summaries, int32_t* h_meta) {
  if (!h) return fail(AFX_ERR_INVALID_ARG, "afx_extract_host_packed: null handle");
  if (n_clips == 0) return AFX_OK;
  if (n_clips < 0 || !h_audio || !h_offsets || !h_lengths || !h_summaries || !h_meta)
    return fail(AFX_ERR_INVALID_ARG, "afx_extract_host_packed: bad argument");
  if (sample_format != AFX_SAMPLE_F32 && sample_format != AFX_SAMPLE_S16)
    return fail(AFX_ERR_INVALID_ARG, "afx_extract_host_packed: unknown sample format %d", sample_format);
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  for (int64_t i = 0; i < n_clips; ++i) {
    if (h_lengths[i] < 0 || h_offsets[i] < 0 || (h_offsets[i] & 7) != 0)
      return fail(AFX_ERR_INVALID_ARG, "afx_extract_host_packed: clip %lld offset must be a non-negative multiple of 8 samples", (long long)i);
    if (h_offsets[i] + h_lengths[i] > h_audio_samples)
      return fail(AFX_ERR_INVALID_ARG, "afx_extract_host_packed: clip %lld runs past the %lld-sample buffer", (long long)i, (long long)h_audio_samples);
  }
  std::vector<HostWave> waves;
  std::vector<long long> rel_off;
  long long max_span = 0;
  if (int rc = plan_waves(h_offsets, h_lengths, n_clips, h->host_wave_samples, waves, rel_off, &max_span)) return rc;
  if (int rc = ensure_host_path(h, (size_t)max_span + 8, sample_format == AFX_SAMPLE_S16, false, n_clips)) return rc;
  const size_t elt = sample_format == AFX_SAMPLE_S16 ? sizeof(int16_t) : sizeof(float);
  const char* base = static_cast<const char*>(h_audio);
  // the last wave's 8-sample rounding may run past the caller's buffer: clamp the copy there
  for (HostWave& w : waves) w.span = std::min<long long>(w.span, std::max<long long>(h_audio_samples - w.base, 0));
  return run_host_waves(h, waves, rel_off, h_lengths, n_clips, feature_mask, sample_format, h_summaries, h_meta,
                        [&](int, const HostWave& w) -> const void* { return base + (size_t)w.base * elt; });
}

int afx_extract_host(afx_handle_t h, const float* const* h_clips, const int32_t* h_lengths, int64_t n_clips,
                     uint32_t feature_mask, float* h_summaries, int32_t* h_meta) {
  if (!h) return fail(AFX_ERR_INVALID_ARG, "afx_extract_host: null handle");
  if (n_clips == 0) return AFX_OK;
  if (n_clips < 0 || !h_clips || !h_lengths || !h_summaries || !h_meta) return fail(AFX_ERR_INVALID_ARG, "afx_extract_host: bad argument");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  std::vector<long long> off(n_clips);
  long long o = 0;
  for (int64_t i = 0; i < n_clips; ++i) {
    if (h_lengths[i] < 0 || (h_lengths[i] > 0 && !h_clips[i])) return fail(AFX_ERR_INVALID_ARG, "afx_extract_host: clip %lld is null or has negative length", (long long)i);
    off[i] = o; o += ((long long)h_lengths[i] + 7) & ~7LL;
  }
  std::vector<HostWave> waves;
  std::vector<long long> rel_off;
  long long max_span = 0;
  if (int rc = plan_waves(reinterpret_cast<const int64_t*>(off.data()), h_lengths, n_clips, std::min<long long>(h->host_wave_samples, 32ll << 20) /* pinned staging slots */, waves, rel_off, &max_span)) return rc;
  if (int rc = ensure_host_path(h, (size_t)max_span + 8, false, true, n_clips)) return rc;
  // pack each wave into a pinned slot with a few host threads (the gather is a plain memcpy per clip)
  auto fill = [&](int slot, const HostWave& w) -> const void* {
    if (cudaEventSynchronize(h->copied[slot]) != cudaSuccess) return nullptr;   // slot's previous H2D has drained
    float* dst = h->h_pin[slot];
    const int nthreads = (int)std::min<int64_t>(std::max(1u, std::min(8u, std::thread::hardware_concurrency())), w.c1 - w.c0);
    auto work = [&](int t) {
      for (int64_t c = w.c0 + t; c < w.c1; c += nthreads) {
        float* d = dst + rel_off[c];
        const long long len = h_lengths[c], padded = (len + 7) & ~7LL;
        if (len > 0) memcpy(d, h_clips[c], (size_t)len * sizeof(float));
        for (long long k = len; k < padded; ++k) d[k] = 0.f;
      }
    };
    if (nthreads <= 1) work(0);
    else { std::vector<std::thread> th; for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t); for (auto& x : th) x.join(); }
    return dst;
  };
  return run_host_waves(h, waves, rel_off, h_lengths, n_clips, feature_mask, AFX_SAMPLE_F32, h_summaries, h_meta, fill);
}

}  // extern "C"

// =====================================================================================================
// afx_extract_static: the fused MFCC + delta + energy pass (afx_extract without the F0 group and without framewise
// output) for callers that must not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  No counterpart in the reference.
//
// The scratch of the pass (batch header, work-item table, ring of per-clip slabs -- what afx_extract takes as
// d_workspace) lives in ONE file-scope device array of AFX_STATIC_WORKSPACE_BYTES that every GPU which loads the library
// holds until exit.  Its name is a host-side stand-in in host code, NOT a device address: the device address comes from
// cudaGetSymbolAddress, once per device with that device current, and is what the kernels are given.
// The call allocates nothing (no cudaMalloc / cudaFree / cudaMallocHost, no event creation): the header arrays are built
// in a stack buffer and copied with cudaMemcpyAsync, the work counter is cleared with cudaMemsetAsync (the array keeps
// the last call's contents; everything else the kernels read is rewritten by this call: the slot owner / arrival
// counters by item_build_kernel, the slabs by the steps before the finisher reads them).
// Every call shares the array: calls whose GPU work overlaps (two streams, two threads, two handles on one device) race.
static __device__ __align__(256) unsigned char g_afx_static_ws[AFX_STATIC_WORKSPACE_BYTES];

namespace {
constexpr int kStaticMaxDevices = 64;
unsigned char* g_static_ws_ptr[kStaticMaxDevices] = {nullptr};   // device address of g_afx_static_ws, per device
std::atomic<int> g_static_ws_lock{0};

// device address of the static scratch on `device` (the current device); resolved once per device
int static_ws_address(int device, unsigned char** out) {
  if (device < 0 || device >= kStaticMaxDevices) return fail(AFX_ERR_UNSUPPORTED, "afx_extract_static: device index %d out of range", device);
  while (g_static_ws_lock.exchange(1, std::memory_order_acquire)) {}
  unsigned char* p = g_static_ws_ptr[device];
  cudaError_t e = cudaSuccess;
  if (!p) {
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_afx_static_ws);
    if (e == cudaSuccess) g_static_ws_ptr[device] = p;
  }
  g_static_ws_lock.store(0, std::memory_order_release);
  if (e != cudaSuccess) return fail(AFX_ERR_CUDA, "afx_extract_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
  *out = p;
  return AFX_OK;
}
}  // namespace

extern "C" {

size_t afx_static_workspace_capacity(void) { return (size_t)AFX_STATIC_WORKSPACE_BYTES; }

int afx_extract_static(afx_handle_t h, const float* d_audio, const int64_t* h_offsets, const int32_t* h_lengths,
                       int64_t n_clips, uint32_t feature_mask, float* d_summaries, int32_t* d_meta, void* stream_v) {
  if (!h) return fail(AFX_ERR_INVALID_ARG, "afx_extract_static: null handle");
  if (n_clips == 0) return AFX_OK;
  if (n_clips < 0 || n_clips > AFX_STATIC_MAX_CLIPS)
    return fail(AFX_ERR_UNSUPPORTED, "afx_extract_static: %lld clips exceed the maximum of %d per call", (long long)n_clips, AFX_STATIC_MAX_CLIPS);
  if (!d_audio || !h_offsets || !h_lengths || !d_summaries || !d_meta) return fail(AFX_ERR_INVALID_ARG, "afx_extract_static: null pointer argument");
  if ((feature_mask & ~(AFX_FEAT_MFCC | AFX_FEAT_ENERGY | AFX_FEAT_PREPROCESSED)) != 0)
    return fail(AFX_ERR_UNSUPPORTED, "afx_extract_static: feature bits 0x%x: only AFX_FEAT_MFCC, AFX_FEAT_ENERGY (and AFX_FEAT_PREPROCESSED) are served", feature_mask);
  const bool raw = (feature_mask & AFX_FEAT_PREPROCESSED) != 0;
  feature_mask &= AFX_FEAT_ALL;
  const float coef = raw ? 0.f : h->cfg.pre_emphasis;
  const float trim_top_db = raw ? 3.0e38f : h->cfg.trim_top_db;
  if ((reinterpret_cast<uintptr_t>(d_audio) & 15u) != 0) return fail(AFX_ERR_INVALID_ARG, "afx_extract_static: d_audio must be 16-byte aligned");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  unsigned char* ws_u = nullptr;
  if (int rc = static_ws_address(h->cfg.device, &ws_u)) return rc;
  char* ws = reinterpret_cast<char*>(ws_u);
  const int n = (int)n_clips;
  const int hop = h->cfg.hop_length;

  // ---- header arrays in a stack buffer (same layout as afx_extract's pinned staging)
  const Layout L = header_layout(n);
  const size_t staged_bytes = L.off_info;
  alignas(16) unsigned char st[40 * AFX_STATIC_MAX_CLIPS + 256];
  if (staged_bytes > sizeof st) return fail(AFX_ERR_UNSUPPORTED, "afx_extract_static: header of %zu bytes exceeds the stack buffer", staged_bytes);
  memset(st, 0, staged_bytes);
  long long* s_offsets = reinterpret_cast<long long*>(st + L.off_offsets);
  long long* s_frame_off = reinterpret_cast<long long*>(st + L.off_frame_off);
  int* s_lengths = reinterpret_cast<int*>(st + L.off_lengths);
  int* s_item_off = reinterpret_cast<int*>(st + L.off_item_off);
  long long fo = 0, max_len = 0, n_items = 0;
  const int isteps = item_steps_for(h, n_clips, h_lengths);
  for (int i = 0; i < n; ++i) {
    if (h_lengths[i] < 0) return fail(AFX_ERR_INVALID_ARG, "afx_extract_static: clip %d has negative length", i);
    if (h_offsets[i] < 0 || (h_offsets[i] & 3) != 0) return fail(AFX_ERR_INVALID_ARG, "afx_extract_static: clip %d offset %lld is not a multiple of 4 samples", i, (long long)h_offsets[i]);
    s_offsets[i] = h_offsets[i];
    s_lengths[i] = h_lengths[i];
    s_frame_off[i] = fo;
    s_item_off[i] = (int)n_items;
    n_items += items_of(h, h_lengths[i], isteps);
    fo += frames_ub(h, h_lengths[i]);
    max_len = std::max<long long>(max_len, h_lengths[i]);
  }
  s_frame_off[n] = fo;
  s_item_off[n] = (int)n_items;
  const bool one_pass = single_pass(h);
  const long long max_trim_blocks = (200 * 1024) / 4 - 8;
  if (!one_pass && ceil_div(max_len, afx::kTrimHop) > max_trim_blocks)
    return fail(AFX_ERR_UNSUPPORTED, "afx_extract_static: clips longer than %lld samples need hop_length dividing 512", max_trim_blocks * afx::kTrimHop);

  const size_t slab_bytes = slab_bytes_for(h, max_len);
  const int n_slots = slots_for(h, n, max_len);
  const size_t items_bytes = align_up((size_t)n_items * sizeof(afx::WorkItem), 256);
  const size_t need = L.header_bytes + items_bytes + (size_t)n_slots * slab_bytes;
  if (need > (size_t)AFX_STATIC_WORKSPACE_BYTES)
    return fail(AFX_ERR_WORKSPACE, "afx_extract_static: this batch needs %zu bytes of scratch, the static array holds %zu (fewer or shorter clips, or afx_extract with a caller workspace)",
                need, (size_t)AFX_STATIC_WORKSPACE_BYTES);

  // pageable source: the copy is staged before the call returns, so the stack buffer may go out of scope
  CUDA_TRY(cudaMemcpyAsync(ws, st, staged_bytes, cudaMemcpyHostToDevice, stream));
  CUDA_TRY(cudaMemsetAsync(ws + L.off_counter, 0, 256, stream));   // the work counter must start at zero
  const long long* d_offsets = reinterpret_cast<const long long*>(ws + L.off_offsets);
  const long long* d_frame_off = reinterpret_cast<const long long*>(ws + L.off_frame_off);
  const int* d_lengths = reinterpret_cast<const int*>(ws + L.off_lengths);
  afx::ClipInfo* d_info = reinterpret_cast<afx::ClipInfo*>(ws + L.off_info);

  if (!one_pass) {
    afx::TrimArgs ta;
    ta.audio = d_audio; ta.offsets = d_offsets; ta.lengths = d_lengths; ta.info = d_info;
    ta.clip_begin = 0; ta.hop = hop; ta.coef = coef; ta.top_db = trim_top_db; ta.raw = raw ? 1 : 0;
    ta.min_frames = (feature_mask & AFX_FEAT_MFCC) ? afx::kDeltaWidth : 1;
    const size_t smem = (size_t)(ceil_div(max_len, afx::kTrimHop) + 4) * sizeof(float);
    afx::trim_kernel<<<n, afx::kTrimThreads, smem, stream>>>(ta);
    ++g_launches;
  }
  afx::StreamArgs sa;
  afx::ClipArgs& ca = sa.c;
  ca.audio = d_audio; ca.offsets = d_offsets; ca.lengths = d_lengths; ca.frame_off = d_frame_off; ca.info = d_info;
  ca.counter = reinterpret_cast<int*>(ws + L.off_counter); ca.n_clips = n;
  ca.slabs = ws + L.header_bytes + items_bytes; ca.slab_bytes = slab_bytes; ca.a_max = slab_rows(h, max_len);
  ca.tab = h->tab; ca.dct_t = h->d_dct_t; ca.hop = hop;
  ca.coef = coef; ca.trim_top_db = trim_top_db; ca.mfcc_top_db = h->cfg.mfcc_top_db; ca.raw = raw ? 1 : 0;
  ca.n_mfcc = h->cfg.n_mfcc; ca.n16 = h->n16; ca.summary_width = afx_summary_width(h); ca.frame_width = afx_frame_width(h);
  ca.mask = feature_mask; ca.summaries = d_summaries; ca.meta = d_meta; ca.frames_out = nullptr;
  ca.pretrimmed = one_pass ? 0 : 1;
  sa.items = reinterpret_cast<const afx::WorkItem*>(ws + L.header_bytes);
  sa.n_items = (int)n_items; sa.n_slots = n_slots; sa.mel_w0_floats = h->mel_w0_floats; sa.l2_mode = h->l2_mode;
  sa.slot_owner = reinterpret_cast<int*>(ws + L.off_slots); sa.slot_done = sa.slot_owner + kMaxSlots;
  const int threads = std::max(n, n_slots);
  afx::item_build_kernel<<<(threads + 255) / 256, 256, 0, stream>>>(sa, reinterpret_cast<const int*>(ws + L.off_item_off),
                                                                     frames_per_step(h) * isteps);
  const long long want = ceil_div(n_items, h->stream_warps);
  const int sgrid = (int)std::max<long long>(1, std::min<long long>(want, h->sm_count));
  h->stream_fn<<<sgrid, h->stream_warps * 32, stream_smem_bytes(h, h->stream_warps), stream>>>(sa);
  g_launches += 2;
  CUDA_TRY(cudaGetLastError());
  return AFX_OK;
}

}  // extern "C"
