// Host ingest: threaded RIFF/WAVE decode straight into the packed (pinned) host buffer the device path reads.
//
// Stands in for the decode half of librosa.load(path, sr=self.sr) as called by load_audio
// (audio_feature_extraction_toolkit/core/feature_extractor.py:52, reference) for every file batch_process (:228-231)
// enumerates: PCM 8/16/24/32-bit and IEEE float32/64, mono mix by channel mean, float32 in [-1, 1] (PCM16 -> x / 32768,
// libsndfile's scaling).  PCM16 mono files can also be delivered as raw 16-bit samples (converted on the device), which
// halves the host->device bytes.  No resampling.  Pure host code, no CUDA.
#pragma once
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

namespace afx {
namespace wav {

constexpr int kMaxChannels = 64;   // channel counts a mono mix-down is sized for

enum Status { OK = 0, OPEN_FAILED = 1, NOT_WAVE = 2, MISSING_CHUNK = 3, UNSUPPORTED = 4, LENGTH_MISMATCH = 5, NOT_PCM16_MONO = 6 };

struct Info {
  int32_t sr = 0, channels = 0, tag = 0, bits = 0;
  int64_t frames = 0;        // samples per channel
  int64_t data_off = 0;      // byte offset of the data chunk's body
  int64_t data_bytes = 0;
};

inline uint32_t rd32(const unsigned char* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
inline uint16_t rd16(const unsigned char* p) { return (uint16_t)(p[0] | (p[1] << 8)); }

// walk the RIFF chunks of an open file; leaves the file position unspecified
inline int parse(FILE* f, Info& info) {
  unsigned char hdr[12];
  if (fread(hdr, 1, 12, f) != 12 || memcmp(hdr, "RIFF", 4) != 0 || memcmp(hdr + 8, "WAVE", 4) != 0) return NOT_WAVE;
  bool have_fmt = false, have_data = false;
  int64_t pos = 12;
  for (;;) {
    unsigned char ch[8];
    if (fseek(f, (long)pos, SEEK_SET) != 0 || fread(ch, 1, 8, f) != 8) break;
    const uint32_t size = rd32(ch + 4);
    if (memcmp(ch, "fmt ", 4) == 0) {
      unsigned char body[40];
      const size_t want = size < sizeof body ? size : sizeof body;
      if (want < 16 || fread(body, 1, want, f) != want) return MISSING_CHUNK;
      info.tag = rd16(body); info.channels = rd16(body + 2); info.sr = (int32_t)rd32(body + 4); info.bits = rd16(body + 14);
      if (info.tag == 0xFFFE && want >= 26) info.tag = rd16(body + 24);   // WAVE_FORMAT_EXTENSIBLE: sub-format GUID starts with the tag
      have_fmt = true;
    } else if (memcmp(ch, "data", 4) == 0) {
      info.data_off = pos + 8;
      info.data_bytes = size;
      have_data = true;   // keep walking: the last data chunk wins, as in the Python reader
    }
    pos += 8 + (int64_t)size + (size & 1);
  }
  if (!have_fmt || !have_data) return MISSING_CHUNK;
  if (info.channels < 1 || info.channels > kMaxChannels) return UNSUPPORTED;   // a crafted fmt chunk must not size a multi-GB decode buffer
  const bool pcm_ok = info.tag == 1 && (info.bits == 8 || info.bits == 16 || info.bits == 24 || info.bits == 32);
  const bool flt_ok = info.tag == 3 && (info.bits == 32 || info.bits == 64);
  if (!pcm_ok && !flt_ok) return UNSUPPORTED;
  // a truncated file: only what is really there
  if (fseek(f, 0, SEEK_END) == 0) {
    const int64_t end = ftell(f);
    if (end >= 0 && info.data_off + info.data_bytes > end) info.data_bytes = end > info.data_off ? end - info.data_off : 0;
  }
  const int64_t bytes_per_frame = (int64_t)info.channels * (info.bits / 8);
  info.frames = info.data_bytes / bytes_per_frame;
  return OK;
}

inline int probe(const char* path, Info& info) {
  FILE* f = fopen(path, "rb");
  if (!f) return OPEN_FAILED;
  const int rc = parse(f, info);
  fclose(f);
  return rc;
}

inline float decode_sample(const unsigned char* p, int tag, int bits) {
  if (tag == 1) {
    switch (bits) {
      case 8:  return ((float)p[0] - 128.0f) / 128.0f;
      case 16: return (float)(int16_t)rd16(p) / 32768.0f;
      case 24: { int32_t v = (int32_t)p[0] | ((int32_t)p[1] << 8) | ((int32_t)p[2] << 16); if (v & 0x800000) v -= 0x1000000; return (float)v / 8388608.0f; }
      default: return (float)((double)(int32_t)rd32(p) / 2147483648.0);
    }
  }
  if (bits == 32) { float v; memcpy(&v, p, 4); return v; }
  double v; memcpy(&v, p, 8); return (float)v;
}

// decode `frames` frames of file `path` into out_f32 (mono mix) or, when out_s16 is given, copy raw PCM16 mono samples
inline int load(const char* path, int64_t frames, float* out_f32, int16_t* out_s16) {
  FILE* f = fopen(path, "rb");
  if (!f) return OPEN_FAILED;
  Info info;
  int rc = parse(f, info);
  if (rc == OK && info.frames != frames) rc = LENGTH_MISMATCH;
  if (rc == OK && out_s16 && !(info.tag == 1 && info.bits == 16 && info.channels == 1)) rc = NOT_PCM16_MONO;
  if (rc != OK) { fclose(f); return rc; }
  fseek(f, (long)info.data_off, SEEK_SET);
  if (out_s16) {   // little-endian host assumed (x86-64 / aarch64)
    const size_t got = fread(out_s16, sizeof(int16_t), (size_t)frames, f);
    fclose(f);
    return got == (size_t)frames ? OK : LENGTH_MISMATCH;
  }
  const int bps = info.bits / 8, ch = info.channels;
  const int64_t chunk_frames = std::max<int64_t>(1, (int64_t)(1 << 18) / ((int64_t)ch * bps));   // 256 KB of file per read, whatever the channel count
  std::vector<unsigned char> buf((size_t)chunk_frames * ch * bps);
  for (int64_t done = 0; done < frames;) {
    const int64_t n = frames - done < chunk_frames ? frames - done : chunk_frames;
    if (fread(buf.data(), (size_t)ch * bps, (size_t)n, f) != (size_t)n) { fclose(f); return LENGTH_MISMATCH; }
    const unsigned char* p = buf.data();
    if (ch == 1) {
      for (int64_t i = 0; i < n; ++i, p += bps) out_f32[done + i] = decode_sample(p, info.tag, info.bits);
    } else {
      for (int64_t i = 0; i < n; ++i) {   // librosa.to_mono: float32 mean over channels
        float s = 0.f;
        for (int c = 0; c < ch; ++c, p += bps) s += decode_sample(p, info.tag, info.bits);
        out_f32[done + i] = s / (float)ch;
      }
    }
    done += n;
  }
  fclose(f);
  return OK;
}

// run fn(i) for i in [0, n) on `threads` host threads (dynamic distribution)
template <typename Fn>
inline void parallel_for(int64_t n, int threads, Fn fn) {
  if (threads < 1) threads = (int)std::thread::hardware_concurrency();
  if (threads < 1) threads = 1;
  if ((int64_t)threads > n) threads = (int)(n > 0 ? n : 1);
  std::atomic<int64_t> next{0};
  auto work = [&]() { for (;;) { const int64_t i = next.fetch_add(1); if (i >= n) break; fn(i); } };   // fn must not throw (callers catch per item)
  if (threads == 1) { work(); return; }
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) pool.emplace_back(work);
  for (auto& t : pool) t.join();
}

}  // namespace wav
}  // namespace afx
