"""Minimal RIFF/WAVE reader + writer (host I/O).

Stands in for the decode half of `librosa.load(path, sr=self.sr)`
(/root/reference/audio_feature_extraction_toolkit/core/feature_extractor.py:52): PCM 8/16/24/32-bit
and IEEE float32/float64, mono mix by channel mean, float32 in [-1, 1] (PCM16 -> x / 32768, as
libsndfile does).  No resampling: that is host-ingest work outside the hot path.
"""
from __future__ import annotations

import struct

import numpy as np

__all__ = ["read_wav", "write_wav", "probe_batch", "load_batch"]


def read_wav(path):
    with open(path, "rb") as f:
        data = f.read()
    if len(data) < 12 or data[:4] != b"RIFF" or data[8:12] != b"WAVE":
        raise ValueError(f"{path}: not a RIFF/WAVE file")
    pos = 12
    fmt = None
    pcm = None
    while pos + 8 <= len(data):
        cid = data[pos:pos + 4]
        size = struct.unpack_from("<I", data, pos + 4)[0]
        body = data[pos + 8:pos + 8 + size]
        if cid == b"fmt ":
            tag, ch, sr, _, _, bits = struct.unpack_from("<HHIIHH", body, 0)
            if tag == 0xFFFE and len(body) >= 26:  # WAVE_FORMAT_EXTENSIBLE: sub-format GUID starts with the tag
                tag = struct.unpack_from("<H", body, 24)[0]
            fmt = (tag, ch, sr, bits)
        elif cid == b"data":
            pcm = body
        pos += 8 + size + (size & 1)
    if fmt is None or pcm is None:
        raise ValueError(f"{path}: missing fmt or data chunk")
    tag, ch, sr, bits = fmt
    if tag == 1:
        if bits == 8:
            x = (np.frombuffer(pcm, dtype=np.uint8).astype(np.float32) - 128.0) / 128.0
        elif bits == 16:
            x = np.frombuffer(pcm, dtype="<i2").astype(np.float32) / 32768.0
        elif bits == 24:
            b = np.frombuffer(pcm[:len(pcm) // 3 * 3], dtype=np.uint8).reshape(-1, 3).astype(np.int32)
            v = b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16)
            v = np.where(v & 0x800000, v - 0x1000000, v)
            x = v.astype(np.float32) / 8388608.0
        elif bits == 32:
            x = (np.frombuffer(pcm, dtype="<i4").astype(np.float64) / 2147483648.0).astype(np.float32)
        else:
            raise ValueError(f"{path}: unsupported PCM width {bits}")
    elif tag == 3:
        x = np.frombuffer(pcm, dtype="<f4" if bits == 32 else "<f8").astype(np.float32)
    else:
        raise ValueError(f"{path}: unsupported WAVE format tag {tag}")
    if ch > 1:
        x = x[:len(x) // ch * ch].reshape(-1, ch).mean(axis=1).astype(np.float32)
    return np.ascontiguousarray(x, dtype=np.float32), int(sr)


def write_wav(path, y, sr, subtype="PCM_16"):
    y = np.asarray(y, dtype=np.float32)
    if subtype == "PCM_16":
        pcm = np.clip(np.round(y * 32768.0), -32768, 32767).astype("<i2").tobytes()
        tag, bits = 1, 16
    elif subtype == "FLOAT":
        pcm = y.astype("<f4").tobytes()
        tag, bits = 3, 32
    else:
        raise ValueError(subtype)
    block = bits // 8
    hdr = b"RIFF" + struct.pack("<I", 36 + len(pcm)) + b"WAVE" + b"fmt " + struct.pack(
        "<IHHIIHH", 16, tag, 1, sr, sr * block, block, bits) + b"data" + struct.pack("<I", len(pcm))
    with open(path, "wb") as f:
        f.write(hdr + pcm)


WAV_STATUS = {1: "cannot open", 2: "not a RIFF/WAVE file", 3: "missing fmt or data chunk", 4: "unsupported WAVE encoding",
              5: "length changed between probe and load", 6: "not 16-bit mono PCM"}


def _c_paths(paths):
    import ctypes as C
    arr = (C.c_char_p * len(paths))(*[str(p).encode() for p in paths])
    return arr


def probe_batch(paths, threads: int = 0):
    """Headers of many WAV files at once (native, threaded: afx_wav_probe_batch).  Returns dict of numpy arrays
    sr, channels, frames, tag, bits, status (0 = ok, else a WAV_STATUS key)."""
    from . import _native as N
    n = len(paths)
    sr = np.zeros(n, np.int32); ch = np.zeros(n, np.int32); fr = np.zeros(n, np.int64)
    fmt = np.zeros(n, np.int32); st = np.zeros(n, np.int32)
    if n:
        cp = _c_paths(paths)
        N.check(N.lib().afx_wav_probe_batch(cp, n, threads, sr.ctypes.data, ch.ctypes.data, fr.ctypes.data,
                                            fmt.ctypes.data, st.ctypes.data))
    return {"sr": sr, "channels": ch, "frames": fr, "tag": fmt >> 8, "bits": fmt & 0xFF, "status": st}


def load_batch(paths, frames, out, offsets, as_pcm16: bool = False, threads: int = 0):
    """Decode many WAV files in parallel straight into `out` (a torch tensor or numpy array, ideally pinned):
    file i -> out[offsets[i] : offsets[i] + frames[i]] as float32 mono, or as raw int16 when as_pcm16 (16-bit mono
    PCM files only).  Returns the per-file status array."""
    from . import _native as N
    n = len(paths)
    st = np.zeros(n, np.int32)
    if n == 0:
        return st
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    frames = np.ascontiguousarray(frames, dtype=np.int64)
    ptr = out.data_ptr() if hasattr(out, "data_ptr") else out.ctypes.data
    cp = _c_paths(paths)
    N.check(N.lib().afx_wav_load_batch(cp, n, threads, N.SAMPLE_S16 if as_pcm16 else N.SAMPLE_F32, ptr,
                                       offsets.ctypes.data, frames.ctypes.data, st.ctypes.data))
    return st
