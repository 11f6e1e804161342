"""B200-native AudioFeatureExtractor: host-side mirror of the reference class.

Same constructor arguments, method names, return keys and error behaviour as
/root/reference/audio_feature_extraction_toolkit/core/feature_extractor.py (cited per method);
the arithmetic runs in hand-written sm_100a CUDA kernels behind the C ABI of libafx_b200.so
(include/afx_b200.h).  PyTorch only provides device memory and streams.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import logging
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence, Tuple

import os

import numpy as np

from .. import _native as N
from ..wav import read_wav

# librosa.note_to_hz('C2'), librosa.note_to_hz('C7') (reference :15-16)
F0_MIN_C2 = 440.0 * 2.0 ** ((36 - 69) / 12.0)
F0_MAX_C7 = 440.0 * 2.0 ** ((96 - 69) / 12.0)

FEATURE_GROUPS = ("f0", "mfcc", "energy")
_MASKS = {"f0": N.AFX_FEAT_F0, "mfcc": N.AFX_FEAT_MFCC, "energy": N.AFX_FEAT_ENERGY}


class FeatureExtractionError(RuntimeError):
    """Raised where librosa would raise ParameterError inside the reference (empty / non-finite /
    too-short audio), so extract_features re-raises and batch_process skips, as in :211-213,:233-235."""


_STATUS_MSG = {
    N.CLIP_EMPTY: "empty signal after silence trim",
    N.CLIP_NONFINITE: "Audio buffer is not finite everywhere",
    N.CLIP_TOO_SHORT: "signal too short: delta(width=9) needs at least 9 frames",
}


def _mask_of(features_to_extract: Optional[Sequence[str]]) -> int:
    if features_to_extract is None:
        return N.AFX_FEAT_ALL
    m = 0
    for f in features_to_extract:
        if f == "preprocessed":   # modifier: the waveforms are outputs of preprocess_audio
            m |= N.AFX_FEAT_PREPROCESSED
            continue
        if f not in _MASKS:
            raise ValueError(f"unknown feature group {f!r}; expected a subset of {FEATURE_GROUPS}")
        m |= _MASKS[f]
    return m


class AudioFeatureExtractor:
    """音頻特徵提取器 (reference :7). Drop-in for the reference class; adds the batch entry points
    `extract_batch` / `extract_batch_device` that the GPU path is built around."""

    def __init__(self, sr: int = 22050, frame_length: int = 1024, hop_length: int = 256, n_mfcc: int = 13,
                 f0_min: float = F0_MIN_C2, f0_max: float = F0_MAX_C7, pre_emphasis: float = 0.97,
                 device: int = 0, extended_stats: bool = False):
        # reference :30-36
        self.sr = sr
        self.frame_length = frame_length
        self.hop_length = hop_length
        self.n_mfcc = n_mfcc
        self.f0_min = f0_min
        self.f0_max = f0_max
        self.pre_emphasis = pre_emphasis
        self.device = device
        self.extended_stats = extended_stats
        logging.basicConfig(level=logging.INFO)  # reference :38
        self.logger = logging.getLogger(__name__)
        # The reference stores whatever it is given (:30-36) and librosa frames at any size.  The kernels exist for
        # frame_length 512 / 1024 / 2048 with any hop_length <= frame_length / 2: say so HERE, not at the first call
        # (host-only query, no device needed).
        if N.lib().afx_clip_kernel_variant(int(sr), int(frame_length), int(hop_length), None) < 0:
            raise ValueError(f"AudioFeatureExtractor: frame_length={frame_length}, hop_length={hop_length} is not supported by the "
                             "B200 kernels (frame_length must be 512, 1024 or 2048 and 1 <= hop_length <= frame_length / 2)")
        if not (1 <= int(n_mfcc) <= 64):
            raise ValueError(f"AudioFeatureExtractor: n_mfcc={n_mfcc} is not supported (1 .. 64)")
        self._handle = None
        self._ws = None
        self._pin_bufs = {}   # reused pinned ingest buffers of batch_process, per sample dtype

    # ------------------------------------------------------------------ native handle
    def _h(self):
        if self._handle is None:
            L = N.lib()
            cfg = N.AfxConfig()
            L.afx_config_default(C.byref(cfg))
            cfg.sr, cfg.frame_length, cfg.hop_length, cfg.n_mfcc = self.sr, self.frame_length, self.hop_length, self.n_mfcc
            cfg.f0_min, cfg.f0_max, cfg.pre_emphasis, cfg.device = self.f0_min, self.f0_max, self.pre_emphasis, self.device
            h = C.c_void_p()
            N.check(L.afx_create(C.byref(cfg), C.byref(h)))
            self._handle = h
            self.summary_width = L.afx_summary_width(h)
            self.frame_width = L.afx_frame_width(h)
        return self._handle

    def close(self):
        if self._handle is not None:
            N.lib().afx_destroy(self._handle)
            self._handle = None
            self._ws = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ device batch path
    def max_frames(self, length: int) -> int:
        return 1 + max(int(length), 0) // self.hop_length

    def extract_batch_device(self, audio, offsets, lengths, features_to_extract=None, return_frames: bool = False,
                             stream=None):
        """Hot path on device-resident audio.

        audio: 1-D float32 CUDA tensor holding all clips; offsets/lengths: host int64/int32 arrays
        (offsets multiples of 4 samples).  Returns (summaries [n, SS] cuda f32, meta [n, 4] cuda i32,
        frames [sum max_frames, FW] cuda f32 or None).  Only enqueues work on the current stream.
        """
        import torch

        L = N.lib()
        h = self._h()
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        lengths = np.ascontiguousarray(lengths, dtype=np.int32)
        n = int(lengths.shape[0])
        assert audio.is_cuda and audio.dtype == torch.float32 and audio.is_contiguous()
        dev = audio.device
        mask = _mask_of(features_to_extract)
        # ragged batch: the device pass takes the longest clips first (the clip a warp finishes last, and the last pYIN
        # decodes, should be short ones: tools/ragged_probe.py, +13 % on a 0.5 - 30 s mix) and the rows go back by index.
        # Not when framewise rows are asked for: their layout follows the caller's clip order.
        order = None
        if not return_frames and n > 1 and int(lengths.min()) != int(lengths.max()):
            order = np.argsort(-lengths.astype(np.int64), kind="stable")
            offsets, lengths = np.ascontiguousarray(offsets[order]), np.ascontiguousarray(lengths[order])
        summaries = torch.empty((n, self.summary_width), dtype=torch.float32, device=dev)
        meta = torch.empty((n, 4), dtype=torch.int32, device=dev)
        frames = None
        if return_frames:
            rows = int(sum(self.max_frames(int(l)) for l in lengths))
            frames = torch.full((rows, self.frame_width), float("nan"), dtype=torch.float32, device=dev)
        need = L.afx_workspace_bytes(h, n, lengths.ctypes.data)
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            self._ws = torch.empty(int(need), dtype=torch.uint8, device=dev)
        s = stream if stream is not None else torch.cuda.current_stream(dev)
        N.check(L.afx_extract(h, audio.data_ptr(), offsets.ctypes.data, lengths.ctypes.data, n, mask,
                              summaries.data_ptr(), meta.data_ptr(), frames.data_ptr() if frames is not None else None,
                              self._ws.data_ptr(), self._ws.numel(), s.cuda_stream))
        if order is not None:
            with torch.cuda.stream(s):
                inv = torch.from_numpy(np.argsort(order)).to(dev)
                summaries, meta = summaries.index_select(0, inv), meta.index_select(0, inv)
        return summaries, meta, frames

    @staticmethod
    def pack_clips(clips: Sequence[np.ndarray], align: int = 4, pinned: bool = True):
        """Pack float32 clips into one (pinned) host tensor with `align`-sample aligned offsets."""
        import torch

        lengths = np.asarray([int(c.shape[0]) for c in clips], dtype=np.int32)
        strides = (lengths.astype(np.int64) + align - 1) // align * align
        offsets = np.zeros(len(clips), dtype=np.int64)
        if len(clips) > 1:
            offsets[1:] = np.cumsum(strides)[:-1]
        total = int(strides.sum()) + align
        buf = torch.zeros(total, dtype=torch.float32, pin_memory=pinned and torch.cuda.is_available())
        nb = buf.numpy()
        for c, o, l in zip(clips, offsets, lengths):
            nb[o:o + l] = np.asarray(c, dtype=np.float32)
        return buf, offsets, lengths

    def extract_batch(self, clips: Sequence[np.ndarray], names: Optional[Sequence[str]] = None,
                      features_to_extract=None, skip_errors: bool = False) -> List[Dict[str, Any]]:
        """extract_features for in-memory waveforms already at self.sr: one H2D copy, one batched
        device pass, one D2H of the summary rows.  Returns dicts in the reference's key order."""
        import torch

        if len(clips) == 0:
            return []
        names = list(names) if names is not None else [f"<clip {i}>" for i in range(len(clips))]
        buf, offsets, lengths = self.pack_clips(clips)
        dev = torch.device("cuda", self.device)
        audio = buf.to(dev, non_blocking=True)
        summaries, meta, _ = self.extract_batch_device(audio, offsets, lengths, features_to_extract)
        s = summaries.cpu().numpy()
        m = meta.cpu().numpy()
        return self.rows_to_dicts(s, m, names, features_to_extract, skip_errors=skip_errors)

    def _extract_wav_group(self, names: Sequence[str], frames: np.ndarray, pcm16: bool, features_to_extract=None):
        """Summary / meta rows (host numpy) and decode status of WAV files already at self.sr.  The files are cut into
        chunks of ~128 M samples; a helper thread decodes chunk i + 1 into one of two reused pinned buffers (native
        threaded reader) while afx_extract_host_packed streams chunk i to the GPU, so file decoding overlaps the H2D
        copies and the kernels (both calls release the GIL).  Raw 16-bit samples when pcm16 (converted on the device)."""
        import threading
        import torch
        from ..wav import load_batch

        L = N.lib()
        h = self._h()
        n = len(names)
        frames = np.asarray(frames, dtype=np.int64)
        hs = np.zeros((n, self.summary_width), np.float32)
        hm = np.zeros((n, 4), np.int32)
        st = np.zeros(n, np.int32)
        if n == 0:
            return hs, hm, st
        strides = (frames + 7) // 8 * 8
        cap = max(int(os.environ.get("AFX_INGEST_CHUNK_SAMPLES", 128 << 20)), int(strides.max()) + 8)
        cuts, acc = [0], 0
        for i in range(n):                                   # greedy cut into chunks of at most `cap` samples
            if acc + int(strides[i]) > cap and i > cuts[-1]:
                cuts.append(i); acc = 0
            acc += int(strides[i])
        cuts.append(n)
        chunks = [(cuts[k], cuts[k + 1]) for k in range(len(cuts) - 1)]
        need = max(int(strides[a:b].sum()) for a, b in chunks) + 8
        dtype = torch.int16 if pcm16 else torch.float32
        bufs = self._pin_bufs.get(dtype)
        if bufs is None or bufs[0].numel() < need:
            # int16 garbage between clips is harmless (never read as signal); float32 gaps must not hold NaN patterns
            mk = torch.empty if pcm16 else torch.zeros
            bufs = [mk(need, dtype=dtype, pin_memory=True) for _ in range(min(2, len(chunks)))]
            self._pin_bufs[dtype] = bufs
        mask = _mask_of(features_to_extract)

        def decode(k):
            a, b = chunks[k]
            off = np.zeros(b - a, np.int64)
            off[1:] = np.cumsum(strides[a:b])[:-1]
            st[a:b] = load_batch(names[a:b], frames[a:b], bufs[k % len(bufs)], off, as_pcm16=pcm16)
            return off

        off = decode(0)
        for k, (a, b) in enumerate(chunks):
            nxt, box = None, {}
            if k + 1 < len(chunks):
                nxt = threading.Thread(target=lambda: box.setdefault("off", decode(k + 1)))
                nxt.start()
            lengths = np.where(st[a:b] == 0, frames[a:b], 0).astype(np.int32)   # a file that failed to decode: empty clip
            ks = np.zeros((b - a, self.summary_width), np.float32)
            km = np.zeros((b - a, 4), np.int32)
            total = int(strides[a:b].sum()) + 8
            N.check(L.afx_extract_host_packed(h, bufs[k % len(bufs)].data_ptr(), N.SAMPLE_S16 if pcm16 else N.SAMPLE_F32, total,
                                              off.ctypes.data, lengths.ctypes.data, b - a, mask, ks.ctypes.data, km.ctypes.data))
            hs[a:b] = ks; hm[a:b] = km
            if nxt is not None:
                nxt.join()
                off = box["off"]
        return hs, hm, st

    def rows_for_files(self, files: Sequence[str], features_to_extract: Optional[Sequence[str]] = None):
        """Summary and meta rows (host numpy, one row per file, in order) of WAV files: probed and decoded by the native
        threaded reader into ONE pinned buffer (raw 16-bit samples when every file is 16-bit mono PCM) that
        afx_extract_host_packed streams to the GPU in waves; files at another sample rate go through load_audio (device
        resampler).  An unreadable file keeps an AFX_CLIP_EMPTY row, so rows stay aligned with `files`."""
        import torch
        from ..wav import probe_batch, WAV_STATUS

        n = len(files)
        L = N.lib()
        h = self._h()   # (creates the handle: summary_width is known from here on)
        hs = np.zeros((n, self.summary_width), np.float32)
        hm = np.zeros((n, 4), np.int32)
        hm[:, 0] = N.CLIP_EMPTY
        if n == 0:
            return hs, hm
        info = probe_batch(list(files))
        keep = [i for i in range(n) if info["status"][i] == 0 and int(info["sr"][i]) == self.sr]
        other = [i for i in range(n) if info["status"][i] == 0 and int(info["sr"][i]) != self.sr]
        for i in range(n):
            if info["status"][i] != 0:
                self.logger.error(f"載入音頻文件失敗: {files[i]}: {WAV_STATUS.get(int(info['status'][i]), 'unreadable')}")
        # longest clips first: the clip a warp finishes last should be a short one (tools/ragged_probe.py: a ragged
        # 0.5 - 30 s batch runs 13 % faster longest-first than in random order); the rows go back by index
        keep.sort(key=lambda i: -int(info["frames"][i]))
        if keep:
            names = [files[i] for i in keep]
            pcm16 = bool(np.all((info["tag"][keep] == 1) & (info["bits"][keep] == 16) & (info["channels"][keep] == 1)))
            ks, km, st = self._extract_wav_group(names, info["frames"][keep], pcm16, features_to_extract)
            for j in np.nonzero(st)[0]:
                self.logger.error(f"載入音頻文件失敗: {names[j]}: {WAV_STATUS.get(int(st[j]), 'unreadable')}")
            hs[keep] = ks; hm[keep] = km
        if other:
            dev = torch.device("cuda", self.device)
            clips, idx = [], []
            for i in other:
                try:
                    clips.append(self.load_audio(files[i])[0]); idx.append(i)
                except Exception:   # load_audio has logged it (reference :54-56)
                    pass
            if clips:
                buf, offsets, lengths = self.pack_clips(clips)
                summ, meta, _ = self.extract_batch_device(buf.to(dev, non_blocking=True), offsets, lengths, features_to_extract)
                hs[idx] = summ.cpu().numpy(); hm[idx] = meta.cpu().numpy()
        return hs, hm

    def batch_process_sharded(self, audio_dir: str, features_to_extract: Optional[Sequence[str]] = None):
        """batch_process over the ranks of an initialised torch.distributed job (one process per GPU): the sorted
        *.wav list is block-partitioned by rank, every rank runs its shard through rows_for_files (native ingest +
        afx_extract_host_packed) on its own GPU, and the per-file rows are gathered to rank 0 -- over NCCL when that
        is the job's backend, else as host tensors -- which returns the list of dicts; the other ranks return None."""
        import torch
        import torch.distributed as dist
        from ..distributed import extract_sharded

        files = sorted(str(p) for p in Path(audio_dir).glob("*.wav"))
        on_gpu = dist.is_available() and dist.is_initialized() and dist.get_backend() == "nccl"
        dev = torch.device("cuda", self.device) if on_gpu else torch.device("cpu")

        def rows(shard):
            hs, hm = self.rows_for_files(shard, features_to_extract)
            return torch.from_numpy(hs).to(dev), torch.from_numpy(hm).to(dev)

        summ, meta = extract_sharded(rows, files)
        if summ is None:
            return None
        results = self.rows_to_dicts(summ.cpu().numpy(), meta.cpu().numpy(), files, features_to_extract, skip_errors=True)
        for r in results:
            self.logger.info(f"成功處理文件: {Path(r['file_path']).name}")
        return results

    def rows_to_dicts(self, s: np.ndarray, m: np.ndarray, names: Sequence[str], features_to_extract=None,
                      skip_errors: bool = False) -> List[Dict[str, Any]]:
        mask = _mask_of(features_to_extract)
        n = self.n_mfcc
        out = []
        rows = np.asarray(s).tolist()          # one bulk float32 -> Python float conversion for the whole batch
        status = np.asarray(m)[:, 0].tolist()
        e0 = 4 + 4 * n
        for i, name in enumerate(names):
            st = status[i]
            if st != N.CLIP_OK:
                err = FeatureExtractionError(f"{name}: {_STATUS_MSG.get(st, 'status %d' % st)}")
                if skip_errors:
                    self.logger.error(f"處理文件 {Path(str(name)).name} 失敗: {err}")  # reference :234
                    continue
                self.logger.error(f"特徵提取失敗: {err}")  # reference :212
                raise err
            row = rows[i]
            d: Dict[str, Any] = {"file_path": name}  # reference :203
            if mask & N.AFX_FEAT_F0:  # reference :110-113
                d.update(f0_mean=row[0], f0_std=row[1], f0_missing_rate=row[2], f0_quality=row[3])
            if mask & N.AFX_FEAT_MFCC:  # reference :147-150
                d.update(mfcc_mean=row[4:4 + n], mfcc_std=row[4 + n:4 + 2 * n],
                         mfcc_delta_mean=row[4 + 2 * n:4 + 3 * n], mfcc_delta2_mean=row[4 + 3 * n:e0])
            if mask & N.AFX_FEAT_ENERGY:  # reference :176-178
                d.update(energy_mean=row[e0], energy_std=row[e0 + 1], energy_range=row[e0 + 2])
            if self.extended_stats:  # README.md:25-28 of the reference (skew / kurtosis)
                if mask & N.AFX_FEAT_MFCC:
                    d.update(mfcc_skew=row[e0 + 3:e0 + 3 + n], mfcc_kurtosis=row[e0 + 3 + n:e0 + 3 + 2 * n])
                if mask & N.AFX_FEAT_ENERGY:
                    d.update(energy_skew=row[e0 + 3 + 2 * n], energy_kurtosis=row[e0 + 4 + 2 * n])
            out.append(d)
        return out

    # ------------------------------------------------------------------ framewise feature set (.npz of the experiment pipeline)
    def extract_framewise(self, clips: Sequence[np.ndarray], names: Optional[Sequence[str]] = None,
                          skip_errors: bool = False) -> List[Dict[str, Any]]:
        """Framewise features of a batch in the layout the reference's experiment pipeline saves with np.savez
        (04_feature_extraction_experiment/feature_extraction.py:191-292, :340-353) and its DTW stage loads:
        'mfcc' [3 n_mfcc, T] (MFCC / delta / delta2), 'f0' [2, T] (nan_to_num(pyin f0) / delta), 'energy' [2, T]
        (rms / delta), 'zcr' [2, T] (zero-crossing rate / delta), computed on the preprocessed (pre-emphasised,
        trimmed) waveform of this class; plus 'short_time_energy' [1 + (L - frame_length) // hop] = sum(frame ** 2) over
        librosa.util.frame's unpadded frames (README.md:33), which is not part of the stored .npz.  One batched device pass + one small kernel (afx_frame_extras)."""
        import torch

        if len(clips) == 0:
            return []
        names = list(names) if names is not None else [f"<clip {i}>" for i in range(len(clips))]
        L = N.lib()
        buf, offsets, lengths = self.pack_clips(clips)
        dev = torch.device("cuda", self.device)
        audio = buf.to(dev, non_blocking=True)
        summaries, meta, frames = self.extract_batch_device(audio, offsets, lengths, None, return_frames=True)
        extras = torch.zeros((frames.shape[0], N.EXTRAS_WIDTH), dtype=torch.float32, device=dev)
        N.check(L.afx_frame_extras(self._h(), audio.data_ptr(), len(clips), meta.data_ptr(), frames.data_ptr(),
                                   extras.data_ptr(), self._ws.data_ptr(), torch.cuda.current_stream(dev).cuda_stream))
        fr, ex, m = frames.cpu().numpy(), extras.cpu().numpy(), meta.cpu().numpy()
        rows = np.concatenate([[0], np.cumsum([self.max_frames(int(l)) for l in lengths])])
        n = self.n_mfcc
        out = []
        for i, name in enumerate(names):
            st = int(m[i, 0])
            if st != N.CLIP_OK:
                err = FeatureExtractionError(f"{name}: {_STATUS_MSG.get(st, 'status %d' % st)}")
                if skip_errors:
                    self.logger.error(f"處理文件 {Path(str(name)).name} 失敗: {err}")
                    continue
                raise err
            T = int(m[i, 3])
            f, e = fr[rows[i]:rows[i] + T], ex[rows[i]:rows[i] + T]
            n_ste = max(0, 1 + (int(m[i, 2]) - int(m[i, 1]) - self.frame_length) // self.hop_length)
            out.append({"file_path": name,
                        "mfcc": np.ascontiguousarray(f[:, :3 * n].T),
                        "f0": np.stack([np.nan_to_num(f[:, 3 * n + 1]), e[:, 3]]),
                        "energy": np.stack([f[:, 3 * n], e[:, 2]]),
                        "zcr": np.stack([e[:, 0], e[:, 1]]),
                        "short_time_energy": e[:n_ste, 4].copy()})
        return out

    # ------------------------------------------------------------------ F0 post-processing (README "中值濾波平滑" / "F0範圍約束")
    def extract_f0_smoothed(self, clips: Sequence[np.ndarray]) -> List[Dict[str, Any]]:
        """pYIN F0 of a batch followed, on the device, by the post-processing the reference advertises (README.md:12-15;
        recipe 04_feature_extraction_experiment/feature_extractor.py:283-301): the voiced frames' F0 through
        scipy.signal.medfilt(5) and savgol_filter(11, 3), clamped to [f0_min, f0_max].  Per clip: 'f0' [T] float32 (NaN
        where unvoiced), 'f0_smoothed' [T] (smoothed values at the voiced frames, NaN elsewhere), 'voiced_flag' [T], and
        the recipe's statistics 'f0_mean', 'f0_global_rmse', 'f0_local_rmse', 'f0_stability', 'f0_missing_rate'."""
        import torch

        if len(clips) == 0:
            return []
        L = N.lib()
        buf, offsets, lengths = self.pack_clips(clips)
        dev = torch.device("cuda", self.device)
        audio = buf.to(dev, non_blocking=True)
        summaries, meta, frames = self.extract_batch_device(audio, offsets, lengths, ("f0",), return_frames=True)
        track = torch.full((frames.shape[0],), float("nan"), dtype=torch.float32, device=dev)
        stats = torch.zeros((len(clips), 4), dtype=torch.float32, device=dev)
        h_len = np.ascontiguousarray(lengths, dtype=np.int32)
        N.check(L.afx_f0_postprocess(self._h(), len(clips), h_len.ctypes.data, meta.data_ptr(), frames.data_ptr(),
                                     track.data_ptr(), stats.data_ptr(), self._ws.data_ptr(),
                                     torch.cuda.current_stream(dev).cuda_stream))
        fr, tr, st, m = frames.cpu().numpy(), track.cpu().numpy(), stats.cpu().numpy(), meta.cpu().numpy()
        rows = np.concatenate([[0], np.cumsum([self.max_frames(int(l)) for l in lengths])])
        c_f0 = 3 * self.n_mfcc + 1
        out = []
        for i in range(len(clips)):
            if int(m[i, 0]) != N.CLIP_OK:
                raise FeatureExtractionError(f"<clip {i}>: {_STATUS_MSG.get(int(m[i, 0]), 'status %d' % int(m[i, 0]))}")
            T = int(m[i, 3])
            f0 = fr[rows[i]:rows[i] + T, c_f0].copy()
            voiced = ~np.isnan(f0)
            out.append({"f0": f0, "f0_smoothed": tr[rows[i]:rows[i] + T].copy(), "voiced_flag": voiced,
                        "f0_mean": float(st[i, 0]), "f0_global_rmse": float(st[i, 1]), "f0_local_rmse": float(st[i, 2]),
                        "f0_stability": float(st[i, 3]),
                        "f0_missing_rate": float(np.sum(~voiced) / T) if T else 1.0})
        return out

    @staticmethod
    def save_features_npz(features: Dict[str, Any], npz_path: str) -> str:
        """np.savez of one clip's framewise features under the reference's keys (feature_extraction.py:351-352)."""
        np.savez(npz_path, **{k: features[k] for k in ("mfcc", "f0", "energy", "zcr")})
        return npz_path

    def batch_export_npz(self, audio_dir: str, out_dir: str) -> List[str]:
        """Every *.wav of a directory -> <out_dir>/<stem>_processed_features.npz (the file-name pattern of
        feature_extraction.py:350, without the timestamp prefix); unreadable / failing files are logged and skipped."""
        out_dir_p = Path(out_dir)
        out_dir_p.mkdir(parents=True, exist_ok=True)
        clips, names = [], []
        for audio_file in sorted(Path(audio_dir).glob("*.wav")):
            try:
                clips.append(self.load_audio(str(audio_file))[0]); names.append(str(audio_file))
            except Exception as e:
                self.logger.error(f"處理文件 {audio_file.name} 失敗: {str(e)}")
        paths = []
        for feats in self.extract_framewise(clips, names, skip_errors=True):
            stem = Path(feats["file_path"]).stem
            paths.append(self.save_features_npz(feats, str(out_dir_p / f"{stem}_processed_features.npz")))
        return paths

    # ------------------------------------------------------------------ reference API
    def _to_device(self, y: np.ndarray):
        import torch
        buf, offsets, lengths = self.pack_clips([np.asarray(y, dtype=np.float32)])
        return buf.to(torch.device("cuda", self.device), non_blocking=True), offsets, lengths

    def preprocess_audio(self, y: np.ndarray) -> np.ndarray:
        """音頻預處理 (reference :58-74): librosa.effects.preemphasis(coef=self.pre_emphasis) then
        librosa.effects.trim(top_db=30), both on the device (afx_preprocess)."""
        import torch
        L = N.lib()
        h = self._h()
        audio, offsets, lengths = self._to_device(y)
        out = torch.zeros_like(audio)
        meta = torch.empty((1, 4), dtype=torch.int32, device=audio.device)
        scratch = torch.empty((1, self.summary_width), dtype=torch.float32, device=audio.device)
        need = L.afx_workspace_bytes(h, 1, lengths.ctypes.data)
        if self._ws is None or self._ws.numel() < need or self._ws.device != audio.device:
            self._ws = torch.empty(int(need), dtype=torch.uint8, device=audio.device)
        N.check(L.afx_preprocess(h, audio.data_ptr(), offsets.ctypes.data, lengths.ctypes.data, 1, out.data_ptr(),
                                 meta.data_ptr(), scratch.data_ptr(), self._ws.data_ptr(), self._ws.numel(),
                                 torch.cuda.current_stream(audio.device).cuda_stream))
        m = meta.cpu().numpy()[0]
        if int(m[0]) == N.CLIP_NONFINITE:   # librosa.util.valid_audio raises ParameterError
            raise FeatureExtractionError(_STATUS_MSG[N.CLIP_NONFINITE])
        return out[int(m[1]):int(m[2])].cpu().numpy()   # empty when every frame is silent, as librosa returns

    def _extract_preprocessed(self, y: np.ndarray, group: str) -> Dict[str, Any]:
        audio, offsets, lengths = self._to_device(y)
        feats = (group, "preprocessed")
        summaries, meta, _ = self.extract_batch_device(audio, offsets, lengths, feats)
        d = self.rows_to_dicts(summaries.cpu().numpy(), meta.cpu().numpy(), ["<array>"], feats)[0]
        d.pop("file_path")
        return d

    def extract_f0(self, y: np.ndarray) -> Dict[str, Any]:
        """提取F0特徵 (reference :76-114): librosa.pyin on an already preprocessed waveform ->
        f0_mean, f0_std, f0_missing_rate, f0_quality."""
        return self._extract_preprocessed(y, "f0")

    def extract_mfcc(self, y: np.ndarray) -> Dict[str, Any]:
        """提取MFCC特徵 (reference :116-151) on an already preprocessed waveform ->
        mfcc_mean, mfcc_std, mfcc_delta_mean, mfcc_delta2_mean."""
        return self._extract_preprocessed(y, "mfcc")

    def extract_energy(self, y: np.ndarray) -> Dict[str, Any]:
        """提取能量特徵 (reference :153-179) on an already preprocessed waveform ->
        energy_mean, energy_std, energy_range."""
        return self._extract_preprocessed(y, "energy")

    def resample_batch_device(self, audio, offsets, lengths, native_sr: int, stream=None):
        """Sample-rate conversion native_sr -> self.sr of device-resident clips (afx_resample): librosa.load's resampling
        step (reference :52) with res_type='polyphase' (scipy.signal.resample_poly defaults; 'soxr_hq' is a C library
        that cannot be restated).  Returns (audio_out cuda f32, offsets int64, lengths int32) in pack_clips' layout."""
        import torch

        L = N.lib()
        h = self._h()
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        lengths = np.ascontiguousarray(lengths, dtype=np.int32)
        n = int(lengths.shape[0])
        out_len = np.asarray([L.afx_resample_length(h, int(native_sr), int(l)) for l in lengths], dtype=np.int64)
        strides = (out_len + 3) // 4 * 4
        out_off = np.zeros(n, dtype=np.int64)
        if n > 1:
            out_off[1:] = np.cumsum(strides)[:-1]
        out = torch.zeros(int(strides.sum()) + 4, dtype=torch.float32, device=audio.device)
        s = stream if stream is not None else torch.cuda.current_stream(audio.device)
        N.check(L.afx_resample(h, audio.data_ptr(), offsets.ctypes.data, lengths.ctypes.data, n, int(native_sr),
                               out.data_ptr(), out_off.ctypes.data, s.cuda_stream))
        return out, out_off, out_len.astype(np.int32)

    def resample(self, y: np.ndarray, native_sr: int) -> np.ndarray:
        """One waveform native_sr -> self.sr on the device; float32 numpy in, float32 numpy out."""
        import torch

        if int(native_sr) == self.sr:
            return np.asarray(y, dtype=np.float32)
        buf, offsets, lengths = self.pack_clips([np.asarray(y, dtype=np.float32)])
        dev = torch.device("cuda", self.device)
        out, _, out_len = self.resample_batch_device(buf.to(dev, non_blocking=True), offsets, lengths, native_sr)
        return out[:int(out_len[0])].cpu().numpy()

    def load_audio(self, audio_path: str) -> Tuple[np.ndarray, int]:
        """載入音頻文件 (reference :41-56): float32 mono at self.sr; a file at another rate is resampled on the device."""
        try:
            y, sr = read_wav(audio_path)
            if sr != self.sr:
                y = self.resample(y, sr)
            return y, self.sr
        except Exception as e:
            self.logger.error(f"載入音頻文件失敗: {str(e)}")
            raise

    def extract_features(self, audio_path: str, features_to_extract: Optional[Sequence[str]] = None) -> Dict[str, Any]:
        """提取所有特徵 (reference :181-213). `features_to_extract` is the README.md:142-145 keyword."""
        try:
            y, _ = self.load_audio(audio_path)
            return self.extract_batch([y], [audio_path], features_to_extract)[0]
        except Exception as e:
            self.logger.error(f"特徵提取失敗: {str(e)}")
            raise

    def _batch_process_mixed_rates(self, files, info, keep, other, features_to_extract):
        """batch_process for a directory that holds files at other sample rates: every readable file is decoded to
        float32 by the native reader, the groups at another rate are converted on the device (afx_resample) and all
        clips go through one device pass, in the enumeration order."""
        import torch
        from ..wav import load_batch, WAV_STATUS

        dev = torch.device("cuda", self.device)
        order = sorted(keep + other)
        names = [files[i] for i in order]
        frames = info["frames"][order].astype(np.int64)
        rates = info["sr"][order].astype(np.int64)
        strides = (frames + 7) // 8 * 8
        offsets = np.zeros(len(order), dtype=np.int64)
        offsets[1:] = np.cumsum(strides)[:-1]
        host = torch.zeros(int(strides.sum()) + 8, dtype=torch.float32, pin_memory=True)
        st = load_batch(names, frames, host, offsets, as_pcm16=False)
        for i in np.nonzero(st)[0]:
            self.logger.error(f"載入音頻文件失敗: {names[i]}: {WAV_STATUS.get(int(st[i]), 'unreadable')}")
        lengths = np.where(st == 0, frames, 0).astype(np.int32)
        raw = host.to(dev, non_blocking=True)
        # target layout: every clip at self.sr
        L = N.lib()
        h = self._h()
        out_len = np.asarray([int(l) if int(r) == self.sr else L.afx_resample_length(h, int(r), int(l))
                              for l, r in zip(lengths, rates)], dtype=np.int64)
        ostr = (out_len + 3) // 4 * 4
        ooff = np.zeros(len(order), dtype=np.int64)
        ooff[1:] = np.cumsum(ostr)[:-1]
        audio = torch.zeros(int(ostr.sum()) + 4, dtype=torch.float32, device=dev)
        s = torch.cuda.current_stream(dev)
        for r in np.unique(rates):
            idx = np.nonzero(rates == r)[0]
            if int(r) == self.sr:
                for i in idx:   # same rate: plain device copies
                    audio[ooff[i]:ooff[i] + out_len[i]].copy_(raw[offsets[i]:offsets[i] + lengths[i]])
                continue
            io, il, oo = (np.ascontiguousarray(offsets[idx]), np.ascontiguousarray(lengths[idx]),
                          np.ascontiguousarray(ooff[idx]))
            N.check(L.afx_resample(h, raw.data_ptr(), io.ctypes.data, il.ctypes.data, len(idx), int(r),
                                   audio.data_ptr(), oo.ctypes.data, s.cuda_stream))
        summ, meta, _ = self.extract_batch_device(audio, ooff, out_len.astype(np.int32), features_to_extract)
        results = self.rows_to_dicts(summ.cpu().numpy(), meta.cpu().numpy(), names, features_to_extract, skip_errors=True)
        for r in results:
            self.logger.info(f"成功處理文件: {Path(r['file_path']).name}")
        return results

    def batch_process(self, audio_dir: str, features_to_extract: Optional[Sequence[str]] = None) -> List[Dict[str, Any]]:
        """批量處理音頻文件 (reference :215-237): every *.wav of the directory in one batched pass.
        Files are probed and decoded by the native threaded reader (afx_wav_*) straight into reused pinned host buffers
        (raw 16-bit samples when every file is 16-bit mono PCM: half the PCIe bytes, converted on the device), chunk by
        chunk, while afx_extract_host_packed streams the previous chunk to the GPU in waves overlapped with the kernels
        (_extract_wav_group).  Files that fail to load or to extract are logged and skipped, in the reference's
        enumeration order."""
        from ..wav import probe_batch, WAV_STATUS

        files = [str(p) for p in Path(audio_dir).glob("*.wav")]
        if not files:
            return []
        L = N.lib()
        h = self._h()
        info = probe_batch(files)
        keep, other = [], []
        for i, f in enumerate(files):
            if info["status"][i] != 0:
                msg = WAV_STATUS.get(int(info["status"][i]), "unreadable")
                self.logger.error(f"載入音頻文件失敗: {f}: {msg}")
                self.logger.error(f"處理文件 {Path(f).name} 失敗: {msg}")
            elif int(info["sr"][i]) != self.sr:
                other.append(i)
            else:
                keep.append(i)
        if other:
            return self._batch_process_mixed_rates(files, info, keep, other, features_to_extract)
        if not keep:
            return []
        names = [files[i] for i in keep]
        pcm16 = bool(np.all((info["tag"][keep] == 1) & (info["bits"][keep] == 16) & (info["channels"][keep] == 1)))
        # the device pass takes the longest clips first (see rows_for_files); rows return to the enumeration order
        order = np.argsort(-info["frames"][keep].astype(np.int64), kind="stable")
        hs_o, hm_o, st_o = self._extract_wav_group([names[k] for k in order], info["frames"][keep][order], pcm16, features_to_extract)
        hs, hm, st = np.empty_like(hs_o), np.empty_like(hm_o), np.empty_like(st_o)
        hs[order] = hs_o; hm[order] = hm_o; st[order] = st_o
        for i in np.nonzero(st)[0]:
            self.logger.error(f"載入音頻文件失敗: {names[i]}: {WAV_STATUS.get(int(st[i]), 'unreadable')}")
        results = self.rows_to_dicts(hs, hm, names, features_to_extract, skip_errors=True)
        for r in results:
            self.logger.info(f"成功處理文件: {Path(r['file_path']).name}")
        return results
