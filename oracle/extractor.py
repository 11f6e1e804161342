"""Oracle restatement of the reference orchestration (TEST INFRASTRUCTURE, parity unpinned).

Follows /root/reference/audio_feature_extraction_toolkit/core/feature_extractor.py method by
method, with `oracle.librosa_restated` standing in for librosa 0.11.0.  Works on in-memory
float32 waveforms (the reference's `load_audio` is a librosa.load call; WAV decoding is host I/O
and is exercised separately).
"""
from __future__ import annotations

import numpy as np
import scipy.stats

from . import librosa_restated as L

GROUPS = ("f0", "mfcc", "energy")


class OracleFeatureExtractor:
    def __init__(self, sr=22050, frame_length=1024, hop_length=256, n_mfcc=13,
                 f0_min=L.note_c2_hz, f0_max=L.note_c7_hz, pre_emphasis=0.97, extended_stats=False,
                 dtype=np.float32):
        # core/feature_extractor.py:30-36
        self.sr = sr
        self.frame_length = frame_length
        self.hop_length = hop_length
        self.n_mfcc = n_mfcc
        self.f0_min = f0_min
        self.f0_max = f0_max
        self.pre_emphasis = pre_emphasis
        self.extended_stats = extended_stats
        self.dtype = dtype

    # core/feature_extractor.py:58-74
    def preprocess_audio(self, y):
        y = np.asarray(y, dtype=self.dtype)
        y_pre = L.preemphasis(y, coef=self.pre_emphasis)
        y_trim, _ = L.trim(y_pre, top_db=30)
        return y_trim

    def trim_indices(self, y):
        y = np.asarray(y, dtype=self.dtype)
        return L.trim_indices(L.preemphasis(y, coef=self.pre_emphasis), top_db=30)

    # core/feature_extractor.py:76-114
    def f0_track(self, y):
        f0, voiced, vprob = L.pyin(y, fmin=self.f0_min, fmax=self.f0_max, sr=self.sr,
                                   frame_length=self.frame_length, hop_length=self.hop_length)
        return f0, voiced, vprob

    def extract_f0(self, y):
        f0, _, _ = self.f0_track(y)
        f0_valid = f0[~np.isnan(f0)]
        if len(f0_valid) > 0:
            f0_mean = np.mean(f0_valid)
            f0_std = np.std(f0_valid)
            f0_missing_rate = np.sum(np.isnan(f0)) / len(f0)
            f0_quality = 1 - f0_missing_rate
        else:
            f0_mean, f0_std, f0_missing_rate, f0_quality = 0, 0, 1, 0
        return {"f0_mean": float(f0_mean), "f0_std": float(f0_std),
                "f0_missing_rate": float(f0_missing_rate), "f0_quality": float(f0_quality)}

    # core/feature_extractor.py:116-151
    def mfcc_frames(self, y):
        m = L.mfcc(y, sr=self.sr, n_mfcc=self.n_mfcc, n_fft=self.frame_length,
                   hop_length=self.hop_length, dtype=self.dtype)
        return m, L.delta(m), L.delta(m, order=2)

    def extract_mfcc(self, y):
        m, d1, d2 = self.mfcc_frames(y)
        out = {"mfcc_mean": np.mean(m, axis=1).tolist(), "mfcc_std": np.std(m, axis=1).tolist(),
               "mfcc_delta_mean": np.mean(d1, axis=1).tolist(), "mfcc_delta2_mean": np.mean(d2, axis=1).tolist()}
        if self.extended_stats:  # README.md:25-28 extension: scipy.stats defaults (biased, Fisher)
            out["mfcc_skew"] = scipy.stats.skew(m.astype(np.float64), axis=1).tolist()
            out["mfcc_kurtosis"] = scipy.stats.kurtosis(m.astype(np.float64), axis=1).tolist()
        return out

    # core/feature_extractor.py:153-179
    def rms_frames(self, y):
        return L.rms(y, frame_length=self.frame_length, hop_length=self.hop_length, dtype=self.dtype)

    def extract_energy(self, y):
        r = self.rms_frames(y)
        out = {"energy_mean": float(np.mean(r)), "energy_std": float(np.std(r)), "energy_range": float(np.ptp(r))}
        if self.extended_stats:
            out["energy_skew"] = float(scipy.stats.skew(r.astype(np.float64), axis=None))
            out["energy_kurtosis"] = float(scipy.stats.kurtosis(r.astype(np.float64), axis=None))
        return out

    # core/feature_extractor.py:181-213 on an in-memory waveform
    def extract_features_array(self, y, name="<array>", features_to_extract=GROUPS):
        y_processed = self.preprocess_audio(y)
        feats = {"file_path": name}
        if "f0" in features_to_extract:
            feats.update(self.extract_f0(y_processed))
        if "mfcc" in features_to_extract:
            feats.update(self.extract_mfcc(y_processed))
        if "energy" in features_to_extract:
            feats.update(self.extract_energy(y_processed))
        return feats

    def framewise(self, y):
        """The framewise feature set the reference's experiment pipeline stores as .npz
        (04_feature_extraction_experiment/feature_extraction.py:191-292, saved :340-353), computed on this class's
        preprocessed waveform: 'mfcc' [3 n_mfcc, T] = MFCC / delta / delta2, 'f0' [2, T] = nan_to_num(pyin f0) / delta,
        'energy' [2, T] = rms / delta, 'zcr' [2, T] = zero-crossing rate / delta."""
        yt = self.preprocess_audio(y)
        m, d1, d2 = self.mfcc_frames(yt)
        f0, _, _ = self.f0_track(yt)
        f0 = np.nan_to_num(f0).reshape(1, -1)
        r = self.rms_frames(yt)
        z = L.zero_crossing_rate(yt, frame_length=self.frame_length, hop_length=self.hop_length)
        # short-time energy of the variant extractor (README.md:33; '04_.../audio_feature_extraction 2/
        # audio_feature_extraction/feature_extractor.py:234-238'): np.sum(librosa.util.frame(y, N, hop) ** 2, axis=0)
        N, hop = self.frame_length, self.hop_length
        if yt.size >= N:
            fr = np.lib.stride_tricks.sliding_window_view(yt, N)[::hop]
            ste = np.sum(fr.astype(np.float64) ** 2, axis=1)
        else:
            ste = np.zeros(0)
        return {"mfcc": np.vstack([m, d1, d2]), "f0": np.vstack([f0, L.delta(f0)]),
                "energy": np.vstack([r, L.delta(r)]), "zcr": np.vstack([z, L.delta(z)]), "short_time_energy": ste}

    def f0_postprocess(self, f0, voiced_flag):
        """F0 post-processing of the reference's experiment recipe (04_feature_extraction_experiment/
        feature_extractor.py:283-301; README.md:12-15 'median smoothing', 'F0 range constraint'), with scipy itself:
        medfilt(min(5, len)) -> savgol_filter(11, 3) from 11 voiced frames on -> clip to [f0_min, f0_max]; returns the
        smoothed voiced values and (mean, global RMSE, local RMSE, stability).  scipy rejects the even kernels of
        2 and 4 voiced frames; the values pass through there."""
        import scipy.signal
        v = np.asarray(f0, dtype=np.float64)[np.asarray(voiced_flag, dtype=bool)]
        if v.size == 0:
            return v, (0.0, np.inf, np.inf, np.inf)
        k = min(5, v.size)
        if k % 2 == 1:
            v = scipy.signal.medfilt(v, kernel_size=k)
        if v.size >= 11:
            v = scipy.signal.savgol_filter(v, window_length=11, polyorder=3)
        v = np.clip(v, self.f0_min, self.f0_max)
        with np.errstate(invalid="ignore"), __import__("warnings").catch_warnings():
            __import__("warnings").simplefilter("ignore")
            local = float(np.sqrt(np.mean(np.diff(v) ** 2)))
        glob = float(np.sqrt(np.mean((v - np.mean(v)) ** 2)))
        return v, (float(np.mean(v)), glob, local, float(np.std(v) / (np.mean(v) + 1e-8)))


def f0_track_task(task):
    """(constructor kwargs, waveform) -> (f0, voiced_flag, voiced_prob) of the preprocessed waveform; a module-level
    function so that multiprocessing pools (tools/f0_parity_sweep.py, tests) can ship it to worker processes."""
    import os
    for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):
        os.environ[k] = "1"
    cfg, y = task
    ora = OracleFeatureExtractor(**cfg)
    return ora.f0_track(ora.preprocess_audio(y))
