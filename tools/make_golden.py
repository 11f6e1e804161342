#!/usr/bin/env python
"""Freeze oracle outputs as golden vectors: tests/golden/afx_golden_<cfg>.npz.

PARITY UNPINNED (see oracle/__init__.py): librosa 0.11.0 cannot be imported here, so these vectors are outputs of the
numpy/scipy restatement in oracle/ on seeded synthetic clips, not of the reference itself.  They pin the oracle against
regressions (CPU test) and give the CUDA path a fixture that does not depend on running the oracle (GPU test).

    python tools/make_golden.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from audio_feature_extraction_b200.synth import make_clip  # noqa: E402
from oracle import OracleFeatureExtractor  # noqa: E402

CONFIGS = {
    "c2": dict(sr=22050, frame_length=1024, hop_length=256, n_mfcc=13),
    "c3": dict(sr=16000, frame_length=512, hop_length=128, n_mfcc=40),
    "c4": dict(sr=44100, frame_length=2048, hop_length=512, n_mfcc=13),
}


def main():
    out_dir = ROOT / "tests" / "golden"
    out_dir.mkdir(parents=True, exist_ok=True)
    for name, cfg in CONFIGS.items():
        kind = "music" if cfg["sr"] == 44100 else "speech"
        f_lo, f_hi = (80.0, 400.0) if cfg["sr"] == 16000 else (110.0, 880.0)
        ora = OracleFeatureExtractor(extended_stats=True, **cfg)
        data = {"cfg": np.asarray([cfg["sr"], cfg["frame_length"], cfg["hop_length"], cfg["n_mfcc"]], np.int64)}
        for i, sec in enumerate((0.9, 0.6)):
            y = make_clip(900 + i, sr=cfg["sr"], seconds=sec, f_lo=f_lo, f_hi=f_hi, kind=kind)
            s, e = ora.trim_indices(y)
            yt = ora.preprocess_audio(y)
            m, d1, d2 = ora.mfcc_frames(yt)
            r = ora.rms_frames(yt)[0]
            f0, voiced, vprob = ora.f0_track(yt)
            feats = ora.extract_features_array(y)
            data[f"clip{i}"] = y
            data[f"trim{i}"] = np.asarray([s, e, m.shape[1]], np.int64)
            data[f"mfcc{i}"] = m.astype(np.float32)
            data[f"delta{i}"] = d1.astype(np.float32)
            data[f"delta2{i}"] = d2.astype(np.float32)
            data[f"rms{i}"] = r.astype(np.float32)
            data[f"f0{i}"] = f0
            data[f"vprob{i}"] = vprob
            keys = [k for k in feats if k != "file_path"]
            data[f"summary_keys{i}"] = np.asarray(keys)
            data[f"summary{i}"] = np.concatenate([np.atleast_1d(np.asarray(feats[k], np.float64)) for k in keys])
        np.savez_compressed(out_dir / f"afx_golden_{name}.npz", **data)
        print(name, "written", (out_dir / f"afx_golden_{name}.npz").stat().st_size, "bytes")


if __name__ == "__main__":
    main()
