"""One-off pYIN parity sweep on a larger seeded set than the test suite carries (the oracle needs seconds per clip):
N clips per configuration, oracle tracks computed on all host cores, device tracks in one batch.  Prints one JSON
line per configuration: voicing agreement, worst deviation in cents, worst voiced-probability difference.
Usage: python tools/f0_parity_sweep.py [N]   (test infrastructure: imports oracle/)"""
import json
import multiprocessing as mp
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

CFGS = [("C2", dict(sr=22050, frame_length=1024, hop_length=256, n_mfcc=13), "speech", 110.0, 880.0),
        ("C3", dict(sr=16000, frame_length=512, hop_length=128, n_mfcc=40), "speech", 80.0, 400.0),
        ("C4", dict(sr=44100, frame_length=2048, hop_length=512, n_mfcc=13), "music", 110.0, 880.0)]


def sweep(n_clips, emit=None):
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from oracle.extractor import f0_track_task
    rng = np.random.default_rng(7)
    out = []
    with mp.get_context("spawn").Pool(len(os.sched_getaffinity(0))) as pool:
        for name, cfg, kind, f_lo, f_hi in CFGS:
            secs = np.round(rng.uniform(0.8, 3.0, size=n_clips), 3)
            clips = [make_clip(5000 + s, sr=cfg["sr"], seconds=float(sec), f_lo=f_lo, f_hi=f_hi, kind=kind) for s, sec in enumerate(secs)]
            clips += [(0.05 * rng.standard_normal(int(1.2 * cfg["sr"]))).astype(np.float32) for _ in range(4)]   # noise only
            ex = AudioFeatureExtractor(**cfg)
            buf, offsets, lengths = ex.pack_clips(clips)
            summ, meta, frames = ex.extract_batch_device(buf.cuda(), offsets, lengths, ("f0",), return_frames=True)
            torch.cuda.synchronize()
            fr, m = frames.cpu().numpy(), meta.cpu().numpy()
            rows = np.concatenate([[0], np.cumsum([ex.max_frames(int(l)) for l in lengths])])
            ref = pool.map(f0_track_task, [(cfg, y) for y in clips], chunksize=1)
            c_f0 = 3 * ex.n_mfcc + 1
            total = agree = voiced_both = 0
            worst_cents = worst_vp = 0.0
            bad_clips = 0
            for i, (f0, voiced, vprob) in enumerate(ref):
                T = len(f0)
                assert m[i, 3] == T, (name, i, m[i], T)
                got = fr[rows[i]:rows[i] + T, c_f0].astype(np.float64)
                gv = ~np.isnan(got)
                total += T; agree += int(np.sum(gv == voiced))
                bad_clips += int(not np.array_equal(gv, voiced))
                both = gv & voiced
                voiced_both += int(both.sum())
                if both.any():
                    worst_cents = max(worst_cents, float(np.max(1200.0 * np.abs(np.log2(got[both] / f0[both])))))
                worst_vp = max(worst_vp, float(np.max(np.abs(fr[rows[i]:rows[i] + T, c_f0 + 1] - vprob))))
            res = {"config": name, "clips": len(clips), "frames": total, "voicing_agreement": agree / total,
                   "clips_with_any_voicing_difference": bad_clips, "voiced_frames_compared": voiced_both,
                   "worst_cents": worst_cents, "worst_voiced_prob_diff": worst_vp}
            out.append(res)
            if emit:
                emit(res)
    return out


def main():
    sweep(int(sys.argv[1]) if len(sys.argv) > 1 else 48, emit=lambda r: print(json.dumps(r), flush=True))


if __name__ == "__main__":
    main()
