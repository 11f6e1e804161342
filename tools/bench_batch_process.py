"""Wall-clock throughput of the reference-facing batch entry on real files: AudioFeatureExtractor.batch_process(dir) over N
16-bit PCM WAV files (written to a tmpfs directory first, untimed), all host work included: glob, header probe, threaded
decode into the pinned buffer, H2D waves, kernels, D2H, dict building, logging.
    python tools/bench_batch_process.py [n_files] [seconds] [features]"""
import json, logging, os, sys, tempfile, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
from audio_feature_extraction_b200 import AudioFeatureExtractor
from audio_feature_extraction_b200.synth import make_batch_torch
from audio_feature_extraction_b200.wav import write_wav

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 10.0
feats = sys.argv[3].split(",") if len(sys.argv) > 3 else ["mfcc", "energy"]
logging.getLogger().setLevel(logging.WARNING)
base = "/dev/shm" if os.path.isdir("/dev/shm") else None
with tempfile.TemporaryDirectory(dir=base) as td:
    audio, off, ln = make_batch_torch(n, sr=22050, seconds=secs, seed=5, device="cuda", align=8)
    a = audio.cpu().numpy(); off = off.numpy(); ln = ln.numpy()
    t0 = time.perf_counter()
    for i in range(n):
        write_wav(os.path.join(td, f"clip_{i:05d}.wav"), a[off[i]:off[i] + ln[i]], 22050)
    t_write = time.perf_counter() - t0
    del audio, a
    ex = AudioFeatureExtractor()
    ex.logger.setLevel(logging.WARNING)
    out = []
    for rep in range(3):
        t0 = time.perf_counter()
        res = ex.batch_process(td, features_to_extract=feats)
        dt = time.perf_counter() - t0
        out.append(dt)
    hours = n * secs / 3600.0
    print(json.dumps({"files": n, "seconds_each": secs, "features": feats, "results": len(res), "wall_s": out,
                      "audio_hours_per_s_best": hours / min(out), "file_bytes": int(n * secs * 22050 * 2),
                      "gb_per_s_best": n * secs * 22050 * 2 / min(out) / 1e9, "write_s": t_write, "dir": td.split("/")[1]}))
