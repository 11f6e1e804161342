"""Phase times of AudioFeatureExtractor.batch_process on a tmpfs directory (same steps, timed one by one)."""
import json, logging, os, sys, tempfile, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
from audio_feature_extraction_b200.core.feature_extractor import _mask_of
from audio_feature_extraction_b200.synth import make_batch_torch
from audio_feature_extraction_b200.wav import write_wav, probe_batch, load_batch
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
with tempfile.TemporaryDirectory(dir="/dev/shm") as td:
    audio, off, ln = make_batch_torch(n, sr=22050, seconds=10.0, seed=5, device="cuda", align=8)
    a = audio.cpu().numpy(); off = off.numpy(); ln = ln.numpy()
    for i in range(n): write_wav(os.path.join(td, f"clip_{i:05d}.wav"), a[off[i]:off[i] + ln[i]], 22050)
    del audio, a
    ex = AudioFeatureExtractor(); ex.logger.setLevel(logging.WARNING); L = N.lib(); h = ex._h()
    ex.batch_process(td, features_to_extract=["mfcc", "energy"])
    for rep in range(2):
        T = {}; t = time.perf_counter()
        def lap(k):
            global t
            t2 = time.perf_counter(); T[k] = round((t2 - t) * 1e3, 2); t = t2
        files = [str(p) for p in Path(td).glob("*.wav")]; lap("glob")
        info = probe_batch(files); lap("probe")
        keep = list(range(n)); names = files
        frames = info["frames"][keep].astype(np.int64)
        strides = (frames + 7) // 8 * 8; offsets = np.zeros(n, np.int64); offsets[1:] = np.cumsum(strides)[:-1]; total = int(strides.sum()) + 8; lap("plan")
        host = torch.zeros(total, dtype=torch.int16, pin_memory=True); lap("pinned_alloc_zero")
        st = load_batch(names, frames, host, offsets, as_pcm16=True); lap("load")
        lengths = np.where(st == 0, frames, 0).astype(np.int32)
        hs = np.zeros((n, ex.summary_width), np.float32); hm = np.zeros((n, 4), np.int32)
        N.check(L.afx_extract_host_packed(h, host.data_ptr(), N.SAMPLE_S16, total, offsets.ctypes.data, lengths.ctypes.data, n, _mask_of(["mfcc", "energy"]), hs.ctypes.data, hm.ctypes.data)); lap("extract_host_packed")
        res = ex.rows_to_dicts(hs, hm, names, ["mfcc", "energy"], skip_errors=True); lap("dicts")
        del host; lap("free")
        print(json.dumps(T))
