"""Default feature set (pYIN + MFCC + energy) at the stock hops and at the 10 ms hops (frame mode of yin_obs_kernel, half width 20
of the Viterbi band): ms per phase and audio-hours/s on 888 x 10 s clips, the filtered Viterbi beside the generic one."""
import json, os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
from audio_feature_extraction_b200.synth import make_batch_torch

dev = torch.device("cuda", 0)
L = N.lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 740
for sr, fl, hops in ((22050, 1024, (256, 200)), (16000, 512, (128, 160)), (44100, 2048, (512, 441))):
    audio, off, ln = make_batch_torch(n, sr=sr, seconds=10.0, seed=5, device=dev, align=8)
    off, ln = off.numpy(), ln.numpy()
    for hop in hops:
        for vit in ("filter", "generic"):
            if vit == "generic": os.environ["AFX_VITERBI"] = "generic"
            else: os.environ.pop("AFX_VITERBI", None)
            ex = AudioFeatureExtractor(sr=sr, frame_length=fl, hop_length=hop, n_mfcc=13)
            groups = ("f0", "mfcc", "energy")
            for _ in range(2): ex.extract_batch_device(audio, off, ln, groups)
            torch.cuda.synchronize()
            import ctypes as C
            h = ex._h()
            L.afx_profile_read(h, None, None, 1); L.afx_profile_enable(h, 1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); s, m, _ = ex.extract_batch_device(audio, off, ln, groups); e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            L.afx_profile_enable(h, 0)
            fm = (C.c_double * N.N_PHASES)(); fn = (C.c_int64 * N.N_PHASES)()
            N.check(L.afx_profile_read(h, fm, fn, 1))
            ph = {N.PHASES[p]: round(fm[p], 2) for p in range(N.N_PHASES) if fn[p]}
            print(json.dumps({"sr": sr, "frame_length": fl, "hop_length": hop, "viterbi": vit, "clips": n, "ms": round(ms, 2),
                              "audio_hours_per_s": round(n * 10 / 3600 / (ms * 1e-3), 1), "phases_ms": ph,
                              "clips_ok": int((m[:, 0] == 0).sum().item()), "voiced_mean": round(float(s[:, 3].mean().item()), 3)}), flush=True)
            del ex
    del audio
