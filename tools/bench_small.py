"""Device timing of the small kernels around the hot path (resampler, framewise extras, F0 post-processing) on resident
inputs; one JSON line each with the algorithmic bytes and the fraction of the measured HBM rate."""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def timed(fn, reps=5):
    import torch
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
    from audio_feature_extraction_b200.synth import make_batch_torch
    hbm = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text()).get("hbm_gbs", 6541.5))
    dev = torch.device("cuda", 0)
    L = N.lib()
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000

    def line(kernel, ms, nbytes, extra):
        print(json.dumps({"kernel": kernel, "ms": ms, "roofline": {"bound": "hbm", "achieved": nbytes / ms / 1e6, "peak": hbm,
                                                                  "unit": "GB/s", "frac": nbytes / ms / 1e6 / hbm}, **extra}), flush=True)

    # ---- resampler: n clips of 10 s at 16 kHz -> 22.05 kHz, and 44.1 kHz -> 22.05 kHz
    ex = AudioFeatureExtractor()
    for native in (16000, 44100, 48000):
        audio, offsets, lengths = make_batch_torch(n, sr=native, seconds=10.0, seed=3, device=dev, align=8)
        offsets, lengths = offsets.numpy(), lengths.numpy()
        out, ooff, olen = ex.resample_batch_device(audio, offsets, lengths, native)
        s = torch.cuda.current_stream(dev).cuda_stream
        h = ex._h()
        lens32 = np.ascontiguousarray(lengths, dtype=np.int32)
        ms = timed(lambda: N.check(L.afx_resample(h, audio.data_ptr(), offsets.ctypes.data, lens32.ctypes.data, n, native,
                                                  out.data_ptr(), ooff.ctypes.data, s)))
        nbytes = 4 * (int(lengths.sum()) + int(olen.sum()))
        line("resample_kernel", ms, nbytes, {"native_sr": native, "clips": n, "audio_hours_per_s": n * 10.0 / 3600.0 / (ms * 1e-3)})
        del audio, out

    # ---- framewise extras and F0 post-processing on n clips of 10 s @ 22.05 kHz
    audio, offsets, lengths = make_batch_torch(n, sr=22050, seconds=10.0, seed=4, device=dev, align=8)
    offsets, lengths = offsets.numpy(), lengths.numpy()
    summ, meta, frames = ex.extract_batch_device(audio, offsets, lengths, None, return_frames=True)
    extras = torch.zeros((frames.shape[0], N.EXTRAS_WIDTH), dtype=torch.float32, device=dev)
    s = torch.cuda.current_stream(dev).cuda_stream
    h = ex._h()
    ms = timed(lambda: N.check(L.afx_frame_extras(h, audio.data_ptr(), n, meta.data_ptr(), frames.data_ptr(), extras.data_ptr(),
                                                  ex._ws.data_ptr(), s)))
    T = int(meta[:, 3].sum().item())
    line("frame_extras_kernel", ms, 4 * int(lengths.sum()) + T * (2 * 4 + 4 * N.EXTRAS_WIDTH),
         {"clips": n, "frames": T, "note": "reads every sample once (zero crossings) + the rms / f0 columns, writes 5 floats per frame"})
    track = torch.full((frames.shape[0],), float("nan"), dtype=torch.float32, device=dev)
    stats = torch.zeros((n, 4), dtype=torch.float32, device=dev)
    lens32 = np.ascontiguousarray(lengths, dtype=np.int32)
    ms = timed(lambda: N.check(L.afx_f0_postprocess(h, n, lens32.ctypes.data, meta.data_ptr(), frames.data_ptr(), track.data_ptr(),
                                                    stats.data_ptr(), ex._ws.data_ptr(), s)))
    line("f0_post_kernel", ms, T * 8 + n * 16, {"clips": n, "frames": T, "note": "one strided f0 column in, smoothed track + 4 statistics out"})


if __name__ == "__main__":
    main()
