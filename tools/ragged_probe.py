"""Which property of the ragged leg costs throughput?  Times the fused MFCC + energy pass on batches cut from the same
synthetic audio: uniform 10 s, random cuts inside one slot (0.5-10 s, no pauses), whole multi-slot clips (20 / 30 s with the
-70 dBFS heads / tails of the slots as pauses), the bench's ragged mix, and the mix sorted by length."""
import json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from audio_feature_extraction_b200 import AudioFeatureExtractor
from audio_feature_extraction_b200.synth import make_batch_torch

dev = torch.device("cuda", 0)
sr, nl = 22050, 2000
audio, off, ln = make_batch_torch(nl, sr=sr, seconds=10.0, seed=1000, device=dev, align=8)
off, ln = off.numpy(), ln.numpy()
stride = int(off[1] - off[0])
ex = AudioFeatureExtractor()

def timed(o, l, reps=5):
    o, l = np.asarray(o, np.int64), np.asarray(l, np.int32)
    for _ in range(2): ex.extract_batch_device(audio, o, l, ("mfcc", "energy"))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): s, m, _ = ex.extract_batch_device(audio, o, l, ("mfcc", "energy"))
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    hours = float(l.astype(np.int64).sum()) / sr / 3600
    return {"ms": round(ms, 3), "audio_hours_per_s": round(hours / (ms * 1e-3), 1), "clips": int(len(l)), "ok": int((m[:, 0] == 0).sum().item())}

rng = np.random.default_rng(7)
res = {}
res["uniform_10s"] = timed(off, ln)
res["cut_in_slot_0.5_10s"] = timed(off, [int(rng.integers(int(0.5 * sr), stride - 8)) for _ in range(nl)])
res["whole_2_slots_20s"] = timed(off[0:nl - 1:2], [2 * stride - 8] * len(off[0:nl - 1:2]))
res["whole_3_slots_30s"] = timed(off[0:nl - 2:3], [3 * stride - 8] * len(off[0:nl - 2:3]))
ro, rl, i = [], [], 0
while i + 3 <= nl:
    k = int(rng.integers(1, 4)); ro.append(int(off[i])); rl.append(int(rng.integers(int(0.5 * sr), k * stride - 8))); i += k
res["bench_ragged_mix"] = timed(ro, rl)
order = np.argsort(-np.asarray(rl))
res["bench_ragged_mix_longest_first"] = timed(np.asarray(ro)[order], np.asarray(rl)[order])
order = np.argsort(np.asarray(rl))
res["bench_ragged_mix_shortest_first"] = timed(np.asarray(ro)[order], np.asarray(rl)[order])
for k, v in res.items(): print(json.dumps({"case": k, **v}))
