"""Host->device copy rate of this box from pinned memory: one stream, then two streams at once (sets the ceiling of the
e2e leg of bench.py, which ships 16-bit PCM)."""
import time
import torch

n = 1 << 30
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); h.random_(0, 255)
h2 = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n, dtype=torch.uint8, device="cuda"); d2 = torch.empty_like(d)
for _ in range(2):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(4):
    d.copy_(h, non_blocking=True)
e1.record(); torch.cuda.synchronize()
print("one stream  : %.1f GB/s" % (4 * n / e0.elapsed_time(e1) / 1e6))
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(4):
    with torch.cuda.stream(s1):
        d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        d2.copy_(h2, non_blocking=True)
torch.cuda.synchronize(); t1 = time.perf_counter()
print("two streams : %.1f GB/s" % (8 * n / (t1 - t0) / 1e9))
for chunk in (16 << 20, 64 << 20, 256 << 20):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for o in range(0, n, chunk):
        d[o:o + chunk].copy_(h[o:o + chunk], non_blocking=True)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print("chunks of %4d MiB: %.1f GB/s" % (chunk >> 20, n / (t1 - t0) / 1e9))

# ---- the same copy while the fused kernel runs on another stream (does the kernel slow the DMA down?)
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
from audio_feature_extraction_b200 import AudioFeatureExtractor
from audio_feature_extraction_b200.synth import make_batch_torch, CONFIGS
cfg = CONFIGS["c2"]
audio, offsets, lengths = make_batch_torch(4000, sr=cfg["sr"], seconds=cfg["seconds"], f_lo=cfg["f_lo"], f_hi=cfg["f_hi"],
                                           kind=cfg["kind"], seed=1, device=torch.device("cuda", 0), align=8)
offsets, lengths = offsets.numpy(), lengths.numpy()
ex = AudioFeatureExtractor()
ex.extract_batch_device(audio, offsets, lengths, ("mfcc", "energy"))
torch.cuda.synchronize()
sk, sc = torch.cuda.Stream(), torch.cuda.Stream()
for busy in (False, True):
    torch.cuda.synchronize()
    if busy:
        with torch.cuda.stream(sk):
            for _ in range(20):                       # ~ 20 x 5.6 ms of kernel time
                ex.extract_batch_device(audio, offsets, lengths, ("mfcc", "energy"), stream=sk)
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(sc):
        a0.record(sc)
        for _ in range(4):
            d.copy_(h, non_blocking=True)
        a1.record(sc)
    torch.cuda.synchronize()
    print("copy with the fused kernel %s: %.1f GB/s" % ("running" if busy else "idle   ", 4 * n / a0.elapsed_time(a1) / 1e6))
