"""Seeded synthetic clips shared by the parity tests, the oracle and bench.py.

speech-like / harmonic clip: f0 ~ U(f_lo, f_hi), 10 harmonics with amplitude 1/k, peak-normalised
to 0.5, optional +-2 % vibrato at 5 Hz, white noise at -30 dBFS RMS, 100 ms raised-cosine fades and
250 ms of -70 dBFS noise at head and tail (so the silence trim has something to trim).
music-like clip: three simultaneous harmonic notes + noise.

`make_clip` is the numpy generator used for parity (bit-identical input on both sides);
`make_batch_torch` builds whole batches on any torch device for the throughput benchmark.
"""
from __future__ import annotations

import numpy as np

__all__ = ["make_clip", "make_clips", "make_batch_torch", "CONFIGS"]

# name -> extractor kwargs + clip recipe (BASELINE.json configs C1..C4)
CONFIGS = {
    "c1": dict(sr=22050, frame_length=1024, hop_length=256, n_mfcc=13, seconds=5.0, f_lo=110.0, f_hi=880.0, kind="speech"),
    "c2": dict(sr=22050, frame_length=1024, hop_length=256, n_mfcc=13, seconds=10.0, f_lo=110.0, f_hi=880.0, kind="speech"),
    "c3": dict(sr=16000, frame_length=512, hop_length=128, n_mfcc=40, seconds=10.0, f_lo=80.0, f_hi=400.0, kind="speech"),
    "c4": dict(sr=44100, frame_length=2048, hop_length=512, n_mfcc=13, seconds=10.0, f_lo=110.0, f_hi=880.0, kind="music"),
}


def _harmonic(t, f0_inst_phase, n_harm=10):
    y = np.zeros_like(t)
    for k in range(1, n_harm + 1):
        y += np.sin(k * f0_inst_phase) / k
    return y


def make_clip(seed: int, sr: int = 22050, seconds: float = 10.0, f_lo: float = 110.0, f_hi: float = 880.0,
              kind: str = "speech", vibrato: bool = True, head_tail_s: float = 0.25) -> np.ndarray:
    """One float32 clip in [-1, 1]; deterministic in (seed, arguments)."""
    rng = np.random.default_rng(seed)
    n = int(round(seconds * sr))
    n_ht = int(round(head_tail_s * sr))
    n_body = max(n - 2 * n_ht, 1)
    t = np.arange(n_body, dtype=np.float64) / sr
    n_notes = 1 if kind == "speech" else 3
    body = np.zeros(n_body)
    for _ in range(n_notes):
        f0 = rng.uniform(f_lo, f_hi)
        dev = 0.02 * f0 if vibrato else 0.0
        # instantaneous phase of f0 + dev*sin(2 pi 5 t)
        phase = 2 * np.pi * f0 * t - (dev / 5.0) * (np.cos(2 * np.pi * 5.0 * t) - 1.0)
        body += _harmonic(t, phase + rng.uniform(0, 2 * np.pi))
    body *= 0.5 / np.max(np.abs(body))
    n_fade = min(int(0.1 * sr), n_body // 2)
    if n_fade > 0:
        ramp = 0.5 - 0.5 * np.cos(np.pi * np.arange(n_fade) / n_fade)
        body[:n_fade] *= ramp
        body[n_body - n_fade:] *= ramp[::-1]
    body += 10 ** (-30 / 20) * rng.standard_normal(n_body)
    y = np.empty(n, dtype=np.float64)
    quiet = 10 ** (-70 / 20)
    y[:n_ht] = quiet * rng.standard_normal(n_ht)
    y[n_ht:n_ht + n_body] = body
    y[n_ht + n_body:] = quiet * rng.standard_normal(n - n_ht - n_body)
    return np.clip(y, -1.0, 1.0).astype(np.float32)


def make_clips(seeds, **kw):
    return [make_clip(s, **kw) for s in seeds]


def make_batch_torch(n_clips: int, sr: int = 22050, seconds: float = 10.0, f_lo: float = 110.0, f_hi: float = 880.0,
                     kind: str = "speech", seed: int = 0, device="cuda", chunk: int = 256, align: int = 4):
    """[n_clips, stride] float32 batch of the same recipe built with torch ops on `device`
    (stride = samples rounded up to `align`); returns (flat_audio, offsets[int64], lengths[int32]).

    Not bit-identical to `make_clip` (different RNG); used where only the workload shape matters.
    """
    import torch

    n = int(round(seconds * sr))
    stride = (n + align - 1) // align * align
    n_ht = int(round(0.25 * sr))
    n_body = n - 2 * n_ht
    n_fade = int(0.1 * sr)
    out = torch.zeros(n_clips * stride, dtype=torch.float32, device=device)
    view = out.view(n_clips, stride)
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    t = torch.arange(n_body, dtype=torch.float32, device=device) / sr
    ramp = 0.5 - 0.5 * torch.cos(torch.pi * torch.arange(n_fade, dtype=torch.float32, device=device) / n_fade)
    n_notes = 1 if kind == "speech" else 3
    for c0 in range(0, n_clips, chunk):
        c1 = min(n_clips, c0 + chunk)
        b = c1 - c0
        body = torch.zeros(b, n_body, dtype=torch.float32, device=device)
        for _ in range(n_notes):
            f0 = f_lo + (f_hi - f_lo) * torch.rand(b, 1, generator=g, device=device)
            ph0 = 2 * torch.pi * torch.rand(b, 1, generator=g, device=device)
            phase = 2 * torch.pi * f0 * t[None, :] - (0.02 * f0 / 5.0) * (torch.cos(2 * torch.pi * 5.0 * t[None, :]) - 1.0) + ph0
            for k in range(1, 11):
                body += torch.sin(k * phase) / k
        body *= 0.5 / body.abs().amax(dim=1, keepdim=True)
        body[:, :n_fade] *= ramp
        body[:, n_body - n_fade:] *= ramp.flip(0)
        body += 10 ** (-30 / 20) * torch.randn(b, n_body, generator=g, device=device)
        quiet = 10 ** (-70 / 20)
        view[c0:c1, :n_ht] = quiet * torch.randn(b, n_ht, generator=g, device=device)
        view[c0:c1, n_ht:n_ht + n_body] = body
        view[c0:c1, n_ht + n_body:n] = quiet * torch.randn(b, n - n_ht - n_body, generator=g, device=device)
    out.clamp_(-1.0, 1.0)
    offsets = torch.arange(n_clips, dtype=torch.int64) * stride
    lengths = torch.full((n_clips,), n, dtype=torch.int32)
    return out, offsets, lengths
