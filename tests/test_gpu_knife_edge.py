"""GPU parity at the knife edges of the two threshold decisions of the path:

* librosa.effects.trim keeps an RMS frame iff 10 log10(r^2) - 10 log10(max r^2) > -30 dB.  The device sums the squares
  of a 2048-sample frame in a different order than numpy (hop-block sums of a warp tree instead of np.mean's pairwise
  sum) and pre-emphasises with one fused multiply-add, so a frame whose level lies within a few float32 ulp of the
  threshold may land on the other side -- and then trim_start / trim_end move by 512 samples.  The tests below pin the
  WIDTH of that band: outside +-2e-6 relative (~ +-17 ulp of the frame's mean square) the indices always match.
* power_to_db(top_db=80) clamps at the clip maximum - 80 dB.  That decision is continuous: a value an ulp on either side
  of the floor moves a log-mel entry by ~1e-5 dB, so parity holds straight through the edge.
Needs a B200: run with `-m gpu`."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _gpu_trim(ex, clips):
    import torch
    buf, offsets, lengths = ex.pack_clips(clips)
    summ, meta, _ = ex.extract_batch_device(buf.cuda(), offsets, lengths, ("energy",))
    torch.cuda.synchronize()
    return meta.cpu().numpy()


def _edge_clip(rng):
    """2 s of noise at -6 dBFS, then 1.5 s of noise decaying by 0.5 dB per 512 samples through the -30 dB trim threshold:
    exactly one RMS frame lies near the threshold, and it is the last candidate for trim_end"""
    sr = 22050
    loud = (0.5 * rng.standard_normal(2 * sr)).astype(np.float32)
    n = int(1.5 * sr)
    env = 10 ** ((-24.0 - 0.5 * np.arange(n) / 512.0) / 20.0)
    quiet = (0.5 * env * rng.standard_normal(n)).astype(np.float32)
    return np.concatenate([loud, quiet]), 2 * sr


def test_trim_decision_matches_outside_a_few_ulp_of_the_threshold(torch_cuda):
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from oracle import OracleFeatureExtractor
    import oracle as O
    ex, ora = AudioFeatureExtractor(), OracleFeatureExtractor()
    rng = np.random.default_rng(11)
    base, n_loud = _edge_clip(rng)
    # the RMS frames through the oracle's own float32 arithmetic: q = the tail frame nearest the threshold
    r = O.rms(O.preemphasis(base, 0.97), 2048, 512)[0].astype(np.float64)
    thr = r.max() * 10 ** (-30 / 20)
    tail = np.arange(n_loud // 512 + 4, len(r) - 3)
    q = int(tail[np.argmin(np.abs(np.log(r[tail] / thr)))])
    g0 = float(thr / r[q])                     # scaling the tail by g0 puts frame q on the threshold
    assert abs(g0 - 1.0) < 0.06
    rels = np.concatenate([np.linspace(-1e-5, 1e-5, 81), np.linspace(-3e-7, 3e-7, 61)])
    clips = []
    for rel in rels:
        y = base.copy()
        y[n_loud:] *= np.float32(g0 * (1.0 + rel))
        clips.append(y)
    meta = _gpu_trim(ex, clips)
    mism, ends = [], set()
    for rel, y, m in zip(rels, clips, meta):
        s, e = ora.trim_indices(y)
        ends.add(e)
        if (int(m[1]), int(m[2])) != (s, e):
            mism.append(float(rel))
    # the scan must see both sides of the decision, and every disagreement must sit inside the ulp band
    assert len(ends) >= 2, "the scan did not cross the threshold"
    assert all(abs(x) <= 2e-6 for x in mism), mism
    print(f"knife edge: {len(mism)} of {len(rels)} scan points disagree, all within |rel| <= {max(map(abs, mism), default=0):.1e}")


def test_trim_flip_rate_on_random_clips(torch_cuda):
    """Random clips with tails within +-3 dB of the trim threshold: the trim indices must match the oracle on all of them
    (a flip needs a frame within ~1e-6 relative of the threshold: expected once per ~10^5 such frames).
    AFX_KNIFE_N scales the sample (default 4000 clips; DESIGN.md quotes the 10^5 run)."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from oracle import OracleFeatureExtractor
    ex, ora = AudioFeatureExtractor(), OracleFeatureExtractor()
    n = int(os.environ.get("AFX_KNIFE_N", "4000"))
    rng = np.random.default_rng(2024)
    flips = total = 0
    for c0 in range(0, n, 2000):
        clips = []
        for _ in range(min(2000, n - c0)):
            loud = (0.3 * rng.standard_normal(int(rng.integers(4000, 9000)))).astype(np.float32)
            lvl = 10 ** ((-30 + rng.uniform(-3, 3)) / 20)
            tail = (0.3 * lvl * rng.standard_normal(int(rng.integers(3000, 9000)))).astype(np.float32)
            head = (0.3 * 10 ** ((-30 + rng.uniform(-3, 3)) / 20) * rng.standard_normal(int(rng.integers(0, 4000)))).astype(np.float32)
            clips.append(np.concatenate([head, loud, tail]))
        meta = _gpu_trim(ex, clips)
        for y, m in zip(clips, meta):
            s, e = ora.trim_indices(y)
            total += 1
            flips += (int(m[1]), int(m[2])) != (s, e)
    print(f"trim flip rate: {flips} of {total} clips")
    assert flips <= max(1, total // 20000), (flips, total)


def test_top_db_clamp_is_continuous_through_its_edge(torch_cuda):
    """A gap whose log-mel minimum crosses (clip max - 80 dB): MFCC parity must hold on both sides and at the edge."""
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor
    import oracle as O
    ex, ora = AudioFeatureExtractor(), OracleFeatureExtractor()
    a = make_clip(60, seconds=1.5)
    rng = np.random.default_rng(5)
    gap = rng.standard_normal(12000).astype(np.float32)
    # find the gap level at which the log-mel range is 80 dB, then scan +-0.02 dB around it
    def span(level):
        y = np.concatenate([a, (level * gap).astype(np.float32), a])
        _, D = O.mfcc(ora.preprocess_audio(y), sr=22050, n_mfcc=13, n_fft=1024, hop_length=256, return_logmel=True)
        return y, float(D.max() - np.min(D))
    lo, hi = 1e-7, 1e-2
    for _ in range(40):
        mid = (lo * hi) ** 0.5
        if span(mid)[1] >= 80.0 - 1e-9: lo = mid
        else: hi = mid
    clips = [span(lo * 10 ** (d / 20))[0] for d in np.linspace(-0.02, 0.02, 9)]
    buf, offsets, lengths = ex.pack_clips(clips)
    summ, meta, frames = ex.extract_batch_device(buf.cuda(), offsets, lengths, ("mfcc", "energy"), return_frames=True)
    torch.cuda.synchronize()
    summ, meta, frames = summ.cpu().numpy(), meta.cpu().numpy(), frames.cpu().numpy()
    rows = np.concatenate([[0], np.cumsum([ex.max_frames(int(l)) for l in lengths])])
    for i, y in enumerate(clips):
        m, d1, d2 = ora.mfcc_frames(ora.preprocess_audio(y))
        T = m.shape[1]
        assert meta[i, 3] == T
        assert np.max(np.abs(frames[rows[i]:rows[i] + T, 0:13].T - m)) <= 1e-3


def test_energy_only_accepts_clips_shorter_than_the_delta_width(torch_cuda):
    """librosa.feature.delta needs 9 frames, librosa.feature.rms does not (core/feature_extractor.py:164-168): without
    the MFCC group a 6-frame clip is extracted, with it the clip is AFX_CLIP_TOO_SHORT."""
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor
    ex, ora = AudioFeatureExtractor(), OracleFeatureExtractor()
    short = make_clip(8, seconds=2.0)[6000:6000 + 1500]
    buf, offsets, lengths = ex.pack_clips([short])
    s1, m1, _ = ex.extract_batch_device(buf.cuda(), offsets, lengths, ("energy",))
    s2, m2, _ = ex.extract_batch_device(buf.cuda(), offsets, lengths, ("mfcc", "energy"))
    torch.cuda.synchronize()
    assert int(m2[0, 0]) == N.CLIP_TOO_SHORT and int(m1[0, 0]) == N.CLIP_OK
    ref = ora.extract_energy(ora.preprocess_audio(short))
    n = ex.n_mfcc
    got = s1.cpu().numpy()[0, 4 + 4 * n:4 + 4 * n + 3]
    np.testing.assert_allclose(got, [ref["energy_mean"], ref["energy_std"], ref["energy_range"]], rtol=1e-5)
