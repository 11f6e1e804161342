"""The CPU oracle (numpy/scipy restatement of librosa 0.11.0 on the AudioFeatureExtractor path) against
independent implementations that are installed here (scipy, torchaudio, transformers.audio_utils) and
against known-answer cases.  Parity with librosa itself is UNPINNED (see oracle/__init__.py): these
tests pin the restatement block by block instead.  CPU only."""
import math

import numpy as np
import pytest
import scipy.fft
import scipy.signal

import oracle as O
from audio_feature_extraction_b200.synth import make_clip
from oracle import OracleFeatureExtractor

CFGS = [(22050, 1024, 256), (16000, 512, 128), (44100, 2048, 512)]


def test_note_constants_match_reference_defaults():
    # core/feature_extractor.py:15-16: librosa.note_to_hz('C2'), ('C7')
    assert O.note_c2_hz == pytest.approx(65.40639132514966, abs=1e-12)
    assert O.note_c7_hz == pytest.approx(2093.004522404789, abs=1e-9)
    assert O.note_c7_hz / O.note_c2_hz == pytest.approx(32.0, rel=1e-15)


def test_preemphasis_first_sample_rule_and_fir():
    rng = np.random.default_rng(0)
    y = rng.standard_normal(1000).astype(np.float32)
    out = O.preemphasis(y, 0.97)
    assert out.dtype == np.float32
    assert out[0] == np.float32(y[0] + (2 * y[0] - y[1]))
    ref = y[1:] + np.float32(-0.97) * y[:-1]
    np.testing.assert_array_equal(out[1:], ref.astype(np.float32))


@pytest.mark.parametrize("L,hop", [(0, 256), (1, 256), (255, 256), (256, 256), (22050, 256), (110250, 256), (160000, 128)])
def test_frame_counts(L, hop):
    y = np.zeros(max(L, 0), dtype=np.float32)
    r = O.rms(y, frame_length=4 * hop, hop_length=hop)
    assert r.shape == (1, 1 + L // hop)


def test_trim_exact_indices_on_digital_silence():
    sr = 22050
    y = np.zeros(5 * sr, dtype=np.float32)
    t = np.arange(2 * sr) / sr
    a, b = 30000, 30000 + 2 * sr
    y[a:b] = 0.5 * np.sin(2 * np.pi * 220 * t)
    s, e = O.trim_indices(y, top_db=30)
    assert s % 512 == 0 and (e % 512 == 0 or e == len(y))
    # frame i covers [512 i - 1024, 512 i + 1024): the first frames touching the tone are far below -30 dB
    assert a - 1536 <= s <= a + 1024 and b - 1024 <= e <= b + 1536
    s0, e0 = O.trim_indices(np.zeros(4096, dtype=np.float32))
    # all-zero signal: every frame is at 0 dB relative to the (zero) max -> nothing is trimmed
    assert (s0, e0) == (0, 4096)


@pytest.mark.parametrize("sr,n_fft,hop", CFGS)
def test_mel_filterbank_matches_torchaudio_and_transformers(sr, n_fft, hop):
    import torch
    import torchaudio
    from transformers.audio_utils import mel_filter_bank

    B = O.mel_filterbank(sr, n_fft)
    assert B.shape == (128, 1 + n_fft // 2) and B.dtype == np.float32
    ta = torchaudio.functional.melscale_fbanks(1 + n_fft // 2, 0.0, sr / 2, 128, sr, norm="slaney", mel_scale="slaney").T.numpy()
    tr = mel_filter_bank(1 + n_fft // 2, 128, 0.0, sr / 2, sr, norm="slaney", mel_scale="slaney").T
    assert np.max(np.abs(B - ta)) < 1e-6
    assert np.max(np.abs(B - tr)) < 1e-6
    assert (B.sum(axis=1) > 0).all()


def test_window_dct_delta_blocks():
    w = O.hamming_periodic(1024)
    k = np.arange(1024)
    np.testing.assert_allclose(w, 0.54 - 0.46 * np.cos(2 * np.pi * k / 1024), atol=1e-15)
    rng = np.random.default_rng(1)
    D = rng.standard_normal((128, 50)).astype(np.float32)
    C = O.dct_ortho(D, 13)
    m = np.arange(128)
    basis = np.stack([(math.sqrt(1 / 128) if c == 0 else math.sqrt(2 / 128)) * np.cos(np.pi * c * (2 * m + 1) / 256) for c in range(13)])
    np.testing.assert_allclose(C, basis @ D.astype(np.float64), atol=2e-5)
    x = rng.standard_normal((3, 40)).astype(np.float32)
    d1, d2 = O.delta(x), O.delta(x, order=2)
    kk = np.arange(-4, 5)
    c2 = np.array([28, 7, -8, -17, -20, -17, -8, 7, 28]) / 462
    for t in range(4, 36):
        np.testing.assert_allclose(d1[:, t], (x[:, t - 4:t + 5] * kk).sum(1) / 60, atol=1e-5)
        np.testing.assert_allclose(d2[:, t], (x[:, t - 4:t + 5] * c2).sum(1), atol=1e-5)
    for t in (0, 1, 2, 3):
        np.testing.assert_allclose(d1[:, t], d1[:, 4], atol=1e-5)
        np.testing.assert_allclose(d2[:, -1 - t], d2[:, -5], atol=1e-5)
    with pytest.raises(ValueError):
        O.delta(x[:, :8])


@pytest.mark.parametrize("sr,n_fft,hop", CFGS)
def test_melspectrogram_db_matches_transformers_pipeline(sr, n_fft, hop):
    from transformers.audio_utils import mel_filter_bank, spectrogram, window_function

    y = make_clip(3, sr=sr, seconds=1.0)
    M = O.melspectrogram(y, sr, n_fft, hop)
    fb = mel_filter_bank(1 + n_fft // 2, 128, 0.0, sr / 2, sr, norm="slaney", mel_scale="slaney")
    S = spectrogram(y, window_function(n_fft, "hamming", periodic=True), frame_length=n_fft, hop_length=hop, power=2.0,
                    center=True, pad_mode="constant", mel_filters=fb)
    assert S.shape == M.shape == (128, 1 + len(y) // hop)
    np.testing.assert_allclose(M, S, rtol=2e-4, atol=1e-9)


def test_mfcc_float32_vs_float64_noise_floor():
    y = OracleFeatureExtractor().preprocess_audio(make_clip(0, seconds=3.0))
    a = O.mfcc(y, 22050, 13, 1024, 256, dtype=np.float32)
    b = O.mfcc(y.astype(np.float64), 22050, 13, 1024, 256, dtype=np.float64)
    assert a.dtype == np.float32
    assert np.max(np.abs(a - b)) < 1e-3   # the float32 path's own rounding is below BASELINE.json's MFCC tolerance


def test_pyin_constants_table():
    # SURVEY.md Appendix A.7 table
    for (sr, fl, hop), exp in zip(CFGS, [(512, 10, 338, 601, 51), (256, 7, 245, 601, 31), (1024, 21, 675, 601, 51)]):
        c = O.pyin_constants(sr, fl, hop, O.note_c2_hz, O.note_c7_hz)
        assert (c["win_length"], c["min_period"], c["max_period"], c["n_pitch_bins"], c["transition_width"]) == exp
    b = O.beta_probs()
    assert b.sum() == pytest.approx(1.0, abs=1e-12)
    np.testing.assert_allclose(b[:5], [0.01527, 0.03934, 0.05534, 0.06513, 0.07021], atol=1e-5)


def test_yin_difference_equals_direct_sum():
    rng = np.random.default_rng(5)
    fl, W = 1024, 512
    x = rng.standard_normal((fl, 3)).astype(np.float32)
    yin = O.yin_cmnd(x, fl, W, 10, 338)
    xd = x.astype(np.float64)
    for t in range(3):
        d = np.array([np.sum((xd[1:W + 1, t] - xd[1 + tau:W + 1 + tau, t]) ** 2) for tau in range(339)])
        cm = np.cumsum(d[1:]) / np.arange(1, 339)
        ref = d[10:339] / cm[9:338]
        np.testing.assert_allclose(yin[:, t], ref, rtol=2e-4)


def test_pyin_220hz_tone_is_exact_bin():
    sr = 22050
    t = np.arange(int(1.0 * sr)) / sr
    y = sum(np.sin(2 * np.pi * 220 * k * t) / k for k in range(1, 6))
    y = (0.3 * y / np.max(np.abs(y))).astype(np.float32)
    y = np.concatenate([1e-4 * np.random.default_rng(0).standard_normal(sr // 4).astype(np.float32), y])
    f0, voiced, vp = O.pyin(y, O.note_c2_hz, O.note_c7_hz, sr=sr, frame_length=1024, hop_length=256)
    assert len(f0) == 1 + len(y) // 256
    body = slice(30, len(f0) - 6)
    assert voiced[body].all()
    expect = O.note_c2_hz * 2 ** (210 / 120)   # bin 210 of 601 = 220.0 Hz
    np.testing.assert_allclose(f0[body], expect, rtol=1e-12)
    assert expect == pytest.approx(220.0, rel=1e-12)
    assert not voiced[:8].any() and np.isnan(f0[:8]).all()


@pytest.mark.parametrize("sr,fl,hop,freq", [(22050, 1024, 256, 110.0), (22050, 1024, 256, 440.0), (16000, 512, 128, 155.563492),
                                            (44100, 2048, 512, 329.627557)])
def test_pyin_analytic_tones_decode_to_the_nearest_pitch_bin(sr, fl, hop, freq):
    """Independent check of the pYIN restatement: a steady harmonic tone must be voiced on every interior
    frame and decode to the 0.1-semitone bin nearest its fundamental -- known without any reference implementation."""
    t = np.arange(int(0.8 * sr)) / sr
    y = sum(np.sin(2 * np.pi * freq * k * t) / k for k in range(1, 5))
    y = (0.3 * y / np.max(np.abs(y))).astype(np.float32)
    f0, voiced, vp = O.pyin(y, O.note_c2_hz, O.note_c7_hz, sr=sr, frame_length=fl, hop_length=hop)
    body = slice(6, len(f0) - 6)
    assert voiced[body].all()
    b = int(round(120 * np.log2(freq / O.note_c2_hz)))
    np.testing.assert_allclose(f0[body], O.note_c2_hz * 2 ** (b / 120), rtol=1e-12)
    assert np.all(vp[body] > 0.5)


def test_pyin_analytic_glide_silence_and_noise():
    sr = 22050
    # a slow glide 200 -> 260 Hz (4.5 semitones per second, far inside the 35.92 octaves/s transition limit): the track
    # must follow within one bin (10 cents) of the instantaneous frequency
    n = int(1.0 * sr)
    t = np.arange(n) / sr
    finst = 200.0 * (260.0 / 200.0) ** t
    phase = 2 * np.pi * np.cumsum(finst) / sr
    y = (0.3 * sum(np.sin(k * phase) / k for k in range(1, 4)) / 1.6).astype(np.float32)
    f0, voiced, _ = O.pyin(y, O.note_c2_hz, O.note_c7_hz, sr=sr, frame_length=1024, hop_length=256)
    idx = np.arange(8, len(f0) - 8)
    assert voiced[idx].all()
    cents = 1200 * np.abs(np.log2(f0[idx] / finst[np.minimum(idx * 256, n - 1)]))
    assert cents.max() <= 15.0, cents.max()
    # digital silence and white noise: no pitch
    f0s, vs, vps = O.pyin(np.zeros(sr // 2, np.float32), O.note_c2_hz, O.note_c7_hz, sr=sr, frame_length=1024, hop_length=256)
    assert not vs.any() and np.isnan(f0s).all() and np.all(vps < 0.5)
    rng = np.random.default_rng(1)
    f0n, vn, _ = O.pyin((0.1 * rng.standard_normal(sr)).astype(np.float32), O.note_c2_hz, O.note_c7_hz, sr=sr, frame_length=1024, hop_length=256)
    assert vn.mean() < 0.2


def test_transition_matrix_structure():
    A = O.pyin_transition(601, 51)
    assert A.shape == (1202, 1202)
    np.testing.assert_allclose(A.sum(axis=1), 1.0, atol=1e-12)
    row = A[300, :601]
    nz = np.flatnonzero(row)
    assert nz[0] == 275 and nz[-1] == 325
    tri = np.concatenate([np.arange(1, 27), np.arange(25, 0, -1)]) / 26.0
    np.testing.assert_allclose(row[275:326], 0.99 * tri / tri.sum(), rtol=1e-12)
    np.testing.assert_allclose(A[300, 601 + 275:601 + 326], 0.01 * tri / tri.sum(), rtol=1e-12)
    # truncated and renormalised at the edge
    r0 = A[0, :601]
    assert np.flatnonzero(r0)[-1] == 25 and r0.sum() == pytest.approx(0.99)


def test_extractor_keys_and_consumer_contract():
    ora = OracleFeatureExtractor()
    y = make_clip(1, seconds=2.0)
    f = ora.extract_features_array(y, features_to_extract=("mfcc", "energy"))
    assert list(f) == ["file_path", "mfcc_mean", "mfcc_std", "mfcc_delta_mean", "mfcc_delta2_mean",
                       "energy_mean", "energy_std", "energy_range"]
    assert len(f["mfcc_mean"]) == 13 and all(isinstance(v, float) for v in f["mfcc_mean"])
    import json
    json.dumps(f)
