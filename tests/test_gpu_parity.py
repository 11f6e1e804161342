"""GPU parity: the CUDA path (through the C ABI of libafx_b200.so) against the CPU oracle on the same
seeded clips, at BASELINE.json's tolerances:
  frame counts / trim indices bit-exact; MFCC <= 1e-3 abs (after the dB step); energy <= 1e-5 rel.
Needs a B200: run with `-m gpu`."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MFCC_ATOL = 1e-3      # BASELINE.json north_star: "MFCC: within 1e-3 absolute after the dB step"
ENERGY_RTOL = 1e-5    # "energy: within 1e-5 relative"
GROUPS = ("mfcc", "energy")


def _framewise(ex, clips, groups=GROUPS):
    import torch
    buf, offsets, lengths = ex.pack_clips(clips)
    audio = buf.cuda()
    summ, meta, frames = ex.extract_batch_device(audio, offsets, lengths, groups, return_frames=True)
    torch.cuda.synchronize()
    rows = np.concatenate([[0], np.cumsum([ex.max_frames(int(l)) for l in lengths])])
    return summ.cpu().numpy(), meta.cpu().numpy(), frames.cpu().numpy(), rows


def _check_clip(ex, ora, y, summ, meta, frames, row0, n):
    s, e = ora.trim_indices(y)
    yt = ora.preprocess_audio(y)
    m, d1, d2 = ora.mfcc_frames(yt)
    r = ora.rms_frames(yt)[0]
    T = m.shape[1]
    # frame counts and indexing: bit-exact
    assert meta.tolist() == [0, s, e, T]
    fr = frames[row0:row0 + T]
    assert np.max(np.abs(fr[:, 0:n].T - m)) <= MFCC_ATOL
    assert np.max(np.abs(fr[:, n:2 * n].T - d1)) <= MFCC_ATOL
    assert np.max(np.abs(fr[:, 2 * n:3 * n].T - d2)) <= MFCC_ATOL
    np.testing.assert_allclose(fr[:, 3 * n], r, rtol=ENERGY_RTOL, atol=1e-12)
    # summary row in the reference's key order (include/afx_b200.h)
    assert np.max(np.abs(summ[4:4 + n] - m.mean(axis=1))) <= MFCC_ATOL
    assert np.max(np.abs(summ[4 + n:4 + 2 * n] - m.std(axis=1))) <= MFCC_ATOL
    assert np.max(np.abs(summ[4 + 2 * n:4 + 3 * n] - d1.mean(axis=1))) <= MFCC_ATOL
    assert np.max(np.abs(summ[4 + 3 * n:4 + 4 * n] - d2.mean(axis=1))) <= MFCC_ATOL
    e0 = 4 + 4 * n
    np.testing.assert_allclose(summ[e0:e0 + 3], [r.mean(), r.std(), np.ptp(r)], rtol=ENERGY_RTOL)


def test_c2_config_framewise_and_summary_parity(torch_cuda):
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor

    secs = [10.0, 5.0, 3.3, 1.0, 0.5, 7.77, 10.0, 2.0]
    clips = [make_clip(s, seconds=sec) for s, sec in enumerate(secs)]
    ex = AudioFeatureExtractor()
    ora = OracleFeatureExtractor()
    summ, meta, frames, rows = _framewise(ex, clips)
    for i, y in enumerate(clips):
        _check_clip(ex, ora, y, summ[i], meta[i], frames, rows[i], ex.n_mfcc)


@pytest.mark.parametrize("cfg", [
    dict(sr=16000, frame_length=512, hop_length=128, n_mfcc=40),     # BASELINE.json configs[2] (C3)
    dict(sr=44100, frame_length=2048, hop_length=512, n_mfcc=13),    # configs[3] (C4)
    dict(sr=22050, frame_length=1024, hop_length=512, n_mfcc=20),
    dict(sr=22050, frame_length=1024, hop_length=128, n_mfcc=13),
    dict(sr=16000, frame_length=512, hop_length=256, n_mfcc=13),
    dict(sr=16000, frame_length=512, hop_length=64, n_mfcc=13),
    dict(sr=44100, frame_length=2048, hop_length=256, n_mfcc=13),
    dict(sr=44100, frame_length=2048, hop_length=1024, n_mfcc=13),   # hop does not divide 512: separate trim pass
    # any hop the reference accepts (core/feature_extractor.py:10-17): the run-time-hop variant of the streaming kernel
    dict(sr=16000, frame_length=512, hop_length=160, n_mfcc=13),     # 04_feature_extraction_experiment's 10 ms hop
    dict(sr=22050, frame_length=1024, hop_length=200, n_mfcc=13),
    dict(sr=44100, frame_length=2048, hop_length=441, n_mfcc=13),
    dict(sr=22050, frame_length=1024, hop_length=33, n_mfcc=13),     # odd hop: tiles at every alignment
    dict(sr=16000, frame_length=512, hop_length=128, n_mfcc=64),     # widest coefficient set (four groups of 16)
    dict(sr=16000, frame_length=512, hop_length=64, n_mfcc=49),
], ids=lambda c: f"{c['sr']}-{c['frame_length']}-{c['hop_length']}-{c['n_mfcc']}")
def test_other_configs_framewise_and_summary_parity(torch_cuda, cfg):
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor

    kind = "music" if cfg["sr"] == 44100 else "speech"
    f_lo, f_hi = (80.0, 400.0) if cfg["sr"] == 16000 else (110.0, 880.0)
    secs = [4.0, 2.5, 0.9, 3.3, 1.7]
    clips = [make_clip(500 + s, sr=cfg["sr"], seconds=sec, f_lo=f_lo, f_hi=f_hi, kind=kind) for s, sec in enumerate(secs)]
    clips.append(np.concatenate([clips[1], np.zeros(30000, np.float32)]))     # trailing digital silence -> tail fix-up + clamp
    ex = AudioFeatureExtractor(**cfg)
    ora = OracleFeatureExtractor(**cfg)
    summ, meta, frames, rows = _framewise(ex, clips)
    for i, y in enumerate(clips):
        _check_clip(ex, ora, y, summ[i], meta[i], frames, rows[i], ex.n_mfcc)


def test_top_db_clamp_redo_pass(torch_cuda):
    """Clips whose log-mel range exceeds 80 dB inside the trimmed region: power_to_db(top_db=80) clamps, so the
    speculative (unclamped) DCT rows must be redone with the clip-global floor (core/feature_extractor.py:127-134)."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor

    rng = np.random.default_rng(5)
    a, b = make_clip(60, seconds=1.5), make_clip(61, seconds=2.0)
    hush = (1e-6 * rng.standard_normal(15000)).astype(np.float32)          # ~ -120 dBFS gap, kept by the trim
    digital_silence = np.zeros(9000, dtype=np.float32)                     # exact zeros -> amin floor, clamped
    tone = (0.5 * np.sin(2 * np.pi * 440.0 * np.arange(44100) / 22050.0)).astype(np.float32)   # pure tone: empty bins
    clips = [np.concatenate([a, hush, b]), np.concatenate([b, digital_silence, a]), tone, make_clip(62, seconds=1.0)]
    ex = AudioFeatureExtractor()
    ora = OracleFeatureExtractor()
    for y in clips[:3]:   # the oracle really clamps on these
        _, D = __import__("oracle").mfcc(ora.preprocess_audio(y), sr=22050, n_mfcc=13, n_fft=1024, hop_length=256, return_logmel=True)
        assert D.min() == pytest.approx(D.max() - 80.0, abs=1e-4)
    summ, meta, frames, rows = _framewise(ex, clips)
    for i, y in enumerate(clips):
        _check_clip(ex, ora, y, summ[i], meta[i], frames, rows[i], ex.n_mfcc)


def test_extended_statistics_match_scipy(torch_cuda):
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor

    clips = [make_clip(40 + s, seconds=4.0) for s in range(3)]
    ex = AudioFeatureExtractor(extended_stats=True)
    ora = OracleFeatureExtractor(extended_stats=True)
    got = ex.extract_batch(clips, features_to_extract=GROUPS)
    for y, g in zip(clips, got):
        ref = ora.extract_features_array(y, features_to_extract=GROUPS)
        np.testing.assert_allclose(g["mfcc_skew"], ref["mfcc_skew"], atol=2e-3)
        np.testing.assert_allclose(g["mfcc_kurtosis"], ref["mfcc_kurtosis"], atol=5e-3)
        assert g["energy_skew"] == pytest.approx(ref["energy_skew"], abs=1e-4)
        assert g["energy_kurtosis"] == pytest.approx(ref["energy_kurtosis"], abs=1e-4)


def test_edge_cases_status_codes(torch_cuda):
    """empty / non-finite / all-silent-after-trim / too-short clips: the librosa exceptions the reference
    re-raises (core/feature_extractor.py:211-213) become per-clip status codes; neighbours are unaffected."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor, FeatureExtractionError, _native as N
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor

    good = make_clip(7, seconds=2.0)
    nan = good.copy(); nan[12345] = np.nan
    inf = good.copy(); inf[-1] = np.inf
    short = make_clip(8, seconds=2.0)[6000:6000 + 1500]       # T = 1 + 1500 // 256 = 6 < 9
    empty = np.zeros(0, dtype=np.float32)
    one = np.array([0.25], dtype=np.float32)
    loud_then_zero = np.concatenate([good, np.zeros(50000, dtype=np.float32)])
    clips = [good, nan, empty, short, inf, one, loud_then_zero, good]
    ex = AudioFeatureExtractor()
    ora = OracleFeatureExtractor()
    summ, meta, frames, rows = _framewise(ex, clips)
    assert meta[:, 0].tolist() == [0, N.CLIP_NONFINITE, N.CLIP_EMPTY, N.CLIP_TOO_SHORT, N.CLIP_NONFINITE,
                                   N.CLIP_TOO_SHORT, 0, 0]
    assert (summ[[1, 2, 3, 4, 5]] == 0).all()
    for i in (0, 6, 7):
        _check_clip(ex, ora, clips[i], summ[i], meta[i], frames, rows[i], ex.n_mfcc)
    with pytest.raises(FeatureExtractionError):
        ex.extract_batch(clips, features_to_extract=GROUPS)
    kept = ex.extract_batch(clips, names=[f"c{i}.wav" for i in range(len(clips))], features_to_extract=GROUPS, skip_errors=True)
    assert [k["file_path"] for k in kept] == ["c0.wav", "c6.wav", "c7.wav"]


def test_ragged_batch_caller_workspace(torch_cuda):
    """Many ragged clips (dynamic clip -> CTA assignment) with a caller-allocated workspace of exactly
    afx_workspace_min_bytes: same rows bit for bit, and a short workspace is refused."""
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
    from audio_feature_extraction_b200.synth import make_clip

    rng = np.random.default_rng(0)
    clips = [make_clip(100 + i, seconds=float(rng.uniform(0.3, 3.0))) for i in range(40)]
    ex = AudioFeatureExtractor()
    buf, offsets, lengths = ex.pack_clips(clips)
    audio = buf.cuda()
    s_ref, m_ref, _ = ex.extract_batch_device(audio, offsets, lengths, GROUPS)
    L, h = N.lib(), ex._h()
    need = L.afx_workspace_min_bytes(h, len(clips), lengths.ctypes.data)
    ws = torch.empty(int(need), dtype=torch.uint8, device="cuda")
    s2 = torch.empty_like(s_ref); m2 = torch.empty_like(m_ref)
    N.check(L.afx_extract(h, audio.data_ptr(), offsets.ctypes.data, lengths.ctypes.data, len(clips), 0x6,
                          s2.data_ptr(), m2.data_ptr(), None, ws.data_ptr(), ws.numel(), None))
    torch.cuda.synchronize()
    assert torch.equal(m_ref, m2)
    assert torch.equal(s_ref, s2)          # same kernels, same order of operations -> bit-identical
    rc = L.afx_extract(h, audio.data_ptr(), offsets.ctypes.data, lengths.ctypes.data, len(clips), 0x6,
                       s2.data_ptr(), m2.data_ptr(), None, ws.data_ptr(), 4096, None)
    assert rc == N.ERR_WORKSPACE


def test_host_entry_points_match_device_entry(torch_cuda):
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
    from audio_feature_extraction_b200.synth import make_clip

    clips = [make_clip(200 + i, seconds=1.0 + 0.37 * i) for i in range(9)]
    ex = AudioFeatureExtractor()
    buf, offsets, lengths = ex.pack_clips(clips, align=8)
    s_ref, m_ref, _ = ex.extract_batch_device(buf.cuda(), offsets, lengths, GROUPS)
    s_ref, m_ref = s_ref.cpu().numpy(), m_ref.cpu().numpy()
    L, h = N.lib(), ex._h()
    n, SS = len(clips), ex.summary_width
    # (1) array of clip pointers
    ptrs = (C.c_void_p * n)(*[c.ctypes.data for c in clips])
    s1 = np.zeros((n, SS), np.float32); m1 = np.zeros((n, 4), np.int32)
    N.check(L.afx_extract_host(h, ptrs, lengths.ctypes.data, n, 0x6, s1.ctypes.data, m1.ctypes.data))
    np.testing.assert_array_equal(m1, m_ref); np.testing.assert_array_equal(s1, s_ref)
    # (2) one packed float32 buffer
    s2 = np.zeros((n, SS), np.float32); m2 = np.zeros((n, 4), np.int32)
    N.check(L.afx_extract_host_packed(h, buf.data_ptr(), N.SAMPLE_F32, buf.numel(), offsets.ctypes.data, lengths.ctypes.data,
                                      n, 0x6, s2.ctypes.data, m2.ctypes.data))
    np.testing.assert_array_equal(m2, m_ref); np.testing.assert_array_equal(s2, s_ref)
    # (3) packed PCM16 == float32 path fed x / 32768
    pcm = np.clip(np.round(buf.numpy() * 32768.0), -32768, 32767).astype(np.int16)
    f_from_pcm = torch.from_numpy(pcm.astype(np.float32) / np.float32(32768.0))
    s_q, m_q, _ = ex.extract_batch_device(f_from_pcm.cuda(), offsets, lengths, GROUPS)
    s3 = np.zeros((n, SS), np.float32); m3 = np.zeros((n, 4), np.int32)
    N.check(L.afx_extract_host_packed(h, pcm.ctypes.data, N.SAMPLE_S16, pcm.size, offsets.ctypes.data, lengths.ctypes.data,
                                      n, 0x6, s3.ctypes.data, m3.ctypes.data))
    np.testing.assert_array_equal(m3, m_q.cpu().numpy()); np.testing.assert_array_equal(s3, s_q.cpu().numpy())


def test_reference_api_files_and_evaluator(torch_cuda, tmp_path):
    """extract_features(path, features_to_extract=[...]) / batch_process(dir) on WAV files, and the
    reference's FeatureEvaluator contract (evaluation/evaluator.py:16-147) on the returned dicts."""
    import json
    from audio_feature_extraction_toolkit import AudioFeatureExtractor, FeatureEvaluator
    from audio_feature_extraction_b200.synth import make_clip
    from audio_feature_extraction_b200.wav import write_wav
    from oracle import OracleFeatureExtractor

    clips = {f"clip{i}.wav": make_clip(300 + i, seconds=1.5 + i) for i in range(3)}
    for name, y in clips.items():
        write_wav(str(tmp_path / name), y, 22050, subtype="FLOAT")
    (tmp_path / "notes.txt").write_text("ignored")
    write_wav(str(tmp_path / "silent.wav"), np.zeros(100, np.float32), 22050, subtype="FLOAT")   # too short -> skipped
    ex = AudioFeatureExtractor(sr=22050, frame_length=1024, hop_length=256, n_mfcc=13, pre_emphasis=0.97)
    ora = OracleFeatureExtractor()
    one = ex.extract_features(str(tmp_path / "clip1.wav"), features_to_extract=["mfcc", "energy"])
    ref = ora.extract_features_array(clips["clip1.wav"], features_to_extract=("mfcc", "energy"))
    assert list(one) == ["file_path", "mfcc_mean", "mfcc_std", "mfcc_delta_mean", "mfcc_delta2_mean",
                         "energy_mean", "energy_std", "energy_range"]
    np.testing.assert_allclose(one["mfcc_mean"], ref["mfcc_mean"], atol=MFCC_ATOL)
    assert one["energy_mean"] == pytest.approx(ref["energy_mean"], rel=ENERGY_RTOL)
    res = ex.batch_process(str(tmp_path), features_to_extract=["mfcc", "energy"])
    assert sorted(r["file_path"] for r in res) == sorted(str(tmp_path / n) for n in clips)
    json.dumps(res)
    ev = FeatureEvaluator()
    rep = ev.generate_evaluation_report(res, str(tmp_path / "report"))
    assert (tmp_path / "report" / "evaluation_detailed.json").exists()
    assert (tmp_path / "report" / "evaluation_summary.csv").exists()
    assert rep["quality_metrics"]["total_files"] == 3


F0_CENTS = 1.0          # BASELINE.json north_star: "F0: within 1 cent on voiced frames"
F0_VOICING = 0.999      # "voicing decisions matching on at least 99.9% of frames"


@pytest.mark.parametrize("cfg,secs", [
    (dict(sr=22050, frame_length=1024, hop_length=256, n_mfcc=13), [3.0, 2.0, 1.2, 4.0]),      # C1 / C2 geometry
    (dict(sr=16000, frame_length=512, hop_length=128, n_mfcc=40), [2.5, 1.5]),                  # C3
    (dict(sr=44100, frame_length=2048, hop_length=512, n_mfcc=13), [2.0, 1.5]),                 # C4
    # hops that do not divide frame_length / 2 (the reference hands any hop to librosa.pyin, core/feature_extractor.py:
    # 87-94): yin_obs_kernel in frame mode (one autocorrelation block per frame), generic Viterbi half width
    (dict(sr=16000, frame_length=512, hop_length=160, n_mfcc=13), [2.0, 1.3, 0.62]),
    (dict(sr=22050, frame_length=1024, hop_length=200, n_mfcc=13), [2.0, 1.1]),
    (dict(sr=44100, frame_length=2048, hop_length=441, n_mfcc=13), [1.5, 0.7]),
    (dict(sr=22050, frame_length=1024, hop_length=64, n_mfcc=13), [1.0]),                       # 16 hops per frame
], ids=["22050-1024-256", "16000-512-128", "44100-2048-512", "16000-512-160", "22050-1024-200", "44100-2048-441", "22050-1024-64"])
def test_pyin_f0_track_and_statistics(torch_cuda, cfg, secs):
    """extract_f0 (core/feature_extractor.py:76-114): the decoded pYIN track frame by frame and the four summary keys."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor

    kind = "music" if cfg["sr"] == 44100 else "speech"
    f_lo, f_hi = (80.0, 400.0) if cfg["sr"] == 16000 else (110.0, 880.0)
    clips = [make_clip(700 + s, sr=cfg["sr"], seconds=sec, f_lo=f_lo, f_hi=f_hi, kind=kind) for s, sec in enumerate(secs)]
    rng = np.random.default_rng(3)
    clips.append((0.05 * rng.standard_normal(int(1.5 * cfg["sr"]))).astype(np.float32))       # noise only: mostly unvoiced
    ex = AudioFeatureExtractor(**cfg)
    ora = OracleFeatureExtractor(**cfg)
    summ, meta, frames, rows = _framewise(ex, clips, groups=("f0", "mfcc", "energy"))
    n = ex.n_mfcc
    total = agree = 0
    for i, y in enumerate(clips):
        yt = ora.preprocess_audio(y)
        f0, voiced, vprob = ora.f0_track(yt)
        T = len(f0)
        assert meta[i, 3] == T and meta[i, 0] == 0
        got = frames[rows[i]:rows[i] + T, 3 * n + 1].astype(np.float64)
        got_v = ~np.isnan(got)
        total += T; agree += int(np.sum(got_v == voiced))
        both = got_v & voiced
        if both.any():
            cents = 1200.0 * np.abs(np.log2(got[both] / f0[both]))
            assert cents.max() <= F0_CENTS, (i, cents.max())
        np.testing.assert_allclose(frames[rows[i]:rows[i] + T, 3 * n + 2], vprob, atol=1e-5)
        ref = ora.extract_f0(yt)
        if np.array_equal(got_v, voiced):
            assert summ[i, 0] == pytest.approx(ref["f0_mean"], rel=1e-5, abs=1e-6)
            assert summ[i, 1] == pytest.approx(ref["f0_std"], rel=1e-4, abs=1e-4)
            assert summ[i, 2] == pytest.approx(ref["f0_missing_rate"], abs=1e-6)
            assert summ[i, 3] == pytest.approx(ref["f0_quality"], abs=1e-6)
        # the other groups are unaffected by the F0 launches
        _check_clip(ex, ora, y, summ[i], meta[i], frames, rows[i], n)
    assert agree / total >= F0_VOICING, (agree, total)


@pytest.mark.parametrize("sr,frame_length,hop_length,half_width", [
    (22050, 1024, 100, 10), (16000, 512, 128, 15), (16000, 512, 160, 20), (22050, 1024, 256, 25), (22050, 1024, 300, 30),
], ids=["hw10", "hw15", "hw20", "hw25", "hw30"])
def test_filtered_viterbi_equals_generic_band_search(torch_cuda, monkeypatch, sr, frame_length, hop_length, half_width):
    """Both band searches of viterbi_kernel are exact restatements of the dense maximum (first index on ties): the filtered
    variant (int32 pre-filter + float64 scan of the surviving chunks, fallback bound, two-pass scan) instantiated for this
    half width must decode the SAME state sequence as the generic one, frame by frame, on tonal, noisy and gliding clips."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    assert 5 * round(35.92 * 12 * hop_length / sr) == half_width
    rng = np.random.default_rng(11)
    clips = [make_clip(900 + s, sr=sr, seconds=sec, f_lo=80.0, f_hi=600.0, kind=k) for s, (sec, k) in
             enumerate([(2.5, "speech"), (1.7, "music"), (3.1, "speech"), (1.2, "music")])]
    clips.append((0.05 * rng.standard_normal(int(1.5 * sr))).astype(np.float32))                       # noise: mostly unvoiced
    glide = np.sin(2 * np.pi * np.cumsum(np.linspace(100.0, 700.0, 2 * sr)) / sr).astype(np.float32) * 0.4  # fast glide: band edges
    clips.append(glide + (0.01 * rng.standard_normal(glide.size)).astype(np.float32))
    kw = dict(sr=sr, frame_length=frame_length, hop_length=hop_length, n_mfcc=13)
    monkeypatch.delenv("AFX_VITERBI", raising=False)
    ex_f = AudioFeatureExtractor(**kw)
    _, meta_f, fr_f, rows = _framewise(ex_f, clips, groups=("f0",))
    monkeypatch.setenv("AFX_VITERBI", "generic")          # read by afx_create
    ex_g = AudioFeatureExtractor(**kw)
    _, meta_g, fr_g, _ = _framewise(ex_g, clips, groups=("f0",))
    n = ex_f.n_mfcc
    assert np.array_equal(meta_f, meta_g) and np.all(meta_f[:, 0] == 0)
    voiced = 0
    for i in range(len(clips)):
        T = int(meta_f[i, 3])
        a, b = fr_f[rows[i]:rows[i] + T, 3 * n + 1], fr_g[rows[i]:rows[i] + T, 3 * n + 1]
        assert np.array_equal(np.isnan(a), np.isnan(b)), i
        assert np.array_equal(a[~np.isnan(a)], b[~np.isnan(b)]), i          # same pitch bin in every voiced frame: same states
        voiced += int(np.sum(~np.isnan(a)))
    assert voiced > 100


@pytest.mark.parametrize("cfg", [dict(), dict(sr=16000, frame_length=512, hop_length=160, n_mfcc=13)], ids=["defaults", "16000-512-160"])
def test_reference_method_api_on_preprocessed_waveforms(torch_cuda, cfg):
    """The reference's public per-stage methods (core/feature_extractor.py:58-179): preprocess_audio(y) returns the
    pre-emphasised + trimmed waveform; extract_f0 / extract_mfcc / extract_energy take THAT waveform (no second
    filter, no second trim) and return the reference's keys.  Also at a hop that takes the run-time-hop kernel and the
    frame mode of the pYIN front end."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor
    ex, ora = AudioFeatureExtractor(**cfg), OracleFeatureExtractor(**cfg)
    sr = cfg.get("sr", 22050)
    for seed in (3, 11):
        y = make_clip(seed, sr=sr, seconds=2.5)
        yt = ex.preprocess_audio(y)
        ref = ora.preprocess_audio(y)
        assert yt.dtype == np.float32 and yt.shape == ref.shape          # trim indices: bit-exact
        np.testing.assert_array_equal(yt, ref.astype(np.float32))         # float32 lfilter arithmetic: bit-exact
        got_m, ref_m = ex.extract_mfcc(yt), ora.extract_mfcc(ref)
        assert list(got_m) == ["mfcc_mean", "mfcc_std", "mfcc_delta_mean", "mfcc_delta2_mean"]
        for k in got_m:
            assert np.max(np.abs(np.asarray(got_m[k]) - np.asarray(ref_m[k]))) <= MFCC_ATOL, k
        got_e, ref_e = ex.extract_energy(yt), ora.extract_energy(ref)
        assert list(got_e) == ["energy_mean", "energy_std", "energy_range"]
        for k in got_e:
            assert got_e[k] == pytest.approx(ref_e[k], rel=ENERGY_RTOL), k
    # F0 on one clip (the pYIN oracle takes seconds)
    y = make_clip(5, sr=sr, seconds=2.0)
    yt = ex.preprocess_audio(y)
    got_f, ref_f = ex.extract_f0(yt), ora.extract_f0(ora.preprocess_audio(y))
    assert list(got_f) == ["f0_mean", "f0_std", "f0_missing_rate", "f0_quality"]
    assert got_f["f0_missing_rate"] == pytest.approx(ref_f["f0_missing_rate"], abs=1e-6)   # float32 summary row
    assert got_f["f0_mean"] == pytest.approx(ref_f["f0_mean"], rel=1e-5)
    assert got_f["f0_std"] == pytest.approx(ref_f["f0_std"], rel=1e-4, abs=1e-3)


def test_preprocess_audio_of_digital_silence_matches_librosa_semantics(torch_cuda):
    """All-zero input: every RMS frame sits at 0 dB re the (zero) maximum, so librosa.effects.trim keeps everything."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from oracle import OracleFeatureExtractor
    ex, ora = AudioFeatureExtractor(), OracleFeatureExtractor()
    y = np.zeros(22050, np.float32)
    got, ref = ex.preprocess_audio(y), ora.preprocess_audio(y)
    assert got.shape == ref.shape
    np.testing.assert_array_equal(got, ref.astype(np.float32))


def test_framewise_npz_feature_set(torch_cuda, tmp_path):
    """The framewise .npz feature set (04_feature_extraction_experiment/feature_extraction.py:191-292,:340-353):
    mfcc 39 x T, f0 / energy / zcr with their deltas, against the oracle on the same clips; then the writer."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from audio_feature_extraction_b200.wav import write_wav
    from oracle import OracleFeatureExtractor
    ex, ora = AudioFeatureExtractor(), OracleFeatureExtractor()
    clips = [make_clip(s, seconds=1.5) for s in (2, 9)]
    got = ex.extract_framewise(clips)
    for y, g in zip(clips, got):
        ref = ora.framewise(y)
        assert g["mfcc"].shape == ref["mfcc"].shape and g["mfcc"].shape[0] == 39          # frame counts: bit-exact
        assert np.max(np.abs(g["mfcc"] - ref["mfcc"])) <= MFCC_ATOL
        np.testing.assert_allclose(g["energy"][0], ref["energy"][0], rtol=ENERGY_RTOL, atol=1e-12)
        np.testing.assert_allclose(g["energy"][1], ref["energy"][1], rtol=1e-4, atol=1e-7)
        np.testing.assert_array_equal(g["zcr"][0], ref["zcr"][0].astype(np.float32))       # integer counts / n_fft: exact
        np.testing.assert_allclose(g["zcr"][1], ref["zcr"][1], rtol=1e-5, atol=1e-7)
        assert g["short_time_energy"].shape == ref["short_time_energy"].shape              # 1 + (L - N) // hop
        np.testing.assert_allclose(g["short_time_energy"], ref["short_time_energy"], rtol=ENERGY_RTOL, atol=1e-12)
        voiced = (ref["f0"][0] > 0)
        assert np.mean((g["f0"][0] > 0) == voiced) >= 0.999                                 # voicing decisions
        both = voiced & (g["f0"][0] > 0)
        cents = 1200 * np.abs(np.log2(g["f0"][0][both] / ref["f0"][0][both]))
        assert cents.size == 0 or cents.max() <= 1.0                                        # F0 within 1 cent
        if np.array_equal(g["f0"][0] > 0, voiced):
            np.testing.assert_allclose(g["f0"][1], ref["f0"][1], rtol=1e-4, atol=1e-3)
    for i, y in enumerate(clips):
        write_wav(str(tmp_path / f"clip{i}.wav"), y, 22050)
    paths = ex.batch_export_npz(str(tmp_path), str(tmp_path / "features"))
    assert [p.split("/")[-1] for p in paths] == ["clip0_processed_features.npz", "clip1_processed_features.npz"]
    z = np.load(paths[0])
    assert sorted(z.files) == ["energy", "f0", "mfcc", "zcr"] and z["mfcc"].shape[0] == 39


def test_batch_process_native_ingest_pcm16_and_mixed(torch_cuda, tmp_path):
    """batch_process through the native threaded WAV reader: a directory of 16-bit mono PCM files takes the raw-PCM
    path (converted on the device), a directory with another encoding the float32 path; both equal extract_features on
    the same file; unreadable files are skipped (core/feature_extractor.py:233-235) and a file at another sample rate
    is resampled on the device (librosa.load's resampling step, :52)."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from audio_feature_extraction_b200.wav import write_wav
    ex = AudioFeatureExtractor()
    d16 = tmp_path / "pcm16"; d16.mkdir()
    for i in range(5):
        write_wav(str(d16 / f"c{i}.wav"), make_clip(40 + i, seconds=1.0 + 0.3 * i), 22050, subtype="PCM_16")
    (d16 / "broken.wav").write_bytes(b"RIFF\x00\x00\x00\x00JUNK")
    write_wav(str(d16 / "rate.wav"), make_clip(1, sr=16000, seconds=1.0), 16000, subtype="PCM_16")
    res = ex.batch_process(str(d16), features_to_extract=["mfcc", "energy"])
    assert sorted(r["file_path"].split("/")[-1] for r in res) == [f"c{i}.wav" for i in range(5)] + ["rate.wav"]
    for r in res:
        one = ex.extract_features(r["file_path"], features_to_extract=["mfcc", "energy"])
        np.testing.assert_allclose(r["mfcc_mean"], one["mfcc_mean"], atol=1e-5)
        assert r["energy_mean"] == pytest.approx(one["energy_mean"], rel=1e-6)
    dmix = tmp_path / "mixed"; dmix.mkdir()
    write_wav(str(dmix / "a.wav"), make_clip(50, seconds=1.2), 22050, subtype="PCM_16")
    write_wav(str(dmix / "b.wav"), make_clip(51, seconds=1.4), 22050, subtype="FLOAT")
    res = ex.batch_process(str(dmix))
    assert len(res) == 2 and all("f0_mean" in r and "mfcc_mean" in r for r in res)


def test_f0_postprocess_matches_scipy(torch_cuda):
    """F0 post-processing (README.md:12-15; 04_feature_extraction_experiment/feature_extractor.py:283-301): medfilt(5) ->
    savgol_filter(11, 3) -> clip on the voiced frames, against scipy through the oracle.  First on the device's own
    pYIN track, then on planted tracks with 0 .. 12 and many voiced frames (every branch of the recipe)."""
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor
    ex, ora = AudioFeatureExtractor(), OracleFeatureExtractor()
    clips = [make_clip(s, seconds=1.0 + 0.25 * s) for s in range(16)]
    got = ex.extract_f0_smoothed(clips)
    for g in got:
        ref, st = ora.f0_postprocess(g["f0"], g["voiced_flag"])
        np.testing.assert_allclose(g["f0_smoothed"][g["voiced_flag"]], ref, rtol=2e-7, atol=0)
        assert np.all(np.isnan(g["f0_smoothed"][~g["voiced_flag"]]))
        np.testing.assert_allclose([g["f0_mean"], g["f0_global_rmse"], g["f0_local_rmse"], g["f0_stability"]], st, rtol=1e-6)
    # planted tracks: overwrite the F0 column of the frame matrix, then post-process again
    buf, offsets, lengths = ex.pack_clips(clips)
    dev = torch.device("cuda", ex.device)
    audio = buf.to(dev)
    _, meta, frames = ex.extract_batch_device(audio, offsets, lengths, ("f0",), return_frames=True)
    m = meta.cpu().numpy()
    rows = np.concatenate([[0], np.cumsum([ex.max_frames(int(l)) for l in lengths])])
    rng = np.random.default_rng(5)
    fr = frames.cpu().numpy().copy()
    c_f0 = 3 * ex.n_mfcc + 1
    planted = []
    for i, nv in enumerate([0, 1, 2, 3, 4, 5, 6, 10, 11, 12, 13, 21, 30, 40, None, None]):
        T = int(m[i, 3])
        nv = T if nv is None else nv
        f0 = np.full(T, np.nan, dtype=np.float32)
        idx = np.sort(rng.choice(T, size=nv, replace=False))
        walk = 220.0 * 2.0 ** np.cumsum(rng.normal(0, 0.02, size=nv))
        walk[rng.random(nv) < 0.1] *= 2.0                          # octave jumps for the median filter
        if i == 15:
            walk *= 12.0                                           # beyond f0_max: the range clamp
        f0[idx] = walk.astype(np.float32)
        fr[rows[i]:rows[i] + T, c_f0] = f0
        planted.append(f0)
    frames.copy_(torch.from_numpy(fr))
    track = torch.full((frames.shape[0],), float("nan"), dtype=torch.float32, device=dev)
    stats = torch.zeros((len(clips), 4), dtype=torch.float32, device=dev)
    h_len = np.ascontiguousarray(lengths, dtype=np.int32)
    N.check(N.lib().afx_f0_postprocess(ex._h(), len(clips), h_len.ctypes.data, meta.data_ptr(), frames.data_ptr(),
                                       track.data_ptr(), stats.data_ptr(), ex._ws.data_ptr(),
                                       torch.cuda.current_stream(dev).cuda_stream))
    tr, st = track.cpu().numpy(), stats.cpu().numpy()
    for i, f0 in enumerate(planted):
        T = f0.size
        voiced = ~np.isnan(f0)
        ref, rst = ora.f0_postprocess(f0, voiced)
        g = tr[rows[i]:rows[i] + T]
        np.testing.assert_allclose(g[voiced], ref, rtol=2e-7, atol=0)
        assert np.all(np.isnan(g[~voiced]))
        np.testing.assert_allclose(st[i], np.asarray(rst, dtype=np.float64), rtol=1e-6, equal_nan=True)
    assert tr[rows[15]:rows[15] + planted[15].size].max() <= np.float32(ex.f0_max)


@pytest.mark.gpu
@pytest.mark.parametrize("native_sr,target_sr", [(16000, 22050), (44100, 22050), (48000, 22050), (8000, 22050),
                                                 (11025, 22050), (22050, 16000), (32000, 22050)])
def test_resample_matches_scipy_resample_poly(torch_cuda, native_sr, target_sr):
    """afx_resample against scipy.signal.resample_poly (= librosa.resample(res_type='polyphase'), the restatable
    stand-in for librosa.load's soxr path, core/feature_extractor.py:52): same filter, same float32 operation order, so
    the comparison is bit for bit; ragged lengths down to 0 and 1 samples."""
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import librosa_restated as LR
    ex = AudioFeatureExtractor(sr=target_sr)
    rng = np.random.default_rng(native_sr)
    clips = [make_clip(90 + i, sr=native_sr, seconds=0.4 + 0.35 * i) for i in range(4)]
    clips += [rng.standard_normal(n).astype(np.float32) * 0.2 for n in (0, 1, 2, 37, 1000)]
    buf, offsets, lengths = ex.pack_clips(clips)
    out, ooff, olen = ex.resample_batch_device(buf.to(torch.device("cuda", ex.device)), offsets, lengths, native_sr)
    out = out.cpu().numpy()
    for y, o, l in zip(clips, ooff, olen):
        ref = LR.resample_polyphase(y, native_sr, target_sr)
        assert int(l) == ref.shape[0] == int(np.ceil(y.shape[0] * target_sr / native_sr))     # lengths: bit-exact
        np.testing.assert_array_equal(out[o:o + l], ref)


@pytest.mark.gpu
def test_load_audio_and_batch_process_resample_other_rates(torch_cuda, tmp_path):
    """load_audio returns (y at self.sr, self.sr) for a file at another rate (reference :52 `librosa.load(path,
    sr=self.sr)`), and batch_process on a mixed-rate directory equals the oracle on the resampled waveforms."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from audio_feature_extraction_b200.wav import write_wav
    from oracle import OracleFeatureExtractor, librosa_restated as LR
    ex, ora = AudioFeatureExtractor(), OracleFeatureExtractor()
    specs = [("a.wav", 16000, "FLOAT"), ("b.wav", 22050, "PCM_16"), ("c.wav", 44100, "FLOAT"), ("d.wav", 48000, "PCM_16")]
    waves = {}
    for i, (name, sr, sub) in enumerate(specs):
        y = make_clip(300 + i, sr=sr, seconds=1.5)
        write_wav(str(tmp_path / name), y, sr, subtype=sub)
        if sub == "PCM_16":
            y = (np.clip(np.round(y * 32768.0), -32768, 32767) / 32768.0).astype(np.float32)
        waves[name] = LR.resample_polyphase(y, sr, 22050)
    y, sr = ex.load_audio(str(tmp_path / "a.wav"))
    assert sr == 22050
    np.testing.assert_array_equal(y, waves["a.wav"])
    res = ex.batch_process(str(tmp_path), features_to_extract=["mfcc", "energy"])
    assert len(res) == 4
    for r in res:
        ref = ora.extract_features_array(waves[r["file_path"].split("/")[-1]], features_to_extract=("mfcc", "energy"))
        assert np.max(np.abs(np.asarray(r["mfcc_mean"]) - np.asarray(ref["mfcc_mean"]))) <= MFCC_ATOL
        assert r["energy_mean"] == pytest.approx(ref["energy_mean"], rel=ENERGY_RTOL)


def test_full_size_c2_batch_properties(torch_cuda):
    """BASELINE.json configs[1] at its full size (10 000 clips x 10 s @ 22.05 kHz, 8.82 GB resident) through properties
    that do not need the oracle at that size: the batch is 625 copies of 16 distinct clips, so (1) every copy must give
    bit-identical rows wherever it sits in the batch and whichever CTA picks it up (a checksum of checksums), (2) the
    rows equal those of the 16 clips extracted alone, which (3) match the oracle; (4) scaling the audio by 1/2 (exact in
    binary floating point) halves the energy rows bit for bit, leaves the trim decisions and frame counts alone and
    shifts MFCC 0 by sqrt(128) * 10 log10(1/4) with every other coefficient unchanged."""
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor
    ex = AudioFeatureExtractor()
    n_base, copies, L = 16, 625, 220500
    base = [make_clip(1000 + s, seconds=10.0) for s in range(n_base)]
    assert all(b.shape[0] == L for b in base)
    dev = torch.device("cuda", ex.device)
    base_dev = torch.from_numpy(np.concatenate(base)).to(dev)
    audio = torch.empty(n_base * copies * L + 8, dtype=torch.float32, device=dev)
    audio[:n_base * copies * L].view(copies, n_base * L).copy_(base_dev.unsqueeze(0).expand(copies, -1))
    n = n_base * copies
    offsets = np.arange(n, dtype=np.int64) * L
    lengths = np.full(n, L, dtype=np.int32)
    groups = ("mfcc", "energy")
    summ, meta, _ = ex.extract_batch_device(audio, offsets, lengths, groups)
    torch.cuda.synchronize()
    assert int(meta[:, 0].abs().max()) == 0
    s3, m3 = summ.view(copies, n_base, -1), meta.view(copies, n_base, 4)
    assert torch.equal(s3, s3[:1].expand_as(s3)) and torch.equal(m3, m3[:1].expand_as(m3))          # (1)
    small_s, small_m, _ = ex.extract_batch_device(base_dev, offsets[:n_base], lengths[:n_base], groups)
    assert torch.equal(small_s, s3[0]) and torch.equal(small_m, m3[0])                               # (2)
    T = m3[0, :, 3].cpu().numpy()
    trimmed = (m3[0, :, 2] - m3[0, :, 1]).cpu().numpy()
    np.testing.assert_array_equal(T, 1 + trimmed // ex.hop_length)                                   # frame counts
    ora = OracleFeatureExtractor()
    rows = s3[0].cpu().numpy()
    nm = ex.n_mfcc
    for i in (0, 9):                                                                                 # (3)
        ref = ora.extract_features_array(base[i], features_to_extract=groups)
        assert np.max(np.abs(rows[i, 4:4 + nm] - np.asarray(ref["mfcc_mean"]))) <= MFCC_ATOL
        assert np.max(np.abs(rows[i, 4 + nm:4 + 2 * nm] - np.asarray(ref["mfcc_std"]))) <= MFCC_ATOL
        assert rows[i, 4 + 4 * nm] == pytest.approx(ref["energy_mean"], rel=ENERGY_RTOL)
    audio.mul_(0.5)                                                                                  # (4)
    summ2, meta2, _ = ex.extract_batch_device(audio, offsets, lengths, groups)
    torch.cuda.synchronize()
    assert torch.equal(meta2, meta)
    e0 = 4 + 4 * nm
    assert torch.equal(summ2[:, e0:e0 + 3], summ[:, e0:e0 + 3] * 0.5)
    shift = float(np.sqrt(128.0) * 10.0 * np.log10(0.25))
    d = (summ2[:, 4:4 + nm] - summ[:, 4:4 + nm]).cpu().numpy()
    assert np.max(np.abs(d[:, 0] - shift)) <= MFCC_ATOL and np.max(np.abs(d[:, 1:])) <= MFCC_ATOL
    dstd = (summ2[:, 4 + nm:4 + 4 * nm] - summ[:, 4 + nm:4 + 4 * nm]).abs().max().item()
    assert dstd <= MFCC_ATOL


def test_full_size_c5_shard_past_2_32_samples(torch_cuda):
    """BASELINE.json configs[4]: 1000 audio-hours = 360 000 clips of C2 sharded by clip over 8 GPUs, i.e. 45 000 clips
    (39.7 GB, 9.9e9 samples: more than 2^32, so every 32-bit sample index on the path would show) resident on ONE GPU.
    The shard is 2 812 copies of 16 distinct clips + 8: every copy must give bit-identical rows wherever it sits, and the
    rows of the clips extracted alone."""
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    ex = AudioFeatureExtractor()
    n_base, n, L = 16, 45000, 220500
    dev = torch.device("cuda", ex.device)
    free, _ = torch.cuda.mem_get_info(dev)
    if free < 48 * 2 ** 30:
        pytest.skip("needs 48 GB of free device memory")
    base_dev = torch.from_numpy(np.concatenate([make_clip(1000 + s, seconds=10.0) for s in range(n_base)])).to(dev)
    assert base_dev.numel() == n_base * L
    audio = torch.empty(n * L + 8, dtype=torch.float32, device=dev)
    copies, rest = divmod(n, n_base)
    audio[:copies * n_base * L].view(copies, n_base * L).copy_(base_dev.unsqueeze(0).expand(copies, -1))
    audio[copies * n_base * L:n * L].copy_(base_dev[:rest * L])
    offsets = np.arange(n, dtype=np.int64) * L
    lengths = np.full(n, L, dtype=np.int32)
    groups = ("mfcc", "energy")
    summ, meta, _ = ex.extract_batch_device(audio, offsets, lengths, groups)
    torch.cuda.synchronize()
    assert int(meta[:, 0].abs().max()) == 0
    small_s, small_m, _ = ex.extract_batch_device(base_dev, offsets[:n_base], lengths[:n_base], groups)
    idx = torch.arange(n, device=dev) % n_base
    assert torch.equal(summ, small_s[idx]) and torch.equal(meta, small_m[idx])
    del audio
    torch.cuda.empty_cache()


@pytest.mark.parametrize("cfg,n_clips,kind,f_lo,f_hi", [
    (dict(sr=16000, frame_length=512, hop_length=128, n_mfcc=40, extended_stats=True), 4096, "speech", 80.0, 400.0),  # C3
    (dict(sr=44100, frame_length=2048, hop_length=512, n_mfcc=13), 2048, "music", 110.0, 880.0),                      # C4
], ids=["C3-full", "C4-full"])
def test_full_size_c3_c4_all_features_batch_invariance(torch_cuda, cfg, n_clips, kind, f_lo, f_hi):
    """BASELINE.json configs[2] / configs[3] at SURVEY.md section 8(d)'s full sizes (4 096 x 10 s @ 16 kHz with the
    skew / kurtosis extension and F0; 2 048 x 10 s @ 44.1 kHz, all features): the batch repeats 8 distinct clips, every
    copy must produce bit-identical rows (pYIN, its chunked scratch and the Viterbi included) and those rows must equal
    the 8 clips extracted alone; frame counts follow 1 + L / hop of the trimmed length."""
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    ex = AudioFeatureExtractor(**cfg)
    n_base = 8
    copies = n_clips // n_base
    base = [make_clip(2000 + s, sr=cfg["sr"], seconds=10.0, f_lo=f_lo, f_hi=f_hi, kind=kind) for s in range(n_base)]
    L = base[0].shape[0]
    assert L % 4 == 0 and all(b.shape[0] == L for b in base)
    dev = torch.device("cuda", ex.device)
    base_dev = torch.from_numpy(np.concatenate(base)).to(dev)
    audio = torch.empty(n_clips * L + 8, dtype=torch.float32, device=dev)
    audio[:n_clips * L].view(copies, n_base * L).copy_(base_dev.unsqueeze(0).expand(copies, -1))
    offsets = np.arange(n_clips, dtype=np.int64) * L
    lengths = np.full(n_clips, L, dtype=np.int32)
    summ, meta, _ = ex.extract_batch_device(audio, offsets, lengths, None)
    torch.cuda.synchronize()
    assert int(meta[:, 0].abs().max()) == 0
    s3, m3 = summ.view(copies, n_base, -1), meta.view(copies, n_base, 4)
    same = (s3 == s3[:1]) | (torch.isnan(s3) & torch.isnan(s3[:1]))
    assert bool(same.all()) and torch.equal(m3, m3[:1].expand_as(m3))
    small_s, small_m, _ = ex.extract_batch_device(base_dev, offsets[:n_base], lengths[:n_base], None)
    assert torch.equal(small_m, m3[0])
    assert bool(((small_s == s3[0]) | (torch.isnan(small_s) & torch.isnan(s3[0]))).all())
    T = m3[0, :, 3].cpu().numpy()
    np.testing.assert_array_equal(T, 1 + (m3[0, :, 2] - m3[0, :, 1]).cpu().numpy() // ex.hop_length)
    f0 = s3[0, :, 0].cpu().numpy()
    # mean F0 inside [f0_min, f0_max], or 0 for a clip without a voiced frame (reference :103-107; the low-pitched
    # speech-like clips at frame_length 512 and the three-note music clips are all-unvoiced for the oracle as well)
    assert np.all(((f0 >= ex.f0_min) & (f0 <= ex.f0_max)) | (f0 == 0))


@pytest.mark.parametrize("cfg,kind,f_lo,f_hi", [
    (dict(sr=22050, frame_length=1024, hop_length=256, n_mfcc=13), "speech", 110.0, 880.0),    # C1 / C2
    (dict(sr=16000, frame_length=512, hop_length=128, n_mfcc=40), "speech", 80.0, 400.0),      # C3
    (dict(sr=44100, frame_length=2048, hop_length=512, n_mfcc=13), "music", 110.0, 880.0),     # C4
], ids=["C2", "C3", "C4"])
def test_parity_set_of_64_seeded_clips(torch_cuda, cfg, kind, f_lo, f_hi):
    """SURVEY.md section 8(d)'s parity set: 64 clips per configuration, seeds 0 .. 63 (lengths 0.6 .. 5 s), compared with
    the oracle stage by stage - trim indices and frame counts bit-exact, framewise MFCC / delta / delta2 within 1e-3,
    RMS within 1e-5 relative, and the summary row."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from oracle import OracleFeatureExtractor
    ex, ora = AudioFeatureExtractor(**cfg), OracleFeatureExtractor(**cfg)
    rng = np.random.default_rng(64)
    secs = np.round(rng.uniform(0.6, 5.0, size=64), 3)
    clips = [make_clip(s, sr=cfg["sr"], seconds=float(sec), f_lo=f_lo, f_hi=f_hi, kind=kind) for s, sec in enumerate(secs)]
    summ, meta, frames, rows = _framewise(ex, clips)
    for i, y in enumerate(clips):
        _check_clip(ex, ora, y, summ[i], meta[i], frames, rows[i], ex.n_mfcc)


def test_pyin_parity_sweep_over_seeded_clips(torch_cuda):
    """tools/f0_parity_sweep.py with 20 + 4 clips per configuration (C2, C3, C4 geometry; the oracle tracks are computed
    on all host cores): BASELINE.json's F0 bar over a few thousand frames per configuration - voicing decisions equal on
    >= 99.9 % of the frames, F0 within 1 cent wherever both are voiced.  The 64-clip run of the same tool is stored in
    profiles/r01h_f0_parity_sweep.jsonl (100 % agreement, 1e-4 cents)."""
    import importlib.util
    from pathlib import Path
    spec = importlib.util.spec_from_file_location("f0_parity_sweep", Path(__file__).resolve().parents[1] / "tools" / "f0_parity_sweep.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    for res in mod.sweep(20):
        assert res["voicing_agreement"] >= F0_VOICING, res
        assert res["worst_cents"] <= F0_CENTS, res
        assert res["frames"] > 1500


def _sharded_worker(rank, world, port, backend, wav_dir, out_path):
    import os
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    ndev = torch.cuda.device_count()
    dev = rank % ndev
    torch.cuda.set_device(dev)
    kw = dict(device_id=torch.device("cuda", dev)) if backend == "nccl" else {}
    dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    try:
        from audio_feature_extraction_b200 import AudioFeatureExtractor
        ex = AudioFeatureExtractor(device=dev)
        res = ex.batch_process_sharded(wav_dir)
        if rank == 0:
            torch.save(res, out_path)
        else:
            assert res is None
    finally:
        dist.destroy_process_group()


def test_batch_process_sharded_two_ranks_matches_single_rank(torch_cuda, tmp_path):
    """BASELINE.json configs[4]: batch_process sharded by clip over the ranks of a torch.distributed job, per-file rows
    gathered to rank 0 (NCCL when the box has two GPUs; on a one-GPU box both ranks share the GPU and gather over gloo)."""
    import socket
    import torch
    import torch.multiprocessing as mp
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import make_clip
    from audio_feature_extraction_b200.wav import write_wav

    d = tmp_path / "wavs"; d.mkdir()
    for i in range(9):
        write_wav(str(d / f"clip_{i:02d}.wav"), make_clip(900 + i, seconds=0.8 + 0.3 * i), 22050)
    (d / "broken.wav").write_bytes(b"RIFFxxxxWAVEjunk")          # logged and skipped (reference :233-235)
    write_wav(str(d / "other_rate.wav"), make_clip(950, sr=16000, seconds=1.5), 16000)   # resampled on the device
    single = {r["file_path"]: r for r in AudioFeatureExtractor().batch_process(str(d))}
    backend = "nccl" if torch.cuda.device_count() >= 2 else "gloo"
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = tmp_path / "rank0.pt"
    mp.spawn(_sharded_worker, args=(2, port, backend, str(d), str(out)), nprocs=2, join=True)
    sharded = torch.load(out, weights_only=False)
    assert [r["file_path"] for r in sharded] == sorted(single)          # enumeration order of the sorted list, broken file skipped
    for r in sharded:
        ref = single[r["file_path"]]
        assert r.keys() == ref.keys()
        for k, v in r.items():
            if k != "file_path":
                np.testing.assert_allclose(np.asarray(v, dtype=np.float64), np.asarray(ref[k], dtype=np.float64), rtol=0, atol=0)
