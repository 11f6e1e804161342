"""Golden vectors under tests/golden/ (made by tools/make_golden.py from the oracle: PARITY UNPINNED, see its header).
CPU: the oracle still reproduces them.  GPU (-m gpu): the CUDA path against the frozen vectors, no oracle involved."""
from pathlib import Path

import numpy as np
import pytest

GOLDEN = Path(__file__).resolve().parent / "golden"
NAMES = ["c2", "c3", "c4"]


def _load(name):
    z = np.load(GOLDEN / f"afx_golden_{name}.npz")
    sr, n_fft, hop, n_mfcc = (int(v) for v in z["cfg"])
    return z, dict(sr=sr, frame_length=n_fft, hop_length=hop, n_mfcc=n_mfcc)


@pytest.mark.parametrize("name", NAMES)
def test_oracle_reproduces_golden(name):
    from oracle import OracleFeatureExtractor
    z, cfg = _load(name)
    ora = OracleFeatureExtractor(extended_stats=True, **cfg)
    for i in range(2):
        y = z[f"clip{i}"]
        s, e = ora.trim_indices(y)
        yt = ora.preprocess_audio(y)
        m, d1, d2 = ora.mfcc_frames(yt)
        assert [s, e, m.shape[1]] == z[f"trim{i}"].tolist()
        np.testing.assert_allclose(m, z[f"mfcc{i}"], atol=2e-4)
        np.testing.assert_allclose(d1, z[f"delta{i}"], atol=2e-4)
        np.testing.assert_allclose(ora.rms_frames(yt)[0], z[f"rms{i}"], rtol=1e-6)
        if i == 0:   # the pYIN restatement takes seconds per clip: one clip per config on CPU
            f0, voiced, vprob = ora.f0_track(yt)
            g = z[f"f0{i}"]
            assert np.array_equal(np.isnan(f0), np.isnan(g))
            np.testing.assert_allclose(f0[~np.isnan(g)], g[~np.isnan(g)], rtol=1e-12)
            np.testing.assert_allclose(vprob, z[f"vprob{i}"], atol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_cuda_path_matches_golden(name, built_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.fail("-m gpu tests need a CUDA device (there is no CPU fallback)")
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    z, cfg = _load(name)
    ex = AudioFeatureExtractor(extended_stats=True, **cfg)
    clips = [z["clip0"], z["clip1"]]
    buf, offsets, lengths = ex.pack_clips(clips)
    summ, meta, frames = ex.extract_batch_device(buf.cuda(), offsets, lengths, None, return_frames=True)
    torch.cuda.synchronize()
    summ, meta, frames = summ.cpu().numpy(), meta.cpu().numpy(), frames.cpu().numpy()
    n = cfg["n_mfcc"]
    row0 = 0
    dicts = ex.rows_to_dicts(summ, meta, ["a", "b"])
    for i in range(2):
        s, e, T = z[f"trim{i}"].tolist()
        assert meta[i].tolist() == [0, s, e, T]                                   # indexing: bit-exact
        fr = frames[row0:row0 + T]
        assert np.max(np.abs(fr[:, :n].T - z[f"mfcc{i}"])) <= 1e-3                 # MFCC 1e-3 abs after the dB step
        assert np.max(np.abs(fr[:, n:2 * n].T - z[f"delta{i}"])) <= 1e-3
        assert np.max(np.abs(fr[:, 2 * n:3 * n].T - z[f"delta2{i}"])) <= 1e-3
        np.testing.assert_allclose(fr[:, 3 * n], z[f"rms{i}"], rtol=1e-5)          # energy 1e-5 rel
        g = z[f"f0{i}"]
        got = fr[:, 3 * n + 1].astype(np.float64)
        assert np.mean(np.isnan(got) == np.isnan(g)) >= 0.999 or np.sum(np.isnan(got) != np.isnan(g)) <= 1
        both = ~np.isnan(got) & ~np.isnan(g)
        if both.any():
            assert np.max(1200 * np.abs(np.log2(got[both] / g[both]))) <= 1.0      # F0 within 1 cent
        keys = [str(k) for k in z[f"summary_keys{i}"]]
        flat = np.concatenate([np.atleast_1d(np.asarray(dicts[i][k], np.float64)) for k in keys])
        ref = z[f"summary{i}"]
        if np.array_equal(np.isnan(got), np.isnan(g)):
            np.testing.assert_allclose(flat[:4], ref[:4], rtol=1e-4, atol=1e-4)    # f0_mean/std/missing/quality
        is_shape = np.zeros(len(ref), bool)   # skew / kurtosis keys: looser (ratios of central moments)
        pos = 0
        for k in keys:
            w = np.atleast_1d(np.asarray(dicts[i][k])).size
            if "skew" in k or "kurt" in k:
                is_shape[pos:pos + w] = True
            pos += w
        np.testing.assert_allclose(flat[4:][~is_shape[4:]], ref[4:][~is_shape[4:]], atol=1e-3, rtol=1e-5)
        np.testing.assert_allclose(flat[is_shape], ref[is_shape], atol=1e-2)
        row0 += ex.max_frames(len(clips[i]))
