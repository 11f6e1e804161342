"""GPU test of afx_extract_static (include/afx_b200.h): the fused MFCC + delta + energy pass with its scratch in a
file-scope static device array, for callers that cannot allocate.  The entry point has no counterpart in the reference;
its oracle is the existing entry point afx_extract on the same inputs.

Collected last on purpose (file name): a failure here must not hide the other tests.
Passes alone in a fresh process and after the other tests (it shares nothing with them but the library).
Needs a B200: run with `-m gpu`."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MFCC_ATOL = 1e-3      # BASELINE.json north_star tolerances, as in tests/test_gpu_parity.py
ENERGY_RTOL = 1e-5


def _call_static(L, N, ex, audio, offsets, lengths, mask, torch):
    """afx_extract_static through the C ABI: outputs are allocated HERE (by the caller), the call itself allocates nothing"""
    n = len(lengths)
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    lengths = np.ascontiguousarray(lengths, dtype=np.int32)
    summ = torch.full((n, ex.summary_width), float("nan"), dtype=torch.float32, device=audio.device)
    meta = torch.full((n, 4), -7, dtype=torch.int32, device=audio.device)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = L.afx_extract_static(ex._h(), audio.data_ptr(), offsets.ctypes.data, lengths.ctypes.data, n, mask,
                              summ.data_ptr(), meta.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.afx_last_error().decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before          # nothing went through torch's allocator either
    return summ.cpu().numpy(), meta.cpu().numpy()


def _check(ex, got_s, got_m, ref_s, ref_m):
    n = ex.n_mfcc
    assert np.array_equal(got_m, ref_m)                      # status, trim indices, frame counts: bit-exact
    ok = ref_m[:, 0] == 0
    assert ok.any()
    e0 = 4 + 4 * n
    assert np.max(np.abs(got_s[ok, 4:e0] - ref_s[ok, 4:e0])) <= MFCC_ATOL
    np.testing.assert_allclose(got_s[ok, e0:e0 + 3], ref_s[ok, e0:e0 + 3], rtol=ENERGY_RTOL)
    assert np.max(np.abs(got_s[ok, e0 + 3:] - ref_s[ok, e0 + 3:])) <= 5e-3      # skew / kurtosis extension
    assert (got_s[~ok] == 0).all()


@pytest.mark.parametrize("cfg", [
    dict(sr=22050, frame_length=1024, hop_length=256, n_mfcc=13),      # speculative single pass
    dict(sr=16000, frame_length=512, hop_length=160, n_mfcc=13),       # run-time hop: the trim kernel runs first
], ids=["1024-256", "512-160"])
def test_static_scratch_entry_matches_afx_extract_on_two_calls_in_a_row(torch_cuda, cfg):
    torch = torch_cuda
    from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
    from audio_feature_extraction_b200.synth import make_clip
    L = N.lib()
    assert L.afx_static_workspace_capacity() == 8 << 20
    ex = AudioFeatureExtractor(**cfg)
    sr = cfg["sr"]
    mask = N.AFX_FEAT_MFCC | N.AFX_FEAT_ENERGY
    # call 1: a ragged batch incl. an empty clip, a clip with a silent tail (fix-ups) and a clamp-heavy clip (redo pass);
    # call 2: fewer, different clips -- the array still holds call 1's items, counters and slabs
    gap = np.zeros(6000, np.float32)
    a, b = make_clip(40, sr=sr, seconds=1.5), make_clip(41, sr=sr, seconds=2.0)
    batches = [
        [make_clip(30 + i, sr=sr, seconds=0.7 + 0.45 * i) for i in range(7)]
        + [np.zeros(0, np.float32), np.concatenate([a, np.zeros(20000, np.float32)]), np.concatenate([a, gap, b])],
        [make_clip(50, sr=sr, seconds=3.1), make_clip(51, sr=sr, seconds=0.9), np.concatenate([b, gap, a])],
    ]
    for clips in batches:
        buf, offsets, lengths = ex.pack_clips(clips)
        audio = buf.cuda()
        ref_s, ref_m, _ = ex.extract_batch_device(audio, offsets, lengths, ("mfcc", "energy"))
        torch.cuda.synchronize()
        got_s, got_m = _call_static(L, N, ex, audio, offsets, lengths, mask, torch)
        _check(ex, got_s, got_m, ref_s.cpu().numpy(), ref_m.cpu().numpy())


def test_static_scratch_entry_rejects_what_does_not_fit(torch_cuda):
    torch = torch_cuda
    from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
    L = N.lib()
    ex = AudioFeatureExtractor()
    h = ex._h()
    audio = torch.zeros(1 << 16, dtype=torch.float32, device="cuda")
    out_s = torch.zeros((300, ex.summary_width), dtype=torch.float32, device="cuda")
    out_m = torch.zeros((300, 4), dtype=torch.int32, device="cuda")

    def call(n, length, mask=N.AFX_FEAT_MFCC | N.AFX_FEAT_ENERGY):
        off = np.zeros(n, np.int64); ln = np.full(n, length, np.int32)
        return L.afx_extract_static(h, audio.data_ptr(), off.ctypes.data, ln.ctypes.data, n, mask, out_s.data_ptr(), out_m.data_ptr(), None)

    assert call(257, 1000) == N.ERR_UNSUPPORTED and b"maximum of 256" in L.afx_last_error()          # too many clips
    assert call(200, 30 * 22050) == N.ERR_WORKSPACE and b"static array holds" in L.afx_last_error()  # scratch beyond 8 MiB
    assert call(4, 1000, N.AFX_FEAT_ALL) == N.ERR_UNSUPPORTED and b"AFX_FEAT_MFCC" in L.afx_last_error()   # no F0 group
    assert call(0, 0) == 0
    torch.cuda.synchronize()
