"""Host ingest (afx_wav_probe_batch / afx_wav_load_batch): the native threaded WAV reader against the Python reader
audio_feature_extraction_b200.wav.read_wav on files of every supported encoding.  Pure host code: runs without a GPU."""
import struct

import numpy as np
import pytest


def _write(path, pcm_bytes, tag, ch, sr, bits, extensible=False, odd_chunk=False):
    block = ch * bits // 8
    if extensible:
        fmt = struct.pack("<HHIIHH", 0xFFFE, ch, sr, sr * block, block, bits) + struct.pack("<HHI", 22, bits, 0) + \
              struct.pack("<H", tag) + b"\x00\x00\x00\x00\x10\x00\x80\x00\x00\xaa\x00\x38\x9b\x71"
    else:
        fmt = struct.pack("<HHIIHH", tag, ch, sr, sr * block, block, bits)
    chunks = b"fmt " + struct.pack("<I", len(fmt)) + fmt
    if odd_chunk:
        chunks += b"LIST" + struct.pack("<I", 3) + b"abc" + b"\x00"      # odd-sized chunk + pad byte
    chunks += b"data" + struct.pack("<I", len(pcm_bytes)) + pcm_bytes
    with open(path, "wb") as f:
        f.write(b"RIFF" + struct.pack("<I", 4 + len(chunks)) + b"WAVE" + chunks)


@pytest.fixture()
def wav_files(tmp_path):
    rng = np.random.default_rng(7)
    n = 3001
    x = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
    files = {}
    def add(name, *a, **k):
        _write(str(tmp_path / name), *a, **k); files[name] = str(tmp_path / name)
    add("pcm16_mono.wav", (x[:, 0] * 32767).astype("<i2").tobytes(), 1, 1, 22050, 16)
    add("pcm16_stereo.wav", (x * 32767).astype("<i2").tobytes(), 1, 2, 22050, 16, odd_chunk=True)
    add("pcm8.wav", ((x[:, 0] * 127) + 128).astype(np.uint8).tobytes(), 1, 1, 8000, 8)
    v24 = (x[:, 0] * 8388607).astype(np.int32)
    add("pcm24.wav", b"".join(struct.pack("<i", int(v))[:3] for v in v24), 1, 1, 44100, 24)
    add("pcm32_ext.wav", (x[:, 0].astype(np.float64) * 2147483647).astype("<i4").tobytes(), 1, 1, 22050, 32, extensible=True)
    add("f32_stereo.wav", x.astype("<f4").tobytes(), 3, 2, 16000, 32)
    add("f64.wav", x[:, 1].astype("<f8").tobytes(), 3, 1, 22050, 64)
    (tmp_path / "broken.wav").write_bytes(b"RIFFxxxxJUNK")
    files["broken.wav"] = str(tmp_path / "broken.wav")
    files["missing.wav"] = str(tmp_path / "missing.wav")
    return files


def test_probe_and_load_match_the_python_reader(built_lib, wav_files):
    from audio_feature_extraction_b200.wav import read_wav, probe_batch, load_batch
    names = sorted(wav_files)
    paths = [wav_files[n] for n in names]
    info = probe_batch(paths, threads=4)
    good = [i for i, n in enumerate(names) if n not in ("broken.wav", "missing.wav")]
    assert info["status"][names.index("broken.wav")] == 2 and info["status"][names.index("missing.wav")] == 1
    assert all(info["status"][i] == 0 for i in good)
    frames = info["frames"].copy()
    frames[[names.index("broken.wav"), names.index("missing.wav")]] = 0
    offsets = np.concatenate([[0], np.cumsum((frames + 7) // 8 * 8)[:-1]]).astype(np.int64)
    out = np.full(int(offsets[-1] + frames[-1] + 8), np.nan, np.float32)
    st = load_batch(paths, frames, out, offsets, threads=3)
    for i in good:
        y, sr = read_wav(paths[i])
        assert sr == info["sr"][i] and len(y) == frames[i] == 3001
        np.testing.assert_array_equal(out[offsets[i]:offsets[i] + frames[i]], y)       # bit-exact decode and mono mix
        assert st[i] == 0
    assert st[names.index("broken.wav")] != 0 and st[names.index("missing.wav")] != 0


def test_raw_pcm16_delivery_and_its_guard(built_lib, wav_files):
    from audio_feature_extraction_b200.wav import read_wav, load_batch
    paths = [wav_files["pcm16_mono.wav"], wav_files["pcm16_stereo.wav"]]
    out = np.zeros(2 * 3008, np.int16)
    st = load_batch(paths, [3001, 3001], out, [0, 3008], as_pcm16=True)
    assert st.tolist() == [0, 6]                                                           # stereo cannot be delivered raw
    y, _ = read_wav(paths[0])
    np.testing.assert_array_equal(out[:3001].astype(np.float32) / 32768.0, y)
    st = load_batch(paths[:1], [3000], np.zeros(3008, np.float32), [0])
    assert st.tolist() == [5]                                                              # length mismatch is reported


@pytest.mark.parametrize("sr,native", [(22050, 16000), (22050, 44100), (22050, 48000), (22050, 8000), (16000, 22050), (22050, 11025)])
def test_resample_filter_design_equals_scipy_firwin(built_lib, sr, native):
    """The host half of afx_resample (no GPU needed): the phase-ordered Kaiser-sinc filter equals, bit for bit, what
    scipy.signal.resample_poly builds (firwin(2 * 10 * max(up, down) + 1, 1 / max(up, down), window=('kaiser', 5.0)) cast
    to float32, times up, zero-padded in front so that the output is centred)."""
    import ctypes as C
    import math
    from scipy.signal import firwin
    up, down, J, npr = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
    count = built_lib.afx_resample_filter(sr, native, None, 0, C.byref(up), C.byref(down), C.byref(J), C.byref(npr))
    g = math.gcd(sr, native)
    assert (up.value, down.value) == (sr // g, native // g) and count == up.value * J.value
    taps = np.zeros(count, dtype=np.float32)
    assert built_lib.afx_resample_filter(sr, native, taps.ctypes.data, count, None, None, None, None) == count
    mr = max(up.value, down.value)
    half = 10 * mr
    h = firwin(2 * half + 1, 1.0 / mr, window=("kaiser", 5.0)).astype(np.float32)
    h *= up.value
    n_pre_pad = down.value - half % down.value
    assert npr.value == (half + n_pre_pad) // down.value
    hp = np.concatenate([np.zeros(n_pre_pad, np.float32), h])
    hp = np.concatenate([hp, np.zeros((-len(hp)) % up.value, np.float32)])
    np.testing.assert_array_equal(taps.reshape(up.value, J.value), hp.reshape(-1, up.value).T)
