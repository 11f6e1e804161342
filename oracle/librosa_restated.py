"""numpy/scipy restatement of the librosa 0.11.0 routines on the AudioFeatureExtractor hot path.

TEST INFRASTRUCTURE (see oracle/__init__.py).  PARITY UNPINNED: librosa itself is absent from
the reference tree and from this image; every function below restates the published librosa
0.11.0 algorithm and names the reference call site that reaches it
(`core/feature_extractor.py` = /root/reference/audio_feature_extraction_toolkit/core/feature_extractor.py).

dtype discipline follows librosa: float32 signal, float64 window product + FFT stored as
complex64, float32 mel / dB / DCT / deltas, float64 pYIN after the float32 energy cumsum.
Pass ``dtype=np.float64`` where offered to get the all-double variant used to measure the
float32 noise floor.
"""
from __future__ import annotations

import math

import numpy as np
import scipy.fft
import scipy.signal
import scipy.stats

try:  # librosa jit-compiles its Viterbi with numba; do the same when numba is importable
    import numba
except Exception:  # pragma: no cover
    numba = None

__all__ = [
    "note_c2_hz", "note_c7_hz", "preemphasis", "frame", "rms", "trim_indices", "trim",
    "hamming_periodic", "stft_power", "hz_to_mel", "mel_to_hz", "mel_filterbank",
    "power_to_db", "dct_ortho", "melspectrogram", "mfcc", "delta", "zero_crossing_rate", "pyin_constants",
    "yin_cmnd", "parabolic_shifts", "pyin_observations", "pyin_transition", "viterbi",
    "pyin", "beta_probs", "TINY",
]

TINY = float(np.finfo(np.float64).tiny)

# librosa.note_to_hz('C2') / ('C7') evaluated in the default arguments of
# core/feature_extractor.py:15-16:  440 * 2**((midi - 69) / 12), midi(C2)=36, midi(C7)=96
note_c2_hz = 440.0 * 2.0 ** ((36 - 69) / 12.0)
note_c7_hz = 440.0 * 2.0 ** ((96 - 69) / 12.0)


# --------------------------------------------------------------------------- pre-emphasis / framing
def preemphasis(y: np.ndarray, coef: float = 0.97) -> np.ndarray:
    """librosa.effects.preemphasis (core/feature_extractor.py:69).

    b=[1,-coef], a=[1] in y.dtype; zi = 2*y[0]-y[1] is handed to scipy.signal.lfilter as the
    filter *state*, so out[0] = y[0] + (2*y[0]-y[1]) and out[n] = y[n] - coef*y[n-1] afterwards.
    """
    y = np.asarray(y)
    b = np.asarray([1.0, -coef], dtype=y.dtype)
    a = np.asarray([1.0], dtype=y.dtype)
    zi = 2 * y[..., 0:1] - y[..., 1:2]
    out, _ = scipy.signal.lfilter(b, a, y, zi=np.asarray(zi, dtype=y.dtype))
    return out.astype(y.dtype, copy=False)


def frame(y: np.ndarray, frame_length: int, hop_length: int) -> np.ndarray:
    """librosa.util.frame on the last axis -> [frame_length, n_frames] view."""
    n = 1 + (y.shape[-1] - frame_length) // hop_length
    if n < 1:
        raise ValueError("input too short for one frame")
    s = y.strides[-1]
    return np.lib.stride_tricks.as_strided(y, shape=(frame_length, n), strides=(s, hop_length * s),
                                           writeable=False)


def rms(y: np.ndarray, frame_length: int = 2048, hop_length: int = 512, dtype=np.float32) -> np.ndarray:
    """librosa.feature.rms(y=...) with center=True, pad_mode='constant' (core/feature_extractor.py:164-168).

    Returns shape [1, T], T = 1 + len(y)//hop_length.
    """
    yp = np.pad(np.asarray(y), (frame_length // 2, frame_length // 2), mode="constant")
    x = frame(yp, frame_length, hop_length)
    power = np.mean(np.square(x, dtype=dtype), axis=-2, keepdims=True)
    return np.sqrt(power)


def trim_indices(y: np.ndarray, top_db: float = 30.0, frame_length: int = 2048, hop_length: int = 512):
    """(start, end) of librosa.effects.trim(y, top_db) (core/feature_extractor.py:72).

    rms(2048/512) -> amplitude_to_db(ref=np.max, amin=1e-5, top_db=None) -> frames with
    dB > -top_db are non-silent; start = hop*first, end = min(len, hop*(last+1)); (0, 0) if none.
    """
    mse = rms(y, frame_length=frame_length, hop_length=hop_length)[0]
    magnitude = np.abs(mse)
    ref_value = np.max(magnitude)
    power = np.square(magnitude)
    amin = 1e-5 ** 2  # python double, demoted to the array dtype by np.maximum (as in librosa)
    log_spec = 10.0 * np.log10(np.maximum(amin, power))
    log_spec -= 10.0 * np.log10(np.maximum(amin, ref_value ** 2))
    nonzero = np.flatnonzero(log_spec > -top_db)
    if nonzero.size == 0:
        return 0, 0
    start = int(nonzero[0]) * hop_length
    end = min(y.shape[-1], (int(nonzero[-1]) + 1) * hop_length)
    return start, end


def trim(y: np.ndarray, top_db: float = 30.0):
    start, end = trim_indices(y, top_db=top_db)
    return y[start:end], (start, end)


# --------------------------------------------------------------------------- STFT / mel / dB / DCT
def hamming_periodic(n_fft: int) -> np.ndarray:
    """scipy.signal.get_window('hamming', n_fft, fftbins=True): 0.54 - 0.46 cos(2 pi k / n_fft), float64."""
    return scipy.signal.get_window("hamming", n_fft, fftbins=True)


def stft_power(y: np.ndarray, n_fft: int, hop_length: int, dtype=np.float32) -> np.ndarray:
    """|librosa.stft(y, n_fft, hop, window='hamming', center=True, pad_mode='constant')|**2.

    The window (float64) times the float32 frame is float64, the FFT (scipy.fft.rfft) runs in
    double and is stored as complex64; np.abs(...)**2 is float32.  Shape [1+n_fft/2, T].
    """
    y = np.asarray(y)
    if y.size == 0:
        raise ValueError("empty signal")
    if not np.all(np.isfinite(y)):
        raise ValueError("Audio buffer is not finite everywhere")
    win = hamming_periodic(n_fft)
    yp = np.pad(y, (n_fft // 2, n_fft // 2), mode="constant")
    frames = frame(yp, n_fft, hop_length)
    spec = scipy.fft.rfft(win[:, None] * frames, axis=0)
    if dtype == np.float32:
        spec = spec.astype(np.complex64)
    return np.abs(spec) ** 2


def hz_to_mel(f):
    """Slaney (htk=False) scale: linear 200/3 Hz per mel below 1 kHz, log above."""
    f = np.asanyarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    with np.errstate(divide="ignore"):
        log_t = min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep
    return np.where(f >= min_log_hz, log_t, mels)


def mel_to_hz(m):
    m = np.asanyarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    freqs = f_sp * m
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), freqs)


def mel_filterbank(sr: int, n_fft: int, n_mels: int = 128, fmin: float = 0.0, fmax: float | None = None,
                   dtype=np.float32) -> np.ndarray:
    """librosa.filters.mel(sr, n_fft, n_mels=128, fmin=0, fmax=sr/2, htk=False, norm='slaney')."""
    if fmax is None:
        fmax = float(sr) / 2
    weights = np.zeros((n_mels, 1 + n_fft // 2), dtype=dtype)
    fftfreqs = np.fft.rfftfreq(n=n_fft, d=1.0 / sr)
    mel_f = mel_to_hz(np.linspace(hz_to_mel(fmin), hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = np.subtract.outer(mel_f, fftfreqs)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    weights *= enorm[:, np.newaxis]
    return weights


def power_to_db(S: np.ndarray, ref: float = 1.0, amin: float = 1e-10, top_db: float | None = 80.0) -> np.ndarray:
    """librosa.power_to_db: 10 log10(max(amin,S)) - 10 log10(max(amin,ref)); clamp at global max - top_db."""
    S = np.asarray(S)
    log_spec = 10.0 * np.log10(np.maximum(S.dtype.type(amin), S))
    log_spec -= 10.0 * np.log10(np.maximum(amin, ref))
    if top_db is not None:
        log_spec = np.maximum(log_spec, log_spec.max() - top_db)
    return log_spec


def dct_ortho(S: np.ndarray, n_out: int) -> np.ndarray:
    """scipy.fftpack.dct(S, axis=-2, type=2, norm='ortho')[:n_out] (librosa.feature.mfcc, lifter=0)."""
    return scipy.fft.dct(S, axis=-2, type=2, norm="ortho")[..., :n_out, :]


def melspectrogram(y, sr, n_fft, hop_length, n_mels=128, dtype=np.float32):
    S = stft_power(y, n_fft, hop_length, dtype=dtype)
    basis = mel_filterbank(sr, n_fft, n_mels=n_mels, dtype=dtype)
    return basis @ S  # einsum('ft,mf->mt') in the array dtype


def mfcc(y, sr, n_mfcc, n_fft, hop_length, n_mels=128, dtype=np.float32, return_logmel=False):
    """librosa.feature.mfcc(y, sr, n_mfcc, n_fft, hop_length, window='hamming') (core/feature_extractor.py:127-134)."""
    M = melspectrogram(y, sr, n_fft, hop_length, n_mels=n_mels, dtype=dtype)
    D = power_to_db(M, top_db=80.0)
    C = dct_ortho(D, n_mfcc)
    if return_logmel:
        return C, D
    return C


def delta(data: np.ndarray, order: int = 1, width: int = 9) -> np.ndarray:
    """librosa.feature.delta(data, order) = savgol_filter(width=9, polyorder=order, deriv=order, mode='interp')
    along the last axis (core/feature_extractor.py:137-138)."""
    data = np.atleast_1d(data)
    if width > data.shape[-1]:
        raise ValueError("when mode='interp', width=%d cannot exceed data.shape[axis]=%d" % (width, data.shape[-1]))
    return scipy.signal.savgol_filter(data, width, polyorder=order, deriv=order, axis=-1, mode="interp")


def zero_crossing_rate(y: np.ndarray, frame_length: int = 2048, hop_length: int = 512, threshold: float = 1e-10) -> np.ndarray:
    """librosa.feature.zero_crossing_rate(y, frame_length, hop_length) (center=True): edge padding by frame_length // 2,
    frames, librosa.zero_crossings(pad=False, zero_pos=True, threshold=1e-10) down each frame, mean over the frame's
    frame_length entries (the first entry of a frame is `pad` = False).  Samples with |x| <= threshold count as +0.
    Called by 04_feature_extraction_experiment/feature_extraction.py:279-283 (the framewise .npz features).  [A9]"""
    y = np.asarray(y)
    yp = np.pad(y, (frame_length // 2, frame_length // 2), mode="edge")
    fr = frame(yp, frame_length, hop_length)                      # [frame_length, T]
    z = np.where(np.abs(fr) <= threshold, 0.0, fr)
    sb = np.signbit(z)
    cross = np.zeros(fr.shape, dtype=bool)
    cross[1:] = sb[1:] != sb[:-1]
    return np.mean(cross, axis=0, keepdims=True)                  # [1, T] float64


# --------------------------------------------------------------------------- pYIN
def pyin_constants(sr: int, frame_length: int, hop_length: int, fmin: float, fmax: float,
                   resolution: float = 0.1, max_transition_rate: float = 35.92):
    win_length = frame_length // 2
    min_period = int(np.floor(sr / fmax))
    max_period = min(int(np.ceil(sr / fmin)), frame_length - win_length - 1)
    n_bins_per_semitone = int(np.ceil(1.0 / resolution))
    n_pitch_bins = int(np.floor(12 * n_bins_per_semitone * np.log2(fmax / fmin))) + 1
    max_semitones_per_frame = round(max_transition_rate * 12 * hop_length / sr)
    transition_width = max_semitones_per_frame * n_bins_per_semitone + 1
    return dict(win_length=win_length, min_period=min_period, max_period=max_period,
                n_bins_per_semitone=n_bins_per_semitone, n_pitch_bins=n_pitch_bins,
                transition_width=transition_width)


def beta_probs(n_thresholds: int = 100, beta_parameters=(2, 18)) -> np.ndarray:
    thresholds = np.linspace(0, 1, n_thresholds + 1)
    cdf = scipy.stats.beta.cdf(thresholds, beta_parameters[0], beta_parameters[1])
    return np.diff(cdf)


def yin_cmnd(y_frames: np.ndarray, frame_length: int, win_length: int, min_period: int, max_period: int,
             acf_dtype=np.float64) -> np.ndarray:
    """librosa.core.pitch._cumulative_mean_normalized_difference.

    y_frames: float32 [frame_length, T].  The autocorrelation goes through np.fft (double under
    the reference's pinned numpy 1.24.4); the energy cumsum stays float32.
    """
    yf = y_frames.astype(acf_dtype, copy=False)
    a = np.fft.rfft(yf, frame_length, axis=-2)
    b = np.fft.rfft(yf[win_length:0:-1, :], frame_length, axis=-2)
    acf_frames = np.fft.irfft(a * b, frame_length, axis=-2)[win_length:, :]
    acf_frames[np.abs(acf_frames) < 1e-6] = 0
    energy_frames = np.cumsum(y_frames ** 2, axis=-2)
    energy_frames = energy_frames[win_length:, :] - energy_frames[:-win_length, :]
    energy_frames[np.abs(energy_frames) < 1e-6] = 0
    yin_frames = energy_frames[:1, :] + energy_frames - 2 * acf_frames
    yin_numerator = yin_frames[min_period:max_period + 1, :]
    tau_range = np.arange(1, max_period + 1)[:, None]
    cumulative_mean = np.cumsum(yin_frames[1:max_period + 1, :], axis=-2) / tau_range
    yin_denominator = cumulative_mean[min_period - 1:max_period, :]
    return yin_numerator / (yin_denominator + TINY)


def parabolic_shifts(x: np.ndarray) -> np.ndarray:
    """librosa.core.pitch._parabolic_interpolation along axis 0 ([lags, T])."""
    shifts = np.zeros_like(x)
    a = x[2:] + x[:-2] - 2 * x[1:-1]
    b = (x[2:] - x[:-2]) / 2
    with np.errstate(divide="ignore", invalid="ignore"):
        s = -b / a
    s[np.abs(b) >= np.abs(a)] = 0
    shifts[1:-1] = s
    return shifts


def _localmin(x: np.ndarray) -> np.ndarray:
    """librosa.util.localmin on a 1-d array: x[i-1] > x[i] <= x[i+1], edges padded with themselves."""
    xp = np.pad(x, 1, mode="edge")
    return (x < xp[:-2]) & (x <= xp[2:])


def pyin_observations(yin_frames, shifts, sr, fmin, min_period, n_pitch_bins, n_bins_per_semitone,
                      bprobs=None, boltzmann_parameter=2.0, no_trough_prob=0.01):
    """librosa.core.pitch.__pyin_helper: trough probabilities -> observation matrix [2*n_pitch_bins, T]."""
    if bprobs is None:
        bprobs = beta_probs()
    n_thresholds = len(bprobs)
    thresholds = np.linspace(0, 1, n_thresholds + 1)
    yin_probs = np.zeros_like(yin_frames)
    for i, yin_frame in enumerate(yin_frames.T):
        is_trough = _localmin(yin_frame)
        is_trough[0] = yin_frame[0] < yin_frame[1]
        (trough_index,) = np.nonzero(is_trough)
        if len(trough_index) == 0:
            continue
        trough_heights = yin_frame[trough_index]
        trough_thresholds = np.less.outer(trough_heights, thresholds[1:])
        trough_positions = np.cumsum(trough_thresholds, axis=0) - 1
        n_troughs = np.count_nonzero(trough_thresholds, axis=0)
        with np.errstate(all="ignore"):
            trough_prior = scipy.stats.boltzmann.pmf(trough_positions, boltzmann_parameter, n_troughs)
        trough_prior[~trough_thresholds] = 0
        probs = trough_prior.dot(bprobs)
        global_min = np.argmin(trough_heights)
        n_thresholds_below_min = np.count_nonzero(~trough_thresholds[global_min, :])
        probs[global_min] += no_trough_prob * np.sum(bprobs[:n_thresholds_below_min])
        yin_probs[trough_index, i] = probs
    yin_period, frame_index = np.nonzero(yin_probs)
    period_candidates = min_period + yin_period
    period_candidates = period_candidates + shifts[yin_period, frame_index]
    f0_candidates = sr / period_candidates
    bin_index = 12 * n_bins_per_semitone * np.log2(f0_candidates / fmin)
    bin_index = np.clip(np.round(bin_index), 0, n_pitch_bins).astype(int)
    observation_probs = np.zeros((2 * n_pitch_bins, yin_frames.shape[1]))
    observation_probs[bin_index, frame_index] = yin_probs[yin_period, frame_index]
    voiced_prob = np.clip(np.sum(observation_probs[:n_pitch_bins, :], axis=0, keepdims=True), 0, 1)
    observation_probs[n_pitch_bins:, :] = (1 - voiced_prob) / n_pitch_bins
    return observation_probs, voiced_prob[0]


def pyin_transition(n_pitch_bins: int, transition_width: int, switch_prob: float = 0.01) -> np.ndarray:
    """kron(transition_loop(2, 1-switch), transition_local(n, width, 'triangle', wrap=False))."""
    n = n_pitch_bins
    w = transition_width
    win = scipy.signal.get_window("triangle", w, fftbins=False)
    local = np.zeros((n, n), dtype=np.float64)
    for i in range(n):
        row = np.zeros(n)
        lpad = (n - w) // 2
        row[lpad:lpad + w] = win
        row = np.roll(row, n // 2 + i + 1)
        row[min(n, i + w // 2 + 1):] = 0
        row[:max(0, i - w // 2)] = 0
        local[i] = row
    local /= local.sum(axis=1, keepdims=True)
    # librosa.sequence.transition_loop(2, 1 - switch_prob): diagonal p, off-diagonal (1 - p) / (n_states - 1)
    p_stay = 1 - switch_prob
    t_switch = np.empty((2, 2))
    t_switch[:] = (1.0 - p_stay) / (2 - 1)
    t_switch[0, 0] = t_switch[1, 1] = p_stay
    return np.kron(t_switch, local)


def _viterbi_dense_py(log_prob, log_trans, log_p_init):  # [T, S], [S, S], [S]
    n_steps, n_states = log_prob.shape
    state = np.zeros(n_steps, dtype=np.uint16)
    value = np.zeros((n_steps, n_states), dtype=np.float64)
    ptr = np.zeros((n_steps, n_states), dtype=np.uint16)
    value[0] = log_prob[0] + log_p_init
    for t in range(1, n_steps):
        trans_out = value[t - 1] + log_trans.T
        for j in range(n_states):
            ptr[t, j] = np.argmax(trans_out[j])
            value[t, j] = log_prob[t, j] + trans_out[j, ptr[t][j]]
    state[-1] = np.argmax(value[-1])
    for t in range(n_steps - 2, -1, -1):
        state[t] = ptr[t + 1, state[t + 1]]
    return state


_viterbi_dense = numba.jit(nopython=True, cache=False)(_viterbi_dense_py) if numba is not None else _viterbi_dense_py


def viterbi(prob: np.ndarray, transition: np.ndarray, p_init: np.ndarray) -> np.ndarray:
    """librosa.sequence.viterbi: log(. + tiny) of all three, dense max/argmax (first index on ties)."""
    log_trans = np.log(transition + TINY)
    log_prob = np.log(prob + TINY)
    log_p_init = np.log(p_init + TINY)
    return _viterbi_dense(np.ascontiguousarray(log_prob.T), log_trans, log_p_init).astype(np.int64)


def pyin(y, fmin, fmax, sr=22050, frame_length=2048, hop_length=None, switch_prob=0.01,
         no_trough_prob=0.01, return_details=False, acf_dtype=np.float64):
    """librosa.pyin(y, fmin=, fmax=, sr=, frame_length=, hop_length=) with library defaults
    (core/feature_extractor.py:87-94).  Returns (f0 [T] float64 with NaN unvoiced, voiced_flag, voiced_prob)."""
    y = np.asarray(y)
    if y.size == 0:
        raise ValueError("empty signal")
    if not np.all(np.isfinite(y)):
        raise ValueError("Audio buffer is not finite everywhere")
    if hop_length is None:
        hop_length = frame_length // 4
    c = pyin_constants(sr, frame_length, hop_length, fmin, fmax)
    yp = np.pad(y, (frame_length // 2, frame_length // 2), mode="constant")
    y_frames = frame(yp, frame_length, hop_length)
    yin_frames = yin_cmnd(y_frames, frame_length, c["win_length"], c["min_period"], c["max_period"],
                          acf_dtype=acf_dtype)
    shifts = parabolic_shifts(yin_frames)
    obs, voiced_prob = pyin_observations(yin_frames, shifts, sr, fmin, c["min_period"], c["n_pitch_bins"],
                                         c["n_bins_per_semitone"], no_trough_prob=no_trough_prob)
    transition = pyin_transition(c["n_pitch_bins"], c["transition_width"], switch_prob)
    p_init = np.zeros(2 * c["n_pitch_bins"])
    p_init[c["n_pitch_bins"]:] = 1 / c["n_pitch_bins"]
    states = viterbi(obs, transition, p_init)
    freqs = fmin * 2 ** (np.arange(c["n_pitch_bins"]) / (12 * c["n_bins_per_semitone"]))
    f0 = freqs[states % c["n_pitch_bins"]]
    voiced_flag = states < c["n_pitch_bins"]
    f0 = f0.copy()
    f0[~voiced_flag] = np.nan
    if return_details:
        return f0, voiced_flag, voiced_prob, dict(yin=yin_frames, shifts=shifts, obs=obs, states=states, consts=c)
    return f0, voiced_flag, voiced_prob


def resample_polyphase(y, orig_sr, target_sr):
    """librosa.resample(y, orig_sr=, target_sr=, res_type='polyphase') (librosa 0.11 core/audio.py): scipy.signal.
    resample_poly(y, target_sr // gcd, orig_sr // gcd) on the float32 signal, fixed to ceil(len * target / orig) samples.
    Stands in for librosa.load's default 'soxr_hq' (the soxr C library, not restatable) on files whose rate != sr."""
    import math
    import scipy.signal
    y = np.asarray(y, dtype=np.float32)
    if int(orig_sr) == int(target_sr):
        return y
    g = math.gcd(int(orig_sr), int(target_sr))
    out = scipy.signal.resample_poly(y, int(target_sr) // g, int(orig_sr) // g)
    n = int(np.ceil(y.shape[-1] * float(target_sr) / float(orig_sr)))
    if out.shape[-1] > n:
        out = out[:n]
    elif out.shape[-1] < n:
        out = np.pad(out, (0, n - out.shape[-1]))
    return np.asarray(out, dtype=np.float32)
