"""CPU oracle for the audio -> log-mel (+MFCC, +SpecAugment masking, +delta) front-end.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: it may be imported
by ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` and by nothing else.  The product path is ``speech_recognition_b200`` (CUDA only).

What it restates
----------------
The hot path of cosmoquester/speech-recognition, ``speech_recognition/data.py:122-328``:

* ``make_spectrogram``          data.py:122-142
* ``make_log_mel_spectrogram``  data.py:145-189
* ``make_mfcc``                 data.py:192-241
* ``spec_augment`` (masking)    data.py:244-307 (time-warp branch :275-280 is out of scope)
* ``delta_accelerate``          data.py:310-328

The reference only *composes* TensorFlow ops (``tf.signal.stft``, ``linear_to_mel_weight_matrix``,
``mfccs_from_log_mel_spectrograms``, ``tf.random.uniform``).  TensorFlow is an un-vendored,
unpinned dependency (``requirements.txt:1``: ``tensorflow>=2``) that is NOT importable in the
build container, so the arithmetic below restates TensorFlow's published algorithm for those ops
(tensorflow/python/ops/signal/{spectral_ops,window_ops,mel_ops,mfcc_ops,dct_ops}.py) and parity is
anchored on the reference's call sites and on its one numeric fixture:

PARITY PIN STATUS: *partially pinned*.  ``tests/data/wav_dataset.tfrecord`` (reference
``tests/test_data.py:53-57``) pins frame count, mel-bin count, dtype, epsilon and the log
(silence in -> log(1e-12) everywhere); the reference's shape tests pin the frame/bin arithmetic
(``tests/test_data.py:60-145``).  Window / FFT / mel-weight / DCT values are NOT pinned by any
reference fixture ("parity unpinned" for those terms); they are pinned here against independent
implementations (``torch.stft``, ``scipy.fft.dct``, analytic known answers) in
``tests/test_oracle.py``.

Every function takes ``dtype``: ``np.float64`` gives the "truth" used for error budgets,
``np.float32`` follows TF's float32 arithmetic order as closely as numpy allows.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import numpy as np

# --------------------------------------------------------------------------------------------
# tf.signal.stft  (reference call sites: data.py:135, 172, 221)
# --------------------------------------------------------------------------------------------


def enclosing_power_of_two(n: int) -> int:
    """tf.signal.stft default ``fft_length`` (data.py:128 docstring): smallest 2**k >= n."""
    return 1 if n <= 1 else 1 << (int(n) - 1).bit_length()


def frame_count(num_samples: int, frame_length: int, frame_step: int) -> int:
    """tf.signal.frame(pad_end=False): max(0, 1 + (L - Lw) // H).

    Equals the reference's own shape assertion ``(L - Lw + H) // H`` (tests/test_data.py:76-78).
    """
    if num_samples < frame_length:
        return 0
    return 1 + (num_samples - frame_length) // frame_step


def hann_window(frame_length: int, dtype=np.float64) -> np.ndarray:
    """tf.signal.hann_window(periodic=True) (window_ops._raised_cosine_window, a=b=0.5).

    even = 1 - Lw % 2 ; n = Lw + even - 1 ; w[i] = 0.5 - 0.5 cos(2 pi i / n); Lw == 1 -> [1].
    """
    if frame_length == 1:
        return np.ones(1, dtype=dtype)
    even = 1 - frame_length % 2
    n = dtype(frame_length + even - 1)
    count = np.arange(frame_length, dtype=dtype)
    cos_arg = dtype(2.0 * np.pi) * count / n
    return (dtype(0.5) - dtype(0.5) * np.cos(cos_arg)).astype(dtype)


def stft(audio: np.ndarray, frame_length: int, frame_step: int, fft_length: Optional[int] = None,
         dtype=np.float64) -> np.ndarray:
    """tf.signal.stft(audio, Lw, H, N) with its defaults (periodic Hann, pad_end=False).

    Frames are windowed, then rfft(length N): zero-padded at the end if Lw < N and CROPPED to the
    first N windowed samples if Lw > N (exercised by the reference at tests/test_data.py:66).
    Returns complex [T, N//2+1].
    """
    if fft_length is None:
        fft_length = enclosing_power_of_two(frame_length)
    audio = np.asarray(audio, dtype=dtype)
    T = frame_count(audio.shape[0], frame_length, frame_step)
    K = fft_length // 2 + 1
    cdtype = np.complex128 if dtype == np.float64 else np.complex64
    if T == 0:
        return np.zeros((0, K), dtype=cdtype)
    idx = np.arange(T)[:, None] * frame_step + np.arange(frame_length)[None, :]
    frames = audio[idx] * hann_window(frame_length, dtype)[None, :]
    frames = frames.astype(dtype)
    # numpy>=2 keeps float32 input in single precision inside pocketfft.
    return np.fft.rfft(frames, n=fft_length, axis=-1).astype(cdtype)


# --------------------------------------------------------------------------------------------
# tf.signal.linear_to_mel_weight_matrix  (reference call sites: data.py:177, 226)
# --------------------------------------------------------------------------------------------

_MEL_BREAK_FREQUENCY_HERTZ = 700.0
_MEL_HIGH_FREQUENCY_Q = 1127.0


def _linspace_tf(start, stop, num: int, dtype) -> np.ndarray:
    """tf.linspace in `dtype`: start + delta * i, with the exact `stop` as last element."""
    start = dtype(start)
    stop = dtype(stop)
    if num == 1:
        return np.array([start], dtype=dtype)
    delta = dtype((stop - start) / dtype(num - 1))
    body = (start + delta * np.arange(1, num - 1, dtype=dtype)).astype(dtype)
    return np.concatenate([[start], body, [stop]]).astype(dtype)


def _hertz_to_mel(f, dtype):
    """mel_ops._hertz_to_mel.  The float32 log is taken as the correctly rounded value (float64 log
    rounded to float32): numpy's SIMD float32 log, glibc logf and Eigen's plog (what TF runs) each
    differ in the last ulp, and no fixture pins which one the reference saw."""
    f = np.asarray(f, dtype=dtype)
    arg = (dtype(1.0) + f / dtype(_MEL_BREAK_FREQUENCY_HERTZ)).astype(dtype)
    log = np.log(arg.astype(np.float64)).astype(dtype)
    return (dtype(_MEL_HIGH_FREQUENCY_Q) * log).astype(dtype)


def validate_mel_arguments(num_mel_bins, sample_rate, lower_edge_hertz, upper_edge_hertz):
    """mel_ops._validate_arguments: the conditions under which TF raises ValueError."""
    if num_mel_bins <= 0:
        raise ValueError("num_mel_bins must be positive. Got: %s" % num_mel_bins)
    if lower_edge_hertz < 0.0:
        raise ValueError("lower_edge_hertz must be non-negative. Got: %s" % lower_edge_hertz)
    if lower_edge_hertz >= upper_edge_hertz:
        raise ValueError("lower_edge_hertz %.1f >= upper_edge_hertz %.1f" %
                         (lower_edge_hertz, upper_edge_hertz))
    if sample_rate <= 0.0:
        raise ValueError("sample_rate must be positive. Got: %s" % sample_rate)
    if upper_edge_hertz > sample_rate / 2:
        raise ValueError("upper_edge_hertz must not be larger than the Nyquist frequency "
                         "(sample_rate / 2). Got %s for sample_rate: %s" %
                         (upper_edge_hertz, sample_rate))


def linear_to_mel_weight_matrix(num_mel_bins: int, num_spectrogram_bins: int, sample_rate,
                                lower_edge_hertz, upper_edge_hertz,
                                dtype=np.float32) -> np.ndarray:
    """HTK mel filterbank [K, M] in TF's op order; TF computes it in float32 (dtype default)."""
    validate_mel_arguments(num_mel_bins, sample_rate, lower_edge_hertz, upper_edge_hertz)
    bands_to_zero = 1
    nyquist = dtype(dtype(sample_rate) / dtype(2.0))
    linear_frequencies = _linspace_tf(0.0, nyquist, num_spectrogram_bins, dtype)[bands_to_zero:]
    spectrogram_bins_mel = _hertz_to_mel(linear_frequencies, dtype)[:, None]
    edges = _linspace_tf(_hertz_to_mel(lower_edge_hertz, dtype),
                         _hertz_to_mel(upper_edge_hertz, dtype), num_mel_bins + 2, dtype)
    lower_edge_mel = edges[None, :-2]
    center_mel = edges[None, 1:-1]
    upper_edge_mel = edges[None, 2:]
    lower_slopes = (spectrogram_bins_mel - lower_edge_mel) / (center_mel - lower_edge_mel)
    upper_slopes = (upper_edge_mel - spectrogram_bins_mel) / (upper_edge_mel - center_mel)
    w = np.maximum(dtype(0.0), np.minimum(lower_slopes, upper_slopes)).astype(dtype)
    return np.pad(w, [[bands_to_zero, 0], [0, 0]]).astype(dtype)


# --------------------------------------------------------------------------------------------
# feature extractors
# --------------------------------------------------------------------------------------------


def make_spectrogram(audio, frame_length, frame_step, fft_length=None, dtype=np.float64):
    """data.py:122-142: |stft| with trailing channel axis -> [T, K, 1]."""
    return np.abs(stft(audio, frame_length, frame_step, fft_length, dtype)).astype(dtype)[:, :, None]


def mel_energies(audio, sample_rate, frame_length, frame_step, fft_length, num_mel_bins=80,
                 lower_edge_hertz=80.0, upper_edge_hertz=7600.0, dtype=np.float64,
                 mel_dtype=np.float32):
    """data.py:172-182: square(abs(stft)) @ mel_matrix -> [T, M] (before +eps and log)."""
    mag = np.abs(stft(audio, frame_length, frame_step, fft_length, dtype)).astype(dtype)
    W = linear_to_mel_weight_matrix(num_mel_bins, fft_length // 2 + 1, sample_rate,
                                    lower_edge_hertz, upper_edge_hertz, mel_dtype).astype(dtype)
    return np.matmul(np.square(mag), W).astype(dtype)


def make_log_mel_spectrogram(audio, sample_rate, frame_length, frame_step, fft_length,
                             num_mel_bins=80, lower_edge_hertz=80.0, upper_edge_hertz=7600.0,
                             epsilon=1e-12, dtype=np.float64, mel_dtype=np.float32):
    """data.py:145-189 -> [T, M, 1].  `mel_dtype` is the precision the filterbank is BUILT in
    (TF: float32) - the largest single error term between a float64 and a float32 pipeline."""
    mel = mel_energies(audio, sample_rate, frame_length, frame_step, fft_length, num_mel_bins,
                       lower_edge_hertz, upper_edge_hertz, dtype, mel_dtype)
    return np.log(mel + dtype(epsilon)).astype(dtype)[:, :, None]


def dct2_matrix(num_mel_bins: int, num_mfcc: int, dtype=np.float64) -> np.ndarray:
    """tf.signal.dct(type=2, norm=None) * rsqrt(2 M) as an [M, num_mfcc] matrix:
    D[m, c] = 2 cos(pi c (2m+1) / (2M)) / sqrt(2M)   (mfcc_ops.mfccs_from_log_mel_spectrograms)."""
    m = np.arange(num_mel_bins, dtype=np.float64)[:, None]
    c = np.arange(num_mfcc, dtype=np.float64)[None, :]
    D = 2.0 * np.cos(np.pi * c * (2.0 * m + 1.0) / (2.0 * num_mel_bins)) / math.sqrt(2.0 * num_mel_bins)
    return D.astype(dtype)


def make_mfcc(audio, sample_rate, frame_length, frame_step, fft_length, num_mel_bins=80,
              num_mfcc=40, lower_edge_hertz=80.0, upper_edge_hertz=7600.0, epsilon=1e-12,
              dtype=np.float64, mel_dtype=np.float32):
    """data.py:192-241 -> [T, num_mfcc, 1]  (num_mfcc > M just returns M columns, like TF's slice)."""
    logmel = make_log_mel_spectrogram(audio, sample_rate, frame_length, frame_step, fft_length,
                                      num_mel_bins, lower_edge_hertz, upper_edge_hertz, epsilon,
                                      dtype, mel_dtype)[:, :, 0]
    n = min(num_mfcc, num_mel_bins)
    return np.matmul(logmel, dct2_matrix(num_mel_bins, n, dtype)).astype(dtype)[:, :, None]


def delta_accelerate(feat: np.ndarray) -> np.ndarray:
    """data.py:310-328: [T, F, 1] -> [T, F, 3] = (x, x - shift(x), delta - shift(delta))."""
    zero_head = np.zeros_like(feat[:1])
    delta = feat - np.concatenate([zero_head, feat[:-1]], axis=0)
    deltas = delta - np.concatenate([zero_head, delta[:-1]], axis=0)
    return np.concatenate([feat, delta, deltas], axis=2)


# --------------------------------------------------------------------------------------------
# SpecAugment masking  (data.py:282-301)
# --------------------------------------------------------------------------------------------

PHILOX_M0 = 0xD2511F53
PHILOX_M1 = 0xCD9E8D57
PHILOX_W0 = 0x9E3779B9
PHILOX_W1 = 0xBB67AE85


def philox4x32_10(counter: Sequence[int], key: Sequence[int]) -> Tuple[int, int, int, int]:
    """Philox4x32-10 (Salmon et al., SC'11) - the generator family TF's tf.random uses."""
    c0, c1, c2, c3 = [int(c) & 0xFFFFFFFF for c in counter]
    k0, k1 = [int(k) & 0xFFFFFFFF for k in key]
    for _ in range(10):
        p0 = PHILOX_M0 * c0
        p1 = PHILOX_M1 * c2
        hi0, lo0 = p0 >> 32, p0 & 0xFFFFFFFF
        hi1, lo1 = p1 >> 32, p1 & 0xFFFFFFFF
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & 0xFFFFFFFF, lo1, (hi0 ^ c3 ^ k1) & 0xFFFFFFFF, lo0
        k0 = (k0 + PHILOX_W0) & 0xFFFFFFFF
        k1 = (k1 + PHILOX_W1) & 0xFFFFFFFF
    return c0, c1, c2, c3


def specaug_raw_draw(seed: int, utterance_id: int, draw_index: int) -> int:
    """The framework's counter-based draw: u32 = Philox(counter=(draw, 0, utt_lo, utt_hi),
    key=(seed_lo, seed_hi))[0].  A pure function of (seed, global utterance id, draw index), so
    masks do not depend on batch composition, sharding or GPU count."""
    seed &= 0xFFFFFFFFFFFFFFFF
    utterance_id &= 0xFFFFFFFFFFFFFFFF
    return philox4x32_10((draw_index, 0, utterance_id & 0xFFFFFFFF, utterance_id >> 32),
                         (seed & 0xFFFFFFFF, seed >> 32))[0]


def _uniform_int(u32: int, lo: int, hi: int) -> int:
    """tf.random.uniform((), lo, hi, tf.int32): lo + (u32 mod (hi - lo)); TF raises if hi <= lo."""
    if hi <= lo:
        raise ValueError("Need minval < maxval, got %d >= %d" % (lo, hi))
    return lo + (u32 % (hi - lo))


def specaug_draw_masks(num_time: int, v: int, F=None, m_F=None, T=None, p=None, m_T=None,
                       raw: Optional[Sequence[int]] = None, seed: int = 0, utterance_id: int = 0
                       ) -> List[Tuple[int, int]]:
    """Draw the mask table for one utterance, in the reference's draw order (data.py:282-301):
    (f, f_0) x m_F then (t, t_0) x m_T.  Returns [(start, width)] * (m_F + m_T); the first m_F
    entries are frequency masks.  `raw` = explicit u32 draws (e.g. recorded from TF); otherwise
    the framework's Philox stream."""
    use_f = all([F, m_F])
    use_t = all([T, p, m_T])
    out: List[Tuple[int, int]] = []
    n = [0]

    def nxt():
        i = n[0]
        n[0] += 1
        return int(raw[i]) if raw is not None else specaug_raw_draw(seed, utterance_id, i)

    if use_f:
        for _ in range(m_F):
            f = _uniform_int(nxt(), 0, F)
            f0 = _uniform_int(nxt(), 0, v - f)
            out.append((f0, f))
    if use_t:
        applied = 0
        max_maskable = int(np.int32(np.float32(num_time) * np.float32(p)))
        for _ in range(m_T):
            t = _uniform_int(nxt(), 0, T)
            t = min(t, max_maskable - applied)
            applied += t
            t0 = _uniform_int(nxt(), 0, num_time - t)
            out.append((t0, t))
    return out


def spec_augment_apply(feat: np.ndarray, freq_masks: Sequence[Tuple[int, int]],
                       time_masks: Sequence[Tuple[int, int]]) -> np.ndarray:
    """Apply masks exactly as the reference does: multiply by a 0/1 mask (data.py:286-288,
    299-301), so masked negative values become -0.0.  sequence_mask(a+b) == sequence_mask(a) is
    False exactly on [a, a+b)."""
    out = np.array(feat, copy=True)
    num_time, v = out.shape[0], out.shape[1]
    for f0, f in freq_masks:
        idx = np.arange(v)
        mask = ((idx < f0 + f) == (idx < f0)).astype(out.dtype)
        out = out * mask[None, :, None]
    for t0, t in time_masks:
        idx = np.arange(num_time)
        mask = ((idx < t0 + t) == (idx < t0)).astype(out.dtype)
        out = out * mask[:, None, None]
    return out


def spec_augment(feat: np.ndarray, v: int, W=None, F=None, m_F=None, T=None, p=None, m_T=None,
                 raw=None, seed: int = 0, utterance_id: int = 0) -> np.ndarray:
    """data.py:244-307 without the time-warp branch (W must be falsy)."""
    if W:
        raise NotImplementedError("time warping (data.py:275-280) is outside the hot-path scope")
    table = specaug_draw_masks(feat.shape[0], v, F, m_F, T, p, m_T, raw, seed, utterance_id)
    nf = m_F if all([F, m_F]) else 0
    return spec_augment_apply(feat, table[:nf], table[nf:])


# --------------------------------------------------------------------------------------------
# batched front-end = what the CUDA path is compared against
# --------------------------------------------------------------------------------------------


def batched_features(utterances: Sequence[np.ndarray], feature_type: str, sample_rate: int,
                     frame_length: int, frame_step: int, fft_length: int, num_mel_bins=80,
                     num_mfcc=40, lower_edge_hertz=80.0, upper_edge_hertz=7600.0, epsilon=1e-12,
                     use_delta=False, spec_aug: Optional[dict] = None, seed: int = 0,
                     first_utterance_id: int = 0, mask_tables=None, t_max: Optional[int] = None,
                     dtype=np.float64, mel_dtype=np.float32):
    """Per-utterance reference pipeline in the reference's stage order (train.py:88-116: feature ->
    spec_augment -> delta), zero-padded to [B, Tmax, F, C] like padded_batch (train.py:192-197).
    Returns (features, n_frames)."""
    feats = []
    for i, x in enumerate(utterances):
        if feature_type == "spectrogram":
            f = make_spectrogram(x, frame_length, frame_step, fft_length, dtype)
        elif feature_type == "log-mel-spectrogram":
            f = make_log_mel_spectrogram(x, sample_rate, frame_length, frame_step, fft_length,
                                         num_mel_bins, lower_edge_hertz, upper_edge_hertz, epsilon,
                                         dtype, mel_dtype)
        elif feature_type == "mfcc":
            f = make_mfcc(x, sample_rate, frame_length, frame_step, fft_length, num_mel_bins,
                          num_mfcc, lower_edge_hertz, upper_edge_hertz, epsilon, dtype, mel_dtype)
        else:
            raise ValueError(feature_type)
        if mask_tables is not None:
            fm, tm = mask_tables[i]
            f = spec_augment_apply(f, fm, tm)
        elif spec_aug:
            f = spec_augment(f, f.shape[1], None, spec_aug.get("F"), spec_aug.get("m_F"),
                             spec_aug.get("T"), spec_aug.get("p"), spec_aug.get("m_T"),
                             seed=seed, utterance_id=first_utterance_id + i)
        if use_delta:
            f = delta_accelerate(f)
        feats.append(f)
    n_frames = np.array([f.shape[0] for f in feats], dtype=np.int32)
    tm = int(n_frames.max()) if t_max is None else t_max
    Fdim = feats[0].shape[1]
    C = feats[0].shape[2]
    out = np.zeros((len(feats), tm, Fdim, C), dtype=dtype)
    for i, f in enumerate(feats):
        out[i, :f.shape[0]] = f
    return out, n_frames
