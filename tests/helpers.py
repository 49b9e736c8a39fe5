"""Shared test helpers: synthetic audio, the stated parity tolerance."""
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CONFIG_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "resources", "configs")

LIBRI = dict(sample_rate=16000, frame_length=320, frame_step=160, fft_length=320, num_mel_bins=80,
             lower_edge_hertz=80.0, upper_edge_hertz=7600.0)

# The reference's own parameter grids (tests/test_data.py:60-68, 81-89, 114-122)
REF_SPECTROGRAM_GRID = [(1024, 1024, 1024), (128, 64, 256), (128, 80, None), (512, 512, 256)]
REF_LOGMEL_GRID = [
    (22050, 1024, 1024, 1024, 80, 10, 10000),
    (16000, 128, 64, 256, 123, 12, 88),
    (32000, 128, 80, 128, 321, 32, 16000),
    (44100, 512, 512, 256, 333, 333, 3333),
]
REF_MFCC_GRID = [
    (22050, 1024, 1024, 1024, 80, 40, 10, 10000),
    (16000, 128, 64, 256, 123, 33, 12, 88),
    (32000, 128, 80, 128, 321, 100, 32, 16000),
    (44100, 512, 512, 256, 333, 333, 333, 3333),
]

# Stated tolerance (DESIGN.md "Parity tolerance"):
#   |delta log-mel| <= LOG_TOL where the float64 mel energy >= max(MEL_FLOOR, REL_FLOOR * E_frame),
#   E_frame = sum of the frame's mel energies; below that floor the bound is linear:
#   |delta mel| <= LIN_ABS + LIN_REL * mel + REL_LIN * E_frame.
# REL_FLOOR expresses float32 dynamic range: a float32 FFT (TF's, numpy's or ours) carries an error
# ~1e-7 * |frame|, so bins > ~80 dB below the frame's energy cannot hold 1e-3 in the log domain
# (measured: numpy/pocketfft float32 vs float64 on a pure tone = 2.7e-3 without the relative floor,
# 1.0e-4 with it; the survey measured 0.115 for the same effect).
LOG_TOL = 1e-3
MEL_FLOOR = 1e-7
REL_FLOOR = 1e-8
LIN_ABS, LIN_REL, REL_LIN = 1e-9, 1e-3, 1e-10


def noise_int16(n, seed=0, sigma=0.1):
    """BASELINE config 2 audio: clip(N(0, sigma^2), -1, 1) quantised to int16."""
    rng = np.random.default_rng(seed)
    x = np.clip(rng.normal(0.0, sigma, n), -1.0, 1.0)
    return np.clip(np.round(x * 32768.0), -32768, 32767).astype(np.int16)


def tone_int16(n, freq, sr=16000, amp=0.5):
    t = np.arange(n) / sr
    return np.round(amp * np.sin(2 * np.pi * freq * t) * 32767).astype(np.int16)


def coloured_int16(n, seed=0):
    """speech-like: 1/f-ish coloured noise with a slow amplitude envelope."""
    rng = np.random.default_rng(seed)
    w = rng.normal(0, 1, n)
    spec = np.fft.rfft(w)
    f = np.arange(spec.size) + 1.0
    x = np.fft.irfft(spec / np.sqrt(f), n)
    x = x / np.abs(x).max()
    env = 0.05 + 0.95 * (0.5 + 0.5 * np.sin(2 * np.pi * np.arange(n) / 8000.0)) ** 2
    return np.round(x * env * 30000).astype(np.int16)


def assert_logmel_close(got, want64, eps=1e-12, what=""):
    """Stated tolerance: log domain above the energy floor, linear domain below it."""
    got = np.asarray(got, dtype=np.float64)
    want64 = np.asarray(want64, dtype=np.float64)
    assert got.shape == want64.shape, (what, got.shape, want64.shape)
    mel_want = np.exp(want64) - eps
    mel_got = np.exp(got) - eps
    e_frame = np.maximum(mel_want, 0.0).sum(axis=-2 if want64.ndim >= 3 else -1, keepdims=True)
    hi = mel_want >= np.maximum(MEL_FLOOR, REL_FLOOR * e_frame)
    if hi.any():
        err = np.abs(got - want64)[hi].max()
        assert err <= LOG_TOL, "%s: max |dlog| %.3e > %.1e above the floor" % (what, err, LOG_TOL)
    lo = ~hi
    if lo.any():
        d = np.abs(mel_got - mel_want)
        bound = LIN_ABS + LIN_REL * np.abs(mel_want) + REL_LIN * e_frame
        assert (d[lo] <= np.broadcast_to(bound, d.shape)[lo]).all(), \
            "%s: linear-domain error %.3e above bound" % (what, (d - bound)[lo].max())
