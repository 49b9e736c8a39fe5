"""Drop-in mirror of the reference's feature factories (speech_recognition/data.py:122-328).

Same names, positional order, defaults and return conventions: each factory returns a per-example
``_wrapper(audio, text=None)`` that yields ``feat`` or ``(feat, text)``, so a pipeline written for
the reference maps these exactly where it mapped the originals (run/train.py:88-116,
run/evaluate.py:59-66, run/inference.py:57-66, run/make_tfrecord.py:52-56).  Tensors are torch
CUDA tensors; the arithmetic runs in libb200mel.so (sm_100a kernels).  There is no CPU path: host
arrays are copied to the current CUDA device, and without one every call raises.

The fused, batched form of the same pipeline is ``FeatureFrontend`` (frontend.py), reachable from a
wrapper as ``wrapper.frontend()``.
"""
import itertools
import threading
from typing import Optional

import numpy as np
import torch

from . import frontend as _fe

# ----------------------------------------------------------------------------------------------
# RNG state: the reference seeds a process-global TF generator (utils.py:123-127).  Here the
# "generator" is a counter: utterance k of the process draws Philox(seed, k), so results are
# reproducible and independent of batching / sharding.
# ----------------------------------------------------------------------------------------------
_rng_lock = threading.Lock()
_seed = 0
_utterance_counter = itertools.count()


def set_random_seed(seed: int):
    """Mirror of utils.set_random_seed (utils.py:123-127) for the SpecAugment draw stream."""
    global _seed, _utterance_counter
    with _rng_lock:
        _seed = int(seed)
        _utterance_counter = itertools.count()


def _next_utterance():
    with _rng_lock:
        return _seed, next(_utterance_counter)


def _to_device_audio(audio) -> torch.Tensor:
    if not torch.cuda.is_available():
        raise RuntimeError("speech_recognition_b200 needs a CUDA device: there is no CPU fallback")
    if isinstance(audio, torch.Tensor):
        t = audio
    else:
        t = torch.from_numpy(np.ascontiguousarray(audio))
    if t.dtype not in (torch.float32, torch.int16):
        t = t.to(torch.float32)
    if t.dim() != 1:
        raise ValueError("audio must be shaped [TimeStep]")
    return t if t.is_cuda else t.cuda(non_blocking=True)


def _to_device_feature(feat) -> torch.Tensor:
    if not torch.cuda.is_available():
        raise RuntimeError("speech_recognition_b200 needs a CUDA device: there is no CPU fallback")
    t = feat if isinstance(feat, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(feat))
    t = t.to(torch.float32)
    if t.dim() != 3:
        raise ValueError("feature must be shaped [TimeStep, AudioDim, Channel]")
    return t if t.is_cuda else t.cuda(non_blocking=True)


class _LazyFrontend:
    """Creates the device plan on first use (like @tf.function traces on first call), one per
    CUDA device, so factories can be constructed on a machine without a GPU."""

    def __init__(self, **kwargs):
        self.kwargs = kwargs
        self._cache = {}
        self._lock = threading.Lock()

    def get(self, device=None) -> _fe.FeatureFrontend:
        if not torch.cuda.is_available():
            raise RuntimeError("speech_recognition_b200 needs a CUDA device: there is no CPU fallback")
        idx = torch.cuda.current_device() if device is None else torch.device(device).index
        with self._lock:
            if idx not in self._cache:
                self._cache[idx] = _fe.FeatureFrontend(device=torch.device("cuda", idx), **self.kwargs)
            return self._cache[idx]


def _feature_wrapper(lazy: _LazyFrontend):
    def _wrapper(audio, text=None):
        audio = _to_device_audio(audio)
        feat, _ = lazy.get(audio.device)(audio[None])
        feat = feat[0]
        if text is None:
            return feat
        return feat, text

    _wrapper.frontend = lazy.get
    return _wrapper


def make_spectrogram(frame_length: int, frame_step: int, fft_length=None):
    """data.py:122-142.  Returns a wrapper: audio [TimeStep] -> [NumFrame, NumFFTUniqueBins, 1]."""
    return _feature_wrapper(_LazyFrontend(audio_feature_type="spectrogram", sample_rate=0,
                                          frame_length=frame_length, frame_step=frame_step,
                                          fft_length=fft_length))


def make_log_mel_spectrogram(
    sample_rate: int,
    frame_length: int,
    frame_step: int,
    fft_length: int,
    num_mel_bins: int = 80,
    lower_edge_hertz: float = 80.0,
    upper_edge_hertz: float = 7600.0,
    epsilon: float = 1e-12,
):
    """data.py:145-189.  Returns a wrapper: audio [TimeStep] -> [NumFrame, NumMelFilterbank, 1]."""
    return _feature_wrapper(_LazyFrontend(audio_feature_type="log-mel-spectrogram", sample_rate=sample_rate,
                                          frame_length=frame_length, frame_step=frame_step,
                                          fft_length=fft_length, num_mel_bins=num_mel_bins,
                                          lower_edge_hertz=lower_edge_hertz, upper_edge_hertz=upper_edge_hertz,
                                          epsilon=epsilon))


def make_mfcc(
    sample_rate: int,
    frame_length: int,
    frame_step: int,
    fft_length: int,
    num_mel_bins: int = 80,
    num_mfcc: int = 40,
    lower_edge_hertz: float = 80.0,
    upper_edge_hertz: float = 7600.0,
    epsilon: float = 1e-12,
):
    """data.py:192-241.  Returns a wrapper: audio [TimeStep] -> [NumFrame, NumMFCC, 1]."""
    return _feature_wrapper(_LazyFrontend(audio_feature_type="mfcc", sample_rate=sample_rate,
                                          frame_length=frame_length, frame_step=frame_step,
                                          fft_length=fft_length, num_mel_bins=num_mel_bins, num_mfcc=num_mfcc,
                                          lower_edge_hertz=lower_edge_hertz, upper_edge_hertz=upper_edge_hertz,
                                          epsilon=epsilon))


def spec_augment(
    v: int,
    W: Optional[int] = None,
    F: Optional[int] = None,
    m_F: Optional[int] = None,
    T: Optional[int] = None,
    p: Optional[float] = None,
    m_T: Optional[int] = None,
):
    """data.py:244-307 (time warping, frequency and time masking).  The wrapper takes a feature
    [TimeStep, v, Channel] and returns the augmented copy.  Draw order and integer mapping follow the
    reference ((f, f_0) x m_F then (t, t_0) x m_T; lo + u32 % range); the stream is
    Philox4x32-10 keyed by (set_random_seed seed, running utterance index); the two draws of the time
    warp (``W``, data.py:275-280: tfa.image.sparse_image_warp, restated in oracle/time_warp_oracle.py) come
    from a second stream of the same key, so the masks are the same with or without the warp.  A draw that
    TF would reject (``maxval <= minval``) raises ValueError."""

    def _wrapper(audio, text=None, utterance_id: Optional[int] = None, seed: Optional[int] = None):
        feat = _to_device_feature(audio)
        g_seed, g_utt = _next_utterance()
        seed_ = g_seed if seed is None else seed
        utt = g_utt if utterance_id is None else utterance_id
        if W:
            warped, st_w = _fe.time_warp(feat[None], [feat.shape[0]], int(W), seed_, utt)
            if int(st_w.item()) != 0:
                raise ValueError("spec_augment: time warp draw had maxval <= minval (num_time - W <= W; "
                                 "TF raises InvalidArgumentError here)")
            feat = warped[0]
        table, nf, nt, status = _fe.specaug_draw([feat.shape[0]], v, F, m_F, T, p, m_T, seed_, utt, feat.device)
        if nf + nt > 0:
            feat = _fe.specaug_apply(feat[None], table, nf, nt)[0]
            if int(status.item()) != 0:
                raise ValueError("spec_augment: a uniform draw had maxval <= minval "
                                 "(mask wider than the feature; TF raises InvalidArgumentError here)")
        if text is None:
            return feat
        return feat, text

    return _wrapper


def delta_accelerate(audio, text=None):
    """data.py:310-328: [TimeStep, AudioDim, 1] -> [TimeStep, AudioDim, 3]."""
    feat = _to_device_feature(audio)
    out = _fe.delta(feat[None])[0]
    if text is None:
        return out
    return out, text
