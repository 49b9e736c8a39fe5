"""speech_recognition_b200 - the audio -> log-mel / MFCC / spectrogram (+SpecAugment, +delta)
front-end of cosmoquester/speech-recognition (speech_recognition/data.py:122-328) as hand-written
sm_100a CUDA kernels behind a C ABI (include/b200mel.h, lib/libb200mel.so).

There is no CPU implementation in this package: every entry point runs on a CUDA device and fails
loudly if the library or a device is missing.
"""
from . import batching, data, feature_cache, ingest  # noqa: F401
from .configs import DataConfig, SpecAugmentConfig  # noqa: F401
from .data import (  # noqa: F401
    delta_accelerate,
    make_log_mel_spectrogram,
    make_mfcc,
    make_spectrogram,
    set_random_seed,
    spec_augment,
)
from .frontend import FeatureFrontend  # noqa: F401
from .ingest import load_audio_file  # noqa: F401

__all__ = [
    "DataConfig",
    "SpecAugmentConfig",
    "FeatureFrontend",
    "make_spectrogram",
    "make_log_mel_spectrogram",
    "make_mfcc",
    "spec_augment",
    "delta_accelerate",
    "set_random_seed",
    "load_audio_file",
    "ingest",
    "batching",
    "feature_cache",
]
