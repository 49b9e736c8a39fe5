"""DataConfig / SpecAugmentConfig: the config SCHEMA IS REPRODUCED FROM THE REFERENCE
(speech_recognition/configs/data_config.py:10-106) key for key - field names, defaults, validation and
properties are the drop-in contract of SURVEY 8(b), not original work; only the callables they return
(this package's CUDA-backed front-end) are.

Same YAML keys, validation and properties, so ``resources/configs/{libri,kspon,clovacall}_config.yml``
load unmodified.  ``audio_feature_fn`` returns this package's CUDA-backed per-example callables;
``frontend(training)`` returns the fused batched front-end for the same keys.
"""
from typing import Optional

import yaml
from pydantic import ValidationError
from pydantic.dataclasses import dataclass
from typing_extensions import Literal

from ..data import make_log_mel_spectrogram, make_mfcc, make_spectrogram, spec_augment


def _missing_args_raise_type_error(cls):
    """The reference runs pydantic v1, whose dataclasses raise TypeError for missing arguments
    (asserted at tests/configs/test_data_config.py:17-18); pydantic v2 would raise ValidationError.
    Bind against the signature first so the reference's behaviour is kept on either version."""
    import dataclasses
    import functools

    init = cls.__init__
    names = [f.name for f in dataclasses.fields(cls)]
    required = [f.name for f in dataclasses.fields(cls)
                if f.default is dataclasses.MISSING and f.default_factory is dataclasses.MISSING]

    @functools.wraps(init)
    def __init__(self, *args, **kwargs):
        if len(args) > len(names):
            raise TypeError("__init__() takes %d positional arguments but %d were given" % (len(names), len(args)))
        given = set(names[: len(args)]) | set(kwargs)
        missing = [n for n in required if n not in given]
        if missing:
            raise TypeError("__init__() missing %d required positional argument%s: %s"
                            % (len(missing), "s" if len(missing) > 1 else "", ", ".join(repr(m) for m in missing)))
        try:
            init(self, *args, **kwargs)
        except ValidationError as e:
            # pydantic v1 lets __post_init__'s AssertionError escape as-is
            errs = e.errors()
            if errs and all(err.get("type") == "assertion_error" for err in errs):
                raise AssertionError(errs[0].get("msg", "")) from None
            raise

    cls.__init__ = __init__
    return cls


@dataclass
class SpecAugmentConfig:
    """Parameters for SpecAugment (configs/data_config.py:10-20)."""

    enable: bool
    W: Optional[int] = None
    F: Optional[int] = None
    m_F: Optional[int] = None
    T: Optional[int] = None
    p: Optional[float] = None
    m_T: Optional[int] = None


@_missing_args_raise_type_error
@dataclass
class DataConfig:
    """Config for audio data processing (configs/data_config.py:23-54)."""

    file_format: Literal["pcm", "wav", "flac", "mp3"]
    audio_feature_type: Literal["spectrogram", "log-mel-spectrogram", "mfcc"]
    sample_rate: int
    frame_length: int
    frame_step: int
    fft_length: int
    max_audio_length: int
    max_token_length: int
    use_delta_accelerate: bool
    spec_augment: SpecAugmentConfig
    num_mel_bins: Optional[int] = None
    num_mfcc: Optional[int] = None
    lower_edge_hertz: Optional[float] = None
    upper_edge_hertz: Optional[float] = None

    def __post_init__(self):
        # configs/data_config.py:56-62
        if self.audio_feature_type in ["log-mel-spectrogram", "mfcc"]:
            assert all(
                [self.num_mel_bins, self.lower_edge_hertz, self.upper_edge_hertz]
            ), '"num_mel_bins", "lower_edge_hertz", "upper_edge_hertz" is required'
        if self.audio_feature_type == "mfcc":
            assert self.num_mfcc, '"num_mfcc" is required'

    @property
    def feature_dim(self):
        return 3 if self.use_delta_accelerate else 1

    @property
    def frequency_dim(self):
        if self.audio_feature_type == "spectrogram":
            return self.fft_length // 2 + 1
        if self.audio_feature_type == "log-mel-spectrogram":
            return self.num_mel_bins
        if self.audio_feature_type == "mfcc":
            return self.num_mfcc

    @property
    def audio_feature_fn(self):
        # configs/data_config.py:77-101
        if self.audio_feature_type == "spectrogram":
            return make_spectrogram(self.frame_length, self.frame_step, self.fft_length)
        if self.audio_feature_type == "log-mel-spectrogram":
            return make_log_mel_spectrogram(
                self.sample_rate,
                self.frame_length,
                self.frame_step,
                self.fft_length,
                self.num_mel_bins,
                self.lower_edge_hertz,
                self.upper_edge_hertz,
            )
        if self.audio_feature_type == "mfcc":
            return make_mfcc(
                self.sample_rate,
                self.frame_length,
                self.frame_step,
                self.fft_length,
                self.num_mel_bins,
                self.num_mfcc,
                self.lower_edge_hertz,
                self.upper_edge_hertz,
            )

    @property
    def spec_augment_fn(self):
        """The callable run/train.py:100-110 builds from this config."""
        sa = self.spec_augment
        return spec_augment(self.frequency_dim, sa.W, sa.F, sa.m_F, sa.T, sa.p, sa.m_T)

    def frontend(self, training: bool = False, device=None):
        """Fused batched front-end (feature -> SpecAugment if training and enabled -> delta)."""
        from ..frontend import FeatureFrontend

        return FeatureFrontend.from_config(self, training=training, device=device)

    @classmethod
    def from_yaml(cls, file_path) -> "DataConfig":
        with open(file_path) as f:
            return cls(**yaml.load(f, yaml.SafeLoader))
