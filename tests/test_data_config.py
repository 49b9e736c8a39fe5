"""DataConfig mirror: same behaviour as the reference's tests/configs/test_data_config.py:13-26,
on the three shipped YAMLs."""
import os
from dataclasses import asdict

import pytest
import yaml
from pydantic import ValidationError

from speech_recognition_b200 import DataConfig

from .helpers import CONFIG_DIR


def _load(name):
    with open(os.path.join(CONFIG_DIR, name)) as f:
        return yaml.load(f, yaml.SafeLoader)


def test_data_config():
    correct_config = _load("libri_config.yml")
    with pytest.raises(TypeError):
        DataConfig()
    with pytest.raises(ValidationError):
        DataConfig(**{**correct_config, **{"file_format": "hello"}})
    config = DataConfig(**correct_config)
    assert asdict(config) == correct_config
    assert config.feature_dim == 3


@pytest.mark.parametrize("name,ftype,fdim", [("libri_config.yml", "log-mel-spectrogram", 80),
                                             ("kspon_config.yml", "mfcc", 40),
                                             ("clovacall_config.yml", "mfcc", 40)])
def test_shipped_configs_load_unmodified(name, ftype, fdim):
    config = DataConfig.from_yaml(os.path.join(CONFIG_DIR, name))
    assert config.audio_feature_type == ftype
    assert config.frequency_dim == fdim
    assert (config.sample_rate, config.frame_length, config.frame_step, config.fft_length) == (16000, 320, 160, 320)
    assert config.spec_augment.enable is False and config.spec_augment.W is None
    assert callable(config.audio_feature_fn)       # constructing the factory needs no GPU
    assert callable(config.spec_augment_fn)


def test_required_keys():
    cfg = _load("libri_config.yml")
    with pytest.raises(AssertionError):
        DataConfig(**{**cfg, "num_mel_bins": None})
    with pytest.raises(AssertionError):
        DataConfig(**{**cfg, "audio_feature_type": "mfcc", "num_mfcc": None})
    spec = DataConfig(**{**cfg, "audio_feature_type": "spectrogram", "num_mel_bins": None})
    assert spec.frequency_dim == 161


def test_time_warp_factory_is_accepted():
    """data.py:244-280: W is a parameter of the factory like the others (the wrapper needs a GPU: tests/test_time_warp.py)."""
    from speech_recognition_b200 import spec_augment

    assert callable(spec_augment(80, W=40, F=27, m_F=1, T=100, p=1.0, m_T=1))
