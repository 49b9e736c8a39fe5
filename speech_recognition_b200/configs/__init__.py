from .data_config import DataConfig, SpecAugmentConfig  # noqa: F401
