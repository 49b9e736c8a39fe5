"""The C-ABI library loads and exports every symbol include/b200mel.h declares (no compute)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    with open(os.path.join(ROOT, "include", "b200mel.h")) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200mel_[a-z_0-9]+)\s*\(", text)))


def test_header_declares_expected_entry_points():
    syms = header_symbols()
    for name in ["b200mel_create", "b200mel_destroy", "b200mel_forward", "b200mel_forward_host",
                 "b200mel_num_frames", "b200mel_last_error", "b200mel_specaug_draw", "b200mel_specaug_apply",
                 "b200mel_delta"]:
        assert name in syms


def test_library_exports_every_declared_symbol(built_lib):
    lib = ctypes.CDLL(built_lib)
    for name in header_symbols():
        assert hasattr(lib, name), "libb200mel.so does not export %s" % name


def test_python_binding_covers_header(built_lib):
    from speech_recognition_b200 import _lib

    assert sorted(_lib.SIGNATURES) == header_symbols()
    lib = _lib.load()
    assert lib.b200mel_version() == 100


def test_config_struct_matches_header():
    from speech_recognition_b200 import _lib

    with open(os.path.join(ROOT, "include", "b200mel.h")) as f:
        text = f.read()
    body = re.search(r"typedef struct b200mel_config \{(.*?)\} b200mel_config;", text, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        names += [n.strip() for n in decl.split(None, 1)[1].split(",")]
    assert names == [n for n, _ in _lib.B200MelConfig._fields_]


def test_argument_errors_without_gpu(built_lib):
    """Status codes and messages that need no device: null / bad arguments; and on a machine
    without CUDA the create call reports NO_DEVICE instead of silently computing on the CPU."""
    from speech_recognition_b200 import _lib

    lib = _lib.load()
    assert lib.b200mel_create(None, 0, None) == _lib.ERR_INVALID_ARGUMENT
    cfg = _lib.B200MelConfig()
    cfg.struct_size = 4
    h = ctypes.c_void_p()
    assert lib.b200mel_create(ctypes.byref(cfg), 0, ctypes.byref(h)) == _lib.ERR_INVALID_ARGUMENT
    assert b"struct_size" in lib.b200mel_last_error()
    import torch

    if not torch.cuda.is_available():
        cfg.struct_size = ctypes.sizeof(_lib.B200MelConfig)
        cfg.audio_feature_type, cfg.frame_length, cfg.frame_step, cfg.fft_length = 0, 320, 160, 320
        assert lib.b200mel_create(ctypes.byref(cfg), 0, ctypes.byref(h)) == _lib.ERR_NO_DEVICE
        with pytest.raises(RuntimeError):
            from speech_recognition_b200 import FeatureFrontend

            FeatureFrontend("spectrogram", 16000, 320, 160, 320)
