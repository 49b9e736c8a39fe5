"""ctypes binding of lib/libb200mel.so (C ABI declared in include/b200mel.h).

The library is loaded lazily on first use so that configuration code can be imported on a machine
without a GPU; any attempt to *compute* without the library or a CUDA device raises - there is no
CPU fallback anywhere in this package.
"""
import ctypes
import os
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
# B200MEL_LIB: developer override used by experiments/run_variants.sh to A/B kernel variants
LIB_PATH = os.environ.get("B200MEL_LIB") or os.path.join(HERE, "lib", "libb200mel.so")

OK, ERR_INVALID_ARGUMENT, ERR_UNSUPPORTED, ERR_CUDA, ERR_NO_DEVICE = 0, 1, 2, 3, 4
FEATURE_TYPES = {"spectrogram": 0, "log-mel-spectrogram": 1, "mfcc": 2}
PCM_I16, PCM_F32 = 0, 1
KERNELS = {"auto": 0, "generic": 1, "tc": 2}


class B200MelConfig(ctypes.Structure):
    _fields_ = [
        ("struct_size", ctypes.c_int32),
        ("audio_feature_type", ctypes.c_int32),
        ("sample_rate", ctypes.c_int32),
        ("frame_length", ctypes.c_int32),
        ("frame_step", ctypes.c_int32),
        ("fft_length", ctypes.c_int32),
        ("num_mel_bins", ctypes.c_int32),
        ("num_mfcc", ctypes.c_int32),
        ("lower_edge_hertz", ctypes.c_float),
        ("upper_edge_hertz", ctypes.c_float),
        ("epsilon", ctypes.c_float),
        ("use_delta_accelerate", ctypes.c_int32),
        ("spec_augment_enable", ctypes.c_int32),
        ("sa_F", ctypes.c_int32),
        ("sa_m_F", ctypes.c_int32),
        ("sa_T", ctypes.c_int32),
        ("sa_m_T", ctypes.c_int32),
        ("sa_p", ctypes.c_float),
    ]


_c = ctypes
_vp, _i32, _i64, _u64 = _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_uint64

# name -> (restype, argtypes); must list every symbol include/b200mel.h declares
SIGNATURES = {
    "b200mel_version": (_c.c_int, []),
    "b200mel_last_error": (_c.c_char_p, []),
    "b200mel_create": (_c.c_int, [_c.POINTER(B200MelConfig), _c.c_int, _c.POINTER(_vp)]),
    "b200mel_destroy": (_c.c_int, [_vp]),
    "b200mel_num_frames": (_c.c_int, [_vp, _i64, _c.POINTER(_i64)]),
    "b200mel_feature_shape": (_c.c_int, [_vp, _c.POINTER(_i32), _c.POINTER(_i32), _c.POINTER(_i32)]),
    "b200mel_mel_weights": (_c.c_int, [_vp, _vp]),
    "b200mel_tc_supported": (_c.c_int, [_vp]),
    "b200mel_launch_count": (_i64, [_vp]),
    "b200mel_forward": (_c.c_int, [_vp, _vp, _i32, _vp, _vp, _i32, _i64, _vp, _i64, _vp, _vp, _u64, _u64,
                                   _vp, _i32, _vp]),
    "b200mel_forward_host": (_c.c_int, [_vp, _vp, _i32, _vp, _vp, _i32, _i64, _i64, _vp, _i64, _vp, _u64,
                                        _u64, _i32]),
    "b200mel_forward_host_submit": (_c.c_int, [_vp, _vp, _i32, _vp, _vp, _i32, _i64, _i64, _vp, _i64, _vp, _u64,
                                               _u64, _i32, _i32]),
    "b200mel_forward_host_wait": (_c.c_int, [_vp, _i32]),
    # (no allocation inside the call; no counterpart in the reference - see include/b200mel.h)
    "b200mel_forward_mfcc_noalloc": (_c.c_int, [_vp, _vp, _vp, _vp, _i32, _i64, _vp, _i64, _vp, _vp, _u64, _u64,
                                                _vp, _vp]),
    "b200mel_specaug_draw": (_c.c_int, [_vp, _i32, _i32, _i32, _i32, _i32, _c.c_float, _i32, _u64, _u64,
                                        _vp, _vp, _vp]),
    "b200mel_specaug_apply": (_c.c_int, [_vp, _vp, _i32, _i64, _i32, _i32, _vp, _i32, _i32, _vp]),
    "b200mel_delta": (_c.c_int, [_vp, _vp, _vp, _i32, _i64, _i32, _vp]),
    "b200mel_time_warp": (_c.c_int, [_vp, _vp, _vp, _i32, _i64, _i32, _i32, _i32, _vp, _u64, _u64, _vp, _vp]),
    "b200mel_wav_parse": (_c.c_int, [_vp, _i64, _c.POINTER(_i32), _c.POINTER(_i32), _c.POINTER(_i32),
                                     _c.POINTER(_i32), _c.POINTER(_i64), _c.POINTER(_i64)]),
    "b200mel_ingest_pcm16": (_c.c_int, [_vp, _vp, _vp, _vp, _i32, _vp, _i32, _vp, _i64, _vp, _vp]),
    "b200mel_tfrecord_masked_crc": (_c.c_uint32, [_vp, _i64]),
    "b200mel_audio_mask_length": (_c.c_int, [_i64, _vp, _vp, _i32, _c.POINTER(_i64), _c.POINTER(_i32)]),
    "b200mel_audio_mask": (_c.c_int, [_vp, _i32, _i64, _i64, _i32, _i64, _vp, _vp]),
}

_lock = threading.Lock()
_lib = None


def load():
    """Load libb200mel.so (once).  Raises RuntimeError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    "%s is missing: build it with `python -m speech_recognition_b200.build` "
                    "(or __graft_entry__.build()).  There is no CPU fallback." % LIB_PATH)
            lib = ctypes.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(lib, name)
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


def last_error():
    msg = load().b200mel_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(status):
    """Map a C status to the exception the reference's stack would raise for the same condition:
    bad parameters -> ValueError (TF: ValueError / InvalidArgumentError); the rest -> RuntimeError."""
    if status == OK:
        return
    msg = last_error()
    if status == ERR_INVALID_ARGUMENT:
        raise ValueError(msg)
    if status == ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    if status == ERR_NO_DEVICE:
        raise RuntimeError("b200mel: " + msg)
    raise RuntimeError("b200mel: CUDA failure: " + msg)
