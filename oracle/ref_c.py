"""ctypes wrapper + build recipe for oracle/logmel_ref.c.  TEST INFRASTRUCTURE ONLY (see that file).

Build output goes to oracle/_ref/ (git-ignored, travels to the GPU box with the snapshot).
"""
import ctypes
import os
import shutil
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "logmel_ref.c")
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "liblogmel_ref.so")

FEATURE_TYPES = {"spectrogram": 0, "log-mel-spectrogram": 1, "mfcc": 2}


def build(force=False):
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= os.path.getmtime(SRC):
        return OUT
    gcc = shutil.which("gcc")
    if gcc is None:
        if os.path.exists(OUT):
            return OUT
        raise RuntimeError("gcc not found and %s is not built" % OUT)
    os.makedirs(OUT_DIR, exist_ok=True)
    # x86-64-v3 (AVX2+FMA) rather than -march=native: the .so is built here and run on the GPU box.
    # -ffp-contract=off keeps float32 products and sums separately rounded, like TF's CPU kernels.
    cmd = [gcc, "-O3", "-march=x86-64-v3", "-ffp-contract=off", "-fopenmp", "-shared", "-fPIC", "-o", OUT, SRC, "-lm"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("gcc failed:\n" + res.stdout + res.stderr)
    return OUT


_lib = None


def load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(build())
        lib.logmel_ref_plan_create.restype = ctypes.c_void_p
        lib.logmel_ref_plan_create.argtypes = [ctypes.c_int] * 7 + [ctypes.c_float] * 3 + [ctypes.c_int]
        lib.logmel_ref_plan_destroy.argtypes = [ctypes.c_void_p]
        lib.logmel_ref_feature_dim.argtypes = [ctypes.c_void_p]
        lib.logmel_ref_num_threads.restype = ctypes.c_int
        lib.logmel_ref_batch.restype = ctypes.c_int
        lib.logmel_ref_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_void_p,
                                         ctypes.c_int64, ctypes.c_void_p, ctypes.c_int]
        _lib = lib
    return _lib


class RefPlan:
    def __init__(self, audio_feature_type, sample_rate, frame_length, frame_step, fft_length=None, num_mel_bins=80,
                 num_mfcc=40, lower_edge_hertz=80.0, upper_edge_hertz=7600.0, epsilon=1e-12,
                 use_delta_accelerate=False):
        self.lib = load()
        self.frame_length, self.frame_step = frame_length, frame_step
        self.C = 3 if use_delta_accelerate else 1
        self.h = self.lib.logmel_ref_plan_create(FEATURE_TYPES[audio_feature_type], sample_rate, frame_length,
                                                 frame_step, fft_length or 0, num_mel_bins or 0, num_mfcc or 0,
                                                 lower_edge_hertz or 0.0, upper_edge_hertz or 0.0, epsilon,
                                                 1 if use_delta_accelerate else 0)
        self.F = self.lib.logmel_ref_feature_dim(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.logmel_ref_plan_destroy(self.h)
            self.h = None

    def num_threads(self):
        return self.lib.logmel_ref_num_threads()

    def __call__(self, pcm, lengths=None, offsets=None, t_max=None, n_threads=0, out=None):
        pcm = np.ascontiguousarray(pcm)
        assert pcm.dtype in (np.int16, np.float32)
        if offsets is None:
            if pcm.ndim == 1:
                pcm = pcm[None]
            B, L = pcm.shape
            lens = np.full(B, L, np.int32) if lengths is None else np.ascontiguousarray(lengths, np.int32)
            offs, stride = None, L
        else:
            offs = np.ascontiguousarray(offsets, np.int64)
            lens = np.ascontiguousarray(lengths, np.int32)
            B, stride = len(offs), 0
        if t_max is None:
            lm = int(lens.max()) if B else 0
            t_max = 0 if lm < self.frame_length else 1 + (lm - self.frame_length) // self.frame_step
        if out is None:
            out = np.empty((B, t_max, self.F, self.C), np.float32)
        nf = np.zeros(B, np.int32)
        vp = ctypes.c_void_p
        used = self.lib.logmel_ref_batch(self.h, pcm.ctypes.data_as(vp), int(pcm.dtype == np.float32),
                                         offs.ctypes.data_as(vp) if offs is not None else None,
                                         lens.ctypes.data_as(vp), B, stride, out.ctypes.data_as(vp), t_max,
                                         nf.ctypes.data_as(vp), n_threads)
        self.threads_used = used
        return out, nf
