"""PCM / WAV ingest straight to the front-end's int16 device entry.

The reference's ``load_audio_file`` (speech_recognition/data.py:82-119) decodes a file into
``float32 [TimeStep, NumChannel] = int16 / 32768`` on the host and averages the channels.  For the
two formats whose payload already *is* little-endian int16 - ``wav`` (data.py:96-99) and KsponSpeech raw
``pcm`` (data.py:100-105, with its zero byte appended to odd-length files) - nothing needs decoding: the
file bytes go to the device as they are and ``b200mel_ingest_pcm16`` turns the payload ranges into the
``[B, L]`` int16 batch (+ ``lengths``) that ``FeatureFrontend`` consumes, or into the reference's float32
channel mean.  ``flac`` / ``mp3`` decoding and ``tfio.audio.resample`` stay with the caller's host decoder
(out of scope).

torch is used for pinned / device memory and streams only; header parsing and the byte work are in
lib/libb200mel.so.  There is no CPU path.
"""
import ctypes
from typing import Callable, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from . import _lib

FILE_FORMATS = ("flac", "wav", "pcm", "mp3")  # data.py:90
_DEVICE_FORMATS = ("wav", "pcm")


class WavHeader:
    """Position and shape of the sample data of a RIFF/WAVE file image."""

    __slots__ = ("format", "channels", "sample_rate", "bits_per_sample", "data_offset", "data_bytes")

    def __init__(self, format, channels, sample_rate, bits_per_sample, data_offset, data_bytes):
        self.format, self.channels, self.sample_rate = format, channels, sample_rate
        self.bits_per_sample, self.data_offset, self.data_bytes = bits_per_sample, data_offset, data_bytes

    def __repr__(self):
        return "WavHeader(format=%d, channels=%d, sample_rate=%d, bits_per_sample=%d, data_offset=%d, data_bytes=%d)" % (
            self.format, self.channels, self.sample_rate, self.bits_per_sample, self.data_offset, self.data_bytes)


def parse_wav_header(file_bytes: Union[bytes, bytearray, memoryview, np.ndarray]) -> WavHeader:
    """Walk the RIFF chunks of a WAV file image (host, ``b200mel_wav_parse``).  Raises ``ValueError`` for
    anything that is not a RIFF/WAVE image with a fmt and a data chunk."""
    buf = np.frombuffer(file_bytes, dtype=np.uint8) if not isinstance(file_bytes, np.ndarray) else file_bytes
    if buf.dtype != np.uint8 or buf.ndim != 1 or not buf.flags["C_CONTIGUOUS"]:
        raise ValueError("file image must be contiguous bytes")
    fmt, ch, sr, bits = (ctypes.c_int32() for _ in range(4))
    off, nbytes = ctypes.c_int64(), ctypes.c_int64()
    ptr = ctypes.c_void_p(buf.ctypes.data) if buf.size else None
    _lib.check(_lib.load().b200mel_wav_parse(ptr, int(buf.size), ctypes.byref(fmt), ctypes.byref(ch), ctypes.byref(sr),
                                             ctypes.byref(bits), ctypes.byref(off), ctypes.byref(nbytes)))
    return WavHeader(fmt.value, ch.value, sr.value, bits.value, off.value, nbytes.value)


def _payload(file_format: str, image: np.ndarray) -> Tuple[int, int, int]:
    """(payload offset, payload bytes, channels) of one file image."""
    if file_format == "pcm":
        return 0, int(image.size), 1
    h = parse_wav_header(image)
    if h.format != 1 or h.bits_per_sample != 16:
        raise ValueError("wav ingest handles 16-bit integer PCM only (got format tag %d, %d bits): decode on the host"
                         % (h.format, h.bits_per_sample))
    if h.channels < 1 or h.channels > 256:
        raise ValueError("wav: %d channels" % h.channels)
    # whole sample frames only (a trailing partial frame of a damaged file is dropped, like a WAV decoder does;
    # the zero-byte padding of odd lengths is the raw-pcm rule of data.py:102-103, not a WAV rule)
    frame = 2 * h.channels
    return h.data_offset, h.data_bytes // frame * frame, h.channels


def _ptr(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def ingest_file_images(images: Sequence[Union[bytes, np.ndarray]], file_format: str, device=None,
                       dtype: Optional[torch.dtype] = None, row_stride: Optional[int] = None,
                       ragged: bool = False):
    """File images (the bytes of whole ``.wav`` / ``.pcm`` files) -> samples on the device.

    Returns ``(pcm, lengths)`` - dense ``[B, row_stride]`` - or ``(pcm, offsets, lengths)`` when ``ragged``
    (1-D ``pcm``, ``offsets`` int64 element offsets rounded up to 8 samples).  ``dtype`` defaults to int16
    when every file is mono (the front-end's native input) and to float32 - the reference's channel mean of
    ``int16 / 32768`` (data.py:99, 116) - otherwise.  Asynchronous on the current torch stream."""
    if file_format not in FILE_FORMATS:
        raise ValueError(f"File Format: {file_format} is not valid!")  # data.py:109
    if file_format not in _DEVICE_FORMATS:
        raise NotImplementedError("%s files need a host decoder (tfio / ffmpeg): out of this front-end's scope; "
                                  "pass the decoded int16 samples to FeatureFrontend instead" % file_format)
    if not torch.cuda.is_available():
        raise RuntimeError("speech_recognition_b200 needs a CUDA device: there is no CPU fallback")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    imgs = [np.frombuffer(im, dtype=np.uint8) if not isinstance(im, np.ndarray) else im for im in images]
    B = len(imgs)
    meta = [_payload(file_format, im) for im in imgs]
    channels = np.array([m[2] for m in meta], dtype=np.int32)
    mono = bool((channels == 1).all())
    if dtype is None:
        dtype = torch.int16 if mono else torch.float32
    if dtype not in (torch.int16, torch.float32):
        raise ValueError("dtype must be torch.int16 or torch.float32")
    if dtype == torch.int16 and not mono:
        raise ValueError("multi-channel files are channel means (data.py:116) and need float32 output")
    # one pinned staging buffer, payloads 16-byte aligned where the header allows nothing better
    starts = np.zeros(B, dtype=np.int64)
    total = 0
    for i, im in enumerate(imgs):
        starts[i] = total
        total += (int(im.size) + 15) // 16 * 16
    stage = torch.empty(max(total, 16), dtype=torch.uint8, pin_memory=True)
    stage_np = stage.numpy()
    for i, im in enumerate(imgs):
        stage_np[starts[i]:starts[i] + im.size] = im
    byte_off = starts + np.array([m[0] for m in meta], dtype=np.int64)
    byte_len = np.array([m[1] for m in meta], dtype=np.int64)
    n16 = (byte_len + 1) // 2
    n_samples = n16 // np.maximum(channels, 1)
    d_bytes = stage.to(dev, non_blocking=True)
    small = np.concatenate([byte_off, byte_len]).astype(np.int64)
    if ragged:
        out_off = np.zeros(B, dtype=np.int64)
        acc = 0
        for i in range(B):
            out_off[i] = acc
            acc += (int(n_samples[i]) + 7) // 8 * 8
        small = np.concatenate([small, out_off])
        pcm = torch.zeros(max(acc, 8), dtype=dtype, device=dev)
        row_stride = 0
    else:
        if row_stride is None:
            row_stride = (int(n_samples.max()) + 7) // 8 * 8 if B else 0
        pcm = torch.empty((B, int(row_stride)), dtype=dtype, device=dev)
    d_small = torch.from_numpy(small).to(dev, non_blocking=True)
    d_off, d_len = d_small[:B], d_small[B:2 * B]
    d_out_off = d_small[2 * B:3 * B] if ragged else None
    d_ch = None if mono else torch.from_numpy(channels).to(dev, non_blocking=True)
    lengths = torch.empty(B, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev)
    _lib.check(_lib.load().b200mel_ingest_pcm16(
        _ptr(d_bytes), _ptr(d_off), _ptr(d_len), _ptr(d_ch), B, _ptr(pcm),
        _lib.PCM_I16 if dtype == torch.int16 else _lib.PCM_F32, _ptr(d_out_off), int(row_stride), _ptr(lengths),
        ctypes.c_void_p(stream.cuda_stream)))
    for t in (d_bytes, d_small, d_ch):
        if t is not None:
            t.record_stream(stream)
    if ragged:
        return pcm, d_out_off, lengths
    return pcm, lengths


def ingest_files(paths: Sequence[str], file_format: str, **kwargs):
    """``ingest_file_images`` over files read from disk."""
    images: List[np.ndarray] = [np.fromfile(p, dtype=np.uint8) for p in paths]
    return ingest_file_images(images, file_format, **kwargs)


def load_audio_file(sample_rate: int, file_format: str, resample: Optional[float] = None
                    ) -> Callable[[str], torch.Tensor]:
    """Mirror of ``load_audio_file`` (data.py:82-119): returns ``_wrapper(audio_file_path)`` giving the
    float32 ``[TimeStep]`` audio tensor (int16 / 32768, channel mean) on the current CUDA device."""
    if file_format not in FILE_FORMATS:
        raise ValueError(f"File Format: {file_format} is not valid!")
    if resample is not None:
        raise NotImplementedError("tfio.audio.resample (data.py:112-113) is outside this front-end's scope")

    def _wrapper(audio_file_path) -> torch.Tensor:
        if isinstance(audio_file_path, bytes):
            audio_file_path = audio_file_path.decode("utf-8")
        pcm, lengths = ingest_files([str(audio_file_path)], file_format, dtype=torch.float32)
        return pcm[0, :int(lengths[0])]

    return _wrapper
