"""CPU restatement of the rows either side of the hot path.  TEST INFRASTRUCTURE ONLY:
imported by tests/ only, never by the product path.

* ``load_audio_pcm16`` follows ``load_audio_file`` (speech_recognition/data.py:94-117) for the ``wav`` and
  ``pcm`` formats: int16 little-endian payload -> float32 / 32768 -> mean over channels.
* ``las_audio_mask`` follows ``LAS Listener._audio_mask`` (speech_recognition/models/las.py:205-217) line by
  line; ``conv_audio_mask`` is the same reduction for an arbitrary list of (time kernel, time stride) layers
  (speech_recognition/models/deepspeech2.py:68-78).

Parity pin status: the reference pins none of these numerically (its only audio fixtures are digital
silence, tests/data/audio_files/*); the WAV case is cross-checked against Python's ``wave`` module in
tests/test_ingest.py.
"""
import struct

import numpy as np


def wav_payload(image: bytes):
    """(channels, sample_rate, bits, payload bytes) of a RIFF/WAVE image (chunk walk)."""
    assert image[:4] == b"RIFF" and image[8:12] == b"WAVE"
    pos, fmt = 12, None
    while pos + 8 <= len(image):
        cid = image[pos:pos + 4]
        (size,) = struct.unpack("<I", image[pos + 4:pos + 8])
        if cid == b"fmt ":
            fmt = struct.unpack("<HHIIHH", image[pos + 8:pos + 24])
        elif cid == b"data":
            return fmt[1], fmt[2], fmt[5], image[pos + 8:pos + 8 + size]
        pos += 8 + size + (size & 1)
    raise ValueError("no data chunk")


def load_audio_pcm16(image: bytes, file_format: str) -> np.ndarray:
    """data.py:94-117 for ``wav`` / ``pcm``: float32 [TimeStep]."""
    if file_format == "wav":
        channels, _, bits, payload = wav_payload(image)
        assert bits == 16
        n = len(payload) // (2 * channels)
        audio = np.frombuffer(payload[:n * 2 * channels], "<i2").reshape(n, channels).astype(np.float32) / np.float32(32768.0)
    elif file_format == "pcm":
        if len(image) % 2 == 1:
            image = image + b"\x00"  # data.py:102-103
        audio = np.frombuffer(image, "<i2").astype(np.float32)[:, None] / np.float32(32768.0)  # data.py:104-105
    else:
        raise ValueError(f"File Format: {file_format} is not valid!")
    # tf.reduce_mean(audio, 1) (data.py:116): float32 sum over channels, then one division
    return (audio.sum(axis=1, dtype=np.float32) / np.float32(audio.shape[1])).astype(np.float32)


def las_audio_mask(audio: np.ndarray, kernel_size: int, strides: int) -> np.ndarray:
    """models/las.py:205-217, statement by statement (AUDIO_PAD_VALUE = 0)."""
    batch_size, sequence_length = audio.shape[:2]
    mask = np.any(audio.reshape(batch_size, sequence_length, -1) != 0.0, axis=2)
    sequence_length -= kernel_size - strides
    sequence_length = sequence_length // strides
    sequence_length -= kernel_size - strides
    sequence_length = sequence_length // strides
    sequence_length *= strides ** 2
    sequence_length = max(sequence_length, 0)
    mask = mask[:, :sequence_length].reshape(batch_size, -1, strides ** 2)
    return np.any(mask, axis=2)


def conv_audio_mask(audio: np.ndarray, time_kernel_sizes, time_strides) -> np.ndarray:
    """models/deepspeech2.py:68-78 with ``stride_complex`` read as the product of the time strides."""
    batch_size, sequence_length = audio.shape[:2]
    mask = np.any(audio.reshape(batch_size, sequence_length, -1) != 0.0, axis=2)
    group = 1
    for k, s in zip(time_kernel_sizes, time_strides):
        sequence_length -= k - s
        sequence_length = sequence_length // s
        group *= s
    sequence_length = max(sequence_length, 0)
    mask = mask[:, :sequence_length * group].reshape(batch_size, sequence_length, group)
    return np.any(mask, axis=2)
