"""Regenerate tests/golden/*.npz from the reference checkout (run in the build container only).

    python tests/golden/make_golden.py [/root/reference]

Reads the reference's only numeric fixture for the hot path,
``tests/data/wav_dataset.tfrecord`` (compared with exact equality by the reference at
``tests/test_data.py:53-57``) plus the audio it was made from, and stores them as small npz files
so the parity tests can run where /root/reference does not exist (the GPU box).
"""
import os
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))

from oracle.tfrecord_io import read_feature_cache  # noqa: E402


def read_wav_int16(path):
    with open(path, "rb") as f:
        data = f.read()
    assert data[:4] == b"RIFF" and data[8:12] == b"WAVE"
    pos = 12
    fmt = None
    while pos < len(data):
        cid, size = data[pos:pos + 4], struct.unpack("<I", data[pos + 4:pos + 8])[0]
        body = data[pos + 8:pos + 8 + size]
        if cid == b"fmt ":
            fmt = struct.unpack("<HHIIHH", body[:16])
        elif cid == b"data":
            assert fmt is not None and fmt[0] == 1 and fmt[5] == 16
            pcm = np.frombuffer(body, "<i2").reshape(-1, fmt[1])
            return pcm, fmt[2]
        pos += 8 + size + (size & 1)
    raise ValueError("no data chunk")


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    examples = read_feature_cache(os.path.join(ref, "tests", "data", "wav_dataset.tfrecord"))
    pcm, rate = read_wav_int16(os.path.join(ref, "tests", "data", "audio_files", "test.wav"))
    out = {"wav_pcm_int16": pcm[:, 0], "wav_rate": np.int32(rate)}
    for i, (feat, tokens) in enumerate(examples):
        out["feature_%d" % i] = feat
        out["tokens_%d" % i] = tokens
    np.savez_compressed(os.path.join(HERE, "ref_wav_dataset_tfrecord.npz"), **out)
    # data fixtures for the rows either side of the path: the reference's feature cache
    # file itself (480 bytes; byte-identity pin of feature_cache.py) and the header of its test.wav
    with open(os.path.join(ref, "tests", "data", "wav_dataset.tfrecord"), "rb") as f:
        cache = f.read()
    with open(os.path.join(HERE, "ref_wav_dataset.tfrecord"), "wb") as f:
        f.write(cache)
    with open(os.path.join(ref, "tests", "data", "audio_files", "test.wav"), "rb") as f:
        wav = f.read()
    with open(os.path.join(HERE, "ref_test_wav_header.bin"), "wb") as f:
        f.write(wav[:44])
    assert not any(wav[44:]) and len(wav) == 132344  # the payload is digital silence: only the header is kept
    print("records:", len(examples), "feature shape:", examples[0][0].shape,
          "unique values:", np.unique(examples[0][0]), "pcm:", pcm.shape, "rate:", rate,
          "pcm min/max:", pcm.min(), pcm.max())


if __name__ == "__main__":
    main()
