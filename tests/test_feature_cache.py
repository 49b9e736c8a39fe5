"""Feature cache: the TF-free writer / reader against the reference's own golden
cache file tests/data/wav_dataset.tfrecord (committed as tests/golden/ref_wav_dataset.tfrecord) - byte
identity of the uncompressed stream - plus round trips, CRC known answers and corruption detection."""
import gzip
import os
import struct

import numpy as np
import pytest

from oracle import tfrecord_io

from .helpers import GOLDEN_DIR

GOLDEN = os.path.join(GOLDEN_DIR, "ref_wav_dataset.tfrecord")


@pytest.fixture(scope="module")
def fc(built_lib):
    from speech_recognition_b200 import feature_cache

    return feature_cache


def _unmask(m):
    r = (m - 0xA282EAD8) & 0xFFFFFFFF
    return ((r >> 17) | (r << 15)) & 0xFFFFFFFF


def test_crc32c_known_answers(fc):
    """RFC 3720 appendix B.4 test vectors of CRC-32C."""
    from speech_recognition_b200 import _lib

    crc = lambda b: _unmask(_lib.load().b200mel_tfrecord_masked_crc(b, len(b)))  # noqa: E731
    assert crc(b"123456789") == 0xE3069283
    assert crc(bytes(32)) == 0x8A9136AA
    assert crc(b"\xff" * 32) == 0x62A8AB43
    assert crc(bytes(range(32))) == 0x46DD794E
    assert crc(bytes(range(31, -1, -1))) == 0x113FDB5C
    # unaligned start and length: same value through a shifted buffer
    buf = b"x" + bytes(range(32)) + b"yz"
    assert crc(buf[1:33]) == 0x46DD794E


def test_reads_reference_golden(fc):
    """The reference asserts its features equal this file's (tests/test_data.py:53-57)."""
    ex = list(fc.read_feature_cache(GOLDEN))
    g = np.load(os.path.join(GOLDEN_DIR, "ref_wav_dataset_tfrecord.npz"))
    assert len(ex) == 2
    for i, (feat, tok) in enumerate(ex):
        assert feat.dtype == np.float32 and feat.shape == (412, 80, 1)
        assert np.array_equal(feat, g["feature_%d" % i]) and np.array_equal(tok, g["tokens_%d" % i])
        assert np.all(feat == np.float32(-27.631021))
    assert bytes(ex[0][1].astype(np.uint8)) == b"Hello World Good night"
    # and agrees with the oracle's independent reader
    ref = tfrecord_io.read_feature_cache(GOLDEN)
    for (f0, t0), (f1, t1) in zip(ex, ref):
        assert np.array_equal(f0, f1) and np.array_equal(t0, t1)


def test_writer_is_byte_identical_to_reference(fc, tmp_path):
    """Serialising the decoded examples again reproduces the reference's uncompressed stream bit for bit:
    field order, varints, shapes, tensor_content and both masked CRCs."""
    want = gzip.open(GOLDEN, "rb").read()
    ex = list(fc.read_feature_cache(GOLDEN))
    path = str(tmp_path / "again.tfrecord")
    with fc.FeatureCacheWriter(path) as w:
        for feat, tok in ex:
            w.write(feat, tok)
    assert w.records == 2
    assert gzip.open(path, "rb").read() == want
    assert len(want) == 263926


@pytest.mark.parametrize("shape", [(0, 80, 1), (1, 80, 3), (37, 40, 3), (5, 161, 1), (1499, 80, 1)])
def test_round_trip(fc, tmp_path, shape):
    rng = np.random.default_rng(sum(shape))
    feat = rng.normal(0, 3, shape).astype(np.float32)
    if feat.size:
        feat.flat[0] = -0.0
    tok = rng.integers(-5, 40000, size=shape[0] % 7, dtype=np.int32)
    for comp in ("GZIP", ""):
        path = str(tmp_path / ("rt%s.tfrecord" % comp))
        with fc.FeatureCacheWriter(path, compression=comp) as w:
            w.write(feat, tok)
            w.write(feat[: shape[0] // 2], tok[:1])
        got = list(fc.read_feature_cache(path, compression=comp))
        assert len(got) == 2
        assert got[0][0].shape == shape and got[0][0].tobytes() == feat.tobytes()
        assert np.array_equal(got[0][1], tok) and got[0][1].dtype == np.int32
        assert got[1][0].shape == (shape[0] // 2,) + shape[1:]
        # the oracle's reader agrees on what was written
        ref = tfrecord_io.read_feature_cache(path) if comp else None
        if ref is not None:
            assert np.array_equal(ref[0][0], feat) and np.array_equal(ref[0][1], tok)


def test_write_batch_unpads(fc, tmp_path):
    feats = np.arange(2 * 5 * 4 * 1, dtype=np.float32).reshape(2, 5, 4, 1)
    path = str(tmp_path / "b.tfrecord")
    with fc.FeatureCacheWriter(path) as w:
        w.write_batch(feats, np.array([3, 9], np.int32), [np.array([1, 2]), np.array([], np.int64)])
    got = list(fc.read_feature_cache(path))
    assert got[0][0].shape == (3, 4, 1) and got[1][0].shape == (5, 4, 1)
    assert np.array_equal(got[0][0], feats[0, :3]) and got[1][1].size == 0


def test_corruption_is_detected(fc, tmp_path):
    raw = bytearray(gzip.open(GOLDEN, "rb").read())
    raw[5000] ^= 1
    bad = str(tmp_path / "bad.tfrecord")
    with gzip.open(bad, "wb") as f:
        f.write(bytes(raw))
    with pytest.raises(ValueError, match="corrupted TFRecord payload"):
        list(fc.read_feature_cache(bad))
    assert len(list(fc.read_feature_cache(bad, verify=False))) == 2
    raw = bytearray(gzip.open(GOLDEN, "rb").read())
    raw[9] ^= 0x10
    with gzip.open(bad, "wb") as f:
        f.write(bytes(raw))
    with pytest.raises(ValueError, match="corrupted TFRecord length"):
        list(fc.read_feature_cache(bad))
    with gzip.open(bad, "wb") as f:
        f.write(bytes(raw[:1000]))
    with pytest.raises(ValueError, match="truncated"):
        list(fc.read_feature_cache(bad))


def test_parse_tensor_val_fields(fc):
    """TensorProtos that use float_val / int_val (small tensors, 'last value repeats') decode too."""
    p = b"\x08\x01" + b"\x12\x04\x12\x02\x08\x03" + b"\x2a\x04" + struct.pack("<f", 2.5)
    assert np.array_equal(fc.parse_tensor(p), np.full(3, 2.5, np.float32))
    p = b"\x08\x03" + b"\x12\x04\x12\x02\x08\x02" + b"\x3a\x02\x07\x09"
    assert np.array_equal(fc.parse_tensor(p), np.array([7, 9], np.int32))
    with pytest.raises(ValueError):
        fc.parse_tensor(p, fc.DT_FLOAT)
    with pytest.raises(ValueError):
        fc.serialize_tensor(np.zeros(3, np.float64))
