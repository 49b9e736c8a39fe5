"""TensorFlow-free writer / reader of the reference's offline feature cache.

Wire format (written by run/make_tfrecord.py:39-58, read by data.py:64-79): a GZIP-compressed TFRecord
file; each record is ``serialize_tensor(stack([serialize_tensor(feature), serialize_tensor(tokens)]))``:

    outer  TensorProto{ dtype = DT_STRING, tensor_shape = [2], string_val = [inner_0, inner_1] }
    inner_0 TensorProto{ dtype = DT_FLOAT, tensor_shape = [T, F, C], tensor_content = float32 LE }
    inner_1 TensorProto{ dtype = DT_INT32, tensor_shape = [N],       tensor_content = int32 LE   }

TFRecord framing: ``u64 length | u32 masked_crc32c(length) | payload | u32 masked_crc32c(payload)``.
The writer emits the fields in TensorFlow's serialisation order, so the uncompressed stream of a cache
written here is byte-identical to the reference's for the same features (pinned on the reference's golden
``tests/data/wav_dataset.tfrecord``, see tests/test_feature_cache.py).  The CRC runs in libb200mel.so.
"""
import gzip
import struct
from typing import Iterable, Iterator, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib

DT_FLOAT, DT_INT32, DT_STRING = 1, 3, 7
_NP = {DT_FLOAT: np.dtype("<f4"), DT_INT32: np.dtype("<i4")}


def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf, pos: int) -> Tuple[int, int]:
    result = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7


def _shape_proto(shape: Sequence[int]) -> bytes:
    body = b"".join(b"\x12" + _varint(len(d)) + d for d in ((b"\x08" + _varint(int(s))) if int(s) else b""
                                                             for s in shape))
    return b"\x12" + _varint(len(body)) + body


def serialize_tensor(array: np.ndarray) -> bytes:
    """``tf.io.serialize_tensor`` for float32 / int32 arrays: dtype, tensor_shape, tensor_content."""
    if array.dtype == np.float32:
        dt = DT_FLOAT
    elif array.dtype == np.int32:
        dt = DT_INT32
    else:
        raise ValueError("feature cache holds float32 features and int32 tokens, got %s" % array.dtype)
    content = np.ascontiguousarray(array, dtype=_NP[dt]).tobytes()
    out = b"\x08" + _varint(dt) + _shape_proto(array.shape)
    if content:
        out += b"\x22" + _varint(len(content)) + content
    return out


def serialize_example(feature: np.ndarray, tokens: np.ndarray) -> bytes:
    """One cache record (run/make_tfrecord.py:39-43)."""
    inner = [serialize_tensor(feature), serialize_tensor(tokens)]
    return (b"\x08" + _varint(DT_STRING) + _shape_proto([2]) +
            b"".join(b"\x42" + _varint(len(s)) + s for s in inner))


def _masked_crc(data: bytes) -> int:
    return int(_lib.load().b200mel_tfrecord_masked_crc(data, len(data)))


def frame_record(payload: bytes) -> bytes:
    head = struct.pack("<Q", len(payload))
    return head + struct.pack("<I", _masked_crc(head)) + payload + struct.pack("<I", _masked_crc(payload))


class FeatureCacheWriter:
    """``tf.data.experimental.TFRecordWriter(path, "GZIP")`` for (feature, tokens) examples."""

    def __init__(self, path: str, compression: str = "GZIP", compresslevel: int = 6):
        if compression not in ("GZIP", ""):
            raise ValueError("compression must be 'GZIP' (what the reference writes) or ''")
        self._f = gzip.open(path, "wb", compresslevel=compresslevel) if compression else open(path, "wb")
        self.records = 0

    def write(self, feature: np.ndarray, tokens: np.ndarray):
        self._f.write(frame_record(serialize_example(np.asarray(feature), np.asarray(tokens))))
        self.records += 1

    def write_batch(self, features, n_frames, tokens: Iterable[np.ndarray]):
        """Un-pad a front-end batch: ``features`` [B, Tmax, F, C] (host array or CUDA tensor, copied once),
        ``n_frames`` [B]; example b is ``features[b, :min(n_frames[b], Tmax)]``."""
        feats = features.detach().cpu().numpy() if hasattr(features, "detach") else np.asarray(features)
        nf = n_frames.detach().cpu().numpy() if hasattr(n_frames, "detach") else np.asarray(n_frames)
        for b, tok in enumerate(tokens):
            self.write(feats[b, :min(int(nf[b]), feats.shape[1])], np.asarray(tok, dtype=np.int32))

    def close(self):
        self._f.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def _fields(buf) -> Iterator[Tuple[int, int, object]]:
    pos, n = 0, len(buf)
    while pos < n:
        tag, pos = _read_varint(buf, pos)
        num, wire = tag >> 3, tag & 7
        if wire == 0:
            val, pos = _read_varint(buf, pos)
        elif wire == 2:
            ln, pos = _read_varint(buf, pos)
            val, pos = buf[pos:pos + ln], pos + ln
        elif wire == 5:
            val, pos = buf[pos:pos + 4], pos + 4
        elif wire == 1:
            val, pos = buf[pos:pos + 8], pos + 8
        else:
            raise ValueError("unsupported protobuf wire type %d" % wire)
        yield num, wire, val


def parse_tensor(buf, out_type: Optional[int] = None):
    """``tf.io.parse_tensor``: TensorProto bytes -> ndarray (DT_FLOAT / DT_INT32) or list of bytes
    (DT_STRING).  Accepts tensor_content as well as the repeated ``*_val`` fields."""
    buf = memoryview(buf)
    dtype, shape, content = None, [], None
    fvals: List[float] = []
    ivals: List[int] = []
    svals: List[bytes] = []
    for num, wire, val in _fields(buf):
        if num == 1:
            dtype = val
        elif num == 2:
            for n2, _, dim in _fields(val):
                if n2 == 2:
                    size = 0
                    for n3, _, v in _fields(dim):
                        if n3 == 1:
                            size = v if v < (1 << 63) else v - (1 << 64)
                    shape.append(size)
        elif num == 4:
            content = val
        elif num == 5:
            if wire == 2:
                fvals.extend(np.frombuffer(val, "<f4").tolist())
            else:
                fvals.append(struct.unpack("<f", val)[0])
        elif num == 7:
            if wire == 2:
                pos = 0
                while pos < len(val):
                    v, pos = _read_varint(val, pos)
                    ivals.append(v)
            else:
                ivals.append(val)
        elif num == 8:
            svals.append(bytes(val))
    if out_type is not None and dtype != out_type:
        raise ValueError("TensorProto dtype %s, expected %s" % (dtype, out_type))
    if dtype == DT_STRING:
        return svals
    if dtype not in _NP:
        raise ValueError("unsupported TensorProto dtype %s" % dtype)
    count = int(np.prod(shape)) if shape else 1
    if content is not None:
        arr = np.frombuffer(content, _NP[dtype])
    else:
        vals = fvals if dtype == DT_FLOAT else [v - (1 << 64) if v >= (1 << 63) else v for v in ivals]
        arr = np.asarray(vals, dtype=_NP[dtype])
        if arr.size == 1 and count > 1:
            arr = np.repeat(arr, count)  # TensorProto's "last value repeats" compression
        elif arr.size == 0:
            arr = np.zeros(count, dtype=_NP[dtype])
    if arr.size != count:
        raise ValueError("TensorProto holds %d values for shape %s" % (arr.size, shape))
    return arr.reshape(shape).copy()


def read_records(path: str, compression: str = "GZIP", verify: bool = True) -> Iterator[bytes]:
    """TFRecord payloads of one file; both CRCs are checked unless ``verify`` is False."""
    opener = gzip.open if compression else open
    with opener(path, "rb") as f:
        while True:
            head = f.read(12)
            if not head:
                return
            if len(head) < 12:
                raise ValueError("truncated TFRecord header in %s" % path)
            (length,) = struct.unpack("<Q", head[:8])
            payload = f.read(length)
            tail = f.read(4)
            if len(payload) != length or len(tail) != 4:
                raise ValueError("truncated TFRecord in %s" % path)
            if verify:
                if struct.unpack("<I", head[8:])[0] != _masked_crc(head[:8]):
                    raise ValueError("corrupted TFRecord length in %s" % path)
                if struct.unpack("<I", tail)[0] != _masked_crc(payload):
                    raise ValueError("corrupted TFRecord payload in %s" % path)
            yield payload


def read_feature_cache(path: str, compression: str = "GZIP", verify: bool = True
                       ) -> Iterator[Tuple[np.ndarray, np.ndarray]]:
    """``get_tfrecord_dataset`` (data.py:64-79) for one file: yields (feature float32, tokens int32)."""
    for payload in read_records(path, compression, verify):
        pair = parse_tensor(payload, DT_STRING)
        if len(pair) != 2:
            raise ValueError("cache record holds %d tensors, expected 2" % len(pair))
        yield parse_tensor(pair[0], DT_FLOAT), parse_tensor(pair[1], DT_INT32)
