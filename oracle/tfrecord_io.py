"""TensorFlow-free reader for the reference's feature cache format.  TEST INFRASTRUCTURE ONLY.

Format written by ``speech_recognition/run/make_tfrecord.py:39-58`` and read back by
``speech_recognition/data.py:64-79``: a GZIP TFRecord file whose records are
``tf.io.serialize_tensor(tf.stack([serialize_tensor(feature), serialize_tensor(tokens)]))``, i.e.
an outer ``TensorProto{dtype=DT_STRING, shape=[2], string_val x2}`` holding two inner TensorProtos.

TFRecord framing: u64 length, u32 masked-crc(length), payload, u32 masked-crc(payload).
Only what the golden fixture needs is implemented (DT_FLOAT / DT_INT32 / DT_STRING, tensor_content
or repeated *_val fields).
"""
from __future__ import annotations

import gzip
import struct
from typing import Dict, Iterator, List, Tuple

import numpy as np

DT_FLOAT, DT_INT32, DT_STRING = 1, 3, 7


def read_tfrecords(path: str, compression: str = "GZIP") -> Iterator[bytes]:
    opener = gzip.open if compression == "GZIP" else open
    with opener(path, "rb") as f:
        while True:
            head = f.read(12)
            if len(head) < 12:
                return
            (length,) = struct.unpack("<Q", head[:8])
            data = f.read(length)
            f.read(4)
            yield data


def _varint(buf: bytes, pos: int) -> Tuple[int, int]:
    result = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7


def _fields(buf: bytes) -> Iterator[Tuple[int, int, object]]:
    pos = 0
    while pos < len(buf):
        tag, pos = _varint(buf, pos)
        num, wire = tag >> 3, tag & 7
        if wire == 0:
            val, pos = _varint(buf, pos)
        elif wire == 1:
            val, pos = buf[pos:pos + 8], pos + 8
        elif wire == 2:
            ln, pos = _varint(buf, pos)
            val, pos = buf[pos:pos + ln], pos + ln
        elif wire == 5:
            val, pos = buf[pos:pos + 4], pos + 4
        else:
            raise ValueError("unsupported wire type %d" % wire)
        yield num, wire, val


def parse_tensor_proto(buf: bytes):
    """TensorProto: 1=dtype, 2=tensor_shape{2=dim{1=size}}, 4=tensor_content, 5=float_val,
    7=int_val, 8=string_val."""
    dtype = None
    shape: List[int] = []
    content = None
    float_val: List[float] = []
    int_val: List[int] = []
    string_val: List[bytes] = []
    for num, wire, val in _fields(buf):
        if num == 1:
            dtype = val
        elif num == 2:
            for n2, _, dim in _fields(val):
                if n2 == 2:
                    size = 0
                    for n3, _, v3 in _fields(dim):
                        if n3 == 1:
                            size = v3
                    shape.append(size)
        elif num == 4:
            content = val
        elif num == 5:
            if wire == 2:
                float_val.extend(np.frombuffer(val, "<f4").tolist())
            else:
                float_val.append(struct.unpack("<f", val)[0])
        elif num == 7:
            if wire == 2:
                p = 0
                while p < len(val):
                    x, p = _varint(val, p)
                    int_val.append(x)
            else:
                int_val.append(val)
        elif num == 8:
            string_val.append(val)
    if dtype == DT_STRING:
        return string_val
    np_dtype = {DT_FLOAT: "<f4", DT_INT32: "<i4"}[dtype]
    n = int(np.prod(shape)) if shape else 1
    if content is not None:
        arr = np.frombuffer(content, np_dtype)
    else:
        vals = float_val if dtype == DT_FLOAT else int_val
        arr = np.array(vals, dtype=np_dtype)
        if arr.size == 1 and n > 1:
            arr = np.full(n, arr[0], dtype=np_dtype)
    return arr.reshape(shape)


def read_feature_cache(path: str) -> List[Tuple[np.ndarray, np.ndarray]]:
    """Decode every (feature, tokens) example of a reference feature-cache file."""
    out = []
    for rec in read_tfrecords(path):
        strings = parse_tensor_proto(rec)
        out.append((parse_tensor_proto(strings[0]), parse_tensor_proto(strings[1])))
    return out
