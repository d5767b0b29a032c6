"""TENF ("tensorfile") container reader/writer.

BEAR data files are TENF containers (reference: bear/tensorfile.py:21-58 for the
format description, visr_bear/src/tensorfile.cpp:187-215 for the C++ reader whose
error behaviour `bear_b200/csrc/tenf.cpp` mirrors).  Layout:

    0   4  magic "TENF"
    4   4  version, u32 LE, must be 0
    8   8  data length D, u64 LE
    16  8  metadata length J, u64 LE
    24  D  raw array bytes
    24+D J UTF-8 JSON; arrays appear as objects
           {"_tenf_type": "array", "dtype": "<f4", "shape": [...],
            "strides": [... bytes ...], "base": <offset from file start>}

This module is an independent implementation used by the host side (synthetic
data files, tests); the product's loader is the C++ one behind the C ABI.
"""
import json
import struct

import numpy as np

MAGIC = b"TENF"
HEADER = struct.Struct("<4sIQQ")
ARRAY_TAG = "_tenf_type"


class FormatError(RuntimeError):
    def __init__(self, what):
        super().__init__("tensorfile format error: " + what)


def _encode(node, chunks, pos, byte_order, alignment):
    """Replace ndarrays in `node` by descriptors, appending their bytes to `chunks`."""
    if isinstance(node, np.ndarray):
        dt = node.dtype.newbyteorder(byte_order)
        arr = np.ascontiguousarray(node, dtype=dt)
        pad = (-pos[0]) % alignment
        chunks.append(b"\0" * pad)
        pos[0] += pad
        desc = {
            ARRAY_TAG: "array",
            "dtype": dt.str,
            "strides": [int(s) for s in arr.strides],
            "shape": [int(s) for s in arr.shape],
            "base": pos[0],
        }
        raw = arr.tobytes()
        chunks.append(raw)
        pos[0] += len(raw)
        return desc
    if isinstance(node, dict):
        return {k: _encode(v, chunks, pos, byte_order, alignment) for k, v in node.items()}
    if isinstance(node, (list, tuple)):
        return [_encode(v, chunks, pos, byte_order, alignment) for v in node]
    if isinstance(node, np.generic):
        return node.item()
    return node


def dumps(tree, byte_order="<", alignment=32):
    """Serialise a nested dict/list/ndarray structure to TENF bytes."""
    chunks, pos = [], [HEADER.size]
    meta = _encode(tree, chunks, pos, byte_order, alignment)
    data = b"".join(chunks)
    meta_b = json.dumps(meta).encode("utf8")
    return HEADER.pack(MAGIC, 0, len(data), len(meta_b)) + data + meta_b


def write(path, tree, byte_order="<", alignment=32):
    with open(path, "wb") as f:
        f.write(dumps(tree, byte_order, alignment))


def _decode(node, buf):
    if isinstance(node, dict):
        if ARRAY_TAG in node:
            if node[ARRAY_TAG] != "array":
                raise FormatError("unknown type")
            shape, strides = node["shape"], node["strides"]
            if len(shape) != len(strides):
                raise FormatError("mismatched strides and shape")
            dt = np.dtype(node["dtype"])
            if any(s % dt.itemsize for s in strides):
                raise FormatError("stride is not a multiple of the type size")
            base = int(node["base"])
            span = sum(st * (sh - 1) for st, sh in zip(strides, shape)) + dt.itemsize
            if base + span > len(buf):
                raise FormatError("file not long enough")
            flat = np.frombuffer(buf, dtype=np.uint8, count=span, offset=base)
            return np.lib.stride_tricks.as_strided(
                flat.view(np.uint8), shape=tuple(shape) + (dt.itemsize,),
                strides=tuple(strides) + (1,), writeable=False
            ).copy().view(dt).reshape(shape)
        return {k: _decode(v, buf) for k, v in node.items()}
    if isinstance(node, list):
        return [_decode(v, buf) for v in node]
    return node


def loads(buf):
    """Parse TENF bytes; returns the JSON tree with arrays materialised (native copies)."""
    buf = memoryview(buf)
    if len(buf) < HEADER.size:
        raise FormatError("file not long enough")
    magic, version, dlen, jlen = HEADER.unpack_from(buf, 0)
    if magic != MAGIC:
        raise FormatError("magic number not found")
    if version != 0:
        raise FormatError("unknown version number")
    if len(buf) < HEADER.size + dlen + jlen:
        raise FormatError("file not long enough")
    if jlen == 0:
        raise FormatError("no JSON metadata found")
    try:
        meta = json.loads(bytes(buf[HEADER.size + dlen: HEADER.size + dlen + jlen]).decode("utf8"))
    except ValueError as e:
        raise FormatError("could not parse JSON metadata: %s" % e)
    return _decode(meta, buf)


def read(path):
    with open(path, "rb") as f:
        return loads(f.read())
