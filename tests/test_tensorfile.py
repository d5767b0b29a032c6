"""TENF reader KATs: the reference's fixtures (visr_bear/test/test_tensorfile.cpp:39-61,
gen_tensorfile_test_files.py) through the Python reader and through the C++ reader behind the C ABI."""
import numpy as np
import pytest

from bear_b200 import renderer, tensorfile
from conftest import GOLDEN

DTYPES = ["float32", "float64", "uint64", "uint32", "uint16", "uint8", "int64", "int32", "int16", "int8"]
FILES = ["tensorfile_le_1.tenf", "tensorfile_le_32.tenf", "tensorfile_be_1.tenf", "tensorfile_be_32.tenf"]
i, j, k = np.mgrid[:2, :4, :8]
EXPECT = (i << 5) + (j << 3) + k


@pytest.mark.parametrize("fname", FILES)
def test_python_reader_kat(fname):
    tree = tensorfile.read(GOLDEN / fname)
    for name in DTYPES:
        a = tree["multi_dim"][name]
        assert a.shape == (2, 4, 8) and a.dtype.kind == np.dtype(name).kind
        assert (a == EXPECT.astype(name)).all()


@pytest.mark.parametrize("fname", FILES)
def test_cxx_reader_kat(fname):
    for name in DTYPES:
        a = renderer.tenf_read_f32(GOLDEN / fname, "multi_dim/" + name)
        assert a.shape == (2, 4, 8)
        assert (a == EXPECT.astype(np.float32)).all()


def test_round_trip_alignment_and_strides(tmp_path):
    rng = np.random.default_rng(0)
    a = rng.standard_normal((3, 5, 7)).astype(np.float32)
    for order in "<>":
        for align in (1, 32):
            p = tmp_path / ("t%s%d.tenf" % ("le" if order == "<" else "be", align))
            tensorfile.write(p, {"x": {"y": a}, "n": 3, "s": "str", "l": [1, 2.5, None]}, order, align)
            t = tensorfile.read(p)
            assert t["n"] == 3 and t["s"] == "str" and t["l"] == [1, 2.5, None]
            assert (t["x"]["y"] == a).all()
            assert (renderer.tenf_read_f32(p, "x/y") == a).all()


@pytest.mark.parametrize("mutate,msg", [
    (lambda b: b[:10], "file not long enough"),
    (lambda b: b"XXXX" + b[4:], "magic number not found"),
    (lambda b: b[:4] + b"\1\0\0\0" + b[8:], "unknown version number"),
    (lambda b: b[:-20], "file not long enough"),
])
def test_format_errors(tmp_path, mutate, msg):
    # error strings of tensorfile::read, visr_bear/src/tensorfile.cpp:187-215
    good = (GOLDEN / "tensorfile_le_1.tenf").read_bytes()
    p = tmp_path / "bad.tenf"
    p.write_bytes(mutate(good))
    with pytest.raises(tensorfile.FormatError, match=msg):
        tensorfile.read(p)
    with pytest.raises(renderer.BearCudaError, match="tensorfile format error: " + msg):
        renderer.tenf_read_f32(p, "multi_dim/float32")


def _with_json(good, edit):
    """The fixture with its trailing JSON metadata edited (header lengths updated)."""
    import json
    import struct
    dlen, jlen = struct.unpack("<QQ", good[8:24])
    meta = json.loads(good[24 + dlen:24 + dlen + jlen])
    edit(meta["multi_dim"]["float32"])
    mb = json.dumps(meta).encode()
    return good[:16] + struct.pack("<Q", len(mb)) + good[24:24 + dlen] + mb


@pytest.mark.parametrize("edit", [
    lambda t: t.__setitem__("shape", [-1, 4, 8]),
    lambda t: t.__setitem__("shape", [2.5, 4, 8]),
    lambda t: t.__setitem__("shape", [2 ** 40, 2 ** 40, 8]),
    lambda t: t.__setitem__("strides", [2 ** 62, 32, 4]),
    lambda t: t.__setitem__("strides", [-128, 32, 4]),
    lambda t: t.__setitem__("base", 2 ** 63),
    lambda t: t.__setitem__("base", -8),
], ids=["negative", "fractional", "huge-count", "huge-stride", "negative-stride", "huge-base", "negative-base"])
def test_crafted_size_fields_are_format_errors_in_the_cxx_reader(tmp_path, edit):
    """Untrusted shape / strides / base: a format error, never a wrapped size or an out-of-bounds read."""
    p = tmp_path / "crafted.tenf"
    p.write_bytes(_with_json((GOLDEN / "tensorfile_le_1.tenf").read_bytes(), edit))
    with pytest.raises(renderer.BearCudaError, match="tensorfile format error"):
        renderer.tenf_read_f32(p, "multi_dim/float32")


def test_crafted_header_lengths_do_not_wrap(tmp_path):
    import struct
    good = (GOLDEN / "tensorfile_le_1.tenf").read_bytes()
    p = tmp_path / "wrap.tenf"
    p.write_bytes(good[:8] + struct.pack("<QQ", 2 ** 64 - 8, 16) + good[24:])   # data + JSON lengths sum to 8 mod 2^64
    with pytest.raises(renderer.BearCudaError, match="file not long enough"):
        renderer.tenf_read_f32(p, "multi_dim/float32")
