"""The C-ABI library loads on a machine without a GPU and exports every symbol that
include/bear_cuda.h declares; the ctypes mirrors have the sizes the header implies."""
import ctypes
import re

import pytest

from bear_b200 import renderer
from conftest import ROOT


def declared_symbols():
    text = (ROOT / "include" / "bear_cuda.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bear_cuda_[a-z0-9_]+)\s*\(", text)))


def test_exports_match_header():
    names = declared_symbols()
    assert len(names) >= 14
    lib = ctypes.CDLL(str(renderer.LIB_PATH))
    for n in names:
        assert hasattr(lib, n), "missing export " + n
    assert sorted(renderer.EXPORTS) == names


def test_struct_sizes():
    assert ctypes.sizeof(renderer.Config) == 17 * 4
    assert ctypes.sizeof(renderer.Data) == 6 * 4 + 4 * 8 + 3 * 8 + 8 + 8
    assert ctypes.sizeof(renderer.Tables) == 8 * 8
    assert ctypes.sizeof(renderer.DeviceIO) == 14 * 8


def test_create_fails_loudly_without_gpu(small_data):
    """No CPU fallback: without a device creation must fail, not silently compute elsewhere."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(renderer.BearCudaError, match="no usable CUDA device"):
        renderer.Renderer(small_data, 1, 0, 0, 512)


def test_argument_errors_before_device(small_data):
    with pytest.raises(ValueError, match="Config: period size must be set"):      # api.cpp:70
        renderer.Renderer(small_data, 1, 0, 0, 0)
    with pytest.raises(ValueError, match="Config: data path must be set"):        # api.cpp:71
        renderer.Renderer("", 1, 0, 0, 512)
    with pytest.raises(renderer.BearCudaError, match="power of two"):
        renderer.Renderer(small_data, 1, 0, 0, 500)


def test_every_entry_point_names_its_reference_counterpart():
    """include/bear_cuda.h: the comment in front of every exported function cites the reference
    interface it replaces (file:line) or says that there is none."""
    import re
    lines = (ROOT / "include" / "bear_cuda.h").read_text().split("\n")
    for name in renderer.EXPORTS:
        at = [i for i, l in enumerate(lines) if re.search(r"\b%s\(" % name, l) and not l.lstrip().startswith("*")]
        assert at, name
        i, doc = at[0], ""
        while not doc.strip() and i > 0:   # declarations that follow each other share one comment
            j = i
            while j > 0 and not re.search(r"\bbear_cuda_\w+\([^)]*\)?.*;\s*$|^#ifdef __cplusplus", lines[j - 1]):
                j -= 1                     # back to the end of the previous declaration
            doc = "\n".join(lines[j:i])
            i = j - 1
        assert re.search(r"\w+\.(cpp|hpp|py):\d+", doc) or "ounterpart" in doc, name
