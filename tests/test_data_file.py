"""Data-file loading: shapes and the reference's error strings (panner.cpp:67-136, data_file.cpp:16-20,
visr_bear/test/test_data_file.py)."""
import numpy as np
import pytest

from bear_b200 import data_file, renderer, tensorfile
from conftest import GOLDEN


def test_unity_fixture_shapes(unity_data):
    d = unity_data
    assert d.brirs.shape == (180, 24, 2, 1) and d.delays.shape == (180, 24, 2)
    assert d.decorrelation_filters.shape == (24, 128) and d.hoa_irs.shape == (16, 2, 128)
    assert d.front_loudspeaker == 2 and d.decorrelation_delay == 63
    assert abs(d.hoa_delay - 110.8633) < 1e-3
    i = renderer.read_data_file_info(GOLDEN / "unity_brirs_decorrelators.tf")
    assert (i.num_views, i.num_virtual_loudspeakers, i.brir_length, i.decorrelator_length) == (180, 24, 1, 128)
    assert (i.num_hoa_decoder_channels, i.hoa_ir_length, i.front_loudspeaker) == (16, 128, 2)
    assert i.has_metadata == 0


def _write(tmp_path, tree, name="f.tf"):
    p = tmp_path / name
    tensorfile.write(p, tree)
    return p


def test_metadata_and_version(tmp_path):
    tree = data_file.synthetic_tree(num_views=2, brir_length=64)
    p = _write(tmp_path, tree)
    i = renderer.read_data_file_info(p)
    assert i.has_metadata == 1 and i.label == b"synthetic" and i.is_released == 0
    assert i.brir_length == 64 and i.fs == 48000.0
    tree["bear_data_version"] = 1
    p = _write(tmp_path, tree, "v1.tf")
    with pytest.raises(data_file.DataFileError, match="data file version mismatch: expected 0, got 1"):
        data_file.load(p)
    with pytest.raises(renderer.BearCudaError, match="data file version mismatch: expected 0, got 1"):
        renderer.read_data_file_info(p)


@pytest.mark.parametrize("edit,msg", [
    (lambda t: t.__setitem__("views", t["views"][:1]), "views axis 0 is wrong size"),
    (lambda t: t.__setitem__("brirs", t["brirs"][:, :, :1]), "brirs axis 2 is wrong size"),
    (lambda t: t.__setitem__("delays", t["delays"][:, :3]), "delays axis 1 is wrong size"),
    (lambda t: t.__setitem__("decorrelation_filters", t["decorrelation_filters"][:5]),
     "decorrelation filters axis 0 is wrong size"),
    (lambda t: t.pop("hoa"), "HOA decoder not present"),
])
def test_shape_errors(tmp_path, edit, msg):
    tree = data_file.synthetic_tree(num_views=2, brir_length=64)
    edit(tree)
    p = _write(tmp_path, tree)
    with pytest.raises(data_file.DataFileError, match="BEAR data file error: " + msg):
        data_file.load(p)
    with pytest.raises(renderer.BearCudaError, match="BEAR data file error: " + msg):
        renderer.read_data_file_info(p)


def test_synthetic_default_shape():
    t = data_file.synthetic_tree(num_views=4)
    assert t["brirs"].shape == (4, 24, 2, 2976)      # dynamic_renderer.cpp:120-123
    assert t["decorrelation_filters"].shape == (24, 128) and t["hoa"]["irs"].shape == (16, 2, 128)
    np.testing.assert_allclose((t["brirs"][0, 2, 0].astype(np.float64) ** 2).sum(), 1.0, rtol=1e-5)
