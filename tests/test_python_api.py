"""bear_b200.api: the visr_bear.api-shaped pybind11 module.  The CPU tests cover what
needs no device (Config, Time, Listener, data-file metadata, loud failure without a GPU); the GPU
tests restate visr_bear/test/test_renderer.py:41-97 (timing API, future blocks, add errors), the
buffer checks of pythonwrappers/array_conversion.hpp:12-44, and render a scene through the module and
through the render_cli.py block loop against the oracle."""
from fractions import Fraction

import numpy as np
import pytest

from bear_b200 import api, control, data_file, tensorfile
from conftest import GOLDEN


def to_frac(t):
    return Fraction(t.numerator, t.denominator)


def test_config_time_listener_without_gpu():
    c = api.Config()
    with pytest.raises(ValueError, match="period size must be set"):      # api.cpp:70
        c.validate()
    c.period_size = 512
    with pytest.raises(ValueError, match="data path must be set"):        # api.cpp:71
        c.validate()
    c.data_path = "x.tf"
    c.validate()
    c.num_objects_channels, c.num_direct_speakers_channels, c.num_hoa_channels = 3, 2, 4
    c2 = api.Config(c)
    assert (c2.num_objects_channels, c2.num_direct_speakers_channels, c2.num_hoa_channels, c2.sample_rate) == (3, 2, 4, 48000)
    assert to_frac(api.Time(512, 48000)) == Fraction(4, 375) and api.Time(3).denominator == 1
    l = api.Listener()
    np.testing.assert_allclose(l.look, [0, 1, 0])
    np.testing.assert_allclose(l.right, [1, 0, 0])
    # test_select_brir.py:84-120 convention: orientation maps world to listener coordinates
    a = np.radians(90)
    l.orientation_quaternion = [np.cos(a / 2), 0, 0, np.sin(a / 2)]
    np.testing.assert_allclose(l.look, [1, 0, 0], atol=1e-12)
    h = api.HOAInput()
    h.channels = [0, 1, 2, 3]
    h.type_metadata.orders, h.type_metadata.degrees, h.type_metadata.normalization = [0, 1, 1, 1], [0, -1, 0, 1], "N3D"
    assert api.HOAInput(h).type_metadata.normalization == "N3D"
    o = api.ObjectsInput()
    o.rtime, o.duration = api.Time(0), api.Time(1, 2)
    o.direct_gains = [0.0] * 24
    assert to_frac(o.duration) == Fraction(1, 2) and o.interpolationLength is None


def test_data_file_metadata_without_gpu(tmp_path):
    m = api.DataFileMetadata.read_from_file(str(GOLDEN / "unity_brirs_decorrelators.tf"))
    assert m.has_metadata in (True, False)
    tree = data_file.synthetic_tree(seed=3, num_views=4, brir_length=300)
    tree["metadata"] = {"label": "test file", "description": "made by a test", "released": False}
    tensorfile.write(tmp_path / "m.tf", tree)
    m = api.DataFileMetadata.read_from_file(str(tmp_path / "m.tf"))
    assert (m.has_metadata, m.label, m.description, m.is_released) == (True, "test file", "made by a test", False)
    with pytest.raises(RuntimeError):
        api.DataFileMetadata.read_from_file(str(tmp_path / "missing.tf"))


def test_renderer_fails_loudly_without_gpu(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    tensorfile.write(tmp_path / "d.tf", data_file.synthetic_tree(seed=3, num_views=4, brir_length=300))
    c = api.Config()
    c.period_size, c.data_path = 256, str(tmp_path / "d.tf")
    with pytest.raises(RuntimeError, match="no usable CUDA device|no CPU fallback"):
        api.Renderer(c)


# ---------------------------------------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def scene_file(tmp_path_factory):
    tree = data_file.synthetic_tree(seed=5, num_views=180, brir_length=300)
    path = tmp_path_factory.mktemp("data") / "scene.tf"
    tensorfile.write(path, tree)
    return str(path), data_file.from_tree(tree)


def basic_config(path, n=1, m=0, h=0, B=512):
    c = api.Config()
    c.num_objects_channels, c.num_direct_speakers_channels, c.num_hoa_channels = n, m, h
    c.period_size, c.data_path = B, path
    return c


def dummy_process_call(r, c):
    buf = lambda n: np.zeros((n, c.period_size), np.float32)
    r.process(buf(c.num_objects_channels), buf(c.num_direct_speakers_channels), buf(c.num_hoa_channels), buf(2))


@pytest.mark.gpu
def test_timing_api(scene_file):        # test_renderer.py:41-61
    c = basic_config(scene_file[0])
    r = api.Renderer(c)
    assert to_frac(r.get_block_start_time()) == 0
    dummy_process_call(r, c)
    assert to_frac(r.get_block_start_time()) == Fraction(c.period_size, c.sample_rate)
    r.set_block_start_time(api.Time(1, 10))
    assert to_frac(r.get_block_start_time()) == Fraction(1, 10)
    dummy_process_call(r, c)
    assert to_frac(r.get_block_start_time()) == Fraction(1, 10) + Fraction(c.period_size, c.sample_rate)


@pytest.mark.gpu
def test_future_blocks_and_add_errors(scene_file):    # test_renderer.py:64-97
    c = basic_config(scene_file[0], m=1)
    r = api.Renderer(c)
    ob, db = api.ObjectsInput(), api.DirectSpeakersInput()
    ob.direct_gains, ob.diffuse_gains, db.gains = [0.0] * r.num_gains, [0.0] * r.num_gains, [0.0] * r.num_gains
    for b in (ob, db):
        b.rtime, b.duration = api.Time(512, 48000), api.Time(512, 48000)
    assert not r.add_objects_block(0, ob) and not r.add_direct_speakers_block(0, db)
    r.set_block_start_time(api.Time(1, 48000))
    assert r.add_objects_block(0, ob) and r.add_direct_speakers_block(0, db)
    with pytest.raises(ValueError):
        r.add_objects_block(2, api.ObjectsInput())
    dummy_process_call(r, c)
    with pytest.raises(ValueError):
        r.add_direct_speakers_block(1, api.DirectSpeakersInput())
    dummy_process_call(r, c)


@pytest.mark.gpu
def test_process_buffer_checks(scene_file):           # array_conversion.hpp:12-44
    c = basic_config(scene_file[0], n=2, B=256)
    r = api.Renderer(c)
    z = lambda n, B=256, dt=np.float32: np.zeros((n, B), dt)
    with pytest.raises(ValueError, match="objects_input has wrong number of channels"):
        r.process(z(1), z(0), z(0), z(2))
    with pytest.raises(ValueError, match="output has wrong length"):
        r.process(z(2), z(0), z(0), z(2, 128))
    with pytest.raises(ValueError, match="hoa_input must have 2 dimensions"):
        r.process(z(2), z(0), np.zeros(256, np.float32), z(2))
    with pytest.raises(ValueError, match="objects_input must have stride of 1 along samples"):
        r.process(np.zeros((2, 512), np.float32)[:, ::2], z(0), z(0), z(2))
    with pytest.raises(TypeError):                    # noconvert: float64 is refused, not copied
        r.process(z(2, dt=np.float64), z(0), z(0), z(2))
    r.process(z(2), z(0), z(0), z(2))


@pytest.mark.gpu
def test_scene_through_the_module_and_the_block_loop_matches_oracle(oracle, scene_file):
    """A scene with timed object blocks, a DirectSpeakers channel and a first-order HOA stream, rendered
    (a) through bear_b200.render_loop.BlockRenderer, the restated bear-render loop, with a short last
    block, and (b) period by period through the oracle with the schedule written out by hand."""
    from bear_b200 import render_loop
    path, data = scene_file
    B, K, L = 256, 7, data.num_virtual_loudspeakers
    nsamp = K * B - 100                                  # the last block is short
    audio = np.concatenate([control.synth_audio(1, K * B, 0), control.synth_audio(1, K * B, 50), control.synth_audio(4, K * B, 90)])
    audio[:, nsamp:] = 0

    def vec(d):
        g = [0.0] * L
        for i, v in d.items():
            g[i] = v
        return g
    a, b = api.ObjectsInput(), api.ObjectsInput()
    a.rtime, a.duration, a.direct_gains, a.diffuse_gains = api.Time(0), api.Time(512, 48000), vec({2: .8, 5: .3}), vec({})
    b.rtime, b.duration, b.direct_gains, b.diffuse_gains = api.Time(512, 48000), api.Time(512, 48000), vec({6: .5}), vec({6: .2, 9: .3})
    d = api.DirectSpeakersInput()
    d.gains = vec({5: 1.0})
    h = api.HOAInput()
    h.channels = [0, 1, 2, 3]
    h.type_metadata.orders, h.type_metadata.degrees = [0, 1, 1, 1], [0, -1, 0, 1]
    br = render_loop.BlockRenderer(path, [[a, b]], [[d]], [[h]], [4], block_size=B)
    out = br.render_all(audio[:1, :nsamp], audio[1:2, :nsamp], audio[2:, :nsamp], nsamp)
    assert out.shape == (2, K * B)

    A = np.array(vec({2: .8, 5: .3}) + vec({}))
    Bg = np.array(vec({6: .5}) + vec({6: .2, 9: .3}))
    ds_raw = np.zeros((L, 1), np.float32); ds_raw[5, 0] = 1.0
    ref = oracle.RefRenderer(data, 1, 1, 4, B)
    want = np.zeros((2, K * B))
    for k in range(K):
        end = (k + 1) * B
        if end <= 512:
            g = A
        elif end < 1024:
            p = (end - 512) / 512
            g = (1 - p) * A + p * Bg
        elif end == 1024:
            g = Bg
        else:
            g = 0 * A
        t = oracle.control(data, 0, g[:, None].astype(np.float32), ds_raw)
        t["brir_index"] = 0
        t["hoa_gains"] = np.eye(16, 4, dtype=np.float32)
        s = slice(k * B, end)
        want[:, s] = ref.process(audio[:1, s], audio[1:2, s], audio[2:, s], t)
    assert np.abs(want).max() > 0.05
    np.testing.assert_allclose(out, want, atol=1e-5, rtol=0)
