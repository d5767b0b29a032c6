"""Control tail: delays, gain normalisation and direct/diffuse split between the gain
calculators and the DSP ports.  The oracle restates the in-tree reference source
(panner.cpp:189-235,277-404; gain_norm.cpp:26-44); the CUDA kernel is compared with it."""
import numpy as np
import pytest

from bear_b200 import control, data_file


@pytest.fixture(scope="module")
def gn_data():
    return data_file.from_tree(data_file.synthetic_tree(seed=3, num_views=4, brir_length=260, gain_norm=True))


def scene(ctl, T, N, seed=0):
    rng = np.random.default_rng(seed)
    az = rng.uniform(-180, 180, size=(T, N))
    dif = np.where(np.arange(N) % 3 == 1, 0.5, 0.0) + np.where(np.arange(N) % 5 == 4, 1.0, 0.0) * 0.5
    raw = ctl.raw_object_gains(az, np.clip(dif, 0, 1), rng.uniform(0.2, 1.0, size=(T, N)))
    raw[:, :, 0] = 0.0                                  # a silent object: falls back to the front loudspeaker delays
    if N > 2:                                           # a dense (extent-like) object
        raw[:, :, 2] = np.abs(rng.standard_normal((T, raw.shape[1]))).astype(np.float32) * 0.2
    L = ctl.L
    ds = np.zeros((T, L, 3), np.float32)
    ds[:, 2, 0] = 1.0; ds[:, 5, 1] = 0.7; ds[:, 6, 1] = 0.3   # last channel silent
    view = (np.arange(T) * 3) % ctl.views.shape[0]
    return raw, ds, view.astype(np.int32)


def numpy_compensation(data, g, dly, view):
    """independent evaluation of Panner::compensation_gain (panner.cpp:277-316,332-346,364-380)"""
    F = data.gain_norm_factors.astype(np.float64)
    S = F.shape[0]
    expected = np.sqrt(sum((g ** 2 * np.diagonal(F[0, view, :, :, e])).sum() for e in (0, 1)))
    total = 0.0
    for e in (0, 1):
        d = min(max(0.0, dly[e] * data.fs - data.decorrelation_delay), S - 1.0)
        di = min(int(np.floor(d)), S - 2)
        for s in (di, di + 1):
            total += (1 - abs(d - s)) * (g @ F[s, view, :, :, e] @ g)
    real = np.sqrt(total)
    return 1.0 if abs(real) < 1e-6 else expected / real


def test_oracle_control_matches_formulas(oracle, gn_data):
    ctl = control.HostControl(gn_data)
    raw, ds, view = scene(ctl, 3, 7)
    L = ctl.L
    for t in range(3):
        out = oracle.control(gn_data, view[t], raw[t], ds[t])
        g = raw[t].astype(np.float64)
        want_d = ctl.mean_delays((g[:L] + g[L:]).T, view[t])                     # [N, 2]
        np.testing.assert_allclose(out["direct_delays"].reshape(-1, 2), want_d, rtol=1e-6)
        np.testing.assert_allclose(out["static_delays"], ctl.static_delays(view[t]), rtol=1e-6)
        for i in range(raw.shape[2]):
            comp = numpy_compensation(gn_data, g[:, i], out["direct_delays"][2 * i:2 * i + 2].astype(np.float64), view[t])
            np.testing.assert_allclose(out["direct_gains"][:, i], comp * g[:L, i], rtol=2e-6, atol=1e-9)
            np.testing.assert_allclose(out["diffuse_gains"][:, i], comp * g[L:, i], rtol=2e-6, atol=1e-9)
        assert np.all(out["direct_gains"][:, 0] == 0)
        assert np.all(out["direct_speakers_gains"][:, 2] == 0)
        comp0 = out["direct_speakers_gains"][2, 0]      # single loudspeaker: expected == real -> 1
        assert abs(comp0 - 1.0) < 1e-6


def test_oracle_control_without_table_is_identity(oracle, small_data):
    ctl = control.HostControl(small_data)
    raw, ds, view = scene(ctl, 1, 4)
    out = oracle.control(small_data, view[0], raw[0], ds[0])
    L = ctl.L
    np.testing.assert_array_equal(out["direct_gains"], raw[0, :L])
    np.testing.assert_array_equal(out["diffuse_gains"], raw[0, L:])
    np.testing.assert_array_equal(out["direct_speakers_gains"], ds[0])


@pytest.mark.gpu
@pytest.mark.parametrize("with_table", [True, False])
def test_device_control_tables_match_oracle(oracle, gn_data, small_data, with_table):
    from bear_b200 import device_io
    from bear_b200.renderer import Renderer
    data = gn_data if with_table else small_data
    ctl = control.HostControl(data)
    T, N, B = 5, 37, 128
    raw, ds, view = scene(ctl, T, N, seed=4)
    with Renderer(data, N, 3, 0, B, max_batch_blocks=T) as r:
        r.control_tables(raw, ds, view)
        r.synchronize()
        io = device_io.DeviceBuffers(r, num_speakers=ctl.L)
        got = {k: v.cpu().numpy() for k, v in io.tables.items()}
    for t in range(T):
        want = oracle.control(data, view[t], raw[t], ds[t])
        for k, w in want.items():
            # double arithmetic summed in a different order, rounded to float at the port: <= 1 ulp
            np.testing.assert_allclose(got[k][t].reshape(w.shape), w, rtol=2.5e-7, atol=1e-12, err_msg="%s period %d" % (k, t))
    assert list(got["brir_index"]) == list(view)


@pytest.mark.gpu
def test_render_with_device_control_equals_render_with_host_tables(oracle, gn_data):
    from bear_b200.renderer import Renderer
    ctl = control.HostControl(gn_data)
    T, N, B = 6, 9, 256
    raw, ds, view = scene(ctl, T, N, seed=5)
    obj, dsa = control.synth_audio(N, T * B, 0), control.synth_audio(3, T * B, 50)
    per = [oracle.control(gn_data, view[t], raw[t], ds[t]) for t in range(T)]
    tb = control.BlockTables({k: np.stack([p[k] for p in per]) for k in per[0]})
    tb["brir_index"] = view
    tb["hoa_gains"] = np.zeros((T, 16, 0), np.float32)
    ref = oracle.RefRenderer(gn_data, N, 3, 0, B).render(obj, dsa, None, tb.blocks())
    with Renderer(gn_data, N, 3, 0, B, max_batch_blocks=T) as r:
        r.control_tables(raw, ds, view)
        out = r.process_batch(obj, dsa, None, None)
    assert np.abs(ref).max() > 0.05
    np.testing.assert_allclose(out, ref, atol=1e-5, rtol=0)
