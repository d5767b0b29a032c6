"""Pins the CPU oracle (oracle/bear_ref.cpp) before anything is compared against it:
component checks against numpy / a naive time-domain model / the reference's own libear
BlockConvolver (oracle/_ref, when it was built in the container), and the properties the
reference's tests assert for this path (test_delays.py:114-136, test_changes.py:341-361)."""
import numpy as np
import pytest

from bear_b200 import control
from conftest import rel_err_db


def sparse_random(n, nnz, seed):
    """libear's KAT input recipe (tests/block_convolver_test_utils.cpp:12-23): nnz unit-variance
    samples at random positions, zeros elsewhere."""
    rng = np.random.default_rng(seed)
    x = np.zeros(n, np.float32)
    x[rng.choice(n, size=min(nnz, n), replace=False)] = rng.standard_normal(min(nnz, n)).astype(np.float32)
    return x


@pytest.mark.parametrize("n", [256, 1024, 4096])
def test_rfft_matches_numpy(oracle, n):
    x = np.random.default_rng(n).standard_normal(n).astype(np.float32)
    ref = np.fft.rfft(x.astype(np.float64))
    assert np.abs(oracle.rfft(x, "f64") - ref).max() < 1e-10
    assert np.abs(oracle.rfft(x, "f32") - ref).max() < 5e-5 * np.sqrt(n / 256)


@pytest.mark.parametrize("period,ntaps", [(128, 1), (128, 128), (128, 300), (512, 2976), (256, 2976), (2048, 2976)])
def test_partitioned_convolver_vs_naive(oracle, period, ntaps):
    # recipe of libear/tests/block_convolver_tests.cpp:83-116 (naive O(N*M) convolution), max_error 1e-6
    n = period * 12
    x = sparse_random(n, 200, 1)
    h = sparse_random(ntaps, 50, 2)
    ref = np.convolve(x.astype(np.float64), h.astype(np.float64))[:n]
    assert np.abs(oracle.upols(x, h, period, "f64") - ref).max() < 1e-12
    assert np.abs(oracle.upols(x, h, period, "f32") - ref).max() < 1e-5


def test_partitioned_convolver_vs_reference_libear(oracle):
    """Steady-state agreement with the reference's own in-tree partitioned convolver."""
    rng = np.random.default_rng(3)
    x = (0.1 * rng.standard_normal(512 * 16)).astype(np.float32)
    h = (rng.standard_normal(2976) * np.exp(-np.arange(2976) / 600) / 17).astype(np.float32)
    y_ref = oracle.ear_block_convolve(x, h, 512)
    if y_ref is None:
        pytest.skip("oracle/_ref not built (reference tree absent)")
    y = oracle.upols(x, h, 512, "f32")
    assert np.abs(y - y_ref).max() < 1e-6
    assert rel_err_db(y, y_ref) < -120


def test_delay_line_is_centred_lagrange3(oracle):
    fs, B = 48000.0, 128
    x = np.random.default_rng(0).standard_normal(B * 8).astype(np.float32)
    # integer nominal delay: realised delay d + 1.5, symmetric 4-tap kernel (-1, 9, 9, -1)/16
    d = 7.0
    y = oracle.delay(x, np.full(8, d / fs, np.float32), B, fs, initial_seconds=d / fs, precision="f64")
    xp = np.concatenate([np.zeros(16), x.astype(np.float64)])
    expect = sum(c * xp[16 - 7 - j: 16 - 7 - j + len(x)] for j, c in enumerate([-1 / 16, 9 / 16, 9 / 16, -1 / 16]))
    # (the delay port carries float32 seconds, so the nominal delay is only exact to ~1e-7 relative)
    assert np.abs(y - expect).max() < 5e-6
    # fractional: polynomial input of degree 3 is delayed exactly by d + 1.5
    t = np.arange(B * 8, dtype=np.float64)
    poly = (1e-3 * t) ** 3 - 0.2 * (1e-3 * t) ** 2 + 0.05
    d = 10.3
    y = oracle.delay(poly.astype(np.float32), np.full(8, d / fs, np.float32), B, fs, initial_seconds=d / fs,
                     precision="f64")
    pd = lambda tt: (1e-3 * tt) ** 3 - 0.2 * (1e-3 * tt) ** 2 + 0.05
    assert np.abs(y[B:] - pd(t[B:] - d - 1.5)).max() < 1e-6


def test_delay_ramp_reaches_target_at_block_end(oracle):
    fs, B = 48000.0, 256
    t = np.arange(B * 3, dtype=np.float64)
    x = (0.01 * t).astype(np.float32)  # linear signal: output = 0.01 (t - delay(t) - 1.5)
    d = np.array([5.0, 20.0, 20.0], np.float32) / fs
    y = oracle.delay(x, d, B, fs, initial_seconds=5.0 / fs, precision="f64", ramp_offset=1)
    n = np.arange(B)
    expect_delay = 5.0 + (n + 1) / B * 15.0
    np.testing.assert_allclose(y[B:2 * B], 0.01 * (t[B:2 * B] - expect_delay - 1.5), atol=1e-6)
    y0 = oracle.delay(x, d, B, fs, initial_seconds=5.0 / fs, precision="f64", ramp_offset=0)
    np.testing.assert_allclose(y0[B:2 * B], 0.01 * (t[B:2 * B] - (5.0 + n / B * 15.0) - 1.5), atol=1e-6)


def _render(oracle, data, N, M, H, B, tb, obj=None, ds=None, hoa=None, precision="f32", **kw):
    r = oracle.RefRenderer(data, N, M, H, B, precision=precision, **kw)
    out = r.render(obj, ds, hoa, tb.blocks())
    r.close()
    return out


def test_path_delays_match_on_unity_fixture(oracle, unity_data):
    """visr_bear/test/test_delays.py:114-136: with unity BRIRs and impulse decorrelators the
    DirectSpeakers, objects-direct and objects-diffuse paths give the same output after two periods
    (atol 2e-4 there)."""
    B, T = 512, 10
    ctl = control.HostControl(unity_data)
    x = np.random.default_rng(0).standard_normal((1, T * B)).astype(np.float32)
    y_ds = _render(oracle, unity_data, 0, 1, 0, B, ctl.tables(T, ds_speakers=[unity_data.front_loudspeaker]), ds=x)
    y_obj = _render(oracle, unity_data, 1, 0, 0, B, ctl.tables(T, obj_az=np.zeros((T, 1))), obj=x)
    y_dif = _render(oracle, unity_data, 1, 0, 0, B, ctl.tables(T, obj_az=np.zeros((T, 1)), obj_diffuse=1.0), obj=x)
    assert np.abs(y_ds).max() > 1.0
    np.testing.assert_allclose(y_ds[:, 2 * B:], y_obj[:, 2 * B:], atol=2e-4)
    np.testing.assert_allclose(y_ds[:, 2 * B:], y_dif[:, 2 * B:], atol=2e-4)


def test_first_block_fades_in_from_zero_gain(oracle, unity_data):
    # gains start at 0 and settle during block 0 (test_delays.py:125-126)
    B, T = 512, 3
    ctl = control.HostControl(unity_data)
    x = np.ones((1, T * B), np.float32)
    y = _render(oracle, unity_data, 1, 0, 0, B, ctl.tables(T, obj_az=np.zeros((T, 1))), obj=x, precision="f64")
    n = np.arange(100, B)
    np.testing.assert_allclose(y[0, 100:B], (n + 1) / B, atol=1e-9)   # ramp (n+1)/B once the delayed step arrived
    assert np.abs(y[:, :60]).max() < 1e-12
    np.testing.assert_allclose(y[:, B + 100:], 1.0, atol=1e-6)


def test_linearity_of_items(oracle, small_data):
    """test_changes.py:341-361: rendering items together equals the sum of individual renders (1e-5)."""
    B, T = 512, 6
    ctl = control.HostControl(small_data)
    rng = np.random.default_rng(5)
    az = rng.uniform(-180, 180, size=(T, 2))
    yaw = 30.0 * np.arange(T)
    obj = control.synth_audio(2, T * B, 0)
    ds = control.synth_audio(1, T * B, 10)
    hoa = control.synth_audio(4, T * B, 20)
    eye = np.eye(16, 4, dtype=np.float32)
    all_tb = ctl.tables(T, obj_az=az, obj_diffuse=[0.0, 0.6], ds_speakers=[5], hoa_gains=eye, H=4, yaw_deg=yaw)
    y_all = _render(oracle, small_data, 2, 1, 4, B, all_tb, obj, ds, hoa)
    parts = [
        _render(oracle, small_data, 1, 0, 0, B, ctl.tables(T, obj_az=az[:, :1], obj_diffuse=[0.0], yaw_deg=yaw), obj[:1]),
        _render(oracle, small_data, 1, 0, 0, B, ctl.tables(T, obj_az=az[:, 1:], obj_diffuse=[0.6], yaw_deg=yaw), obj[1:]),
        _render(oracle, small_data, 0, 1, 0, B, ctl.tables(T, ds_speakers=[5], yaw_deg=yaw), ds=ds),
        _render(oracle, small_data, 0, 0, 4, B, ctl.tables(T, hoa_gains=eye, H=4, yaw_deg=yaw), hoa=hoa),
    ]
    np.testing.assert_allclose(y_all, sum(parts), atol=1e-5)


def test_view_change_is_one_block_output_crossfade(oracle, small_data):
    """rcl::InterpolatingFirFilterMatrix as configured in dsp.cpp:73-87: in the block where the view
    index changes, out = (1-r) conv(old view) + r conv(new view) over the same input history."""
    B, T, kc = 256, 8, 5
    ctl = control.HostControl(small_data)
    obj = control.synth_audio(1, T * B, 3)
    base = ctl.tables(T, obj_az=np.full((T, 1), 40.0))
    mk = lambda views: control.BlockTables({**base, "brir_index": np.array(views, np.int32)})
    y_a = _render(oracle, small_data, 1, 0, 0, B, mk([0] * T), obj, precision="f64")
    y_b = _render(oracle, small_data, 1, 0, 0, B, mk([3] * T), obj, precision="f64")
    y_x = _render(oracle, small_data, 1, 0, 0, B, mk([0] * kc + [3] * (T - kc)), obj, precision="f64")
    s = slice(kc * B, (kc + 1) * B)
    r = (np.arange(B) + 1) / B
    # y_b differs from the switched render only through its own block-0 transition, long gone by block kc
    np.testing.assert_allclose(y_x[:, :kc * B], y_a[:, :kc * B], atol=1e-12)
    np.testing.assert_allclose(y_x[:, s], (1 - r) * y_a[:, s] + r * y_b[:, s], atol=1e-9)
    np.testing.assert_allclose(y_x[:, (kc + 1) * B:], y_b[:, (kc + 1) * B:], atol=1e-9)


def test_f32_oracle_within_tolerance_of_f64(oracle, small_data):
    B, T = 512, 8
    ctl = control.HostControl(small_data)
    N, M, H, tb = control.scenario(ctl, "cfg5_ht", T, B, seed=2)
    tb["brir_index"] = (tb["brir_index"] % small_data.num_views).astype(np.int32)
    obj, ds, hoa = (control.synth_audio(c, T * B, s) for c, s in ((N, 0), (M, 100), (H, 200)))
    y32 = _render(oracle, small_data, N, M, H, B, tb, obj, ds, hoa, "f32")
    y64 = _render(oracle, small_data, N, M, H, B, tb, obj, ds, hoa, "f64")
    assert np.abs(y32 - y64).max() < 1e-5 and rel_err_db(y32, y64) < -100
