"""Host table producers against the in-tree formulas (panner.cpp:172-235, select_brir.cpp:34-53)."""
import numpy as np

from bear_b200 import control


def test_pan_is_unit_power_and_sparse(unity_data):
    ctl = control.HostControl(unity_data)
    g = ctl.pan(np.linspace(-180, 180, 721))
    np.testing.assert_allclose((g ** 2).sum(-1), 1.0, atol=1e-12)
    assert ((g != 0).sum(-1) <= 2).all()
    assert ctl.pan(np.array([0.0]))[0, unity_data.front_loudspeaker] == 1.0


def test_mean_delay_formula(unity_data):
    ctl = control.HostControl(unity_data)
    d = unity_data
    g = np.zeros(24); g[[2, 5]] = [0.3, 0.7]
    got = ctl.mean_delays(g, 7)
    dl = d.delays.astype(np.float64)
    want = [(0.3 * dl[7, 2, e] + 0.7 * dl[7, 5, e] + 63.0) / 48000.0 for e in (0, 1)]
    np.testing.assert_allclose(got, want, rtol=1e-12)
    # fallback to the front loudspeaker when the gains sum to <= 1e-6 (panner.cpp:202-205)
    got = ctl.mean_delays(np.zeros(24), 7)
    np.testing.assert_allclose(got, (dl[7, 2] + 63.0) / 48000.0, rtol=1e-12)
    # defaults (panner.cpp:172-186)
    front = (dl[0, 2, 0] + dl[0, 2, 1]) / 2
    assert abs(d.default_static_delay() - front / 48000.0) < 1e-15
    assert abs(d.default_direct_delay() - (front + 63.0) / 48000.0) < 1e-15


def test_static_delays_and_view_selection(unity_data):
    ctl = control.HostControl(unity_data)
    sd = ctl.static_delays(np.array([3]))
    assert sd.shape == (1, 48)
    np.testing.assert_allclose(sd[0, 2 * 5 + 1], unity_data.delays[3, 5, 1] / 48000.0)
    # views are yaw steps of 2 degrees (fixture): yaw 2k degrees selects view k
    assert list(ctl.select_view(np.array([0.0, 2.0, 90.0, 358.0, 1.2]))) == [0, 1, 45, 179, 1]


def test_tables_shapes(unity_data):
    ctl = control.HostControl(unity_data)
    N, M, H, tb = control.scenario(ctl, "cfg5_ht", 4, 512)
    assert (N, M, H) == (16, 12, 16)
    assert tb["direct_gains"].shape == (4, 24, 16) and tb["direct_delays"].shape == (4, 32)
    assert tb["direct_speakers_gains"].shape == (4, 24, 12) and tb["hoa_gains"].shape == (4, 16, 16)
    assert tb["static_delays"].shape == (4, 48) and list(tb["brir_index"]) == [0, 1, 2, 3]
    assert all(v.dtype == (np.int32 if k == "brir_index" else np.float32) for k, v in tb.items())
