"""Parity of the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
Bar (BASELINE.json north_star): max-abs error <= 1e-5 and relative error energy <= -100 dB."""
import numpy as np
import pytest

from bear_b200 import control, data_file
from conftest import GOLDEN, rel_err_db

pytestmark = pytest.mark.gpu

ATOL = 1e-5
REL_DB = -100.0


def check(out, ref, what=""):
    e = np.abs(out.astype(np.float64) - ref).max()
    db = rel_err_db(out, ref)
    assert e <= ATOL and db <= REL_DB, "%s: max-abs %.3e, rel %.1f dB" % (what, e, db)


def gpu_render(data, N, M, H, B, tb, obj, ds, hoa, mode, **kw):
    from bear_b200.renderer import Renderer
    T = tb.num_blocks
    if mode == "stream":
        with Renderer(data, N, M, H, B, max_batch_blocks=1, **kw) as r:
            out = np.zeros((2, T * B), np.float32)
            for k in range(T):
                s = slice(k * B, (k + 1) * B)
                out[:, s] = r.process(obj[:, s] if N else None, ds[:, s] if M else None,
                                      hoa[:, s] if H else None, tb.block(k))
            return out
    batch = T if mode == "batch" else int(mode[5:])
    with Renderer(data, N, M, H, B, max_batch_blocks=batch, **kw) as r:
        return r.render(obj, ds, hoa, tb, batch=batch)


def ref_render(oracle, data, N, M, H, B, tb, obj, ds, hoa, **kw):
    r = oracle.RefRenderer(data, N, M, H, B, **kw)
    out = r.render(obj, ds, hoa, tb.blocks())
    r.close()
    return out


def case_tables(ctl, name, T, rng):
    az1 = rng.uniform(-180, 180, size=(T, 1))
    az5 = rng.uniform(-180, 180, size=(T, 5))
    yaw = (30.0 * np.arange(T)) % 360.0
    eye = np.eye(16, 4, dtype=np.float32)
    return {
        "obj_static": (dict(obj_az=np.zeros((T, 1))), 1, 0, 0),
        "obj_moving": (dict(obj_az=az1), 1, 0, 0),
        "obj_diffuse": (dict(obj_az=az1, obj_diffuse=1.0), 1, 0, 0),
        "obj_views": (dict(obj_az=az5, obj_diffuse=[0, .5, 0, 1, .2], yaw_deg=yaw), 5, 0, 0),
        "ds": (dict(ds_speakers=[2, 5, -1]), 0, 3, 0),
        "hoa": (dict(hoa_gains=eye, H=4), 0, 0, 4),
        "all_views": (dict(obj_az=az5, obj_diffuse=[0, .5, 0, 1, .2], ds_speakers=[2, 5, -1], hoa_gains=eye, H=4,
                           yaw_deg=yaw), 5, 3, 4),
    }[name]


@pytest.mark.parametrize("mode", ["batch", "batch3", "stream"])
@pytest.mark.parametrize("B", [128, 512, 2048])
@pytest.mark.parametrize("name", ["obj_static", "obj_moving", "obj_diffuse", "obj_views", "ds", "hoa", "all_views"])
def test_paths_match_oracle(oracle, small_data, name, B, mode):
    T = max(6, 3072 // B)
    ctl = control.HostControl(small_data)
    kw, N, M, H = case_tables(ctl, name, T, np.random.default_rng(B))
    tb = ctl.tables(T, **kw)
    obj, ds, hoa = (control.synth_audio(c, T * B, s) for c, s in ((N, 0), (M, 100), (H, 200)))
    ref = ref_render(oracle, small_data, N, M, H, B, tb, obj, ds, hoa)
    out = gpu_render(small_data, N, M, H, B, tb, obj, ds, hoa, mode)
    assert np.abs(ref).max() > 0.05
    check(out, ref, "%s B=%d %s" % (name, B, mode))


@pytest.mark.parametrize("offsets", [(0, 0, 0), (1, 0, 1), (0, 1, 0)])
def test_ramp_convention_switches_follow_oracle(oracle, small_data, offsets):
    B, T = 256, 8
    ctl = control.HostControl(small_data)
    kw, N, M, H = case_tables(ctl, "all_views", T, np.random.default_rng(1))
    tb = ctl.tables(T, **kw)
    obj, ds, hoa = (control.synth_audio(c, T * B, s) for c, s in ((N, 0), (M, 100), (H, 200)))
    ref = ref_render(oracle, small_data, N, M, H, B, tb, obj, ds, hoa, ramp_offsets=offsets)
    out = gpu_render(small_data, N, M, H, B, tb, obj, ds, hoa, "batch", ramp_offsets=offsets)
    check(out, ref, str(offsets))
    other = ref_render(oracle, small_data, N, M, H, B, tb, obj, ds, hoa, ramp_offsets=tuple(1 - o for o in offsets))
    assert np.abs(out - other).max() > 1e-4   # the switch is not a no-op


@pytest.mark.parametrize("splits", [(1, 1), (2, 3), (4, 16)])
def test_partial_sum_splits(oracle, small_data, splits):
    """object_splits / cmac_splits only change how partial sums are grouped."""
    B, T, N = 256, 6, 70
    ctl = control.HostControl(small_data)
    rng = np.random.default_rng(2)
    tb = ctl.tables(T, obj_az=rng.uniform(-180, 180, size=(T, N)), obj_diffuse=np.where(np.arange(N) % 3 == 0, .4, 0.),
                    obj_gain=0.3, ds_speakers=[2, 5], yaw_deg=30.0 * np.arange(T))
    obj, ds = control.synth_audio(N, T * B, 0), control.synth_audio(2, T * B, 500)
    ref = ref_render(oracle, small_data, N, 2, 0, B, tb, obj, ds, None)
    out = gpu_render(small_data, N, 2, 0, B, tb, obj, ds, None, "batch", object_splits=splits[0], cmac_splits=splits[1])
    check(out, ref, str(splits))


@pytest.mark.parametrize("chunk", [1, 2, 5])
def test_chunked_launch_and_overlapped_upload(oracle, small_data, chunk):
    """A call issued in chunks of periods (uploads of later chunks overlap the kernels of earlier
    ones) gives the same result as the one-launch form."""
    B, T = 256, 11
    ctl = control.HostControl(small_data)
    kw, N, M, H = case_tables(ctl, "all_views", T, np.random.default_rng(7))
    tb = ctl.tables(T, **kw)
    obj, ds, hoa = (control.synth_audio(c, T * B, s) for c, s in ((N, 0), (M, 100), (H, 200)))
    ref = ref_render(oracle, small_data, N, M, H, B, tb, obj, ds, hoa)
    whole = gpu_render(small_data, N, M, H, B, tb, obj, ds, hoa, "batch")
    out = gpu_render(small_data, N, M, H, B, tb, obj, ds, hoa, "batch", chunk_blocks=chunk)
    check(out, ref, "chunk %d" % chunk)
    np.testing.assert_array_equal(out, whole)
    out2 = gpu_render(small_data, N, M, H, B, tb, obj, ds, hoa, "batch6", chunk_blocks=chunk)   # 6 + 5 periods
    check(out2, ref, "chunk %d, two calls" % chunk)


VIEW_PATTERNS = {
    "constant": lambda T, rng: np.zeros(T),
    "alternating": lambda T, rng: 30.0 * (np.arange(T) % 2),
    "runs": lambda T, rng: 30.0 * (np.arange(T) // 7),                    # 7 periods per view: items of 6 + 1
    "random4": lambda T, rng: 30.0 * rng.integers(0, 4, size=T),          # many repeats, items of every size
    "cycle": lambda T, rng: 30.0 * (np.arange(T) % 12),                   # every view, 3-4 periods each
}


@pytest.mark.parametrize("mode,kw", [("batch", {}), ("batch17", {}), ("batch", {"chunk_blocks": 9}),
                                     ("batch", {"cmac_splits": 5})])
@pytest.mark.parametrize("pattern", sorted(VIEW_PATTERNS))
def test_view_grouped_convolution(oracle, small_data, pattern, mode, kw):
    """Launches of >= 8 periods group their periods by BRIR view in the cMAC kernel (one filter load
    for up to 6 periods): view sequences with every item size, across calls and chunked launches."""
    B, T = 256, 40
    ctl = control.HostControl(small_data)
    rng = np.random.default_rng(len(pattern))
    yaw = VIEW_PATTERNS[pattern](T, rng)
    tb = ctl.tables(T, obj_az=rng.uniform(-180, 180, size=(T, 3)), obj_diffuse=[0, .5, 1], ds_speakers=[4],
                    hoa_gains=np.eye(16, 4, dtype=np.float32), H=4, yaw_deg=yaw)
    obj, ds, hoa = (control.synth_audio(c, T * B, s) for c, s in ((3, 0), (1, 100), (4, 200)))
    ref = ref_render(oracle, small_data, 3, 1, 4, B, tb, obj, ds, hoa)
    out = gpu_render(small_data, 3, 1, 4, B, tb, obj, ds, hoa, mode, **kw)
    assert np.abs(ref).max() > 0.05
    check(out, ref, "%s %s %s" % (pattern, mode, kw))


def test_calls_of_irregular_length(oracle, small_data):
    """One renderer fed calls of 1, 2, 9, 3, 14, 1, 5, 8 ... periods: state carried between calls
    (delay-line history of P = 6 periods, input/bus histories, previous tables) whatever the call sizes,
    shorter and longer than the number of partitions, grouped and ungrouped launches."""
    from bear_b200.renderer import Renderer
    B, sizes = 128, [1, 2, 9, 3, 14, 1, 5, 8, 1, 1, 11]
    T = sum(sizes)
    ctl = control.HostControl(small_data)
    rng = np.random.default_rng(21)
    tb = ctl.tables(T, obj_az=rng.uniform(-180, 180, size=(T, 4)), obj_diffuse=[0, .3, 1, .6], ds_speakers=[1, 7],
                    hoa_gains=np.eye(16, 4, dtype=np.float32), H=4, yaw_deg=30.0 * rng.integers(0, 5, size=T))
    obj, ds, hoa = (control.synth_audio(c, T * B, s) for c, s in ((4, 0), (2, 100), (4, 200)))
    ref = ref_render(oracle, small_data, 4, 2, 4, B, tb, obj, ds, hoa)
    out = np.empty((2, T * B), np.float32)
    with Renderer(small_data, 4, 2, 4, B, max_batch_blocks=max(sizes)) as r:
        k0 = 0
        for n in sizes:
            a, b = k0 * B, (k0 + n) * B
            out[:, a:b] = r.process_batch(obj[:, a:b], ds[:, a:b], hoa[:, a:b], tb.slice(k0, k0 + n))
            k0 += n
    check(out, ref, "irregular calls")


def test_single_period_calls_mixed_with_batch_calls(oracle, small_data):
    """Single-period calls run as one captured graph that keeps a host copy of the tables so that a
    port passed as NULL means "unchanged"; a batch call in between makes that copy stale, after which
    NULL ports must still mean "as in the previous period" (the batch call's last one)."""
    from bear_b200.renderer import Renderer
    B, T = 128, 26
    ctl = control.HostControl(small_data)
    rng = np.random.default_rng(33)
    yaw = np.repeat(30.0 * rng.integers(0, 6, size=T // 2), 2)          # the view changes every other period
    tb = ctl.tables(T, obj_az=rng.uniform(-180, 180, size=(T, 2)), obj_diffuse=[0.2, .8], ds_speakers=[3],
                    hoa_gains=np.eye(16, 4, dtype=np.float32), H=4, yaw_deg=yaw)
    obj, ds, hoa = (control.synth_audio(c, T * B, s) for c, s in ((2, 0), (1, 100), (4, 200)))
    ref = ref_render(oracle, small_data, 2, 1, 4, B, tb, obj, ds, hoa)
    out = np.empty((2, T * B), np.float32)

    def single(r, k):
        t = tb.block(k)
        if k > 0 and tb["brir_index"][k] == tb["brir_index"][k - 1]:     # ports the reference writes on change only
            t["brir_index"], t["static_delays"] = -1, None
        if k > 0:
            t["hoa_gains"] = None
        s = slice(k * B, (k + 1) * B)
        out[:, s] = r.process(obj[:, s], ds[:, s], hoa[:, s], t)

    with Renderer(small_data, 2, 1, 4, B, max_batch_blocks=8) as r:
        k = 0
        for kind, n in [("single", 5), ("batch", 7), ("single", 6), ("batch", 3), ("single", 5)]:
            if kind == "single":
                for _ in range(n):
                    single(r, k)
                    k += 1
            else:
                a, b = k * B, (k + n) * B
                out[:, a:b] = r.process_batch(obj[:, a:b], ds[:, a:b], hoa[:, a:b], tb.slice(k, k + n))
                k += n
        assert k == T
    check(out, ref, "single-period and batch calls mixed")


@pytest.mark.parametrize("fmt", ["f32", "s16", "s24", "s32"])
def test_interleaved_file_frames(small_data, fmt):
    """bear_cuda_process_batch_frames: interleaved file frames (float or PCM), a track map with gaps,
    a silent input and a permutation; bit-equal to the planar float call on the decoded samples."""
    from bear_b200.renderer import Renderer
    B, T, C_file = 256, 19, 11                      # 19 periods: chunked upload with a ragged last chunk
    N, M, H = 4, 2, 4
    tracks = np.array([7, 0, -1, 3, 10, 9, 1, 2, 4, 5], np.int32)
    ctl = control.HostControl(small_data)
    rng = np.random.default_rng(5)
    tb = ctl.tables(T, obj_az=rng.uniform(-180, 180, size=(T, N)), obj_diffuse=[0, .4, 1, .1], ds_speakers=[2, 6],
                    hoa_gains=np.eye(16, 4, dtype=np.float32), H=4, yaw_deg=30.0 * rng.integers(0, 4, size=T))
    x = 0.25 * rng.standard_normal((T * B, C_file))
    if fmt == "f32":
        frames, decoded = x.astype(np.float32), x.astype(np.float32)
        kw = {}
    elif fmt == "s16":
        frames = np.clip(np.round(x * 32768.0), -32768, 32767).astype(np.int16)
        decoded, kw = frames.astype(np.float32) / np.float32(32768.0), {}
    elif fmt == "s32":
        frames = np.clip(np.round(x * 2147483648.0), -2**31, 2**31 - 1).astype(np.int32)
        decoded, kw = (frames.astype(np.float64) / 2147483648.0).astype(np.float32), {}
        decoded = frames.astype(np.float32) * np.float32(1.0 / 2147483648.0)   # int -> float rounds, then exact scaling
    else:
        v = np.clip(np.round(x * 8388608.0), -2**23, 2**23 - 1).astype(np.int32)
        b = np.empty((T * B, C_file, 3), np.uint8)
        b[..., 0], b[..., 1], b[..., 2] = v & 0xFF, (v >> 8) & 0xFF, (v >> 16) & 0xFF
        frames, decoded = b.reshape(-1), v.astype(np.float32) / np.float32(8388608.0)
        kw = {"fmt": "s24", "frame_channels": C_file}
    planar = np.zeros((N + M + H, T * B), np.float32)
    for i, c in enumerate(tracks):
        if c >= 0:
            planar[i] = decoded[:, c]
    with Renderer(small_data, N, M, H, B, max_batch_blocks=T) as r:
        want = r.process_batch(planar[:N], planar[N:N + M], planar[N + M:], tb)
    with Renderer(small_data, N, M, H, B, max_batch_blocks=T) as r:
        got = r.process_batch_frames(frames, tracks, tb, **kw)
        assert np.abs(want).max() > 0.02
        np.testing.assert_array_equal(got, want)
        with pytest.raises(ValueError, match="track"):
            r.process_batch_frames(frames, np.where(tracks == 10, 11, tracks), tb, **kw)


def test_more_views_than_the_grouping_table_holds(oracle):
    """A data file with more views than the grouping kernel's histogram (2560) renders through the
    ungrouped kernel; an out-of-range brir_index is still refused."""
    data = data_file.from_tree(data_file.synthetic_tree(seed=5, num_views=2600, brir_length=100))
    B, T = 128, 12
    ctl = control.HostControl(data)
    rng = np.random.default_rng(11)
    tb = ctl.tables(T, obj_az=rng.uniform(-180, 180, size=(T, 2)), obj_diffuse=[0, .7], yaw_deg=rng.uniform(0, 360, size=T))
    assert len(set(tb["brir_index"].tolist())) > 6 and tb["brir_index"].max() > 1000
    obj = control.synth_audio(2, T * B, 0)
    none = np.zeros((0, T * B), np.float32)
    ref = ref_render(oracle, data, 2, 0, 0, B, tb, obj, none, none)
    out = gpu_render(data, 2, 0, 0, B, tb, obj, none, none, "batch")
    check(out, ref, "2600 views")


def test_unity_fixture_path_delays_through_file_loader(oracle, unity_data):
    """visr_bear/test/test_delays.py:114-136 on the CUDA path, data file read by the C++ loader."""
    B, T = 512, 10
    path = str(GOLDEN / "unity_brirs_decorrelators.tf")
    ctl = control.HostControl(unity_data)
    x = np.random.default_rng(0).standard_normal((1, T * B)).astype(np.float32)
    none = np.zeros((0, T * B), np.float32)
    fl = unity_data.front_loudspeaker
    tb_ds, tb_obj = ctl.tables(T, ds_speakers=[fl]), ctl.tables(T, obj_az=np.zeros((T, 1)))
    tb_dif = ctl.tables(T, obj_az=np.zeros((T, 1)), obj_diffuse=1.0)
    y_ds = gpu_render(path, 0, 1, 0, B, tb_ds, none, x, none, "stream")
    y_obj = gpu_render(path, 1, 0, 0, B, tb_obj, x, none, none, "stream")
    y_dif = gpu_render(path, 1, 0, 0, B, tb_dif, x, none, none, "batch")
    np.testing.assert_allclose(y_ds[:, 2 * B:], y_obj[:, 2 * B:], atol=2e-4)
    np.testing.assert_allclose(y_ds[:, 2 * B:], y_dif[:, 2 * B:], atol=2e-4)
    np.testing.assert_allclose(y_ds, ref_render(oracle, unity_data, 0, 1, 0, B, tb_ds, None, x, None), atol=2e-5)


def test_single_block_unchanged_ports(oracle, small_data):
    """Ports the reference only writes on change (static_delays, brir_index, hoa_gains): NULL / -1 in
    the single-block call means 'unchanged' (static_delay_calc.cpp:19-30, select_brir.cpp:34-53)."""
    from bear_b200.renderer import Renderer
    B, T = 256, 6
    ctl = control.HostControl(small_data)
    tb = ctl.tables(T, obj_az=np.full((T, 1), 20.0), hoa_gains=np.eye(16, 4, dtype=np.float32), H=4,
                    yaw_deg=np.array([0, 0, 60, 60, 60, 60.0]))
    obj, hoa = control.synth_audio(1, T * B, 0), control.synth_audio(4, T * B, 9)
    ref = ref_render(oracle, small_data, 1, 0, 4, B, tb, obj, None, hoa)
    out = np.zeros((2, T * B), np.float32)
    with Renderer(small_data, 1, 0, 4, B) as r:
        for k in range(T):
            t = tb.block(k)
            if k not in (0, 2):   # listener "changes" only in blocks 0 and 2
                t["static_delays"], t["brir_index"], t["hoa_gains"] = None, None, None
            s = slice(k * B, (k + 1) * B)
            out[:, s] = r.process(obj[:, s], None, hoa[:, s], t)
    check(out, ref)


def test_reset_makes_one_instance_render_a_batch_of_files(oracle, small_data):
    """BASELINE configs[4] (batch of mixed files): one renderer + reset between files == a fresh
    renderer per file (what the reference does, bear/render_cli.py:503-512) == the oracle."""
    from bear_b200 import parallel
    from bear_b200.renderer import Renderer
    B, T = 512, 6
    ctl = control.HostControl(small_data)
    files = []
    for f in range(3):
        N, M, H, tb = control.scenario(ctl, "cfg5_ht", T, B, seed=20 + f)
        tb["brir_index"] = ((tb["brir_index"] + 5 * f) % small_data.num_views).astype(np.int32)
        tb["static_delays"] = ctl.static_delays(tb["brir_index"]).astype(np.float32)
        audio = [control.synth_audio(c, T * B, 1000 * f + s, 0.05) for c, s in ((N, 0), (M, 100), (H, 200))]
        files.append((*audio, tb))
    with Renderer(small_data, N, M, H, B, max_batch_blocks=T) as r:
        got = {}
        for rank in range(2):                       # two "ranks" share the instance here; files 0,2 then 1
            got.update(parallel.render_file_batch(r, files, rank, 2))
    assert sorted(got) == [0, 1, 2]
    for f, (obj, ds, hoa, tb) in enumerate(files):
        check(got[f], ref_render(oracle, small_data, N, M, H, B, tb, obj, ds, hoa), "file %d" % f)
        fresh = gpu_render(small_data, N, M, H, B, tb, obj, ds, hoa, "batch")
        np.testing.assert_array_equal(got[f], fresh)


def _odd_data(L, A, brir_length, num_views=5):
    """data file with fewer loudspeakers / a lower-order HOA decoder than default.tf"""
    tree = data_file.synthetic_tree(seed=7, num_views=num_views, brir_length=brir_length)
    tree["brirs"] = tree["brirs"][:, :L]
    tree["delays"] = tree["delays"][:, :L]
    tree["decorrelation_filters"] = tree["decorrelation_filters"][:L]
    tree["layout"] = dict(tree["layout"], channels=tree["layout"]["channels"][:L])
    tree["hoa"] = dict(tree["hoa"], irs=tree["hoa"]["irs"][:A])
    return data_file.from_tree(tree)


@pytest.mark.parametrize("L,A,brir_length,B,N,M,H", [
    (10, 4, 100, 128, 3, 2, 4),      # single partition, first-order decoder, 10 loudspeakers
    (7, 9, 1500, 1024, 33, 1, 9),    # odd loudspeaker count (gain rows not a multiple of 4), P = 2
    (24, 16, 2976, 4096, 2, 0, 1),   # largest period, P = 1
    (3, 1, 700, 256, 1, 5, 1),       # tiny layout, order-0 decoder
])
def test_unusual_shapes(oracle, L, A, brir_length, B, N, M, H):
    data = _odd_data(L, A, brir_length)
    ctl = control.HostControl(data)
    T = max(4, 2048 // B)
    rng = np.random.default_rng(L * 100 + A)
    ds_spk = [int(rng.integers(-1, L)) for _ in range(M)]
    tb = ctl.tables(T, obj_az=rng.uniform(-180, 180, size=(T, N)), obj_diffuse=rng.uniform(0, 1, size=N),
                    obj_gain=1.0 / np.sqrt(max(N, 1)), ds_speakers=ds_spk, hoa_gains=0.3 * rng.standard_normal((A, H)), H=H,
                    yaw_deg=72.0 * np.arange(T))
    obj, ds, hoa = (control.synth_audio(c, T * B, s) for c, s in ((N, 0), (M, 100), (H, 200)))
    ref = ref_render(oracle, data, N, M, H, B, tb, obj, ds, hoa)
    assert np.abs(ref).max() > 0.02
    for mode in ("batch", "stream"):
        check(gpu_render(data, N, M, H, B, tb, obj, ds, hoa, mode), ref, "L=%d A=%d B=%d %s" % (L, A, B, mode))


def test_shared_filter_bank(oracle, small_data):
    """bear_cuda_create_shared: renderers with different channel counts share one set of spectra; each
    matches the oracle, and the tables outlive the renderer that built them."""
    from bear_b200.renderer import Renderer
    B, T = 256, 6
    ctl = control.HostControl(small_data)
    first = Renderer(small_data, 1, 0, 0, B, max_batch_blocks=T)
    kw, N, M, H = case_tables(ctl, "all_views", T, np.random.default_rng(3))
    second = Renderer(first, N, M, H, B, max_batch_blocks=T)
    third = Renderer(second, 0, 3, 0, B, max_batch_blocks=1)
    assert first.info.shared_bytes == second.info.shared_bytes == third.info.shared_bytes > 0
    with pytest.raises(ValueError, match="period size differs"):
        Renderer(first, 1, 0, 0, 2 * B)
    first.close()                                  # the bank lives on in the other two
    tb = ctl.tables(T, **kw)
    obj, ds, hoa = (control.synth_audio(c, T * B, s) for c, s in ((N, 0), (M, 100), (H, 200)))
    check(second.render(obj, ds, hoa, tb), ref_render(oracle, small_data, N, M, H, B, tb, obj, ds, hoa), "second")
    tb3 = ctl.tables(T, ds_speakers=[2, 5, -1])
    out3 = np.concatenate([third.process(None, ds[:, k * B:(k + 1) * B], None, tb3.block(k)) for k in range(T)], axis=1)
    check(out3, ref_render(oracle, small_data, 0, 3, 0, B, tb3, None, ds, None), "third")
    second.close(); third.close()


def test_argument_errors(small_data):
    from bear_b200.renderer import Renderer
    B = 256
    ctl = control.HostControl(small_data)
    tb = ctl.tables(2, obj_az=np.zeros((2, 1)))
    obj = control.synth_audio(1, 2 * B, 0)
    with Renderer(small_data, 1, 0, 0, B, max_batch_blocks=1) as r:
        with pytest.raises(ValueError, match="max_batch_blocks"):
            r.process_batch(obj, None, None, tb)
        bad = tb.slice(0, 1)
        bad["brir_index"] = np.array([small_data.num_views], np.int32)
        with pytest.raises(ValueError, match="brir_index out of range"):
            r.process_batch(obj[:, :B], None, None, bad)
        with pytest.raises(ValueError, match="shape"):
            r.process(obj[:, :B - 1], None, None, tb.block(0))
        # still usable after errors
        r.process(obj[:, :B], None, None, tb.block(0))


def _bad_tables(small_data, T, B, what):
    ctl = control.HostControl(small_data)
    tb = ctl.tables(T, obj_az=np.zeros((T, 2)), ds_speakers=[2])
    if what == "long":
        tb["direct_delays"][T - 1, 1] = 0.5          # 24000 samples; the reference allows 1 s (per_ear_delay.cpp:25)
    elif what == "negative":
        tb["direct_speakers_delays"][0, 0] = -1e-4
    elif what == "static":
        tb["static_delays"][1, 5] = 0.25
    elif what == "nan":
        tb["direct_delays"][0, 0] = np.nan
    return tb


@pytest.mark.parametrize("what", ["long", "negative", "static", "nan"])
def test_table_delays_out_of_range_are_reported(small_data, what):
    """A delay the history buffers cannot hold is an error of the call (it used to be clamped silently);
    nothing outside the renderer's buffers is touched and the renderer stays usable."""
    from bear_b200.renderer import Renderer
    from bear_b200 import device_io
    B, T = 256, 3
    tb = _bad_tables(small_data, T, B, what)
    good = control.HostControl(small_data).tables(T, obj_az=np.zeros((T, 2)), ds_speakers=[2])
    obj, ds = control.synth_audio(2, T * B, 0), control.synth_audio(1, T * B, 50)
    with Renderer(small_data, 2, 1, 0, B, max_batch_blocks=T) as r:
        with pytest.raises(ValueError, match="table delay"):
            r.process_batch(obj, ds, None, tb)
        r.reset()
        r.process_batch(obj, ds, None, good)
        # one period per call (plain and graph path)
        r.reset()
        with pytest.raises(ValueError, match="table delay"):
            for k in range(T):
                r.process(obj[:, k * B:(k + 1) * B], ds[:, k * B:(k + 1) * B], None, tb.block(k))
        # device-resident tables: reported by the next synchronize
        r.reset()
        io = device_io.DeviceBuffers(r)
        io.load_audio(np.concatenate([obj, ds]))
        io.load_tables(tb)
        r.process_device(T)
        with pytest.raises(ValueError, match="table delay"):
            r.synchronize()
        r.synchronize()   # reported once


def test_view_index_out_of_range_on_the_device_path(small_data):
    from bear_b200.renderer import Renderer
    from bear_b200 import device_io
    B = 256
    for T in (3, 9):      # plain and view-grouped convolution kernel
        ctl = control.HostControl(small_data)
        tb = ctl.tables(T, obj_az=np.zeros((T, 1)))
        tb["brir_index"][T - 2] = small_data.num_views + 7
        obj = control.synth_audio(1, T * B, 0)
        with Renderer(small_data, 1, 0, 0, B, max_batch_blocks=T) as r:
            io = device_io.DeviceBuffers(r)
            io.load_audio(obj)
            io.load_tables(tb)
            r.process_device(T)
            with pytest.raises(ValueError, match="brir_index out of range"):
                r.synchronize()
            assert np.isfinite(io.read_output(T)).all()


def test_dense_and_listed_gain_columns_many_objects(oracle, small_data):
    """Gain columns of every density in one scene of 70 objects (several staging chunks, two object
    splits): pair-wise panned (2 non-zeros), 3 and 4 non-zeros (the record lists of mix_prep_kernel),
    and all 24 non-zero (`extent` objects of the reference benchmark, visr_bear/test/benchmark.cpp:132-139,
    274-277: the table-driven path), direct and diffuse, changing every period."""
    B, T, N = 256, 9, 70
    L = small_data.num_virtual_loudspeakers
    ctl = control.HostControl(small_data)
    rng = np.random.default_rng(5)
    tb = ctl.tables(T, obj_az=rng.uniform(-180, 180, size=(T, N)), obj_diffuse=np.where(np.arange(N) % 3 == 0, .4, 0.),
                    obj_gain=0.3, ds_speakers=[2, 5], yaw_deg=30.0 * np.arange(T))
    for i in range(N):
        kind = i % 5
        if kind in (1, 2, 3):
            nnz = {1: 3, 2: 4, 3: L}[kind]
            for key in ("direct_gains", "diffuse_gains"):
                g = np.zeros((T, L), np.float32)
                for k in range(T):
                    idx = rng.choice(L, size=nnz, replace=False)
                    g[k, idx] = rng.uniform(-0.3, 0.3, size=nnz)
                if key == "diffuse_gains" and i % 2:
                    g[:] = 0
                tb[key][:, :, i] = g
    dsg = tb["direct_speakers_gains"]
    dsg[:, :, 1] = rng.uniform(0, 0.2, size=(T, L))      # a dense DirectSpeakers column as well
    obj, ds = control.synth_audio(N, T * B, 0), control.synth_audio(2, T * B, 500)
    ref = ref_render(oracle, small_data, N, 2, 0, B, tb, obj, ds, None)
    for kw in ({}, {"object_splits": 3}, {"chunk_blocks": 4}):
        out = gpu_render(small_data, N, 2, 0, B, tb, obj, ds, None, "batch", **kw)
        check(out, ref, "dense %s" % kw)
    out = gpu_render(small_data, N, 2, 0, B, tb, obj, ds, None, "stream")
    check(out, ref, "dense stream")


def test_zero_channel_configs(small_data):
    from bear_b200.renderer import Renderer
    B = 128
    ctl = control.HostControl(small_data)
    tb = ctl.tables(3)
    with Renderer(small_data, 0, 0, 0, B, max_batch_blocks=3) as r:   # nothing connected: silence
        assert not r.process_batch(None, None, None, tb).any()
        assert not r.process(None, None, None, tb.block(0)).any()     # NULL arrays allowed (test_renderer.cpp:27,57)


def test_device_resident_path_matches_host_path(small_data):
    """bear_cuda_process_device on the staging areas == bear_cuda_process_batch from host buffers."""
    import torch
    from bear_b200.renderer import Renderer
    from bear_b200 import device_io
    B, T = 512, 5
    ctl = control.HostControl(small_data)
    N, M, H, tb = control.scenario(ctl, "cfg5_ht", T, B, seed=3)
    tb["brir_index"] = (tb["brir_index"] % small_data.num_views).astype(np.int32)
    audio = [control.synth_audio(c, 2 * T * B, s) for c, s in ((N, 0), (M, 100), (H, 200))]
    with Renderer(small_data, N, M, H, B, max_batch_blocks=T) as a, Renderer(small_data, N, M, H, B, max_batch_blocks=T) as b:
        io = device_io.DeviceBuffers(b)
        for call in range(2):
            s = slice(call * T * B, (call + 1) * T * B)
            want = a.process_batch(*[x[:, s] for x in audio], tb)
            io.load_audio(np.concatenate([x[:, s] for x in audio]))
            io.load_tables(tb)
            b.process_device(T)
            b.synchronize()
            got = io.read_output(T)
            np.testing.assert_array_equal(got, want)


@pytest.fixture(scope="module")
def default_shape_data():
    return data_file.from_tree(data_file.synthetic_tree(seed=0, num_views=24))


@pytest.mark.parametrize("name,B,T", [("cfg1", 512, 24), ("cfg2_ht", 512, 24), ("cfg3_ht", 512, 16), ("cfg5_ht", 1024, 8),
                                      ("cfg4_ht", 256, 16)])
def test_baseline_configs_default_shape(oracle, default_shape_data, name, B, T):
    """BASELINE.json configs on a data file of the production shape (BRIR 2976 taps, P = 6/12/3)."""
    data = default_shape_data
    ctl = control.HostControl(data)
    N, M, H, tb = control.scenario(ctl, name, T, B, seed=1)
    tb["brir_index"] = (tb["brir_index"] % data.num_views).astype(np.int32)
    tb["static_delays"] = ctl.static_delays(tb["brir_index"]).astype(np.float32)
    # level the scene so that the rendered mix stays within full scale (the 1e-5 bar is absolute)
    scale = 0.1 / max(1.0, np.sqrt((N + M + H) / 4.0))
    obj, ds, hoa = (control.synth_audio(c, T * B, s, scale) for c, s in ((N, 0), (M, 5000), (H, 6000)))
    ref = ref_render(oracle, data, N, M, H, B, tb, obj, ds, hoa)
    out = gpu_render(data, N, M, H, B, tb, obj, ds, hoa, "batch8")
    assert 0.05 < np.abs(ref).max() < 1.5
    check(out, ref, name)


def test_linearity_at_full_size(default_shape_data):
    """Size-independent property (test_changes.py:341-361): a 1024-object render equals the sum of the
    renders of its 8 object groups -- the property object-group sharding relies on."""
    data = default_shape_data
    B, T, N, G = 256, 24, 1024, 8
    ctl = control.HostControl(data)
    _, _, _, tb = control.scenario(ctl, "cfg4_ht", T, B, seed=4)
    tb["brir_index"] = (tb["brir_index"] % data.num_views).astype(np.int32)
    tb["static_delays"] = ctl.static_delays(tb["brir_index"]).astype(np.float32)
    obj = control.synth_audio(N, T * B, 0, 0.1 / np.sqrt(N / 4.0))
    whole = gpu_render(data, N, 0, 0, B, tb, obj, None, None, "batch")
    acc = np.zeros_like(whole, dtype=np.float64)
    per = N // G
    for g in range(G):
        s = slice(g * per, (g + 1) * per)
        part = control.BlockTables({**tb, "direct_gains": tb["direct_gains"][:, :, s],
                                    "diffuse_gains": tb["diffuse_gains"][:, :, s],
                                    "direct_delays": tb["direct_delays"][:, 2 * s.start:2 * s.stop]})
        acc += gpu_render(data, per, 0, 0, B, part, obj[s], None, None, "batch")
    assert np.abs(whole).max() > 0.05
    np.testing.assert_allclose(whole, acc, atol=1e-5)
