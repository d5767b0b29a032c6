"""End-to-end replay of the reference's own golden vectors (visr_bear/test/test_changes.py:294-361,
no-head-track scenes) through the oracle and -- on the GPU box -- the CUDA path.

The goldens were rendered with default.tf, which is not in the tree.  What the three static scenes
need from it is (a) the HOA decoder and the delay table, which the in-tree fixture carries
unchanged, and (b) one 2976-tap BRIR pair per scene, which oracle/identify_conventions.py fitted on
the steady state of each golden (samples >= 4096) and stored in identified_brirs.npz.  The first
eight periods -- where gains ramp up from zero and delays ramp from their defaults -- were never
seen by that fit, so matching them pins the reference's ramp and delay-line conventions; the
reference's own tolerance for these files is atol 1e-4, the bar here is 2e-5."""
import json

import numpy as np
import pytest

from bear_b200 import control, data_file, tensorfile
from conftest import GOLDEN

B, NBLK = 512, 93
N = B * NBLK
ATOL = 2e-5


@pytest.fixture(scope="module")
def replay_data():
    fx = tensorfile.read(GOLDEN / "unity_brirs_decorrelators.tf")
    fit = np.load(GOLDEN / "identified_brirs.npz")
    L = fx["delays"].shape[1]
    brirs = np.zeros((1, L, 2, 2976), np.float32)
    for e in (0, 1):
        brirs[0, 5, e] = fit["no_head_track_direct_speakers_out_ear%d" % e]      # M+030
        brirs[0, 2, e] = fit["no_head_track_objects_out_simple_ear%d" % e]       # M+000
    irs = np.array(fx["hoa"]["irs"], np.float32)
    # default.tf's decoder is left/right symmetric (the frontal golden has identical ears); the
    # fixture's right ear is not, so mirror the left ear: sign -1 for the m < 0 harmonics
    acn = np.arange(irs.shape[0])
    n = np.floor(np.sqrt(acn)).astype(int)
    m = acn - n * (n + 1)
    irs[:, 1] = irs[:, 0] * np.where(m < 0, -1.0, 1.0)[:, None]
    tree = {"views": np.array(fx["views"][:1]), "brirs": brirs, "delays": np.array(fx["delays"][:1]), "fs": 48000.0,
            "layout": fx["layout"], "decorrelation_filters": np.array(fx["decorrelation_filters"]),
            "decorrelation_delay": fx["decorrelation_delay"], "front_loudspeaker": fx["front_loudspeaker"],
            "hoa": {"irs": irs, "delay": fx["hoa"]["delay"]}}
    return data_file.from_tree(tree)


def scene(name, data):
    """(N, M, H, tables, objects, direct_speakers, hoa, golden) of one test_changes.py scene."""
    x0 = np.load(GOLDEN / "input.npy")
    ctl = control.HostControl(data)
    golden = np.load(GOLDEN / ("no_head_track_%s.npy" % name))[:, :N]
    none = np.zeros((0, N), np.float32)
    if name == "hoa_out":        # get_hoa_input(1, "SN3D"), roll 3000: first order, source in front -> ACN 0 and 3
        x = np.roll(x0, 3000, axis=1)[:, :N]
        audio = np.concatenate([x, 0 * x, 0 * x, x])
        return 0, 0, 4, ctl.tables(NBLK, hoa_gains=np.eye(16, 4, dtype=np.float32), H=4), none, none, audio, golden
    if name == "direct_speakers_out":   # PolarSpeakerPosition(30, 0, 1), roll 2000 -> M+030
        return 0, 1, 0, ctl.tables(NBLK, ds_speakers=[5]), none, np.roll(x0, 2000, axis=1)[:, :N], none, golden
    if name == "objects_out_simple":    # PolarPosition(0, 0, 1), roll 0 -> M+000
        return 1, 0, 0, ctl.tables(NBLK, obj_az=np.zeros((NBLK, 1))), x0[:, :N], none, none, golden
    raise ValueError(name)


SCENES = ["hoa_out", "direct_speakers_out", "objects_out_simple"]


@pytest.mark.parametrize("name", SCENES)
def test_oracle_reproduces_reference_golden(oracle, replay_data, name):
    Nn, M, H, tb, obj, ds, hoa, golden = scene(name, replay_data)
    r = oracle.RefRenderer(replay_data, Nn, M, H, B)
    out = r.render(obj, ds, hoa, tb.blocks())
    assert np.abs(golden).max() > 0.1
    np.testing.assert_allclose(out[:, :8 * B], golden[:, :8 * B], atol=ATOL, rtol=0)   # transient: never fitted
    np.testing.assert_allclose(out, golden, atol=ATOL, rtol=0)


@pytest.mark.parametrize("offsets", [(0, 1, 1), (1, 0, 1)])
def test_other_conventions_miss_the_golden(oracle, replay_data, offsets):
    """The replay has teeth: the other ramp conventions miss the reference by far more than the bar."""
    Nn, M, H, tb, obj, ds, hoa, golden = scene("direct_speakers_out", replay_data)
    out = oracle.RefRenderer(replay_data, Nn, M, H, B, ramp_offsets=offsets).render(obj, ds, hoa, tb.blocks())
    assert np.abs(out[:, :8 * B] - golden[:, :8 * B]).max() > 10 * ATOL


def test_identification_report_matches_defaults():
    rep = json.loads((GOLDEN / "pinned_conventions.json").read_text())
    assert rep["pinned"]["gain_ramp_offset"] == 1 and rep["pinned"]["delay_ramp_offset"] == 1
    assert rep["pinned"]["lagrange_method_delay"] == 1.5
    for k, v in rep["verdict"].items():
        assert "gain_ramp_offset=1" in v, (k, v)
    hoa = rep["cases"]["no_head_track_hoa_out"]["max_abs_error"]
    assert hoa["method_delay=1.5 gain_ramp_offset=1 ear=0"] < 1e-5 < hoa["method_delay=1.5 gain_ramp_offset=0 ear=0"]


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["stream", "batch"])
@pytest.mark.parametrize("name", SCENES)
def test_cuda_path_reproduces_reference_golden(replay_data, name, mode):
    from bear_b200.renderer import Renderer
    Nn, M, H, tb, obj, ds, hoa, golden = scene(name, replay_data)
    if mode == "batch":
        with Renderer(replay_data, Nn, M, H, B, max_batch_blocks=NBLK) as r:
            out = r.process_batch(obj, ds, hoa, tb)
    else:
        out = np.zeros((2, N), np.float32)
        with Renderer(replay_data, Nn, M, H, B) as r:
            for k in range(NBLK):
                s = slice(k * B, (k + 1) * B)
                out[:, s] = r.process(obj[:, s] if Nn else None, ds[:, s] if M else None, hoa[:, s] if H else None,
                                      tb.block(k))
    np.testing.assert_allclose(out, golden, atol=ATOL, rtol=0)
