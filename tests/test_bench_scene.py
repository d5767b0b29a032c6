"""The bench scene is the same scene at every GPU count: the object groups a rank generates for
`bench.py --gpus N` are slices of the single-GPU scene (strong scaling is only meaningful if the work
is), and the bench line keeps the keys the driver reads."""
import importlib.util
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent


def _bench():
    spec = importlib.util.spec_from_file_location("bench_module", ROOT / "bench.py")
    mod = importlib.util.module_from_spec(spec)
    argv, sys.argv = sys.argv, ["bench.py"]
    try:
        spec.loader.exec_module(mod)
    finally:
        sys.argv = argv
    return mod


def test_rank_scenes_are_slices_of_the_whole_scene():
    bench = _bench()
    data, ctl = bench.data_and_control()
    T, B, total = 6, 256, 256
    whole = bench.scene_tables(ctl, T, B, 0, total)
    audio = bench.scene_audio(0, total, T * B, 0.05)
    for world in (2, 4):
        mine = total // world
        for rank in range(world):
            tb = bench.scene_tables(ctl, T, B, rank * mine, mine)
            sl = slice(rank * mine, (rank + 1) * mine)
            assert np.array_equal(tb["direct_gains"], whole["direct_gains"][:, :, sl])
            assert np.array_equal(tb["diffuse_gains"], whole["diffuse_gains"][:, :, sl])
            assert np.array_equal(tb["direct_delays"], whole["direct_delays"][:, 2 * sl.start:2 * sl.stop])
            # what every rank must agree on: the listener
            assert np.array_equal(tb["brir_index"], whole["brir_index"])
            assert np.array_equal(tb["static_delays"], whole["static_delays"])
            assert np.array_equal(bench.scene_audio(rank * mine, mine, T * B, 0.05), audio[sl])


def test_scene_has_the_properties_the_bench_text_claims():
    bench = _bench()
    data, ctl = bench.data_and_control()
    T = 8
    tb = bench.scene_tables(ctl, T, 256, 0, 128)
    views = np.asarray(tb["brir_index"])
    assert np.all(np.diff(views) != 0), "a view change in every period"
    g = np.asarray(tb["direct_gains"])
    assert np.any(g[1:] != g[:-1]), "per-period metadata: the gains move"
    q = np.asarray(tb["diffuse_gains"])
    has_dif = (q != 0).any(axis=(0, 1))
    assert 0.2 <= has_dif.mean() <= 0.3, "every 4th object is half diffuse"
    nnz = ((g != 0) | (q != 0)).sum(axis=1)
    assert nnz.max() <= 3, "point sources: one panning triplet per period"
    dense = bench.scene_tables(ctl, T, 256, 0, 128, dense_every=4)
    nnz_d = (np.asarray(dense["direct_gains"]) != 0).sum(axis=1)
    assert (nnz_d[:, ::4] == ctl.L).all() and nnz_d[:, 1::4].max() <= 3


def test_workload_config_names_the_baseline_configs():
    bench = _bench()

    class A:
        workload = "cfg4_ht"
    c1 = bench.workload_config(A, 1872, 1)
    c8 = bench.workload_config(A, 1872, 8)
    assert c1["objects_total"] == c8["objects_total"] == 1024 and c8["objects_per_gpu"] == 128
    assert "configs[3]" in c1["workload"] and c1["period"] == 256
    assert c1["sharding"] == "single GPU" and "NVLink" in c8["sharding"]
