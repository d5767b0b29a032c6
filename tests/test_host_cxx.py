"""include/bear_b200/host.hpp: the C++ host layer over the C ABI.  The C++ test
program tests/cxx/test_host.cpp restates the reference's tests for config validation, listener
axes, view selection, block timing and SH rotation; here it is compiled and run, and on the GPU a
scene rendered through the C++ Renderer is compared with the oracle."""
import subprocess

import numpy as np
import pytest

from bear_b200 import control, data_file, tensorfile
from conftest import ROOT


def _compile(tmp_path_factory, name):
    exe = tmp_path_factory.mktemp("cxx") / name
    lib_dir = ROOT / "bear_b200"
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-pthread", "-I", str(ROOT / "include"),
                           str(ROOT / "tests" / "cxx" / (name + ".cpp")), "-o", str(exe), "-L", str(lib_dir),
                           "-lbear_cuda", "-Wl,-rpath," + str(lib_dir)])
    return str(exe)


@pytest.fixture(scope="module")
def test_host(tmp_path_factory):
    return _compile(tmp_path_factory, "test_host")


@pytest.fixture(scope="module")
def test_dynamic(tmp_path_factory):
    return _compile(tmp_path_factory, "test_dynamic")


def test_wrappers_without_gpu(test_dynamic):
    """VariableBlockSizeAdapter, BackgroundBuilder, DynamicRenderer (include/bear_b200/dynamic.hpp) around a
    fake renderer: one-period delay, metadata queueing, builder start/swap/error, fade sequence."""
    r = subprocess.run([test_dynamic], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK (0 failures)"), r.stdout + r.stderr


@pytest.mark.gpu
def test_wrappers_on_gpu(test_dynamic, scene_file):
    """the same wrappers around bear_b200::Renderer: every dynamic period is the static render times n/B,
    times 1-n/B, zero or unchanged, in a legal order (test_dynamic_renderer.py:253-291)."""
    r = subprocess.run([test_dynamic, "gpu", scene_file[0]], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "OK (0 failures)" in r.stdout, r.stdout + r.stderr


def test_host_layer_without_gpu(test_host):
    r = subprocess.run([test_host], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK (0 failures)"), r.stdout + r.stderr


@pytest.fixture(scope="module")
def scene_file(tmp_path_factory):
    tree = data_file.synthetic_tree(seed=5, num_views=180, brir_length=300, gain_norm=False)
    tree["gain_norm_quick"] = {"factors": np.random.default_rng(1).random((40, 180, 48, 48, 2), dtype=np.float32) + 0.5}
    path = tmp_path_factory.mktemp("data") / "scene.tf"
    tensorfile.write(path, tree)
    return str(path), data_file.from_tree(tree)


@pytest.mark.gpu
def test_renderer_api_on_gpu(test_host, scene_file):
    r = subprocess.run([test_host, "gpu", scene_file[0]], capture_output=True, text=True)
    assert r.returncode == 0 and "OK (0 failures)" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_cxx_renderer_matches_oracle(oracle, test_host, scene_file, tmp_path):
    path, data = scene_file
    K, B, L = 8, 256, data.num_virtual_loudspeakers
    audio = np.concatenate([control.synth_audio(2, K * B, 0), control.synth_audio(1, K * B, 50), control.synth_audio(4, K * B, 90)])
    (tmp_path / "in.bin").write_bytes(audio.astype(np.float32).tobytes())
    r = subprocess.run([test_host, "render", path, str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), str(K)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    out = np.frombuffer((tmp_path / "out.bin").read_bytes(), np.float32).reshape(2, K * B)

    # the same scene, period by period, through the oracle (schedule semantics written out by hand)
    def vec(d, q):
        g = np.zeros(2 * L)
        for i, v in d.items():
            g[i] = v
        for i, v in q.items():
            g[L + i] = v
        return g
    A, Bg, C = vec({2: .8, 5: .3}, {}), vec({6: .5, 9: .6}, {6: .2, 9: .3}), vec({0: .4}, {0: .4, 1: .1})
    ds_raw = np.zeros((L, 1), np.float32); ds_raw[5, 0] = 1.0
    hoa = np.eye(16, 4, dtype=np.float32)
    per = []
    for k in range(K):
        end = (k + 1) * B
        if end <= 1024:
            g0 = A
        elif end < 2048:
            p = (end - 1024) / 1024
            g0 = (1 - p) * A + p * Bg
        elif end == 2048:
            g0 = Bg
        else:
            g0 = 0 * A
        raw = np.stack([g0, C], 1).astype(np.float32)
        view = 0 if k < 2 else 2          # look vector at +4.6 degrees: nearest view is 4 degrees (index 2 of the 2-degree grid)
        t = oracle.control(data, view, raw, ds_raw)
        t["brir_index"] = view
        t["hoa_gains"] = hoa if k < 2 else None
        per.append((t, view))
    # HOA matrix after the yaw: rotation of the first-order components (checked against host.hpp's own test)
    from math import cos, sin, radians
    a = radians(4.6)
    rot = np.eye(16, dtype=np.float64)
    # M Y(p) = Y(q p): the orientation {cos(a/2),0,0,-sin(a/2)} turns directions by -a in azimuth, so
    # ACN1 = sin(az) -> sin(az - a) = cos(a) ACN1 - sin(a) ACN3, ACN3 = cos(az) -> cos(a) ACN3 + sin(a) ACN1
    rot[1, 1], rot[1, 3], rot[3, 1], rot[3, 3] = cos(a), -sin(a), sin(a), cos(a)
    ref = oracle.RefRenderer(data, 2, 1, 4, B)
    want = np.zeros((2, K * B))
    for k, (t, view) in enumerate(per):
        t["hoa_gains"] = (hoa if k < 2 else (rot @ np.eye(16, 4)).astype(np.float32))
        s = slice(k * B, (k + 1) * B)
        want[:, s] = ref.process(audio[:2, s], audio[2:3, s], audio[3:, s], t)
    assert np.abs(want).max() > 0.05
    np.testing.assert_allclose(out, want, atol=1e-5, rtol=0)
