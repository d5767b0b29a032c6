#!/usr/bin/env python
"""Developer check: CUDA path vs CPU oracle, path by path, on a small synthetic data file.
Prints one line per case; exits non-zero if any case exceeds 1e-5."""
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from bear_b200 import control, data_file  # noqa: E402
from bear_b200.renderer import Renderer  # noqa: E402
from oracle import bear_ref  # noqa: E402


def err(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    d = a - b
    rel = 10 * np.log10((d ** 2).sum() / max((b ** 2).sum(), 1e-300) + 1e-300)
    return np.abs(d).max(), rel


def main():
    data = data_file.from_tree(data_file.synthetic_tree(seed=1, num_views=12, brir_length=700))
    ctl = control.HostControl(data)
    rng = np.random.default_rng(0)
    bad = 0
    cases = []
    for B in (128, 512, 2048):
        T = max(6, 3072 // B)
        yaw = (30.0 * np.arange(T)) % 360.0
        az1 = rng.uniform(-180, 180, size=(T, 1))
        az5 = rng.uniform(-180, 180, size=(T, 5))
        eye = np.eye(16, 4, dtype=np.float32)
        cases += [
            ("obj1 static", B, T, dict(obj_az=np.zeros((T, 1))), 1, 0, 0),
            ("obj1 moving", B, T, dict(obj_az=az1), 1, 0, 0),
            ("obj1 diffuse", B, T, dict(obj_az=az1, obj_diffuse=1.0), 1, 0, 0),
            ("obj5 mixed+views", B, T, dict(obj_az=az5, obj_diffuse=[0, .5, 0, 1, .2], yaw_deg=yaw), 5, 0, 0),
            ("ds3", B, T, dict(ds_speakers=[2, 5, -1]), 0, 3, 0),
            ("hoa4", B, T, dict(hoa_gains=eye, H=4), 0, 0, 4),
            ("all+views", B, T, dict(obj_az=az5, obj_diffuse=[0, .5, 0, 1, .2], ds_speakers=[2, 5, -1],
                                     hoa_gains=eye, H=4, yaw_deg=yaw), 5, 3, 4),
        ]
    for name, B, T, kw, N, M, H in cases:
        tb = ctl.tables(T, **kw)
        obj = control.synth_audio(N, T * B, 0)
        ds = control.synth_audio(M, T * B, 100)
        hoa = control.synth_audio(H, T * B, 200)
        ref = bear_ref.RefRenderer(data, N, M, H, B).render(obj, ds, hoa, tb.blocks())
        for mode, batch in (("batch", T), ("batch3", 3), ("stream", 1)):
            r = Renderer(data, N, M, H, B, max_batch_blocks=batch)
            if mode == "stream":
                out = np.zeros((2, T * B), np.float32)
                for k in range(T):
                    s = slice(k * B, (k + 1) * B)
                    out[:, s] = r.process(obj[:, s] if N else None, ds[:, s] if M else None,
                                          hoa[:, s] if H else None, tb.block(k))
            else:
                out = r.render(obj, ds, hoa, tb, batch=batch)
            e, rel = err(out, ref)
            flag = "" if e <= 1e-5 and rel <= -100 else "  <-- FAIL"
            bad += bool(flag)
            print("B=%4d %-18s %-7s S=%d KS=%d maxabs=%.3e rel=%.1f dB peak=%.3f%s"
                  % (B, name, mode, r.info.object_splits, r.info.cmac_splits, e, rel, np.abs(ref).max(), flag))
            if flag:
                blk = np.abs(out.astype(np.float64) - ref).reshape(2, T, B).max(axis=(0, 2))
                print("      per-block max err:", np.array2string(blk, precision=2))
            r.close()
    # crude timing: cfg2-like, 1 s per call
    full = data_file.from_tree(data_file.synthetic_tree(seed=0, num_views=32))
    ctl = control.HostControl(full)
    B, T = 512, 94
    N, M, H, tb = control.scenario(ctl, "cfg2_ht", T, B)
    tb["brir_index"] = (np.arange(T) % 32).astype(np.int32)
    tb["static_delays"] = ctl.static_delays(tb["brir_index"]).astype(np.float32)
    obj = control.synth_audio(N, T * B, 0)
    r = Renderer(full, N, M, H, B, max_batch_blocks=T)
    r.process_batch(obj, None, None, tb)
    t0 = time.perf_counter()
    for _ in range(5):
        r.process_batch(obj, None, None, tb)
    dt = (time.perf_counter() - t0) / 5
    print("cfg2-like host API: %.3f ms per %.2f s of audio -> RTF %.0f" % (dt * 1e3, T * B / 48000, T * B / 48000 / dt))
    t0 = time.perf_counter()
    for _ in range(20):
        r.process_device(T)
    r.synchronize()
    dt = (time.perf_counter() - t0) / 20
    print("cfg2-like device:   %.3f ms per %.2f s of audio -> RTF %.0f" % (dt * 1e3, T * B / 48000, T * B / 48000 / dt))
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
