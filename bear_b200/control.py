"""Host-side producers of the per-block Control->DSP tables.

In the reference these tables are produced every block by the Control graph
(visr_bear/src/control.cpp) from ADM metadata through libear; that stays on the host
(BASELINE.json north_star) and needs Boost + libear, which this image lacks.  What the
render core consumes is only the *tables* (visr_bear/src/top.cpp:34-51).  This module
produces tables with the reference's formulas where they are in-tree, and a documented
synthetic stand-in for the libear panner:

  direct delays     Panner::get_delays                   visr_bear/src/panner.cpp:189-208
  static delays     Panner::get_diffuse_delays           visr_bear/src/panner.cpp:224-235
  view selection    SelectBRIR::process (argmax views.look) visr_bear/src/select_brir.cpp:34-53
  direct/diffuse    DirectDiffuseSplit                   visr_bear/src/direct_diffuse_split.cpp:24-32
  gains             SYNTHETIC: constant-power pair-wise panning on the horizontal ring of the
                    data file's layout (<=2 non-zeros, unit power), standing in for
                    ear::GainCalculatorObjects (panner.cpp:139-155)

All control math is float64 and is cast to float32 at the port, as in the reference
(e.g. direct_delay_calc.cpp:36-37).
"""
import numpy as np


class BlockTables(dict):
    """Dense per-block tables for T blocks (keys follow the DSP port names, dsp.cpp:30-96):

        direct_gains            f32 [T, L, N]
        diffuse_gains           f32 [T, L, N]
        direct_delays           f32 [T, 2N]   seconds, index 2i+e
        direct_speakers_gains   f32 [T, L, M]
        direct_speakers_delays  f32 [T, 2M]   seconds
        static_delays           f32 [T, 2L]   seconds, index 2l+e
        brir_index              i32 [T]
        hoa_gains               f32 [T, A, H]
    """

    @property
    def num_blocks(self):
        return int(self["brir_index"].shape[0])

    def block(self, k, prev_view=None):
        """Tables of block k in the single-block (reference port) form for the oracle."""
        t = {key: self[key][k] for key in (
            "direct_gains", "diffuse_gains", "direct_delays", "direct_speakers_gains",
            "direct_speakers_delays", "static_delays", "hoa_gains")}
        t["brir_index"] = int(self["brir_index"][k])
        return t

    def blocks(self):
        return [self.block(k) for k in range(self.num_blocks)]

    def slice(self, k0, k1):
        return BlockTables({key: v[k0:k1] for key, v in self.items()})


class HostControl:
    def __init__(self, data):
        self.data = data
        self.L = data.num_virtual_loudspeakers
        self.A = data.n_hoa_channels
        self.fs = data.fs
        self.delays = data.delays.astype(np.float64)  # [V, L, 2] samples
        views = data.views.astype(np.float64)
        self.views = views / np.linalg.norm(views, axis=1, keepdims=True)
        ring = []
        if isinstance(data.layout, dict):
            for idx, ch in enumerate(data.layout["channels"]):
                if abs(ch["polar_position"]["elevation"]) < 1e-6 and idx < self.L:
                    ring.append((float(ch["polar_position"]["azimuth"]) % 360.0, idx))
        if len(ring) < 2:  # no usable layout: spread the first min(L,8) speakers evenly
            n = min(self.L, 8)
            ring = [(360.0 * i / n, i) for i in range(n)]
        ring.sort()
        self.ring_az = np.array([a for a, _ in ring])
        self.ring_idx = np.array([i for _, i in ring])

    # ---- synthetic panner -------------------------------------------------------------
    def pan(self, az_deg):
        """az_deg: array [...]; returns gains [..., L] float64, unit power, <=2 non-zeros."""
        az = np.asarray(az_deg, dtype=np.float64) % 360.0
        R = len(self.ring_az)
        hi = np.searchsorted(self.ring_az, az, side="right") % R
        lo = (hi - 1) % R
        a0, a1 = self.ring_az[lo], self.ring_az[hi]
        span = (a1 - a0) % 360.0
        span = np.where(span == 0, 360.0, span)
        f = ((az - a0) % 360.0) / span
        g = np.zeros(az.shape + (self.L,))
        np.put_along_axis(g, self.ring_idx[lo][..., None], np.cos(f * np.pi / 2)[..., None], -1)
        g_hi = np.zeros_like(g)
        np.put_along_axis(g_hi, self.ring_idx[hi][..., None], np.sin(f * np.pi / 2)[..., None], -1)
        return g + g_hi

    def raw_object_gains(self, obj_az, obj_diffuse=None, obj_gain=None):
        """What GainCalcObjects would emit for T periods: [T, 2L, N] float32, rows 0..L-1 direct,
        L..2L-1 diffuse (synthetic panner; no normalisation, no delays)."""
        obj_az = np.asarray(obj_az, dtype=np.float64)
        T, N = obj_az.shape
        g = self.pan(obj_az)
        if obj_gain is not None:
            g = g * np.broadcast_to(np.asarray(obj_gain, dtype=np.float64), (T, N))[..., None]
        dif = np.zeros((T, N)) if obj_diffuse is None else np.broadcast_to(np.asarray(obj_diffuse, dtype=np.float64), (T, N))
        both = np.concatenate([g * np.sqrt(1.0 - dif)[..., None], g * np.sqrt(dif)[..., None]], -1)   # [T,N,2L]
        return np.ascontiguousarray(both.transpose(0, 2, 1), dtype=np.float32)

    # ---- reference formulas ------------------------------------------------------------
    def select_view(self, yaw_deg):
        """Listener yaw (degrees, anticlockwise seen from above) -> view index; the look vector is
        the ADM front (0,1,0) rotated by yaw; closest view by maximum dot product."""
        yaw = np.radians(np.asarray(yaw_deg, dtype=np.float64))
        look = np.stack([-np.sin(yaw), np.cos(yaw), np.zeros_like(yaw)], -1)
        return np.argmax(look @ self.views.T, axis=-1).astype(np.int32)

    def mean_delays(self, gains, view):
        """gains [..., L] (direct+diffuse or DirectSpeakers), view [...] -> seconds [..., 2]."""
        gains = np.asarray(gains, dtype=np.float64)
        d = self.delays[np.asarray(view)]            # [..., L, 2]
        total = gains.sum(-1)
        num = np.einsum("...l,...le->...e", gains, d)
        fl = self.data.front_loudspeaker
        mean = np.where((total > 1e-6)[..., None], num / np.where(total > 1e-6, total, 1.0)[..., None],
                        d[..., fl, :])
        return (mean + self.data.decorrelation_delay) / self.fs

    def static_delays(self, view):
        """view [...] -> seconds [..., 2L], index 2l+e."""
        d = self.delays[np.asarray(view)] / self.fs
        return d.reshape(d.shape[:-2] + (2 * self.L,))

    # ---- table assembly ----------------------------------------------------------------
    def tables(self, T, obj_az=None, obj_diffuse=None, obj_gain=None, ds_speakers=None, hoa_gains=None,
               yaw_deg=None, H=0):
        """Build BlockTables for T blocks.

        obj_az      [T, N] azimuth (deg) of each object at the END of each block (gains are the
                    block-end targets, gain_calc_objects.cpp:136)
        obj_diffuse [T, N] or [N] diffuseness in [0,1] (direct = sqrt(1-d) g, diffuse = sqrt(d) g)
        obj_gain    [T, N] or None linear gain
        ds_speakers [M] virtual-loudspeaker index per DirectSpeakers channel, -1 = silent (LFE)
        hoa_gains   [A, H] or [T, A, H]
        yaw_deg     [T] listener yaw or None (view 0)
        """
        L, A = self.L, self.A
        view = np.zeros(T, np.int32) if yaw_deg is None else self.select_view(yaw_deg)
        tb = BlockTables()
        tb["brir_index"] = view
        tb["static_delays"] = self.static_delays(view).astype(np.float32)
        if obj_az is not None and np.asarray(obj_az).size:
            obj_az = np.asarray(obj_az, dtype=np.float64).reshape(T, -1)
            N = obj_az.shape[1]
            g = self.pan(obj_az)                         # [T, N, L]
            if obj_gain is not None:
                g = g * np.broadcast_to(np.asarray(obj_gain, dtype=np.float64), (T, N))[..., None]
            dif = np.zeros((T, N)) if obj_diffuse is None else np.broadcast_to(
                np.asarray(obj_diffuse, dtype=np.float64), (T, N))
            g_dir = g * np.sqrt(1.0 - dif)[..., None]
            g_dif = g * np.sqrt(dif)[..., None]
            tb["direct_gains"] = np.ascontiguousarray(g_dir.transpose(0, 2, 1), dtype=np.float32)
            tb["diffuse_gains"] = np.ascontiguousarray(g_dif.transpose(0, 2, 1), dtype=np.float32)
            dl = self.mean_delays(g_dir + g_dif, view[:, None])  # [T, N, 2]
            tb["direct_delays"] = dl.reshape(T, 2 * N).astype(np.float32)
        else:
            tb["direct_gains"] = np.zeros((T, L, 0), np.float32)
            tb["diffuse_gains"] = np.zeros((T, L, 0), np.float32)
            tb["direct_delays"] = np.zeros((T, 0), np.float32)
        if ds_speakers is not None and len(ds_speakers):
            ds_speakers = np.asarray(ds_speakers)
            M = len(ds_speakers)
            g = np.zeros((M, L))
            for j, sp in enumerate(ds_speakers):
                if sp >= 0:
                    g[j, sp] = 1.0
            tb["direct_speakers_gains"] = np.ascontiguousarray(
                np.broadcast_to(g.T[None], (T, L, M)), dtype=np.float32)
            dl = self.mean_delays(np.broadcast_to(g[None], (T, M, L)), view[:, None])
            tb["direct_speakers_delays"] = dl.reshape(T, 2 * M).astype(np.float32)
        else:
            tb["direct_speakers_gains"] = np.zeros((T, L, 0), np.float32)
            tb["direct_speakers_delays"] = np.zeros((T, 0), np.float32)
        if hoa_gains is not None and H:
            hg = np.asarray(hoa_gains, dtype=np.float32)
            tb["hoa_gains"] = np.ascontiguousarray(np.broadcast_to(hg, (T, A, H)))
        else:
            tb["hoa_gains"] = np.zeros((T, A, 0), np.float32)
        return tb


def synth_audio(channels, samples, seed_base=0, scale=0.1):
    """float32 N(0, scale^2) per channel, seed = seed_base + channel index
    (mirrors visr_bear/test/benchmark.cpp:79-86)."""
    out = np.empty((channels, samples), np.float32)
    for c in range(channels):
        out[c] = scale * np.random.default_rng(seed_base + c).standard_normal(samples, dtype=np.float32)
    return out


def scenario(ctl, name, T, period, seed=0):
    """The BASELINE.json configs as (N, M, H, tables).  Azimuths ~U(-180,180) change every block
    (benchmark.cpp:121-147 `update_every_time`); `*_ht` variants step the listener yaw 2 deg per
    block so every block crossfades to a new BRIR view."""
    rng = np.random.default_rng(seed)
    L, A = ctl.L, ctl.A
    ht = name.endswith("_ht")
    base = name[:-3] if ht else name
    yaw = (2.0 * np.arange(T)) % 360.0 if ht else None
    eye = np.eye(A, 16, dtype=np.float32)
    if base == "cfg1":    # 1 static object, az 0
        return 1, 0, 0, ctl.tables(T, obj_az=np.zeros((T, 1)), yaw_deg=yaw)
    if base in ("cfg2", "cfg4"):  # N moving objects, new position every block, some diffuse
        N = 64 if base == "cfg2" else 1024
        az = rng.uniform(-180, 180, size=(T, N))
        dif = np.where(np.arange(N) % 4 == 3, 0.5, 0.0)
        return N, 0, 0, ctl.tables(T, obj_az=az, obj_diffuse=dif, yaw_deg=yaw)
    if base == "cfg3":    # 4+7+0 bed (12 ch, LFE silent) + 3rd-order HOA, identity decode-input matrix
        ds = [2, 5, 6, -1, 8, 9, 3, 4, 10, 11, 14, 15][:12]
        return 0, 12, 16, ctl.tables(T, ds_speakers=[s if s < L else -1 for s in ds], hoa_gains=eye, H=16,
                                    yaw_deg=yaw)
    if base == "cfg5":    # one file of the mixed batch: 16 objects + 12 DS + 16 HOA
        az = rng.uniform(-180, 180, size=(T, 16))
        ds = [2, 5, 6, -1, 8, 9, 3, 4, 10, 11, 14, 15]
        return 16, 12, 16, ctl.tables(T, obj_az=az, obj_diffuse=np.where(np.arange(16) % 4 == 3, 0.5, 0.0),
                                     ds_speakers=[s if s < L else -1 for s in ds], hoa_gains=eye, H=16,
                                     yaw_deg=yaw)
    raise ValueError("unknown scenario " + name)
