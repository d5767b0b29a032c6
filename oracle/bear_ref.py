"""ctypes binding of the CPU oracle (oracle/bear_ref.cpp).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs,
never by the product package."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
_f32p = C.POINTER(C.c_float)


class Setup(C.Structure):
    _fields_ = [
        ("num_objects", C.c_int32), ("num_direct_speakers", C.c_int32), ("num_hoa", C.c_int32),
        ("period", C.c_int32), ("fs", C.c_double),
        ("num_views", C.c_int32), ("num_speakers", C.c_int32), ("brir_length", C.c_int32),
        ("decorrelator_length", C.c_int32), ("num_hoa_decoder_channels", C.c_int32),
        ("hoa_ir_length", C.c_int32),
        ("brirs", _f32p), ("decorrelators", _f32p), ("hoa_irs", _f32p),
        ("default_direct_delay", C.c_double), ("default_static_delay", C.c_double),
        ("hoa_delay", C.c_double), ("max_delay_samples", C.c_int32),
        ("gain_ramp_offset", C.c_int32), ("delay_ramp_offset", C.c_int32),
        ("xfade_ramp_offset", C.c_int32),
    ]


def build():
    subprocess.check_call(["make", "-s", "-C", str(HERE)])


_libs = {}
_native = {"flags": None}


def use_native_build():
    """For the timed CPU baseline (bench.py) only: rebuild the f32 oracle for THIS machine
    (-O3 -march=native, what BASELINE.md section 3 asks for) into oracle/_native/ and make lib("f32")
    use it.  Returns the compiler flags in use; falls back to the shipped -march=x86-64-v3 build when no
    compiler is around.  Parity tests never call this: they run the shipped build."""
    if _native["flags"] is None:
        flags = "-O3 -march=native -std=c++17 -fPIC -shared"
        out = HERE / "_native" / "libbear_ref_f32.so"
        try:
            out.parent.mkdir(exist_ok=True)
            subprocess.check_call(["g++"] + flags.split() + ["-DREAL=float", "-o", str(out), str(HERE / "bear_ref.cpp")],
                                  stderr=subprocess.DEVNULL)
            _native["path"] = out
            _native["flags"] = flags
            _libs.pop("f32", None)
        except (OSError, subprocess.CalledProcessError):
            _native["flags"] = "shipped build (-O3 -march=x86-64-v3)"
    return _native["flags"]



def lib(precision="f32"):
    if precision not in _libs:
        path = HERE / ("libbear_ref_%s.so" % precision)
        if precision == "f32" and _native.get("path"):
            path = _native["path"]
        if not path.exists():
            build()
        l = C.CDLL(str(path))
        l.bear_ref_create.restype = C.c_void_p
        l.bear_ref_create.argtypes = [C.POINTER(Setup)]
        l.bear_ref_destroy.argtypes = [C.c_void_p]
        l.bear_ref_process.argtypes = [C.c_void_p] + [C.c_void_p] * 9 + [C.c_int, C.c_void_p, C.c_void_p]
        l.bear_ref_upols.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        l.bear_ref_delay.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_double, C.c_double,
                                     C.c_int, C.c_int, C.c_void_p]
        l.bear_ref_rfft.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        l.bear_ref_control.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int, C.c_void_p, C.c_void_p,
                                       C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p] + [C.c_void_p] * 6
        _libs[precision] = l
    return _libs[precision]


def _real(precision):
    return np.float32 if precision == "f32" else np.float64


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


class RefRenderer:
    """Block-by-block oracle renderer at the Control->DSP table boundary."""

    def __init__(self, data, num_objects, num_direct_speakers, num_hoa, period, precision="f32",
                 max_delay_samples=256, ramp_offsets=(1, 1, 1)):
        self.l = lib(precision)
        self.real = _real(precision)
        self.N, self.M, self.H, self.B = num_objects, num_direct_speakers, num_hoa, period
        self.L, self.A = data.num_virtual_loudspeakers, data.n_hoa_channels
        self._keep = (_f32(data.brirs), _f32(data.decorrelation_filters), _f32(data.hoa_irs))
        s = Setup()
        s.num_objects, s.num_direct_speakers, s.num_hoa, s.period = num_objects, num_direct_speakers, num_hoa, period
        s.fs = data.fs
        s.num_views, s.num_speakers, s.brir_length = data.num_views, self.L, data.brir_length
        s.decorrelator_length = data.decorrelation_filters.shape[1]
        s.num_hoa_decoder_channels, s.hoa_ir_length = self.A, data.hoa_irs.shape[2]
        s.brirs, s.decorrelators, s.hoa_irs = (a.ctypes.data_as(_f32p) for a in self._keep)
        s.default_direct_delay = data.default_direct_delay()
        s.default_static_delay = data.default_static_delay()
        s.hoa_delay = data.hoa_delay / data.fs
        s.max_delay_samples = max_delay_samples
        s.gain_ramp_offset, s.delay_ramp_offset, s.xfade_ramp_offset = ramp_offsets
        self.h = self.l.bear_ref_create(C.byref(s))

    def close(self):
        if getattr(self, "h", None):
            self.l.bear_ref_destroy(self.h)
            self.h = None

    __del__ = close

    def process(self, obj, ds, hoa, tables):
        """obj [N,B], ds [M,B], hoa [H,B]; tables: dict with the port names of dsp.cpp
        (direct_gains, diffuse_gains, direct_delays, direct_speakers_gains,
        direct_speakers_delays, static_delays|None, brir_index|-1, hoa_gains|None)."""
        out = np.zeros((2, self.B), dtype=self.real)
        a = [_f32(obj), _f32(ds), _f32(hoa)] + [_f32(tables.get(k)) for k in (
            "direct_gains", "diffuse_gains", "direct_delays", "direct_speakers_gains",
            "direct_speakers_delays", "static_delays")]
        hg = _f32(tables.get("hoa_gains"))
        bi = tables.get("brir_index", -1)
        self.l.bear_ref_process(self.h, *[_ptr(x) for x in a], -1 if bi is None else int(bi),
                                _ptr(hg), _ptr(out))
        return out

    def render(self, obj, ds, hoa, block_tables):
        """Render T blocks; obj [N,T*B] etc.; block_tables: list of T dicts."""
        T, B = len(block_tables), self.B
        out = np.zeros((2, T * B), dtype=self.real)
        sl = lambda x, k: None if x is None or x.shape[0] == 0 else x[:, k * B:(k + 1) * B]
        for k, t in enumerate(block_tables):
            out[:, k * B:(k + 1) * B] = self.process(sl(obj, k), sl(ds, k), sl(hoa, k), t)
        return out


def upols(x, taps, period, precision="f32"):
    x, taps = _f32(x), _f32(taps)
    y = np.zeros(len(x), dtype=_real(precision))
    lib(precision).bear_ref_upols(_ptr(x), len(x), _ptr(taps), len(taps), period, _ptr(y))
    return y


def delay(x, delay_seconds, period, fs=48000.0, initial_seconds=0.0, max_delay_samples=256,
          ramp_offset=1, precision="f32"):
    x, d = _f32(x), _f32(delay_seconds)
    y = np.zeros(len(x), dtype=_real(precision))
    lib(precision).bear_ref_delay(_ptr(x), len(x), _ptr(d), period, fs, initial_seconds,
                                  max_delay_samples, ramp_offset, _ptr(y))
    return y


def rfft(x, precision="f32"):
    x = _f32(x)
    out = np.zeros(len(x) + 2, dtype=_real(precision))
    lib(precision).bear_ref_rfft(_ptr(x), len(x), _ptr(out))
    return out[0::2] + 1j * out[1::2]


def ear_block_convolve(x, taps, period):
    """The reference's own libear BlockConvolver (oracle/_ref), if it was built here."""
    path = HERE / "_ref" / "libear_block_convolver.so"
    if not path.exists():
        return None
    l = C.CDLL(str(path))
    x, taps = _f32(x), _f32(taps)
    y = np.zeros(len(x), dtype=np.float32)
    l.ear_block_convolve(_ptr(x), len(x), _ptr(taps), len(taps), period, _ptr(y))
    return y


def control(data, view, gains, ds_gains=None):
    """The Control graph's table producers for one period (oracle restatement, bear_ref_control):
    gains [2L,N] raw object gains (direct rows, then diffuse rows), ds_gains [L,M] -> dict of the DSP
    port tables of that period."""
    L, V = data.num_virtual_loudspeakers, data.num_views
    g = _f32(gains) if gains is not None else np.zeros((2 * L, 0), np.float32)
    dg = _f32(ds_gains) if ds_gains is not None else np.zeros((L, 0), np.float32)
    N, M = g.shape[1], dg.shape[1]
    gn = getattr(data, "gain_norm_factors", None)
    gn = None if gn is None else _f32(gn)
    delays = _f32(data.delays)
    out = {"direct_gains": np.zeros((L, N), np.float32), "diffuse_gains": np.zeros((L, N), np.float32),
           "direct_delays": np.zeros(2 * N, np.float32), "direct_speakers_gains": np.zeros((L, M), np.float32),
           "direct_speakers_delays": np.zeros(2 * M, np.float32), "static_delays": np.zeros(2 * L, np.float32)}
    lib("f32").bear_ref_control(L, V, 0 if gn is None else gn.shape[0], data.fs, data.decorrelation_delay,
                                data.front_loudspeaker, _ptr(delays), _ptr(gn), int(view), N, _ptr(g), M, _ptr(dg),
                                *[_ptr(out[k]) for k in ("direct_gains", "diffuse_gains", "direct_delays",
                                                         "direct_speakers_gains", "direct_speakers_delays", "static_delays")])
    return out
