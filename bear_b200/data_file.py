"""BEAR data-file access for the host side, and the synthetic default-shaped file.

`load()` applies the same shape/version checks and error strings as the reference's
`Panner::Panner` (visr_bear/src/panner.cpp:52-137) and `check_data_file_version`
(visr_bear/src/data_file.cpp:16-20).  The product path loads data files with the C++
reader behind the C ABI (`bear_cuda_create_from_file`); this module exists so that the
Python host side (table generation, tests, bench) sees the same tables.

`make_synthetic()` writes a TENF of the production shape pinned by
visr_bear/src/dynamic_renderer.cpp:120-123 (BRIR 2976 taps, decorrelators 128) because
default.tf is a release download that is not in the tree (visr_bear/data/files.txt:1-3).
"""
from dataclasses import dataclass, field
from pathlib import Path

import numpy as np

from . import tensorfile

REPO = Path(__file__).resolve().parent.parent
UNITY_FIXTURE = REPO / "tests" / "golden" / "unity_brirs_decorrelators.tf"


class DataFileError(RuntimeError):
    pass


def _check(ok, message):
    if not ok:
        raise DataFileError("BEAR data file error: " + message)


@dataclass
class DataFile:
    views: np.ndarray                  # [V,3] f32
    brirs: np.ndarray                  # [V,L,2,T] f32, C order
    delays: np.ndarray                 # [V,L,2] f32, samples
    decorrelation_filters: np.ndarray  # [L,Td] f32
    hoa_irs: np.ndarray                # [A,2,Th] f32
    fs: float
    decorrelation_delay: float
    front_loudspeaker: int
    hoa_delay: float = 0.0
    gain_norm_factors: object = None   # [S,V,2L,2L,2] f32 ("gain_norm_quick"/"factors", panner.cpp:113-123) or None
    layout: object = None
    extra: dict = field(default_factory=dict)

    @property
    def num_views(self):
        return self.brirs.shape[0]

    @property
    def num_virtual_loudspeakers(self):
        return self.brirs.shape[1]

    @property
    def brir_length(self):
        return self.brirs.shape[3]

    @property
    def n_hoa_channels(self):
        return self.hoa_irs.shape[0]

    # Panner::get_default_direct_delay / get_default_static_delay, panner.cpp:172-186 (seconds)
    def default_static_delay(self):
        fl = self.front_loudspeaker
        return (float(self.delays[0, fl, 0]) + float(self.delays[0, fl, 1])) / 2.0 / self.fs

    def default_direct_delay(self):
        fl = self.front_loudspeaker
        front = (float(self.delays[0, fl, 0]) + float(self.delays[0, fl, 1])) / 2.0
        return (front + self.decorrelation_delay) / self.fs


def from_tree(tree):
    version = tree.get("bear_data_version", 0)
    if not isinstance(version, int):
        raise DataFileError("version must be an integer")
    if version > 0:
        raise DataFileError("data file version mismatch: expected 0, got %d" % version)

    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    views, brirs, delays = f32(tree["views"]), f32(tree["brirs"]), f32(tree["delays"])
    dec = f32(tree["decorrelation_filters"])
    _check(views.ndim == 2, "views must have 2 dimensions")
    _check(brirs.ndim == 4, "brirs must have 4 dimensions")
    _check(delays.ndim == 3, "delays must have 3 dimensions")
    _check(dec.ndim == 2, "decorrelation_filters must have 2 dimensions")
    V, L = brirs.shape[0], brirs.shape[1]
    _check(views.shape[0] == V, "views axis 0 is wrong size")
    _check(views.shape[1] == 3, "views axis 1 is wrong size")
    _check(brirs.shape[2] == 2, "brirs axis 2 is wrong size")
    _check(delays.shape[0] == V, "delays axis 0 is wrong size")
    _check(delays.shape[1] == L, "delays axis 1 is wrong size")
    _check(delays.shape[2] == 2, "delays axis 2 is wrong size")
    _check(dec.shape[0] == L, "decorrelation filters axis 0 is wrong size")
    _check("hoa" in tree, "HOA decoder not present")
    hoa = f32(tree["hoa"]["irs"])
    _check(hoa.ndim == 3, "HOA decoder must have 3 dimensions")
    _check(hoa.shape[1] == 2, "HOA decoder axis 1 is wrong size")
    order = int(round(np.sqrt(hoa.shape[0]))) - 1
    if (order + 1) ** 2 != hoa.shape[0]:
        raise DataFileError("bad number of hoa channels")
    gn = None
    if "gain_norm_quick" in tree:
        gn = f32(tree["gain_norm_quick"]["factors"])
        _check(gn.ndim == 5, "gain comp factors must have 5 dimensions")
        _check(gn.shape[1] == V, "gain comp factors axis 1 is wrong size")
        _check(gn.shape[2] == L * 2, "gain comp factors axis 2 is wrong size")
        _check(gn.shape[3] == L * 2, "gain comp factors axis 3 is wrong size")
        _check(gn.shape[4] == 2, "gain comp factors axis 4 is wrong size")
    known = {"views", "brirs", "delays", "decorrelation_filters", "hoa", "fs",
             "decorrelation_delay", "front_loudspeaker", "layout", "gain_norm_quick"}
    return DataFile(
        views=views, brirs=brirs, delays=delays, decorrelation_filters=dec, hoa_irs=hoa,
        fs=float(tree["fs"]), decorrelation_delay=float(int(tree["decorrelation_delay"])),
        front_loudspeaker=int(tree["front_loudspeaker"]),
        hoa_delay=float(tree["hoa"].get("delay", 0.0)), gain_norm_factors=gn, layout=tree.get("layout"),
        extra={k: v for k, v in tree.items() if k not in known},
    )


def load(path):
    return from_tree(tensorfile.read(path))


def gain_norm_factors(brirs, decorrelation_filters, delays, dec_delay):
    """The "gain_norm_quick" factor table of a data file, by the recipe of the reference's authoring
    tool (bear/process_irs/finalise.py:210-257): inner products between all pairs of the delayed
    direct-path BRIRs and the decorrelated, delay-rounded diffuse-path BRIRs, per integer direct
    delay 0..max_delay-1, view and ear.  O(S V (2L)^2 T): only for small synthetic files."""
    V, L, _, T = brirs.shape
    b = brirs.astype(np.float64)
    rd = np.round(delays).astype(int)
    S = int(np.max(np.ceil(delays))) + 2
    dl = decorrelation_filters.shape[1]
    ir_len = max(T + dl - 1 + rd.max(), T + int(dec_delay) + S - 1)
    out = np.zeros((S, V, 2 * L, 2 * L, 2), np.float32)
    for v in range(V):
        dif = np.zeros((L, 2, ir_len))
        for l in range(L):
            for e in range(2):
                c = np.convolve(b[v, l, e], decorrelation_filters[l].astype(np.float64))
                dif[l, e, rd[v, l, e]:rd[v, l, e] + len(c)] = c
        for s in range(S):
            direct = np.zeros((L, 2, ir_len))
            direct[:, :, s + int(dec_delay):s + int(dec_delay) + T] = b[v]
            allb = np.concatenate([direct, dif], 0)
            out[s, v] = np.einsum("les,mes->lme", allb, allb)
    return out


def synthetic_tree(seed=0, num_views=180, brir_length=2976, decorrelator_length=128,
                   hoa_ir_length=128, decay=600.0, gain_norm=False):
    """Synthetic data-file tree of the default.tf shape (SURVEY.md §8d recipe).

    Views are unit yaw vectors every 360/V degrees (as the in-tree fixture); delays,
    layout and scalars come from the reference's unity fixture (copied to tests/golden);
    BRIRs are exponentially decaying noise with the front filter at unit energy;
    decorrelators are unit-energy noise; HOA decode filters are N(0, 0.1^2).
    """
    tpl = tensorfile.read(UNITY_FIXTURE)
    rng = np.random.default_rng(seed)
    L = tpl["delays"].shape[1]
    yaw = np.arange(num_views) * (2.0 * np.pi / num_views)
    views = np.stack([-np.sin(yaw), np.cos(yaw), np.zeros_like(yaw)], 1).astype(np.float32)
    t = np.arange(brir_length, dtype=np.float64)
    brirs = rng.standard_normal((num_views, L, 2, brir_length)) * np.exp(-t / decay)
    fl = int(tpl["front_loudspeaker"])
    brirs /= np.sqrt(np.sum(brirs[0, fl, 0] ** 2))
    if num_views == tpl["delays"].shape[0]:
        delays = np.array(tpl["delays"], dtype=np.float32)
    else:  # sub-sample the fixture's views
        idx = (np.arange(num_views) * tpl["delays"].shape[0]) // num_views
        delays = np.array(tpl["delays"][idx], dtype=np.float32)
    dec = rng.standard_normal((L, decorrelator_length))
    dec /= np.sqrt(np.sum(dec ** 2, axis=1, keepdims=True))
    hoa = 0.1 * rng.standard_normal((16, 2, hoa_ir_length))
    extra = {}
    if gain_norm:
        extra["gain_norm_quick"] = {"factors": gain_norm_factors(brirs.astype(np.float32), dec.astype(np.float32),
                                                                 delays, int(tpl["decorrelation_delay"]))}
    return {
        **extra,
        "views": views,
        "brirs": brirs.astype(np.float32),
        "delays": delays,
        "fs": 48000.0,
        "layout": tpl["layout"],
        "decorrelation_filters": dec.astype(np.float32),
        "decorrelation_delay": int(tpl["decorrelation_delay"]),
        "front_loudspeaker": fl,
        "hoa": {"irs": hoa.astype(np.float32), "delay": float(tpl["hoa"]["delay"])},
        "metadata": {"label": "synthetic", "released": False,
                     "description": "synthetic default-shaped BEAR data file (seed %d)" % seed},
    }


def make_synthetic(path, **kw):
    path = Path(path)
    if not path.exists():
        tmp = path.with_suffix(".tmp%d" % np.random.randint(1 << 30))
        tensorfile.write(tmp, synthetic_tree(**kw))
        tmp.replace(path)
    return str(path)
