"""ctypes binding of libbear_cuda.so (include/bear_cuda.h) and a thin host-side renderer.

The class mirrors the part of `bear::Renderer` (visr_bear/include/bear/api.hpp:155-256) that sits
on the hot path -- construction from a Config + data file, `process()` per period -- at the
Control->DSP table boundary: where the reference's `process()` runs its Control graph and then the
DSP graph, `Renderer.process()` here takes the Control outputs (bear_b200.control.BlockTables)
and runs the CUDA DSP.  There is no CPU fallback: a missing library or device raises.
"""
import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "libbear_cuda.so"

OK, ERR_INVALID_ARGUMENT, ERR_DATA_FILE, ERR_CUDA, ERR_UNSUPPORTED = 0, -1, -2, -3, -4

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)


class Config(C.Structure):
    """bear_cuda_config; field meaning follows bear::Config (api.hpp:16-61)."""
    _fields_ = [
        ("num_objects_channels", C.c_uint32), ("num_direct_speakers_channels", C.c_uint32),
        ("num_hoa_channels", C.c_uint32), ("period_size", C.c_uint32), ("sample_rate", C.c_uint32),
        ("max_batch_blocks", C.c_uint32), ("device", C.c_int32), ("max_delay_samples", C.c_uint32),
        ("object_splits", C.c_uint32), ("cmac_splits", C.c_uint32),
        ("gain_ramp_offset", C.c_uint32), ("delay_ramp_offset", C.c_uint32),
        ("xfade_ramp_offset", C.c_uint32), ("chunk_blocks", C.c_uint32), ("reserved", C.c_uint32 * 3),
    ]


class Data(C.Structure):
    _fields_ = [
        ("num_views", C.c_uint32), ("num_virtual_loudspeakers", C.c_uint32), ("brir_length", C.c_uint32),
        ("decorrelator_length", C.c_uint32), ("num_hoa_decoder_channels", C.c_uint32),
        ("hoa_ir_length", C.c_uint32),
        ("brirs", _f32p), ("delays", _f32p), ("decorrelation_filters", _f32p), ("hoa_irs", _f32p),
        ("fs", C.c_double), ("decorrelation_delay", C.c_double), ("hoa_delay", C.c_double),
        ("front_loudspeaker", C.c_uint32), ("num_gain_norm_delays", C.c_uint32),
        ("gain_norm_factors", _f32p),
    ]


class Tables(C.Structure):
    _fields_ = [
        ("direct_gains", _f32p), ("diffuse_gains", _f32p), ("direct_delays", _f32p),
        ("direct_speakers_gains", _f32p), ("direct_speakers_delays", _f32p), ("static_delays", _f32p),
        ("brir_index", _i32p), ("hoa_gains", _f32p),
    ]


class DeviceIO(C.Structure):
    _fields_ = [
        ("audio_in", C.c_void_p), ("audio_stride", C.c_size_t), ("audio_offset", C.c_size_t),
        ("direct_gains", C.c_void_p), ("diffuse_gains", C.c_void_p), ("direct_delays", C.c_void_p),
        ("direct_speakers_gains", C.c_void_p), ("direct_speakers_delays", C.c_void_p),
        ("static_delays", C.c_void_p), ("brir_index", C.c_void_p), ("hoa_gains", C.c_void_p),
        ("audio_out", C.c_void_p), ("out_stride", C.c_size_t), ("stream", C.c_void_p),
    ]


class Info(C.Structure):
    _fields_ = [
        ("num_partitions", C.c_uint32), ("object_splits", C.c_uint32), ("cmac_splits", C.c_uint32),
        ("history_samples", C.c_uint32), ("kernels_per_call", C.c_uint32), ("device_bytes", C.c_uint64),
        ("launches_total", C.c_uint64), ("default_direct_delay", C.c_double),
        ("default_static_delay", C.c_double), ("shared_bytes", C.c_uint64),
        ("cmac_items_last", C.c_uint32), ("cmac_splits_last", C.c_uint32),
        ("world", C.c_uint32), ("rank", C.c_uint32), ("speakers_begin", C.c_uint32), ("speakers_end", C.c_uint32),
    ]


class DataFileInfo(C.Structure):
    _fields_ = [
        ("has_metadata", C.c_int32), ("is_released", C.c_int32), ("label", C.c_char * 256),
        ("description", C.c_char * 1024),
        ("num_views", C.c_uint32), ("num_virtual_loudspeakers", C.c_uint32), ("brir_length", C.c_uint32),
        ("decorrelator_length", C.c_uint32), ("num_hoa_decoder_channels", C.c_uint32),
        ("hoa_ir_length", C.c_uint32), ("front_loudspeaker", C.c_uint32), ("reserved", C.c_uint32),
        ("fs", C.c_double), ("decorrelation_delay", C.c_double), ("hoa_delay", C.c_double),
    ]


EXPORTS = (
    "bear_cuda_create", "bear_cuda_create_from_file", "bear_cuda_destroy", "bear_cuda_last_error",
    "bear_cuda_process", "bear_cuda_process_batch", "bear_cuda_get_device_io",
    "bear_cuda_process_device", "bear_cuda_synchronize", "bear_cuda_get_info",
    "bear_cuda_host_alloc", "bear_cuda_host_free", "bear_cuda_read_data_file_info",
    "bear_cuda_tenf_read_f32", "bear_cuda_set_profiling", "bear_cuda_get_profile",
    "bear_cuda_process_device_static_scratch", "bear_cuda_control_tables", "bear_cuda_reset",
    "bear_cuda_create_shared", "bear_cuda_process_batch_frames",
    "bear_cuda_comm_blob", "bear_cuda_comm_connect", "bear_cuda_comm_speaker_range",
    "bear_cuda_process_device_pipelined", "bear_cuda_result_stream",
)
STATIC_SCRATCH_BYTES = 4 << 20
NUM_KERNELS = 11
COMM_BLOB_BYTES = 256


class KernelProfile(C.Structure):
    _fields_ = [("name", C.c_char * 32), ("total_ms", C.c_double), ("launches", C.c_uint64)]


def build(force=False):
    """Compile libbear_cuda.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    if force:
        subprocess.check_call(["make", "-s", "-C", str(HERE / "csrc"), "clean"])
    subprocess.check_call(["make", "-s", "-C", str(HERE / "csrc")])


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = Path(os.environ.get("BEAR_CUDA_LIB", LIB_PATH))    # (a differently tuned build, for experiments)
        if not path.exists():
            raise RuntimeError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)" % path)
        l = C.CDLL(str(path))
        vp, u32, sz = C.c_void_p, C.c_uint32, C.c_size_t
        l.bear_cuda_create.argtypes = [C.POINTER(Config), C.POINTER(Data), C.POINTER(vp)]
        l.bear_cuda_create_from_file.argtypes = [C.POINTER(Config), C.c_char_p, C.POINTER(vp)]
        l.bear_cuda_create_shared.argtypes = [C.POINTER(Config), vp, C.POINTER(vp)]
        l.bear_cuda_destroy.argtypes = [vp]
        l.bear_cuda_destroy.restype = None
        l.bear_cuda_last_error.argtypes = [vp]
        l.bear_cuda_last_error.restype = C.c_char_p
        l.bear_cuda_process.argtypes = [vp, vp, vp, vp, C.POINTER(Tables), vp]
        l.bear_cuda_process_batch.argtypes = [vp, u32, vp, sz, vp, sz, vp, sz, C.POINTER(Tables), vp, sz]
        l.bear_cuda_process_batch_frames.argtypes = [vp, u32, vp, u32, u32, _i32p, C.POINTER(Tables), vp, sz]
        l.bear_cuda_get_device_io.argtypes = [vp, C.POINTER(DeviceIO)]
        l.bear_cuda_process_device.argtypes = [vp, u32]
        l.bear_cuda_process_device_static_scratch.argtypes = [vp, u32]
        l.bear_cuda_process_device_pipelined.argtypes = [vp, u32]
        l.bear_cuda_result_stream.argtypes = [vp, C.POINTER(C.c_void_p)]
        l.bear_cuda_synchronize.argtypes = [vp]
        l.bear_cuda_reset.argtypes = [vp]
        l.bear_cuda_get_info.argtypes = [vp, C.POINTER(Info)]
        l.bear_cuda_host_alloc.argtypes = [sz]
        l.bear_cuda_host_alloc.restype = vp
        l.bear_cuda_host_free.argtypes = [vp]
        l.bear_cuda_host_free.restype = None
        l.bear_cuda_read_data_file_info.argtypes = [C.c_char_p, C.POINTER(DataFileInfo)]
        l.bear_cuda_tenf_read_f32.argtypes = [C.c_char_p, C.c_char_p, vp, sz, C.POINTER(sz),
                                              C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        l.bear_cuda_control_tables.argtypes = [vp, u32, vp, vp, vp, vp]
        l.bear_cuda_comm_blob.argtypes = [vp, u32, vp]
        l.bear_cuda_comm_connect.argtypes = [vp, u32, u32, vp, C.c_int32]
        l.bear_cuda_comm_speaker_range.argtypes = [u32, u32, u32, C.POINTER(u32), C.POINTER(u32)]
        l.bear_cuda_comm_speaker_range.restype = None
        l.bear_cuda_set_profiling.argtypes = [vp, C.c_int]
        l.bear_cuda_get_profile.argtypes = [vp, C.POINTER(KernelProfile), C.c_int]
        _lib = l
    return _lib


class BearCudaError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(message)
        self.code = code


def _raise(code, message):
    # same exception classes as the reference throws at this boundary (api.cpp:68-72,158-159)
    if code == ERR_INVALID_ARGUMENT:
        raise ValueError(message)
    raise BearCudaError(code, message)


def read_data_file_info(path):
    """bear::DataFileMetadata::read_from_file (api.hpp:261-287) + table dimensions; host only."""
    info = DataFileInfo()
    rc = lib().bear_cuda_read_data_file_info(str(path).encode(), C.byref(info))
    if rc != OK:
        _raise(rc, lib().bear_cuda_last_error(None).decode())
    return info


def tenf_read_f32(path, key):
    """Read one array of a TENF file through the C++ reader; host only."""
    l = lib()
    count, ndim, shape = C.c_size_t(), C.c_uint32(), (C.c_uint32 * 8)()
    rc = l.bear_cuda_tenf_read_f32(str(path).encode(), key.encode(), None, 0, C.byref(count), shape, C.byref(ndim))
    if rc != OK:
        _raise(rc, l.bear_cuda_last_error(None).decode())
    out = np.empty(count.value, np.float32)
    rc = l.bear_cuda_tenf_read_f32(str(path).encode(), key.encode(), C.c_void_p(out.ctypes.data), out.size,
                                   C.byref(count), shape, C.byref(ndim))
    if rc != OK:
        _raise(rc, l.bear_cuda_last_error(None).decode())
    return out.reshape([shape[i] for i in range(ndim.value)])


def pinned_empty(shape, dtype=np.float32):
    """numpy array backed by page-locked host memory (bear_cuda_host_alloc)."""
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    ptr = lib().bear_cuda_host_alloc(max(n, 16))
    if not ptr:
        raise MemoryError("bear_cuda_host_alloc failed")
    buf = (C.c_char * max(n, 16)).from_address(ptr)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _pinned_keep[arr.__array_interface__["data"][0]] = ptr
    return arr


_pinned_keep = {}


def _as_f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class Renderer:
    """One renderer instance bound to one GPU."""

    def __init__(self, data, num_objects=0, num_direct_speakers=0, num_hoa=0, period=0, max_batch_blocks=1,
                 device=0, max_delay_samples=0, object_splits=0, cmac_splits=0, ramp_offsets=(1, 1, 1),
                 sample_rate=48000, chunk_blocks=0):
        """`data`: path to a BEAR data file, a bear_b200.data_file.DataFile, or another Renderer whose
        filter spectra and tables are to be shared (bear_cuda_create_shared)."""
        self._l = lib()
        self._h = C.c_void_p()
        cfg = Config()
        cfg.num_objects_channels, cfg.num_direct_speakers_channels = num_objects, num_direct_speakers
        cfg.num_hoa_channels, cfg.period_size, cfg.sample_rate = num_hoa, period, sample_rate
        cfg.max_batch_blocks, cfg.device, cfg.max_delay_samples = max_batch_blocks, device, max_delay_samples
        cfg.object_splits, cfg.cmac_splits, cfg.chunk_blocks = object_splits, cmac_splits, chunk_blocks
        cfg.gain_ramp_offset, cfg.delay_ramp_offset, cfg.xfade_ramp_offset = ramp_offsets
        if isinstance(data, Renderer):   # share the filter spectra / data-file tables of an existing renderer
            rc = self._l.bear_cuda_create_shared(C.byref(cfg), data._h, C.byref(self._h))
        elif isinstance(data, (str, os.PathLike)):
            rc = self._l.bear_cuda_create_from_file(C.byref(cfg), str(data).encode(), C.byref(self._h))
        else:
            keep = [_as_f32(data.brirs), _as_f32(data.delays), _as_f32(data.decorrelation_filters),
                    _as_f32(data.hoa_irs)]
            d = Data()
            d.num_views, d.num_virtual_loudspeakers = data.num_views, data.num_virtual_loudspeakers
            d.brir_length, d.decorrelator_length = data.brir_length, data.decorrelation_filters.shape[1]
            d.num_hoa_decoder_channels, d.hoa_ir_length = data.hoa_irs.shape[0], data.hoa_irs.shape[2]
            d.brirs, d.delays, d.decorrelation_filters, d.hoa_irs = (a.ctypes.data_as(_f32p) for a in keep)
            d.fs, d.decorrelation_delay, d.hoa_delay = data.fs, data.decorrelation_delay, data.hoa_delay
            d.front_loudspeaker = data.front_loudspeaker
            gn = getattr(data, "gain_norm_factors", None)
            if gn is not None:
                keep.append(_as_f32(gn))
                d.gain_norm_factors = keep[-1].ctypes.data_as(_f32p)
                d.num_gain_norm_delays = gn.shape[0]
            rc = self._l.bear_cuda_create(C.byref(cfg), C.byref(d), C.byref(self._h))
        if rc != OK:
            _raise(rc, self._l.bear_cuda_last_error(None).decode())
        self.N, self.M, self.H, self.B = num_objects, num_direct_speakers, num_hoa, period
        self.max_batch_blocks = max_batch_blocks
        self.info = self.get_info()

    # ------------------------------------------------------------------ lifetime
    def close(self):
        if getattr(self, "_h", None):
            self._l.bear_cuda_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc != OK:
            _raise(rc, self._l.bear_cuda_last_error(self._h).decode())

    def get_info(self):
        i = Info()
        self._check(self._l.bear_cuda_get_info(self._h, C.byref(i)))
        return i

    # ------------------------------------------------------------------ tables
    @staticmethod
    def _tables_struct(tb, keep):
        t = Tables()
        for name, _ in Tables._fields_:
            v = tb.get(name) if tb is not None else None
            if v is None:
                continue
            if name == "brir_index":
                a = np.ascontiguousarray(np.atleast_1d(v), dtype=np.int32)
                keep.append(a)
                t.brir_index = a.ctypes.data_as(_i32p)
            else:
                a = _as_f32(v)
                keep.append(a)
                if a.size:
                    setattr(t, name, a.ctypes.data_as(_f32p))
        return t

    # ------------------------------------------------------------------ processing
    def process(self, objects_input, direct_speakers_input, hoa_input, tables, output=None):
        """One period, the buffer contract of bear::Renderer::process (api.hpp:165-177):
        objects_input [N,B], direct_speakers_input [M,B], hoa_input [H,B] float32 (None when the
        count is 0); `tables`: mapping with the DSP port names, entries may be None = unchanged."""
        keep = []
        B = self.B

        def ptrs(x, count):
            if count == 0:
                return None
            x = _as_f32(x)
            if x.shape != (count, B):
                raise ValueError("input must have shape (%d, %d)" % (count, B))
            keep.append(x)
            arr = (C.c_void_p * count)(*[x[i].ctypes.data for i in range(count)])
            keep.append(arr)
            return C.cast(arr, C.c_void_p)

        out = np.empty((2, B), np.float32) if output is None else output
        outp = (C.c_void_p * 2)(out[0].ctypes.data, out[1].ctypes.data)
        t = self._tables_struct(tables, keep)
        self._check(self._l.bear_cuda_process(self._h, ptrs(objects_input, self.N),
                                              ptrs(direct_speakers_input, self.M), ptrs(hoa_input, self.H),
                                              C.byref(t), C.cast(outp, C.c_void_p)))
        return out

    def control_tables(self, object_gains, direct_speakers_gains, brir_index, hoa_gains=None):
        """bear_cuda_control_tables: raw gain-calculator outputs of T periods -> the renderer's device
        tables (delays, gain normalisation, direct/diffuse split, static delays).  object_gains
        [T,2L,N], direct_speakers_gains [T,L,M], brir_index [T], hoa_gains [T,A,H]."""
        bi = np.ascontiguousarray(brir_index, dtype=np.int32)
        og = None if object_gains is None or self.N == 0 else _as_f32(object_gains)
        dg = None if direct_speakers_gains is None or self.M == 0 else _as_f32(direct_speakers_gains)
        hg = None if hoa_gains is None or self.H == 0 else _as_f32(hoa_gains)
        p = lambda a: None if a is None else C.c_void_p(a.ctypes.data)
        self._check(self._l.bear_cuda_control_tables(self._h, len(bi), p(og), p(dg), p(bi), p(hg)))
        self._pending_T = len(bi)

    def process_batch(self, objects, direct_speakers, hoa, tables, output=None):
        """T periods in one call: objects [N,T*B], direct_speakers [M,T*B], hoa [H,T*B]
        (row-contiguous float32, row stride free), tables: BlockTables with T blocks, or None to
        use the tables prepared by the preceding control_tables() call."""
        keep = []
        T = self._pending_T if tables is None else int(np.asarray(tables["brir_index"]).shape[0])
        TB = T * self.B

        def arr(x, count):
            if count == 0:
                return None, 0
            if x.dtype != np.float32 or x.strides[1] != 4:
                x = _as_f32(x)
            if x.shape[0] != count or x.shape[1] < TB:
                raise ValueError("input must have shape (%d, >=%d)" % (count, TB))
            keep.append(x)
            return C.c_void_p(x.ctypes.data), x.strides[0] // 4

        if output is False:       # a rank of a shared scene that does not receive the mix
            out, out_ptr, out_stride = None, None, 0
        else:
            out = np.empty((2, TB), np.float32) if output is None else output
            out_ptr, out_stride = C.c_void_p(out.ctypes.data), out.strides[0] // 4
        o, os_ = arr(objects, self.N)
        d, ds_ = arr(direct_speakers, self.M)
        h, hs_ = arr(hoa, self.H)
        t = self._tables_struct(tables, keep)
        self._check(self._l.bear_cuda_process_batch(self._h, T, o, os_, d, ds_, h, hs_,
                                                    None if tables is None else C.byref(t),
                                                    out_ptr, out_stride))
        return out

    FRAME_FORMATS = {"f32": 0, "s16": 1, "s24": 2, "s32": 3}

    def process_batch_frames(self, frames, tracks, tables, fmt=None, frame_channels=None, output=None):
        """T periods straight from a file's interleaved frames: `frames` is [T*B, channels] float32 /
        int16 / int32, or a flat uint8 array of packed 3-byte samples with fmt="s24" and
        frame_channels given; `tracks` [N+M+H] = file channel of every input (-1 = silence)."""
        frames = np.ascontiguousarray(frames)
        if fmt is None:
            fmt = {np.dtype(np.float32): "f32", np.dtype(np.int16): "s16", np.dtype(np.int32): "s32"}[frames.dtype]
        if frame_channels is None:
            frame_channels = frames.shape[1]
        T = int(np.asarray(tables["brir_index"]).shape[0])
        bps = {"f32": 4, "s16": 2, "s24": 3, "s32": 4}[fmt]
        if frames.nbytes < T * self.B * frame_channels * bps:
            raise ValueError("frames hold fewer than %d periods" % T)
        tr = np.ascontiguousarray(tracks, dtype=np.int32)
        if tr.shape != (self.N + self.M + self.H,):
            raise ValueError("tracks must have one entry per input channel")
        keep = [frames, tr]
        if output is False:       # a rank of a shared scene that does not receive the mix
            out, out_ptr, out_stride = None, None, 0
        else:
            out = np.empty((2, T * self.B), np.float32) if output is None else output
            out_ptr, out_stride = C.c_void_p(out.ctypes.data), out.strides[0] // 4
        t = self._tables_struct(tables, keep)
        self._check(self._l.bear_cuda_process_batch_frames(
            self._h, T, C.c_void_p(frames.ctypes.data), self.FRAME_FORMATS[fmt], int(frame_channels),
            tr.ctypes.data_as(_i32p), C.byref(t), out_ptr, out_stride))
        return out

    def render(self, objects, direct_speakers, hoa, tables, batch=None, output=None):
        """Render all blocks of `tables`, `batch` blocks per call (default: max_batch_blocks).
        output=False: this rank of a shared scene does not receive the mix (returns None)."""
        T = tables.num_blocks
        batch = batch or self.max_batch_blocks
        out = None if output is False else np.empty((2, T * self.B), np.float32)
        sl = lambda x, a, b: None if x is None or x.shape[0] == 0 else x[:, a * self.B:b * self.B]
        for k0 in range(0, T, batch):
            k1 = min(T, k0 + batch)
            res = self.process_batch(sl(objects, k0, k1), sl(direct_speakers, k0, k1), sl(hoa, k0, k1),
                                     tables.slice(k0, k1), output=False if out is None else None)
            if out is not None:
                out[:, k0 * self.B:k1 * self.B] = res
        return out

    # ------------------------------------------------------------------ device-resident path
    def device_io(self):
        io = DeviceIO()
        self._check(self._l.bear_cuda_get_device_io(self._h, C.byref(io)))
        return io

    def process_device(self, T):
        self._check(self._l.bear_cuda_process_device(self._h, T))

    def process_device_pipelined(self, T):
        """bear_cuda_process_device that leaves the exchange + convolution of its periods queued on the
        renderer's second stream (shared scenes; see include/bear_cuda.h for the rules).  The output is
        complete on result_stream()."""
        self._check(self._l.bear_cuda_process_device_pipelined(self._h, T))

    def result_stream(self):
        """cudaStream_t (an int) on which the output of a pipelined call completes."""
        st = C.c_void_p()
        self._check(self._l.bear_cuda_result_stream(self._h, C.byref(st)))
        return st.value or 0

    def set_profiling(self, enable):
        self._check(self._l.bear_cuda_set_profiling(self._h, int(enable)))

    def get_profile(self, reset=True):
        """{kernel name: (total ms, launches)} from CUDA events on the renderer's stream."""
        arr = (KernelProfile * NUM_KERNELS)()
        self._check(self._l.bear_cuda_get_profile(self._h, arr, int(reset)))
        return {e.name.decode(): (e.total_ms, e.launches) for e in arr}

    def process_device_static_scratch(self, T):
        """bear_cuda_process_device without any allocation: partial spectra go to the library's
        built-in device array (see include/bear_cuda.h for the limit and the sharing rules)."""
        self._check(self._l.bear_cuda_process_device_static_scratch(self._h, T))

    # ------------------------------------------------------------------ one scene on several GPUs
    def comm_blob(self, rank):
        """bear_cuda_comm_blob: what this rank tells the others (bytes of length COMM_BLOB_BYTES)."""
        buf = (C.c_ubyte * COMM_BLOB_BYTES)()
        self._check(self._l.bear_cuda_comm_blob(self._h, rank, buf))
        return bytes(buf)

    def comm_connect(self, rank, world, blobs, root=0):
        """bear_cuda_comm_connect: `blobs` = the comm_blob() of every rank, ordered by rank (list of
        bytes or one bytes object).  root < 0: every rank receives the mix."""
        data = b"".join(blobs) if not isinstance(blobs, (bytes, bytearray)) else bytes(blobs)
        if len(data) != world * COMM_BLOB_BYTES:
            raise ValueError("need one blob of %d bytes per rank" % COMM_BLOB_BYTES)
        buf = (C.c_ubyte * len(data)).from_buffer_copy(data)
        self._check(self._l.bear_cuda_comm_connect(self._h, rank, world, buf, root))
        self.info = self.get_info()

    def reset(self):
        """Back to the freshly constructed state (bear_cuda_reset): next file of a batch."""
        self._check(self._l.bear_cuda_reset(self._h))

    def synchronize(self):
        self._check(self._l.bear_cuda_synchronize(self._h))
