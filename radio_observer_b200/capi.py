"""ctypes binding of include/ro_cuda.h (libro_cuda.so).

This is the binding the tests and bench.py use; it adds nothing to the C ABI.
There is deliberately no fallback: if the shared library is missing or a call
fails, RoError is raised.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence, Tuple

import numpy as np

PREC_FP32, PREC_FP64 = 0, 1
TUNE_NO_FUSED, TUNE_NO_PAIR, TUNE_NO_CHAIN = 1, 2, 4
IN_INT16, IN_FLOAT32, IN_COMPLEX64 = 0, 1, 2
MAX_BANDS = 4

_HERE = os.path.dirname(os.path.abspath(__file__))


class RoError(RuntimeError):
    pass


class WfConfig(C.Structure):
    _fields_ = [
        ("bins", C.c_int32), ("overlap", C.c_int32), ("sample_rate", C.c_int32),
        ("in_format", C.c_int32), ("iq_gain", C.c_double), ("iq_phase_shift", C.c_int32),
        ("n_bands", C.c_int32), ("band_lo", C.c_int32 * MAX_BANDS), ("band_hi", C.c_int32 * MAX_BANDS),
        ("det_lo", C.c_int32), ("det_w", C.c_int32), ("noise_lo", C.c_int32), ("noise_w", C.c_int32),
        ("avg_bins", C.c_int32), ("device", C.c_int32), ("max_batch_frames", C.c_int32),
        ("row_avg", C.c_int32), ("bin_decim", C.c_int32), ("precision", C.c_int32),
        ("tuning", C.c_int32), ("scratch_mb", C.c_int32), ("reserved", C.c_int32 * 1),
    ]


class DetRec(C.Structure):
    _fields_ = [("noise", C.c_float), ("peak", C.c_int32), ("avg", C.c_float), ("pad", C.c_int32)]


DET_DTYPE = np.dtype([("noise", "<f4"), ("peak", "<i4"), ("avg", "<f4"), ("pad", "<i4")])

_lib = None


def library_path() -> str:
    return os.environ.get("RO_CUDA_LIB", os.path.join(_HERE, "libro_cuda.so"))


def load_library():
    """dlopen libro_cuda.so and declare every prototype of include/ro_cuda.h."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        raise RoError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
    lib = C.CDLL(path)
    vp, sz, u64 = C.c_void_p, C.c_size_t, C.c_uint64
    P = C.POINTER
    protos = {
        "ro_wf_create": (C.c_int, [P(WfConfig), P(vp)]),
        "ro_wf_destroy": (None, [vp]),
        "ro_wf_row_width": (C.c_int, [vp]),
        "ro_wf_overlap": (C.c_int, [vp]),
        "ro_wf_out_bins": (C.c_int, [vp]),
        "ro_wf_row_count": (sz, [vp, sz]),
        "ro_wf_reduce_resident": (C.c_int, [vp, vp, sz, vp, P(sz), C.c_int]),
        "ro_wf_push": (C.c_int, [vp, vp, sz]),
        "ro_wf_poll": (C.c_int, [vp, vp, sz, vp, P(sz), P(u64)]),
        "ro_wf_flush": (C.c_int, [vp]),
        "ro_wf_reset": (C.c_int, [vp]),
        "ro_wf_frame_count": (sz, [vp, sz]),
        "ro_wf_process_resident": (C.c_int, [vp, vp, sz, vp, vp, sz, P(sz), C.c_int]),
        "ro_wf_process_host": (C.c_int, [vp, vp, sz, vp, vp, sz, P(sz)]),
        "ro_wf_detect_resident": (C.c_int, [vp, vp, sz, vp, C.c_int]),
        "ro_wf_set_stream": (C.c_int, [vp, vp]),
        "ro_wf_sync": (C.c_int, [vp]),
        "ro_wf_set_profiling": (C.c_int, [vp, C.c_int]),
        "ro_wf_kernel_name": (C.c_char_p, [vp]),
        "ro_wf_long_rows_static": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_double, vp, sz, vp, sz, P(sz), vp, C.c_int]),
        "ro_wf_kernel_times": (C.c_int, [vp, P(C.c_double), P(u64)]),
        "ro_wf_launch_count": (u64, [vp]),
        "ro_dev_count": (C.c_int, [P(C.c_int)]),
        "ro_dev_alloc": (C.c_int, [C.c_int, sz, P(vp)]),
        "ro_dev_free": (C.c_int, [C.c_int, vp]),
        "ro_dev_upload": (C.c_int, [C.c_int, vp, vp, sz]),
        "ro_dev_download": (C.c_int, [C.c_int, vp, vp, sz]),
        "ro_dev_memset": (C.c_int, [C.c_int, vp, C.c_int, sz]),
        "ro_host_alloc_pinned": (C.c_int, [sz, P(vp)]),
        "ro_host_free_pinned": (C.c_int, [vp]),
        "ro_window_table": (C.c_int, [C.c_int32, vp]),
        "ro_twiddle_table": (C.c_int, [C.c_int32, vp]),
        "ro_last_error": (C.c_char_p, []),
        "ro_version": (C.c_char_p, []),
    }
    for name, (res, args) in protos.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    lib._ro_protos = tuple(protos)
    _lib = lib
    return lib


def _check(rc: int):
    if rc != 0:
        raise RoError(f"ro_cuda error {rc}: {load_library().ro_last_error().decode()}")


def window_table(bins: int) -> np.ndarray:
    out = np.empty(bins, dtype=np.float32)
    _check(load_library().ro_window_table(bins, out.ctypes.data))
    return out


def twiddle_table(n: int) -> np.ndarray:
    out = np.empty(2 * n, dtype=np.float32)
    _check(load_library().ro_twiddle_table(n, out.ctypes.data))
    return out.view(np.complex64)


def device_count() -> int:
    n = C.c_int(0)
    _check(load_library().ro_dev_count(C.byref(n)))
    return n.value


STATIC_SCRATCH_BYTES = 64 << 20        # RO_STATIC_SCRATCH_BYTES
STATIC_MAX_BINS = 131072               # RO_STATIC_MAX_BINS


def long_rows_static(device: int, bins: int, overlap: int, in_format: int, iq_gain: float, d_iq: int,
                     n_samples: int, d_rows: int, max_frames: int, cuda_stream: Optional[int] = None,
                     sync: bool = True) -> int:
    """ro_wf_long_rows_static: the long-frame transform on the library's static device arrays (no
    handle, no allocation).  Device pointers as integers; returns the number of rows written."""
    n = C.c_size_t(0)
    _check(load_library().ro_wf_long_rows_static(device, bins, overlap, in_format, float(iq_gain), d_iq, n_samples,
                                                 d_rows, max_frames, C.byref(n), cuda_stream, 1 if sync else 0))
    return n.value


class DeviceBuffer:
    """cudaMalloc'd memory owned through the C ABI (no torch needed)."""

    def __init__(self, nbytes: int, device: int = 0):
        self.lib = load_library()
        self.device, self.nbytes = device, int(nbytes)
        p = C.c_void_p()
        _check(self.lib.ro_dev_alloc(device, self.nbytes, C.byref(p)))
        self.ptr = p.value

    def upload(self, arr: np.ndarray, offset: int = 0):
        arr = np.ascontiguousarray(arr)
        assert offset + arr.nbytes <= self.nbytes
        _check(self.lib.ro_dev_upload(self.device, self.ptr + offset, arr.ctypes.data, arr.nbytes))

    def download(self, dtype, count: int, offset: int = 0) -> np.ndarray:
        out = np.empty(count, dtype=dtype)
        assert offset + out.nbytes <= self.nbytes
        _check(self.lib.ro_dev_download(self.device, out.ctypes.data, self.ptr + offset, out.nbytes))
        return out

    def free(self):
        if self.ptr:
            self.lib.ro_dev_free(self.device, self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class PinnedArray:
    """numpy view over cudaMallocHost memory."""

    def __init__(self, shape, dtype):
        self.lib = load_library()
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        p = C.c_void_p()
        _check(self.lib.ro_host_alloc_pinned(max(n, 1), C.byref(p)))
        self.ptr = p.value
        buf = (C.c_char * max(n, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            self.lib.ro_host_free_pinned(self.ptr)
            self.ptr = None


class Waterfall:
    """One stream on one GPU: thin object wrapper over ro_wf_*."""

    def __init__(self, bins: int, overlap: int, *, sample_rate: int = 48000, in_format: int = IN_INT16,
                 iq_gain: float = 0.0, iq_phase_shift: int = 0,
                 bands: Sequence[Tuple[int, int]] = (), detector: Optional[dict] = None,
                 device: int = 0, max_batch_frames: int = 0, row_avg: int = 1, bin_decim: int = 1,
                 precision: int = 0, tuning: int = 0, scratch_mb: int = 0):
        self.lib = load_library()
        cfg = WfConfig()
        cfg.bins, cfg.overlap, cfg.sample_rate, cfg.in_format = bins, overlap, sample_rate, in_format
        cfg.iq_gain, cfg.iq_phase_shift = iq_gain, iq_phase_shift
        cfg.n_bands = len(bands)
        for i, (lo, hi) in enumerate(bands):
            cfg.band_lo[i], cfg.band_hi[i] = lo, hi
        if detector:
            cfg.det_lo, cfg.det_w = detector["det_lo"], detector["det_w"]
            cfg.noise_lo, cfg.noise_w = detector["noise_lo"], detector["noise_w"]
            cfg.avg_bins = detector["avg_bins"]
        cfg.device, cfg.max_batch_frames = device, max_batch_frames
        cfg.row_avg, cfg.bin_decim, cfg.precision = row_avg, bin_decim, precision
        cfg.tuning, cfg.scratch_mb = tuning, scratch_mb
        self.cfg = cfg
        h = C.c_void_p()
        _check(self.lib.ro_wf_create(C.byref(cfg), C.byref(h)))
        self.h = h
        self.bins = bins
        self.overlap = self.lib.ro_wf_overlap(h)
        self.hop = bins - self.overlap
        self.row_width = self.lib.ro_wf_row_width(h)
        self.out_bins = self.lib.ro_wf_out_bins(h)
        self.row_avg = max(1, row_avg)
        self.device = device
        self.det_on = bool(detector)

    def close(self):
        if self.h:
            self.lib.ro_wf_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def frame_count(self, n_samples: int) -> int:
        return int(self.lib.ro_wf_frame_count(self.h, n_samples))

    def row_count(self, n_samples: int) -> int:
        """output rows: frame_count // row_avg"""
        return int(self.lib.ro_wf_row_count(self.h, n_samples))

    def _n_samples(self, iq: np.ndarray) -> int:
        if self.cfg.in_format == IN_COMPLEX64:
            assert iq.dtype == np.complex128
            return iq.size
        want = np.int16 if self.cfg.in_format == IN_INT16 else np.float32
        assert iq.dtype == want, (iq.dtype, want)
        return iq.size // 2

    # -- streaming path --------------------------------------------------
    def push(self, iq: np.ndarray):
        iq = np.ascontiguousarray(iq)
        _check(self.lib.ro_wf_push(self.h, iq.ctypes.data, self._n_samples(iq)))

    def flush(self):
        _check(self.lib.ro_wf_flush(self.h))

    def reset(self):
        _check(self.lib.ro_wf_reset(self.h))

    def poll(self, max_rows: int = 1 << 12):
        rows = np.empty((max_rows, self.row_width), dtype=np.float32)
        det = np.zeros(max_rows, dtype=DET_DTYPE)
        n = C.c_size_t(0)
        first = C.c_uint64(0)
        _check(self.lib.ro_wf_poll(self.h, rows.ctypes.data, max_rows, det.ctypes.data, C.byref(n), C.byref(first)))
        return rows[:n.value], det[:n.value], first.value

    def poll_all(self):
        rs, ds, first = [], [], None
        while True:
            r, d, f = self.poll()
            if len(r) == 0:
                break
            if first is None:
                first = f
            rs.append(r.copy())
            ds.append(d.copy())
        if not rs:
            return np.empty((0, self.row_width), np.float32), np.zeros(0, DET_DTYPE), 0
        return np.concatenate(rs), np.concatenate(ds), first

    # -- batch paths -------------------------------------------------------
    def process_host(self, iq: np.ndarray, rows_out: Optional[np.ndarray] = None,
                     det_out: Optional[np.ndarray] = None, want_rows: bool = True):
        iq = np.ascontiguousarray(iq) if not isinstance(iq, np.ndarray) or not iq.flags.c_contiguous else iq
        ns = self._n_samples(iq)
        F = self.row_count(ns)
        if rows_out is None and want_rows:
            rows_out = np.empty((F, self.row_width), dtype=np.float32)
        if det_out is None and self.det_on:
            det_out = np.zeros(F, dtype=DET_DTYPE)
        n = C.c_size_t(0)
        _check(self.lib.ro_wf_process_host(
            self.h, iq.ctypes.data, ns,
            rows_out.ctypes.data if rows_out is not None else None,
            det_out.ctypes.data if det_out is not None else None, F, C.byref(n)))
        return rows_out, det_out, n.value

    def process_resident(self, d_iq: int, n_samples: int, d_rows: int, d_det: Optional[int],
                         max_frames: Optional[int] = None, sync: bool = True) -> int:
        n = C.c_size_t(0)
        mf = self.row_count(n_samples) if max_frames is None else max_frames
        _check(self.lib.ro_wf_process_resident(self.h, d_iq, n_samples, d_rows, d_det, mf, C.byref(n), int(sync)))
        return n.value

    def reduce_resident(self, d_rows: int, n_in: int, d_out: int, sync: bool = True) -> int:
        n = C.c_size_t(0)
        _check(self.lib.ro_wf_reduce_resident(self.h, d_rows, n_in, d_out, C.byref(n), int(sync)))
        return n.value

    def detect_resident(self, d_rows: int, n_rows: int, d_det: int, sync: bool = True):
        _check(self.lib.ro_wf_detect_resident(self.h, d_rows, n_rows, d_det, int(sync)))

    # -- control -----------------------------------------------------------
    def set_stream(self, cuda_stream: Optional[int]):
        _check(self.lib.ro_wf_set_stream(self.h, cuda_stream))

    def sync(self):
        _check(self.lib.ro_wf_sync(self.h))

    def set_profiling(self, on: bool):
        _check(self.lib.ro_wf_set_profiling(self.h, int(on)))

    def kernel_times(self):
        ms = (C.c_double * 2)()
        ln = (C.c_uint64 * 2)()
        _check(self.lib.ro_wf_kernel_times(self.h, ms, ln))
        return [ms[0], ms[1]], [ln[0], ln[1]]

    def launch_count(self) -> int:
        return int(self.lib.ro_wf_launch_count(self.h))

    def kernel_name(self) -> str:
        """What the last transform launch of this handle ran."""
        return self.lib.ro_wf_kernel_name(self.h).decode()
