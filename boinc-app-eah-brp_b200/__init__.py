"""brp_b200 -- B200-native Einstein@Home BRP per-template search loop.

Python is only the test/benchmark harness here: everything below is a ctypes view of the C ABI in
``include/brp_b200.h`` (``libbrp_b200.so``, hand-written sm_100a CUDA + C++ host code).  There is no
Python or CPU implementation of the hot path in this package: if the shared library is missing the
import fails, and without a CUDA device every compute call returns a ``RADPUL_CUDA_*`` error.

Two layers, mirroring the header:

* :class:`LegacyBackend` -- the reference's own backend functions with the reference's names and
  argument meaning (``set_up_resampling`` ... ``tear_down_harmonic_summing``), one template per call.
* :class:`Search` -- the additive batched API (``brp_search_*``) used for all throughput numbers.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from . import synth  # noqa: F401  (numpy-only helpers: synthetic work-units and banks)

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "libbrp_b200.so"

RADPUL_EMEM, RADPUL_EFILE, RADPUL_EIO, RADPUL_EVAL, RADPUL_EMISC = 1, 2, 3, 4, 5
RADPUL_CUDA_DEVICE_FIND = 1001
RADPUL_CUDA_FFT_PLAN = 1011

N_CAND_5, N_CAND = 100, 500


class BrpError(RuntimeError):
    def __init__(self, code: int, where: str):
        super().__init__(f"{where} failed with RADPUL code {code}")
        self.code = code


def _check(rc: int, where: str) -> None:
    if rc != 0:
        raise BrpError(rc, where)


class Geometry(C.Structure):
    """``brp_geometry``: derived sizes of demod_binary.c:782,1087-1099."""
    _fields_ = [
        ("nsamples_unpadded", C.c_uint32), ("nsamples", C.c_uint32), ("fft_size", C.c_uint32),
        ("window_2", C.c_uint32), ("fundamental_idx_hi", C.c_uint32), ("harmonic_idx_hi", C.c_uint32),
        ("t_obs", C.c_float), ("dt", C.c_float), ("step_inv", C.c_float),
    ]

    @property
    def nr_pages(self) -> int:
        return (self.fundamental_idx_hi >> 10) + 1


class Template(C.Structure):
    """``brp_template``: per-template constants of demod_binary.c:1208-1230."""
    _fields_ = [("P", C.c_float), ("tau", C.c_float), ("Psi0", C.c_float), ("Omega", C.c_float), ("S0", C.c_float)]


class ResampParams(C.Structure):
    """``RESAMP_PARAMS`` (structs.h:151-161)."""
    _fields_ = [
        ("nsamples", C.c_uint), ("nsamples_unpadded", C.c_uint), ("fft_size", C.c_uint),
        ("tau", C.c_float), ("Omega", C.c_float), ("Psi0", C.c_float), ("dt", C.c_float),
        ("step_inv", C.c_float), ("S0", C.c_float),
    ]


class DIfloatPtr(C.Union):
    """``DIfloatPtr`` (diptr.h:38-41)."""
    _fields_ = [("host_ptr", C.POINTER(C.c_float)), ("device_ptr", C.c_ulonglong)]


HIT_DTYPE = np.dtype([("power", "<f4"), ("bin", "<u4"), ("tmpl", "<u4"), ("harm_idx", "<u4")])
CAND_DTYPE = np.dtype(
    [("power", "<f8"), ("P_b", "<f8"), ("tau", "<f8"), ("Psi", "<f8"), ("fA", "<f8"), ("n_harm", "<u4"), ("f0", "<u4")]
)
assert HIT_DTYPE.itemsize == 16 and CAND_DTYPE.itemsize == 48

# every symbol include/brp_b200.h declares (tests check that the library exports exactly these)
EXPORTED_SYMBOLS = [
    "initialize_cuda", "set_up_resampling", "run_resampling", "tear_down_resampling", "set_up_fft", "run_fft",
    "tear_down_fft", "set_up_harmonic_summing", "run_harmonic_summing", "tear_down_harmonic_summing", "shutdown_cuda",
    "brp_derive_geometry", "brp_derive_template", "brp_base_thresholds", "brp_chisq_Q", "brp_chisq_Qinv",
    "brp_bank_read", "brp_bank_free", "brp_wu_read", "brp_series_free", "brp_toplist_from_hits",
    "brp_toplist_merge", "brp_finalize_candidates", "brp_format_candidates", "brp_write_candidates", "brp_checkpoint_write",
    "brp_checkpoint_read", "brp_search_create", "brp_search_destroy", "brp_search_set_series",
    "brp_search_set_series_device", "brp_search_reset", "brp_search_run", "brp_search_sync", "brp_search_collect",
    "brp_search_last_run_ms", "brp_search_stream", "brp_search_kernel_launches", "brp_search_profile_batch",
    "brp_stage_resample", "brp_stage_resample_compare", "brp_stage_spectrum", "brp_stage_spectrum_of", "brp_stage_harmonic_summing",
    "brp_stage_sin_lut", "brp_stage_round_index", "brp_resample_noalloc", "brp_whiten_series", "brp_whiten_series_zap",
    "brp_zaplist_read", "brp_zaplist_free", "brp_zap_noise",
]


def _load() -> C.CDLL:
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). brp_b200 has no CPU fallback."
        )
    return C.CDLL(str(LIB_PATH))


lib = _load()

_f32 = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64 = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_i32 = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")

lib.brp_derive_geometry.argtypes = [C.c_uint32, C.c_double, C.c_float, C.c_float, C.c_uint32, C.POINTER(Geometry)]
lib.brp_derive_template.argtypes = [C.c_double, C.c_double, C.c_double, C.c_float, C.POINTER(Template)]
lib.brp_base_thresholds.argtypes = [C.c_float, C.c_uint32, _f32]
lib.brp_chisq_Q.argtypes = [C.c_double, C.c_int]
lib.brp_chisq_Q.restype = C.c_double
lib.brp_chisq_Qinv.argtypes = [C.c_double, C.c_int]
lib.brp_chisq_Qinv.restype = C.c_double
lib.brp_bank_read.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_double)), C.POINTER(C.c_uint32)]
lib.brp_bank_free.argtypes = [C.POINTER(C.c_double)]
lib.brp_bank_free.restype = None
lib.brp_wu_read.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_uint32), C.POINTER(C.c_double)]
lib.brp_series_free.argtypes = [C.POINTER(C.c_float)]
lib.brp_series_free.restype = None
lib.brp_toplist_from_hits.argtypes = [C.c_void_p, C.c_uint32, _f64, C.c_uint32, C.c_void_p, C.c_void_p]
lib.brp_toplist_merge.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
lib.brp_finalize_candidates.argtypes = [C.c_void_p, C.c_void_p]
lib.brp_format_candidates.argtypes = [C.c_void_p, C.POINTER(Geometry), C.c_double, C.c_char_p, C.c_int]
lib.brp_write_candidates.argtypes = [C.c_char_p, C.c_void_p, C.POINTER(Geometry), C.c_double]
lib.brp_checkpoint_write.argtypes = [C.c_char_p, C.c_uint32, C.c_char_p, C.c_void_p]
lib.brp_checkpoint_read.argtypes = [C.c_char_p, C.POINTER(C.c_uint32), C.c_char_p, C.c_void_p]
lib.brp_search_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(Geometry), C.c_float, C.c_uint32]
lib.brp_search_destroy.argtypes = [C.c_void_p]
lib.brp_search_set_series.argtypes = [C.c_void_p, C.c_void_p]
lib.brp_search_set_series_device.argtypes = [C.c_void_p, C.c_void_p]
lib.brp_search_reset.argtypes = [C.c_void_p]
lib.brp_search_run.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
lib.brp_search_sync.argtypes = [C.c_void_p]
lib.brp_search_collect.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
lib.brp_search_last_run_ms.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
lib.brp_search_stream.argtypes = [C.c_void_p]
lib.brp_search_stream.restype = C.c_void_p
lib.brp_search_kernel_launches.argtypes = [C.c_void_p]
lib.brp_search_kernel_launches.restype = C.c_uint64
lib.brp_search_profile_batch.argtypes = [C.c_void_p, _f64, C.c_uint32, C.c_uint32, _f32]
lib.brp_stage_resample.argtypes = [C.c_void_p, C.POINTER(Template), C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_float)]
lib.brp_stage_resample_compare.argtypes = [C.c_void_p, _f64, C.c_uint32, C.POINTER(C.c_uint64)]
lib.brp_stage_spectrum.argtypes = [C.c_void_p, _f32, C.c_uint32]
lib.brp_stage_spectrum_of.argtypes = [C.c_void_p, _f32, _f32, C.c_uint32]
lib.brp_stage_harmonic_summing.argtypes = [C.c_void_p, _f32, _f32, C.c_uint32, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
lib.brp_stage_sin_lut.argtypes = [C.c_int, _f32, _f32, C.c_uint32]
lib.brp_stage_round_index.argtypes = [C.c_int, _f32, _i32, C.c_uint32]

lib.brp_resample_noalloc.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(Template), C.c_float, C.c_float, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p]

lib.brp_whiten_series.argtypes = [C.c_int, _f32, C.c_uint32, C.c_float, C.c_uint32, C.c_uint32, C.POINTER(C.c_float)]
lib.brp_whiten_series_zap.argtypes = [C.c_int, _f32, C.c_uint32, C.c_float, C.c_uint32, C.c_double, C.c_void_p, C.c_uint32,
                                      C.POINTER(C.c_float)]
lib.brp_zaplist_read.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_double)), C.POINTER(C.c_uint32)]
lib.brp_zaplist_free.argtypes = [C.POINTER(C.c_double)]
lib.brp_zaplist_free.restype = None
lib.brp_zap_noise.argtypes = [C.c_uint32, C.c_float, C.c_uint32, C.c_double, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32,
                              C.POINTER(C.c_uint32)]

lib.initialize_cuda.argtypes = [C.c_int, C.POINTER(C.c_int)]
lib.set_up_resampling.argtypes = [DIfloatPtr, C.POINTER(DIfloatPtr), C.POINTER(ResampParams), C.c_void_p, C.c_void_p]
lib.run_resampling.argtypes = [DIfloatPtr, DIfloatPtr, C.POINTER(ResampParams)]
lib.tear_down_resampling.argtypes = [DIfloatPtr]
lib.set_up_fft.argtypes = [DIfloatPtr, C.POINTER(DIfloatPtr), C.c_uint32, C.c_uint]
lib.run_fft.argtypes = [DIfloatPtr, DIfloatPtr, C.c_uint32, C.c_uint, C.c_float]
lib.tear_down_fft.argtypes = [DIfloatPtr]
lib.set_up_harmonic_summing.argtypes = [C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.c_uint), C.c_uint, C.c_uint]
lib.run_harmonic_summing.argtypes = [C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.POINTER(C.c_int32)), C.c_uint, DIfloatPtr,
                                     C.c_uint, C.c_uint, C.c_uint, C.POINTER(C.c_float)]
lib.tear_down_harmonic_summing.argtypes = [C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.POINTER(C.c_int32))]
lib.shutdown_cuda.argtypes = []


# ------------------------------------------------------------------------------------------------ host helpers

def derive_geometry(nsamples_unpadded: int, tsample_us: float, padding: float, f0: float, window: int) -> Geometry:
    g = Geometry()
    _check(lib.brp_derive_geometry(nsamples_unpadded, tsample_us, padding, f0, window, C.byref(g)), "brp_derive_geometry")
    return g


def derive_template(P: float, tau: float, psi0: float, step_inv: float) -> Template:
    t = Template()
    _check(lib.brp_derive_template(P, tau, psi0, step_inv, C.byref(t)), "brp_derive_template")
    return t


def base_thresholds(fA: float, fft_size: int) -> np.ndarray:
    thr = np.empty(5, np.float32)
    _check(lib.brp_base_thresholds(fA, fft_size, thr), "brp_base_thresholds")
    return thr


def bank_read(path) -> np.ndarray:
    p = C.POINTER(C.c_double)()
    n = C.c_uint32(0)
    _check(lib.brp_bank_read(str(path).encode(), C.byref(p), C.byref(n)), "brp_bank_read")
    out = np.ctypeslib.as_array(p, shape=(max(n.value, 1) * 3,))[: n.value * 3].copy().reshape(n.value, 3)
    lib.brp_bank_free(p)
    return out


def wu_read(path):
    p = C.POINTER(C.c_float)()
    n = C.c_uint32(0)
    ts = C.c_double(0)
    _check(lib.brp_wu_read(str(path).encode(), C.byref(p), C.byref(n), C.byref(ts)), "brp_wu_read")
    out = np.ctypeslib.as_array(p, shape=(max(n.value, 1),))[: n.value].copy()
    lib.brp_series_free(p)
    return out, ts.value


def toplist_from_hits(hits: np.ndarray, bank: np.ndarray):
    hits = np.ascontiguousarray(hits, HIT_DTYPE)
    bank = np.ascontiguousarray(bank, np.float64)
    cands = np.zeros(N_CAND, CAND_DTYPE)
    tidx = np.full(N_CAND, -1, np.int32)
    _check(lib.brp_toplist_from_hits(hits.ctypes.data, len(hits), bank, bank.shape[0], cands.ctypes.data, tidx.ctypes.data),
           "brp_toplist_from_hits")
    return cands, tidx


def toplist_merge(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    a, b = np.ascontiguousarray(a, CAND_DTYPE), np.ascontiguousarray(b, CAND_DTYPE)
    out = np.zeros(N_CAND, CAND_DTYPE)
    _check(lib.brp_toplist_merge(a.ctypes.data, b.ctypes.data, out.ctypes.data), "brp_toplist_merge")
    return out


def finalize_candidates(cands: np.ndarray, tidx: np.ndarray):
    cands, tidx = cands.copy(), tidx.copy()
    _check(lib.brp_finalize_candidates(cands.ctypes.data, tidx.ctypes.data), "brp_finalize_candidates")
    return cands, tidx


def format_candidates(cands_final: np.ndarray, g: Geometry, tsample_us: float) -> str:
    c = cands_final.copy()
    buf = C.create_string_buffer(1 << 16)
    n = lib.brp_format_candidates(c.ctypes.data, C.byref(g), tsample_us, buf, len(buf))
    if n < 0:
        raise BrpError(RADPUL_EIO, "brp_format_candidates")
    return buf.raw[:n].decode()


def write_candidates(path, cands_final: np.ndarray, g: Geometry, tsample_us: float) -> None:
    c = cands_final.copy()
    _check(lib.brp_write_candidates(str(path).encode(), c.ctypes.data, C.byref(g), tsample_us), "brp_write_candidates")


def checkpoint_write(path, n_template: int, originalfile: str, cands: np.ndarray) -> None:
    c = np.ascontiguousarray(cands, CAND_DTYPE)
    _check(lib.brp_checkpoint_write(str(path).encode(), n_template, originalfile.encode(), c.ctypes.data), "brp_checkpoint_write")


def checkpoint_read(path):
    n = C.c_uint32(0)
    name = C.create_string_buffer(256)
    cands = np.zeros(N_CAND, CAND_DTYPE)
    _check(lib.brp_checkpoint_read(str(path).encode(), C.byref(n), name, cands.ctypes.data), "brp_checkpoint_read")
    return n.value, name.value.decode(), cands


# ------------------------------------------------------------------------------------------------ batched search

class Search:
    """One work-unit on one GPU (``brp_search_*``)."""

    def __init__(self, g: Geometry, fA: float = 0.08, max_batch: int = 16, device: int = 0):
        self.g = g
        self.h = C.c_void_p()
        _check(lib.brp_search_create(C.byref(self.h), device, C.byref(g), fA, max_batch), "brp_search_create")
        self.max_batch = max_batch
        self._keep = None

    def close(self) -> None:
        if self.h:
            lib.brp_search_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_series(self, series: np.ndarray) -> None:
        series = np.ascontiguousarray(series, np.float32)
        assert series.size == self.g.nsamples_unpadded
        self._keep = series
        _check(lib.brp_search_set_series(self.h, series.ctypes.data), "brp_search_set_series")

    def set_series_ptr(self, host_ptr: int) -> None:
        """Host pointer (e.g. pinned memory) to Nu floats."""
        _check(lib.brp_search_set_series(self.h, host_ptr), "brp_search_set_series")

    def set_series_device(self, dev_ptr: int) -> None:
        _check(lib.brp_search_set_series_device(self.h, dev_ptr), "brp_search_set_series_device")

    def reset(self) -> None:
        _check(lib.brp_search_reset(self.h), "brp_search_reset")

    def run(self, bank: np.ndarray, first_global_index: int = 0) -> None:
        bank = np.ascontiguousarray(bank, np.float64)
        _check(lib.brp_search_run(self.h, bank.ctypes.data, bank.shape[0], first_global_index), "brp_search_run")

    def sync(self) -> None:
        _check(lib.brp_search_sync(self.h), "brp_search_sync")

    def collect(self) -> np.ndarray:
        cap = 4096
        while True:
            hits = np.zeros(cap, HIT_DTYPE)
            n = C.c_uint32(0)
            rc = lib.brp_search_collect(self.h, hits.ctypes.data, cap, C.byref(n))
            if rc == RADPUL_EMEM:
                cap = int(n.value)
                continue
            _check(rc, "brp_search_collect")
            return hits[: n.value].copy()

    def last_run_ms(self) -> float:
        ms = C.c_float(0)
        _check(lib.brp_search_last_run_ms(self.h, C.byref(ms)), "brp_search_last_run_ms")
        return ms.value

    @property
    def stream(self) -> int:
        return int(lib.brp_search_stream(self.h) or 0)

    @property
    def kernel_launches(self) -> int:
        return int(lib.brp_search_kernel_launches(self.h))

    def profile_batch(self, bank: np.ndarray, batch: int, repeats: int = 5) -> np.ndarray:
        bank = np.ascontiguousarray(bank[:batch], np.float64)
        ms = np.zeros(4, np.float32)
        _check(lib.brp_search_profile_batch(self.h, bank, batch, repeats, ms), "brp_search_profile_batch")
        return ms

    # ---- stage taps (parity tests)
    def stage_resample(self, t: Template, want_series: bool = True, want_index: bool = True):
        out = np.empty(self.g.nsamples, np.float32) if want_series else None
        idx = np.full(self.g.nsamples_unpadded, -1, np.int32) if want_index else None
        ns, mean = C.c_uint32(0), C.c_float(0)
        _check(lib.brp_stage_resample(self.h, C.byref(t), out.ctypes.data if want_series else None,
                                      idx.ctypes.data if want_index else None, C.byref(ns), C.byref(mean)), "brp_stage_resample")
        return out, idx, ns.value, mean.value

    def stage_resample_compare(self, bank: np.ndarray) -> int:
        """Number of index positions where the run-detecting resampler differs from per-sample evaluation (must be 0)."""
        bank = np.ascontiguousarray(bank, np.float64)
        bad = C.c_uint64(0)
        _check(lib.brp_stage_resample_compare(self.h, bank, bank.shape[0], C.byref(bad)), "brp_stage_resample_compare")
        return bad.value

    def stage_spectrum(self, count: int | None = None) -> np.ndarray:
        count = self.g.fft_size if count is None else count
        ps = np.empty(count, np.float32)
        _check(lib.brp_stage_spectrum(self.h, ps, count), "brp_stage_spectrum")
        return ps

    def stage_spectrum_of(self, series_n: np.ndarray, count: int | None = None) -> np.ndarray:
        count = self.g.fft_size if count is None else count
        series_n = np.ascontiguousarray(series_n, np.float32)
        assert series_n.size == self.g.nsamples
        ps = np.empty(count, np.float32)
        _check(lib.brp_stage_spectrum_of(self.h, series_n, ps, count), "brp_stage_spectrum_of")
        return ps

    def stage_harmonic_summing(self, ps: np.ndarray, thr: np.ndarray, tmpl_index: int = 0) -> np.ndarray:
        ps = np.ascontiguousarray(ps, np.float32)
        assert ps.size >= self.g.harmonic_idx_hi
        cap = 5 * max(self.g.fundamental_idx_hi, 1)
        hits = np.zeros(cap, HIT_DTYPE)
        n = C.c_uint32(0)
        _check(lib.brp_stage_harmonic_summing(self.h, ps, np.asarray(thr, np.float32), tmpl_index, hits.ctypes.data, cap, C.byref(n)),
               "brp_stage_harmonic_summing")
        return hits[: n.value].copy()


def stage_sin_lut(x: np.ndarray, device: int = 0) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(x)
    _check(lib.brp_stage_sin_lut(device, x, out, x.size), "brp_stage_sin_lut")
    return out


def stage_round_index(f: np.ndarray, device: int = 0) -> np.ndarray:
    f = np.ascontiguousarray(f, np.float32)
    out = np.empty(f.shape, np.int32)
    _check(lib.brp_stage_round_index(device, f, out, f.size), "brp_stage_round_index")
    return out


NOALLOC_MAX_SAMPLES = 1 << 24


def resample_noalloc(d_series: int, g: Geometry, t: Template, d_out: int, d_n_steps: int, d_mean: int, stream: int = 0) -> None:
    """``brp_resample_noalloc``: every argument named ``d_*`` is a raw device address supplied by the caller (for instance
    ``tensor.data_ptr()`` of buffers the caller owns); nothing is allocated by the call, the work is only enqueued."""
    _check(lib.brp_resample_noalloc(d_series, g.nsamples_unpadded, g.nsamples, C.byref(t), g.dt, g.step_inv, d_out, d_n_steps,
                                    d_mean, stream), "brp_resample_noalloc")


def whiten_series(series: np.ndarray, padding: float, window: int = 1000, device: int = 0, zap: np.ndarray | None = None,
                  tsample_us: float = 0.0):
    """``brp_whiten_series_zap``: returns (whitened copy, device milliseconds).  zap = float64[n, 2] (fmin, fmax) in Hz."""
    out = np.ascontiguousarray(series, np.float32).copy()
    ms = C.c_float(0)
    z = np.zeros((0, 2)) if zap is None else np.ascontiguousarray(zap, np.float64).reshape(-1, 2)
    _check(lib.brp_whiten_series_zap(device, out, out.size, padding, window, tsample_us, z.ctypes.data if len(z) else None, len(z),
                                     C.byref(ms)), "brp_whiten_series_zap")
    return out, ms.value


def read_zaplist(path) -> np.ndarray:
    """``brp_zaplist_read``: float64[n, 2] (fmin, fmax) in Hz, in file order."""
    ptr, n = C.POINTER(C.c_double)(), C.c_uint32(0)
    _check(lib.brp_zaplist_read(str(path).encode(), C.byref(ptr), C.byref(n)), "brp_zaplist_read")
    try:
        return np.ctypeslib.as_array(ptr, shape=(max(n.value, 1) * 2,))[: 2 * n.value].reshape(-1, 2).copy()
    finally:
        lib.brp_zaplist_free(ptr)


def zap_noise(seed_bits: int, padding: float, nsamples: int, tsample_us: float, zap: np.ndarray):
    """``brp_zap_noise``: (bins uint32[n], values float32[n, 2]) of every draw of the zap step, in draw order (host only)."""
    z = np.ascontiguousarray(zap, np.float64).reshape(-1, 2)
    n = C.c_uint32(0)
    _check(lib.brp_zap_noise(seed_bits, padding, nsamples, tsample_us, z.ctypes.data, len(z), None, None, 0, C.byref(n)), "brp_zap_noise")
    bins, vals = np.empty(n.value, np.uint32), np.empty((n.value, 2), np.float32)
    _check(lib.brp_zap_noise(seed_bits, padding, nsamples, tsample_us, z.ctypes.data, len(z), bins.ctypes.data, vals.ctypes.data,
                             n.value, C.byref(n)), "brp_zap_noise")
    return bins, vals


def search_bank(series: np.ndarray, g: Geometry, bank: np.ndarray, fA: float = 0.08, max_batch: int = 16, device: int = 0):
    """Whole loop through the batched API: returns (cands[500] as right after the loop, template index per slot)."""
    s = Search(g, fA, max_batch, device)
    try:
        s.set_series(series)
        s.run(bank)
        hits = s.collect()
    finally:
        s.close()
    return toplist_from_hits(hits, bank)


# ------------------------------------------------------------------------------------------------ legacy backend

class LegacyBackend:
    """The reference's backend interface (cuda/app/demod_binary_cuda.h:42-52, demod_binary_hs_cuda.cuh:37-39) with the
    reference's call sequence: ``initialize_cuda`` -> ``set_up_*`` -> per template ``run_*`` -> ``tear_down_*`` -> ``shutdown_cuda``."""

    def __init__(self, device: int = 0):
        dev = C.c_int(device)
        _check(lib.initialize_cuda(1, C.byref(dev)), "initialize_cuda")
        self.device = dev.value
        self.params = ResampParams()
        self.resamp = DIfloatPtr()
        self.ps = DIfloatPtr()
        self.sumspec = (C.POINTER(C.c_float) * 5)()
        self.dirty = (C.POINTER(C.c_int32) * 5)()
        self.nr_pages = C.c_uint(0)
        self._series = None
        self._up = [False, False, False]

    def set_up(self, series: np.ndarray, g: Geometry) -> None:
        self.g = g
        self._series = np.ascontiguousarray(series, np.float32)
        self.params.nsamples = g.nsamples
        self.params.nsamples_unpadded = g.nsamples_unpadded
        self.params.fft_size = g.fft_size
        self.series_ptr = DIfloatPtr()
        self.series_ptr.host_ptr = self._series.ctypes.data_as(C.POINTER(C.c_float))
        _check(lib.set_up_resampling(self.series_ptr, C.byref(self.resamp), C.byref(self.params), None, None), "set_up_resampling")
        self._up[0] = True
        _check(lib.set_up_fft(self.resamp, C.byref(self.ps), g.nsamples, g.fft_size), "set_up_fft")
        self._up[1] = True
        _check(lib.set_up_harmonic_summing(self.sumspec, self.dirty, C.byref(self.nr_pages), g.fundamental_idx_hi, g.harmonic_idx_hi),
               "set_up_harmonic_summing")
        self._up[2] = True

    def run_template(self, t: Template, thrA: np.ndarray):
        """One loop iteration (demod_binary.c:1233-1293). Returns (sumspec[5] host views, dirty[5])."""
        g, p = self.g, self.params
        p.tau, p.Omega, p.Psi0, p.dt, p.step_inv, p.S0 = t.tau, t.Omega, t.Psi0, g.dt, g.step_inv, t.S0
        _check(lib.run_resampling(self.series_ptr, self.resamp, C.byref(p)), "run_resampling")
        norm = np.float32(1.0 / g.nsamples)
        _check(lib.run_fft(self.resamp, self.ps, g.nsamples, g.fft_size, norm), "run_fft")
        thr = (C.c_float * 5)(*[float(v) for v in thrA])
        _check(lib.run_harmonic_summing(self.sumspec, self.dirty, self.nr_pages, self.ps, g.window_2, g.fundamental_idx_hi,
                                        g.harmonic_idx_hi, thr), "run_harmonic_summing")
        fu = g.fundamental_idx_hi
        ss = [np.ctypeslib.as_array(self.sumspec[k], shape=(fu,)) for k in range(5)]
        dd = [np.ctypeslib.as_array(self.dirty[k], shape=(self.nr_pages.value,)) for k in range(5)]
        return ss, dd

    def tear_down(self) -> None:
        if self._up[2]:
            lib.tear_down_harmonic_summing(self.sumspec, self.dirty)
        if self._up[1]:
            lib.tear_down_fft(self.ps)
        if self._up[0]:
            lib.tear_down_resampling(self.resamp)
        self._up = [False, False, False]

    def shutdown(self) -> None:
        self.tear_down()
        lib.shutdown_cuda()
