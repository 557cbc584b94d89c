"""TEST INFRASTRUCTURE ONLY: ctypes front-ends for the two CPU checkers in this directory.

  * ``Oracle``    -> oracle/libbrp_oracle.so, our restatement (brp_oracle.cpp); built by ``make -C oracle``.
  * ``Reference`` -> oracle/_ref/libbrp_ref.so, the reference's own translation units compiled in place
                     from /root/reference (only buildable where that tree exists; the .so travels).

Nothing in the product package imports this module.  Only tests/, __graft_entry__.smoke() and the CPU
legs of bench.py do, and only as the checker / the timed CPU baseline.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ORACLE_SO = HERE / "libbrp_oracle.so"
REF_SO = HERE / "_ref" / "libbrp_ref.so"
REF_BIN = HERE / "_ref" / "demod_binary_ref"

f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def build(ref: bool = True) -> None:
    """Compile the checkers (gcc/g++ only).  Building the checker is not using it."""
    target = ["all"] if ref else ["libbrp_oracle.so"]
    subprocess.run(["make", "-s", "-C", str(HERE)] + target, check=True)


class Geometry(C.Structure):
    _fields_ = [
        ("nsamples_unpadded", C.c_uint32),
        ("nsamples", C.c_uint32),
        ("fft_size", C.c_uint32),
        ("window_2", C.c_uint32),
        ("fundamental_idx_hi", C.c_uint32),
        ("harmonic_idx_hi", C.c_uint32),
        ("t_obs", C.c_float),
        ("dt", C.c_float),
        ("step_inv", C.c_float),
    ]

    @property
    def nr_pages(self) -> int:
        return (self.fundamental_idx_hi >> 10) + 1


class Template(C.Structure):
    _fields_ = [("P", C.c_float), ("tau", C.c_float), ("Psi0", C.c_float), ("Omega", C.c_float), ("S0", C.c_float)]


CAND_DTYPE = np.dtype(
    [("power", "<f8"), ("P_b", "<f8"), ("tau", "<f8"), ("Psi", "<f8"), ("fA", "<f8"), ("n_harm", "<u4"), ("f0", "<u4")]
)
assert CAND_DTYPE.itemsize == 48


class Oracle:
    def __init__(self) -> None:
        if not ORACLE_SO.exists():
            build(ref=False)
        L = C.CDLL(str(ORACLE_SO))
        self.L = L
        L.brp_oracle_derive_geometry.argtypes = [C.c_uint32, C.c_double, C.c_float, C.c_float, C.c_uint32, C.POINTER(Geometry)]
        L.brp_oracle_derive_template.argtypes = [C.c_double, C.c_double, C.c_double, C.c_float, C.POINTER(Template)]
        L.brp_oracle_sin_lut.argtypes = [C.c_float]
        L.brp_oracle_sin_lut.restype = C.c_float
        L.brp_oracle_resample.argtypes = [f32p, C.POINTER(Geometry), C.POINTER(Template), f32p, C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_float)]
        L.brp_oracle_power_spectrum.argtypes = [f32p, C.c_uint32, C.c_float, f32p]
        L.brp_oracle_harmonic_summing.argtypes = [f32p, C.POINTER(Geometry), f32p, f32p, i32p]
        L.brp_oracle_harmonic_maxima.argtypes = [f32p, C.POINTER(Geometry), f32p]
        L.brp_oracle_chisq_Q.argtypes = [C.c_double, C.c_int]
        L.brp_oracle_chisq_Q.restype = C.c_double
        L.brp_oracle_chisq_Qinv.argtypes = [C.c_double, C.c_int]
        L.brp_oracle_chisq_Qinv.restype = C.c_double
        L.brp_oracle_base_thresholds.argtypes = [C.c_float, C.c_uint32, f32p]
        L.brp_oracle_finalize.argtypes = [C.c_void_p, C.c_void_p]
        L.brp_oracle_format_candidates.argtypes = [C.c_void_p, C.POINTER(Geometry), C.c_double, C.c_char_p, C.c_int]
        L.brp_oracle_whiten.argtypes = [f32p, C.c_uint32, C.c_float, C.c_uint32]
        L.brp_oracle_whiten_zap.argtypes = [f32p, C.c_uint32, C.c_float, C.c_uint32, C.c_double, C.c_void_p, C.c_uint32]
        L.brp_oracle_rngmed.argtypes = [f32p, C.c_uint32, C.c_uint32, f32p]
        L.brp_oracle_search.argtypes = [f32p, C.POINTER(Geometry), C.c_float, f64p, C.c_uint32, C.c_void_p, C.c_void_p]

    def geometry(self, nu: int, tsample_us: float, padding: float, f0: float, window: int) -> Geometry:
        g = Geometry()
        self.L.brp_oracle_derive_geometry(nu, tsample_us, padding, f0, window, C.byref(g))
        return g

    def template(self, P: float, tau: float, psi0: float, step_inv: float) -> Template:
        t = Template()
        self.L.brp_oracle_derive_template(P, tau, psi0, step_inv, C.byref(t))
        return t

    def sin_lut(self, x: np.ndarray) -> np.ndarray:
        x = np.asarray(x, np.float32)
        return np.array([self.L.brp_oracle_sin_lut(float(v)) for v in x.ravel()], np.float32).reshape(x.shape)

    def resample(self, series: np.ndarray, g: Geometry, t: Template, want_idx: bool = True):
        out = np.empty(g.nsamples, np.float32)
        idx = np.full(g.nsamples_unpadded, -1, np.int32) if want_idx else None
        ns, mean = C.c_uint32(0), C.c_float(0)
        rc = self.L.brp_oracle_resample(series, C.byref(g), C.byref(t), out, idx.ctypes.data if want_idx else None, C.byref(ns), C.byref(mean))
        assert rc == 0
        return out, idx, ns.value, mean.value

    def power_spectrum(self, resampled: np.ndarray, nsamples: int) -> np.ndarray:
        ps = np.empty(nsamples // 2 + 1, np.float32)
        rc = self.L.brp_oracle_power_spectrum(resampled, nsamples, np.float32(1.0 / nsamples), ps)
        assert rc == 0
        return ps

    def harmonic_summing(self, ps: np.ndarray, g: Geometry, thr: np.ndarray):
        fu = g.fundamental_idx_hi
        ss = np.empty(4 * fu, np.float32)
        dirty = np.empty(5 * g.nr_pages, np.int32)
        self.L.brp_oracle_harmonic_summing(ps, C.byref(g), np.asarray(thr, np.float32), ss, dirty)
        return ss.reshape(4, fu), dirty.reshape(5, g.nr_pages)

    def harmonic_maxima(self, ps: np.ndarray, g: Geometry) -> np.ndarray:
        fu = g.fundamental_idx_hi
        mx = np.empty(5 * fu, np.float32)
        self.L.brp_oracle_harmonic_maxima(ps, C.byref(g), mx)
        return mx.reshape(5, fu)

    def base_thresholds(self, fA: float, fft_size: int) -> np.ndarray:
        thr = np.empty(5, np.float32)
        self.L.brp_oracle_base_thresholds(fA, fft_size, thr)
        return thr

    def whiten(self, series: np.ndarray, padding: float, window: int) -> np.ndarray:
        out = np.ascontiguousarray(series, np.float32).copy()
        rc = self.L.brp_oracle_whiten(out, out.size, padding, window)
        assert rc == 0
        return out

    def whiten_zap(self, series: np.ndarray, padding: float, window: int, tsample_us: float, zap: np.ndarray) -> np.ndarray:
        """whitening with zap-list noise; zap = float64[n, 2] (fmin, fmax) in Hz in file order"""
        out = np.ascontiguousarray(series, np.float32).copy()
        zap = np.ascontiguousarray(zap, np.float64).reshape(-1, 2)
        rc = self.L.brp_oracle_whiten_zap(out, out.size, padding, window, tsample_us, zap.ctypes.data if len(zap) else None, len(zap))
        assert rc == 0
        return out

    def rngmed(self, data: np.ndarray, window: int) -> np.ndarray:
        data = np.ascontiguousarray(data, np.float32)
        out = np.empty(len(data) - window + 1, np.float32)
        assert self.L.brp_oracle_rngmed(data, len(data), window, out) == 0
        return out

    def search(self, series: np.ndarray, g: Geometry, fA: float, bank: np.ndarray):
        bank = np.ascontiguousarray(bank, np.float64)
        cands = np.zeros(500, CAND_DTYPE)
        tidx = np.full(500, -1, np.int32)
        rc = self.L.brp_oracle_search(series, C.byref(g), fA, bank, bank.shape[0], cands.ctypes.data, tidx.ctypes.data)
        assert rc == 0
        return cands, tidx

    def finalize(self, cands: np.ndarray, tidx: np.ndarray):
        cands, tidx = cands.copy(), tidx.copy()
        self.L.brp_oracle_finalize(cands.ctypes.data, tidx.ctypes.data)
        return cands, tidx

    def format_candidates(self, cands_final: np.ndarray, g: Geometry, tsample_us: float) -> str:
        c = cands_final.copy()
        buf = C.create_string_buffer(1 << 16)
        n = self.L.brp_oracle_format_candidates(c.ctypes.data, C.byref(g), tsample_us, buf, len(buf))
        assert n >= 0
        return buf.raw[:n].decode()


class Reference:
    """The reference's own CPU code (oracle/_ref).  FFT arithmetic behind it is oracle/shim_fftw.c."""

    def __init__(self) -> None:
        if not REF_SO.exists():
            raise FileNotFoundError(f"{REF_SO} not built (needs /root/reference; run make -C oracle)")
        L = C.CDLL(str(REF_SO))
        self.L = L
        L.brp_ref_sin_lut.argtypes = [C.c_float]
        L.brp_ref_sin_lut.restype = C.c_float
        L.brp_ref_resample.argtypes = [f32p, C.c_uint32, C.c_uint32] + [C.c_float] * 6 + [f32p]
        L.brp_ref_resamp_indices.argtypes = [C.c_uint32] + [C.c_float] * 6 + [i32p, C.POINTER(C.c_uint32), C.c_void_p]
        L.brp_ref_fft_power.argtypes = [f32p, C.c_uint32, C.c_float, f32p]
        L.brp_ref_harmonic_summing.argtypes = [f32p, C.c_uint32, C.c_uint32, C.c_uint32, f32p, f32p, i32p, C.POINTER(C.c_uint32)]
        L.brp_ref_rngmed.argtypes = [f32p, C.c_uint, C.c_uint, f32p]

    @staticmethod
    def available() -> bool:
        return REF_SO.exists()

    def sin_lut(self, x: np.ndarray) -> np.ndarray:
        x = np.asarray(x, np.float32)
        return np.array([self.L.brp_ref_sin_lut(float(v)) for v in x.ravel()], np.float32).reshape(x.shape)

    def resample(self, series, g: Geometry, t: Template) -> np.ndarray:
        out = np.empty(g.nsamples, np.float32)
        rc = self.L.brp_ref_resample(series, g.nsamples, g.nsamples_unpadded, t.tau, t.Omega, t.Psi0, g.dt, g.step_inv, t.S0, out)
        assert rc == 0
        return out

    def resamp_indices(self, g: Geometry, t: Template):
        idx = np.full(g.nsamples_unpadded, -1, np.int32)
        ns = C.c_uint32(0)
        rc = self.L.brp_ref_resamp_indices(g.nsamples_unpadded, t.tau, t.Omega, t.Psi0, g.dt, g.step_inv, t.S0, idx, C.byref(ns), None)
        assert rc == 0
        return idx, ns.value

    def fft_power(self, resampled: np.ndarray, nsamples: int) -> np.ndarray:
        ps = np.empty(nsamples // 2 + 1, np.float32)
        buf = np.ascontiguousarray(resampled, np.float32).copy()
        rc = self.L.brp_ref_fft_power(buf, nsamples, np.float32(1.0 / nsamples), ps)
        assert rc == 0
        return ps

    def harmonic_summing(self, ps: np.ndarray, g: Geometry, thr: np.ndarray):
        fu = g.fundamental_idx_hi
        ss = np.empty(4 * fu, np.float32)
        dirty = np.empty(5 * g.nr_pages, np.int32)
        npg = C.c_uint32(0)
        buf = np.ascontiguousarray(ps, np.float32).copy()
        rc = self.L.brp_ref_harmonic_summing(buf, g.window_2, fu, g.harmonic_idx_hi, np.asarray(thr, np.float32).copy(), ss, dirty, C.byref(npg))
        assert rc == 0 and npg.value == g.nr_pages
        return ss.reshape(4, fu), dirty.reshape(5, g.nr_pages)

    def rngmed(self, data: np.ndarray, window: int) -> np.ndarray:
        data = np.ascontiguousarray(data, np.float32)
        out = np.empty(len(data) - window + 1, np.float32)
        self.L.brp_ref_rngmed(data, len(data), window, out)
        return out

    @staticmethod
    def run_cli(args: list[str], cwd: str | None = None, timeout: float = 3600) -> subprocess.CompletedProcess:
        """Run the unmodified reference driver (demod_binary.c MAIN) with the reference's own CLI."""
        return subprocess.run([str(REF_BIN)] + args, cwd=cwd, capture_output=True, text=True, timeout=timeout)
