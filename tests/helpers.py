"""Shared helpers for the parity tests (inputs are regenerated from seeds; see tests/golden/make_golden.py)."""
import numpy as np

import brp_b200  # noqa: F401
from brp_b200 import synth

GOLDEN_TEMPLATES = np.array([
    [1000.0, 0.0, 0.0],
    [1462.994097917309, 0.192481315985, 1.753485476554],
    [31.25, 0.0123, 4.5],
    [12.5, 0.004, 0.3],
])


def golden_series_AB():
    return synth.make_series(1 << 14, seed=synth.SEED + 7)[0]


def golden_case_C():
    series, q = synth.make_series(1 << 16, seed=synth.SEED + 11,
                                  pulsar=synth.Pulsar(f_spin=173.7, P_orb=40.0, tau=0.01, Psi0=1.0, summed_power=200.0))
    return series, q


def small_bank(n, seed):
    """Short-period orbits so that a seconds-long test series sees whole orbital cycles."""
    bank = synth.make_bank(n, seed=seed)
    bank[2:, 0] = np.round(np.linspace(8.0, 90.0, n - 2), 12)
    bank[2:, 1] = np.round(bank[2:, 1] * 0.05, 12)
    return bank


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def power_close(a, b, rtol=1e-4):
    """|a-b| <= rtol * max(|b|, <P>) (SURVEY.md 8c tap 2: chi^2 bins can be arbitrarily close to 0, hence the floor)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if b.size == 0:
        return 0.0, np.zeros(0)
    floor = max(float(np.mean(np.abs(b))), 1e-30)
    err = np.abs(a - b) / np.maximum(np.abs(b), floor)
    return float(err.max(initial=0.0)), err
