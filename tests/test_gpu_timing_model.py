"""GPU parity, tap 1 pieces: the device restatement of the reference's timing model must be BIT-EXACT
(erp_utilities.cpp:176-209, demod_binary_resamp_cpu.c:116), since resampling indices are integer work."""
import numpy as np
import pytest

import brp_b200
from helpers import bits

pytestmark = pytest.mark.gpu


def test_sin_lut_bit_exact(oracle):
    rng = np.random.default_rng(11)
    x = np.concatenate([
        rng.uniform(-100, 100, 400000), rng.uniform(0, 12, 400000),           # the physical range: Omega*t + Psi0 < ~10
        rng.uniform(-1e-3, 1e-3, 20000), rng.uniform(-3e7, 3e7, 20000),        # tiny and huge (slow path) arguments
        np.arange(-64, 65) * (6.283185 / 64.0), np.arange(-64, 65) * (6.283185 / 128.0),
        np.array([0.0, -0.0, 6.283185, -6.283185, 3.1415925, 1e-38, -1e-38, 2.6e7, -2.6e7, 4194304 * 6.283185]),
    ]).astype(np.float32)
    got = brp_b200.stage_sin_lut(x)
    want = oracle.sin_lut(x)
    bad = np.nonzero(bits(got) != bits(want))[0]
    assert bad.size == 0, f"{bad.size} mismatches, first x={x[bad[0]]!r}: gpu {got[bad[0]]!r} oracle {want[bad[0]]!r}"


def test_round_index_bit_exact():
    rng = np.random.default_rng(12)
    base = rng.integers(0, 1 << 22, 300000).astype(np.float64)
    f = np.concatenate([
        base + rng.random(300000), base + 0.5, base + 0.25, base + 0.75, np.nextafter((base + 0.5).astype(np.float32), 0),
        rng.uniform(-1.4, 1.0, 50000), rng.uniform(0, 64, 50000), rng.uniform(4194300, 8388607.5, 50000),
        rng.uniform(8388608, 16777215, 50000),                                  # >= 2^23: generic path
        np.array([0.0, -0.0, 0.5, 0.49999997, 1.5, 2.5, -0.5, -0.49999997, 8388607.5, 8388607.0, 4194303.5]),
    ]).astype(np.float32)
    got = brp_b200.stage_round_index(f)
    want = (f.astype(np.float64) + 0.5).astype(np.int64)      # C truncation toward zero == numpy astype
    assert np.array_equal(got.astype(np.int64), np.maximum(want, 0))
