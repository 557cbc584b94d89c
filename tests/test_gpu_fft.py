"""GPU parity, tap 2a: hand-written two-pass FFT + power epilogue against the oracle's double-precision DFT.
Tolerance (north_star / SURVEY.md 8c): |a-b| <= 1e-4 * max(|b|, <P>).  Full-size checks use size-independent
properties (impulse, tone, Parseval, linearity of the amplitude spectrum)."""
import numpy as np
import pytest

import brp_b200
from brp_b200 import synth
from helpers import GOLDEN_TEMPLATES, golden_series_AB, power_close

pytestmark = pytest.mark.gpu

RTOL = 1e-4


def _search(n, padding=1.0, window=100):
    nu = int(round(n / padding))
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, padding, 400.0, window)
    assert g.nsamples == n
    return brp_b200.Search(g, max_batch=1), g


@pytest.mark.parametrize("n", [1 << 9, 1 << 10, 1 << 12, 1 << 13, 1 << 14, 1 << 15, 1 << 16, 1 << 17, 1 << 18, 1 << 19, 1 << 20, 1 << 21])
def test_spectrum_matches_oracle_pow2(oracle, n):
    s, g = _search(n, window=10)
    x = np.random.default_rng(n).standard_normal(n).astype(np.float32)
    got = s.stage_spectrum_of(x)
    want = oracle.power_spectrum(x, n)
    err, _ = power_close(got, want, RTOL)
    assert err <= RTOL, err
    assert got[0] == 0.0
    # much tighter in the mean: a float32 FFT should sit near 1e-6
    rel = np.linalg.norm(got.astype(np.float64) - want) / np.linalg.norm(want)
    assert rel < 2e-6, rel
    s.close()


@pytest.mark.parametrize("n", [3 << 9, 3 << 11, 3 << 12, 3 << 14, 3 << 15, 3 << 17, 3 << 20])
def test_spectrum_matches_oracle_3x(oracle, n):
    s, g = _search(n, padding=3.0, window=10)
    x = np.random.default_rng(n).standard_normal(n).astype(np.float32)
    got = s.stage_spectrum_of(x)
    want = oracle.power_spectrum(x, n)
    err, _ = power_close(got, want, RTOL)
    assert err <= RTOL, err
    rel = np.linalg.norm(got.astype(np.float64) - want) / np.linalg.norm(want)
    assert rel < 2e-6, rel
    s.close()


def test_unsupported_length_is_rejected():
    g = brp_b200.derive_geometry(5 << 12, synth.TSAMPLE_US, 1.0, 400.0, 10)
    with pytest.raises(brp_b200.BrpError) as e:
        brp_b200.Search(g, max_batch=1)
    assert e.value.code == brp_b200.RADPUL_CUDA_FFT_PLAN


@pytest.mark.parametrize("tag,padding", [("A", 1.0), ("B", 3.0)])
def test_resample_then_spectrum_matches_golden(golden, tag, padding):
    """Mean padding applied on load by the FFT's first pass == the reference's materialised padding."""
    series = golden_series_AB()
    g = brp_b200.derive_geometry(1 << 14, synth.TSAMPLE_US, padding, 400.0, 200)
    s = brp_b200.Search(g, max_batch=2)
    s.set_series(series)
    for it, row in enumerate(GOLDEN_TEMPLATES):
        t = brp_b200.derive_template(*row, g.step_inv)
        s.stage_resample(t, want_series=False, want_index=False)
        got = s.stage_spectrum()
        err, _ = power_close(got, golden[f"{tag}{it}_ps"], RTOL)
        assert err <= RTOL, (tag, it, err)
    s.close()


@pytest.mark.parametrize("log2nu,padding", [(20, 1.0), (20, 3.0), (21, 1.0), (21, 3.0)])
def test_short_work_units_resample_then_spectrum(oracle, log2nu, padding):
    """Work-units of 2^20 and 2^21 samples run a 2048-row (3 x 2048 with -P 3) pass 1 on the persistent tensor-copy kernel and a
    short pass 2 (brp_fft_plan_create): resample on the device, then the spectrum with the padding applied on load, against the
    oracle's transform of the materialised padded series."""
    nu = 1 << log2nu
    series, _ = synth.make_series(nu, seed=log2nu)
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, padding, 400.0, 1000)
    s = brp_b200.Search(g, max_batch=1)
    s.set_series(series)
    t = brp_b200.derive_template(1462.994097917309, 0.192481315985, 1.753485476554, g.step_inv)
    out, _, n_steps, mean = s.stage_resample(t, want_series=True, want_index=False)
    got = s.stage_spectrum()
    padded = out.copy()
    padded[n_steps:] = np.float32(mean)
    want = oracle.power_spectrum(padded, g.nsamples)
    err, _ = power_close(got, want, RTOL)
    assert err <= RTOL, err
    rel = np.linalg.norm(got.astype(np.float64) - want) / np.linalg.norm(want)
    assert rel < 2e-6, rel
    s.close()


@pytest.mark.parametrize("n", [1 << 22, 3 << 22, 1 << 23])
def test_full_size_properties(n):
    padding = 3.0 if n % 3 == 0 else 1.0
    s, g = _search(n, padding=padding, window=1000)
    f = g.fft_size
    # impulse: |X_k| = 1 for every k  ->  P = 1/N
    x = np.zeros(n, np.float32)
    x[12345] = 1.0
    got = s.stage_spectrum_of(x)
    assert got[0] == 0.0 and np.allclose(got[1:], 1.0 / n, rtol=2e-5, atol=0)
    # pure tone on a bin: all power in that bin, P = A^2 N / 4
    k0 = 1234567
    t = np.arange(n, dtype=np.float64)
    x = (0.5 * np.cos(2 * np.pi * k0 * t / n + 0.3)).astype(np.float32)
    got = s.stage_spectrum_of(x)
    assert abs(got[k0] - 0.25 * n / 4) <= 1e-4 * 0.25 * n / 4
    rest = np.delete(got, k0)
    assert rest.max() < 1e-6 * got[k0]
    # Parseval on noise: sum_k P_k (k=1..F-1, interior bins counted twice) = sum x^2 - (DC and Nyquist terms)
    rng = np.random.default_rng(3)
    x = rng.standard_normal(n).astype(np.float32)
    x -= np.float32(x.mean(dtype=np.float64))
    got = s.stage_spectrum_of(x).astype(np.float64)
    total = 2.0 * got[1:f - 1].sum() + got[f - 1]
    want = float(np.sum(x.astype(np.float64) ** 2)) - float(np.sum(x.astype(np.float64))) ** 2 / n
    assert abs(total - want) <= 2e-6 * want
    # spot-check 64 bins against a direct double-precision DFT
    ks = rng.integers(1, f, 64)
    xd = x.astype(np.float64)
    for k in ks[:16]:
        ph = np.exp(-2j * np.pi * ((k * np.arange(n, dtype=np.int64)) % n) / n)
        want_k = abs(np.dot(xd, ph)) ** 2 / n
        assert abs(got[k] - want_k) <= 1e-4 * max(want_k, 1.0)
    s.close()


_TMA_SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, sys.argv[1])
import brp_b200
from brp_b200 import synth
nu = 1 << 22
series, _ = synth.make_series(nu, seed=77)
g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, 1.0, 400.0, 1000)
s = brp_b200.Search(g, max_batch=1)
s.set_series(series)
t = brp_b200.derive_template(1462.994097917309, 0.192481315985, 1.753485476554, g.step_inv)
s.stage_resample(t, want_series=False, want_index=False)
np.save(sys.argv[2], s.stage_spectrum(g.harmonic_idx_hi))
s.close()
"""


def test_tensor_copy_kernels_equal_the_plain_kernels_bit_for_bit(tmp_path):
    """The production FFT kernels (tiles moved by TMA tensor copies, fft_pass1_tma_kernel / fft_pass2_tma_kernel) do the same
    arithmetic as the plain kernels every other shape uses; BRP_FFT_TMA=0 selects the plain ones.  Same work-unit, same
    template, two processes: the power spectra must be identical bit patterns."""
    import os
    import subprocess
    import sys
    from pathlib import Path
    root = str(Path(__file__).resolve().parent.parent)
    outs = []
    for flag in ("1", "0"):
        out = tmp_path / f"ps_{flag}.npy"
        env = dict(os.environ, BRP_FFT_TMA=flag)
        r = subprocess.run([sys.executable, "-c", _TMA_SCRIPT, root, str(out)], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(np.load(out))
    assert outs[0].size > 1_000_000 and np.all(np.isfinite(outs[0])) and outs[0][1:].min() >= 0 and outs[0].max() > 0
    assert np.array_equal(outs[0].view(np.uint32), outs[1].view(np.uint32))
