"""CPU-only: pins the oracle (oracle/brp_oracle.cpp) against (a) golden vectors produced by the reference's
own code, (b) that code itself when oracle/_ref is present, (c) independent numerics (scipy)."""
import tempfile

import numpy as np
import pytest

from helpers import GOLDEN_TEMPLATES, bits, golden_case_C, golden_series_AB, power_close, small_bank

import brp_b200  # noqa: F401
from brp_b200 import synth


@pytest.mark.parametrize("tag,padding", [("A", 1.0), ("B", 3.0)])
def test_oracle_matches_golden_stages(oracle, golden, tag, padding):
    series = golden_series_AB()
    g = oracle.geometry(1 << 14, synth.TSAMPLE_US, padding, 400.0, 200)
    thr = golden[f"{tag}_thr"]
    for it, row in enumerate(GOLDEN_TEMPLATES):
        t = oracle.template(*row, g.step_inv)
        res, idx, ns, _ = oracle.resample(series, g, t)
        assert ns == int(golden[f"{tag}{it}_nsteps"])
        assert np.array_equal(idx, golden[f"{tag}{it}_idx"])                      # bit-exact indices
        assert np.array_equal(bits(res), bits(golden[f"{tag}{it}_resampled"]))    # incl. serial-mean padding
        ps = oracle.power_spectrum(res, g.nsamples)
        assert np.array_equal(bits(ps), bits(golden[f"{tag}{it}_ps"]))
        ss, dirty = oracle.harmonic_summing(ps, g, thr)
        assert np.array_equal(bits(ss), bits(golden[f"{tag}{it}_sumspec"]))
        assert np.array_equal(dirty, golden[f"{tag}{it}_dirty"])
        assert dirty.sum() > 0


@pytest.mark.parametrize("tag,padding", [("C1", 1.0), ("C3", 3.0)])
def test_oracle_search_matches_golden_candidate_file(oracle, golden, tag, padding):
    series, _ = golden_case_C()
    bank = golden["C_bank"]
    g = oracle.geometry(1 << 16, synth.TSAMPLE_US, padding, 400.0, 200)
    cands, tidx = oracle.search(series, g, 0.08, bank)
    cf, _ = oracle.finalize(cands, tidx)
    text = oracle.format_candidates(cf, g, synth.TSAMPLE_US)
    ref_text = bytes(golden[f"{tag}_candfile"]).decode()
    assert text == ref_text
    assert len(text.splitlines()) > 5


def test_sin_lut_matches_reference(oracle, reference):
    x = np.concatenate([np.linspace(-50, 50, 40001), np.array([0.0, -0.0, 1e-9, -1e-9, 6.283185, -6.283185, 1e4, -1e4])]).astype(np.float32)
    assert np.array_equal(bits(oracle.sin_lut(x)), bits(reference.sin_lut(x)))


@pytest.mark.parametrize("padding", [1.0, 3.0, 2.0])
def test_oracle_matches_reference_stages(oracle, reference, padding):
    nu = 1 << 15
    series, _ = synth.make_series(nu, seed=99)
    g = oracle.geometry(nu, synth.TSAMPLE_US, padding, 400.0, 100)
    thr = oracle.base_thresholds(0.08, g.fft_size) * np.float32(0.6)
    for row in small_bank(6, 5):
        t = oracle.template(*row, g.step_inv)
        res, idx, ns, _ = oracle.resample(series, g, t)
        ridx, rns = reference.resamp_indices(g, t)
        assert ns == rns and np.array_equal(idx, ridx)
        rres = reference.resample(series, g, t)
        assert np.array_equal(bits(res), bits(rres))
        assert np.array_equal(series[idx[:ns]], rres[:ns])          # the gather really uses those indices
        ps, rps = oracle.power_spectrum(res, g.nsamples), reference.fft_power(rres, g.nsamples)
        assert np.array_equal(bits(ps), bits(rps))
        ss, d = oracle.harmonic_summing(ps, g, thr)
        rss, rd = reference.harmonic_summing(rps, g, thr)
        assert np.array_equal(bits(ss), bits(rss)) and np.array_equal(d, rd)


def test_oracle_search_matches_reference_driver(oracle, reference):
    """Whole loop vs the unmodified reference MAIN() run through its own command line."""
    nu = 1 << 15
    series, q = synth.make_series(nu, seed=4242, pulsar=synth.Pulsar(P_orb=25.0, tau=0.008, Psi0=2.0, summed_power=150.0))
    bank = small_bank(10, 77)
    bank[1] = (25.0, 0.008, 2.0)
    with tempfile.TemporaryDirectory() as d:
        synth.write_binary_wu(d + "/wu.binary", q)
        synth.write_bank(d + "/bank.txt", bank)
        r = reference.run_cli(["-i", d + "/wu.binary", "-o", d + "/o.cand", "-t", d + "/bank.txt", "-c", d + "/c.cpt",
                               "-f", "300", "-A", "0.08", "-P", "1.0", "-B", "100"])
        assert r.returncode == 0, r.stderr
        ref_text = open(d + "/o.cand").read()
    g = oracle.geometry(nu, synth.TSAMPLE_US, 1.0, 300.0, 100)
    cands, tidx = oracle.search(series, g, 0.08, bank)
    cf, _ = oracle.finalize(cands, tidx)
    assert oracle.format_candidates(cf, g, synth.TSAMPLE_US) == ref_text


@pytest.mark.parametrize("n", [1 << 12, 3 << 12, 1 << 17, 3 << 16, 9 << 10])
def test_oracle_fft_against_scipy_float64(oracle, n):
    """FFTW is absent: the DFT definition is the contract, cross-checked against pocketfft in float64."""
    import scipy.fft
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n).astype(np.float32)
    ps = oracle.power_spectrum(x, n)
    want = np.abs(scipy.fft.rfft(x.astype(np.float64))) ** 2 / n
    want[0] = 0.0
    err, _ = power_close(ps, want, rtol=1e-6)
    assert err < 1e-6   # float32 rounding of the spectrum only (~1.2e-7 on amplitude)
    assert ps[0] == 0.0


def test_chisq_against_scipy(oracle):
    from scipy.stats import chi2
    for n in (1, 2, 4, 8, 16):
        for q in (1e-3, 4e-8, 1e-12):
            x = oracle.L.brp_oracle_chisq_Qinv(q, n)
            assert abs(x - chi2.isf(q, 2 * n)) <= 1e-9 * x
            assert abs(oracle.L.brp_oracle_chisq_Q(x, n) - q) <= 1e-9 * q
    thr = oracle.base_thresholds(0.08, 6291457)   # the reference bench config
    assert np.allclose(2 * thr, [18.139 * 2, 21.241 * 2, 26.269 * 2, 34.648 * 2, 48.958 * 2], rtol=2e-4)


def test_geometry_table(oracle):
    """Derived sizes of SURVEY.md section 8(d)."""
    rows = [(1 << 22, 1.0, (4194304, 2097153, 500, 109851, 1757613)),
            (1 << 22, 3.0, (12582912, 6291457, 500, 329552, 5272839)),
            (1 << 22, 2.0, (8388608, 4194305, 500, 219702, 3515226)),
            (1 << 20, 1.0, (1048576, 524289, 500, 27463, 439403))]
    for nu, pad, want in rows:
        g = oracle.geometry(nu, synth.TSAMPLE_US, pad, 400.0, 1000)
        assert (g.nsamples, g.fft_size, g.window_2, g.fundamental_idx_hi, g.harmonic_idx_hi) == want


@pytest.mark.parametrize("tag,padding", [("C1W", 1.0), ("C3W", 3.0)])
def test_oracle_whitening_matches_golden_candidate_file(oracle, golden, tag, padding):
    """Whitening pre-step (demod_binary.c:858-1079, empty zap list) restated in the oracle: the reference driver run with
    -W produced these candidate files."""
    series, _ = golden_case_C()
    bank = golden["C_bank"]
    g = oracle.geometry(1 << 16, synth.TSAMPLE_US, padding, 400.0, 200)
    white = oracle.whiten(series, padding, 200)
    cands, tidx = oracle.search(white, g, 0.08, bank)
    cf, _ = oracle.finalize(cands, tidx)
    assert oracle.format_candidates(cf, g, synth.TSAMPLE_US) == bytes(golden[f"{tag}_candfile"]).decode()


def test_oracle_running_median_matches_reference(oracle, reference):
    rng = np.random.default_rng(8)
    x = rng.exponential(1.0, 6000).astype(np.float32)
    x[100:140] = x[100]                       # ties
    for window in (3, 10, 11, 100, 1000, 1001):
        assert np.array_equal(bits(oracle.rngmed(x, window)), bits(reference.rngmed(x, window)))
