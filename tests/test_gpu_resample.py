"""GPU parity, tap 1: resampling indices and n_steps bit-exact against the oracle and the reference-generated golden
vectors; gathered samples bit-exact; mean within 1e-6 (the reference's serial float32 sum is documented, not reproduced)."""
import numpy as np
import pytest

import brp_b200
from brp_b200 import synth
from helpers import GOLDEN_TEMPLATES, bits, golden_series_AB, small_bank

pytestmark = pytest.mark.gpu


def _check(series, g, t, got, ref_idx, ref_ns, ref_res):
    out, idx, ns, mean = got
    assert ns == ref_ns
    assert np.array_equal(idx[:ns], ref_idx[:ns])                       # bit-exact index array
    assert np.array_equal(bits(out[:ns]), bits(ref_res[:ns]))            # gathered samples
    assert np.array_equal(out[:ns], series[idx[:ns]])
    assert np.all(idx[ns:] == -1)                                        # nothing written past n_steps
    ref_mean = float(ref_res[ns]) if ns < g.nsamples else float(np.mean(ref_res[:ns], dtype=np.float64))
    assert abs(mean - ref_mean) <= 1e-6 * max(1.0, abs(ref_mean))
    assert np.all(out[ns:] == np.float32(mean))                          # mean padding [n_steps, N)
    exact = float(np.mean(series[idx[:ns]].astype(np.float64)))
    assert abs(mean - exact) <= 2e-7 * max(1.0, abs(exact))


@pytest.mark.parametrize("tag,padding", [("A", 1.0), ("B", 3.0)])
def test_resample_matches_golden(golden, tag, padding):
    series = golden_series_AB()
    g = brp_b200.derive_geometry(1 << 14, synth.TSAMPLE_US, padding, 400.0, 200)
    s = brp_b200.Search(g, max_batch=2)
    s.set_series(series)
    for it, row in enumerate(GOLDEN_TEMPLATES):
        t = brp_b200.derive_template(*row, g.step_inv)
        _check(series, g, t, s.stage_resample(t), golden[f"{tag}{it}_idx"], int(golden[f"{tag}{it}_nsteps"]), golden[f"{tag}{it}_resampled"])
    s.close()


@pytest.mark.parametrize("log2n,padding", [(12, 1.0), (17, 1.0), (18, 3.0), (22, 1.0)])
def test_resample_matches_oracle(oracle, log2n, padding):
    nu = 1 << log2n
    series, _ = synth.make_series(nu, seed=100 + log2n)
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, padding, 400.0, 1000 if log2n > 14 else 100)
    go = oracle.geometry(nu, synth.TSAMPLE_US, padding, 400.0, 1000 if log2n > 14 else 100)
    bank = np.concatenate([synth.make_bank(5, seed=log2n), small_bank(5, seed=log2n)[2:],
                           np.array([[700.0, 0.335, 6.2], [2231.0, 0.3349, 3.14159], [660.0, 0.2, 4.71]])])
    if log2n == 22:
        bank = bank[[0, 1, 2, 8, 9]]
    s = brp_b200.Search(g, max_batch=2)
    s.set_series(series)
    for row in bank:
        t = brp_b200.derive_template(*row, g.step_inv)
        to = oracle.template(*row, go.step_inv)
        res, idx, ns, _ = oracle.resample(series, go, to)
        _check(series, g, t, s.stage_resample(t), idx, ns, res)
    s.close()


def test_resample_unaligned_length_and_dc_offset(oracle):
    """Nu not a multiple of the block size or of 4; series with a large DC level (un-whitened data):
    indices stay bit-exact, the mean is the exact one (the reference's serial float32 sum drifts here)."""
    nu = (1 << 15) + 1234 + 2
    rng = np.random.default_rng(5)
    series = (rng.integers(0, 16, nu) / 1.4645e-4).astype(np.float32)
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, 1.0, 100.0, 100)
    go = oracle.geometry(nu, synth.TSAMPLE_US, 1.0, 100.0, 100)
    # FFT lengths must be 2^a or 3*2^a: this odd length is only valid for the resampler, so go through a padded geometry
    g2 = brp_b200.derive_geometry(1 << 16, synth.TSAMPLE_US, 1.0, 100.0, 100)
    g2.nsamples_unpadded = nu
    s = brp_b200.Search(g2, max_batch=1)
    s.set_series(series)
    for row in small_bank(6, seed=9):
        t = brp_b200.derive_template(*row, g.step_inv)
        to = oracle.template(*row, go.step_inv)
        _, idx, ns, _ = oracle.resample(series, go, to)
        out, gidx, gns, mean = s.stage_resample(t)
        assert gns == ns and np.array_equal(gidx[:ns], idx[:ns])
        exact = float(np.mean(series[idx[:ns]].astype(np.float64)))
        assert abs(mean - exact) <= 2e-7 * abs(exact)
    s.close()


@pytest.mark.parametrize("nu", [(1 << 22) + 4096, 1 << 23])
def test_resample_long_series_uses_the_literal_double_index_arithmetic(oracle, nu):
    """Series longer than 2^22 samples: i - del_t can reach 2^23, where the FP32 rounding shortcut is no longer exact, so the
    resampler switches to the reference's literal (int)((double)f + 0.5) (resample_kernel<0>, demod_binary_resamp_cpu.c:116).
    Indices, n_steps and gathered samples bit-exact against the oracle."""
    series, _ = synth.make_series(nu, seed=23, pulsar=None)
    go = oracle.geometry(nu, synth.TSAMPLE_US, 1.0, 400.0, 1000)
    g = brp_b200.derive_geometry(1 << 23, synth.TSAMPLE_US, 1.0, 400.0, 1000)     # FFT lengths must be 2^a or 3*2^a
    g.nsamples_unpadded = nu
    s = brp_b200.Search(g, max_batch=1)
    s.set_series(series)
    for row in ([1462.994097917309, 0.192481315985, 1.753485476554], [660.0, 0.335, 6.2], [1000.0, 0.0, 0.0], [40.0, 3.0, 5.5]):
        t = brp_b200.derive_template(*row, g.step_inv)
        res, idx, ns, _ = oracle.resample(series, go, oracle.template(*row, go.step_inv))
        out, gidx, gns, mean = s.stage_resample(t)
        assert gns == ns and np.array_equal(gidx[:ns], idx[:ns]), row
        assert np.array_equal(bits(out[:ns]), bits(res[:ns]))
        exact = float(np.mean(series[idx[:ns]].astype(np.float64)))
        assert abs(mean - exact) <= 2e-7 * max(1.0, abs(exact))
    s.close()


ADVERSARIAL = np.array([
    [1000.0, 0.0, 0.0],          # null template: del_t == 0 everywhere
    [660.0, 0.335, 6.2],         # corner of the production bank
    [120.0, 1.0, 1.0],           # amplitude 15 000 samples, 2.3 orbits per work-unit
    [40.0, 3.0, 5.5],            # amplitude 46 000 samples, slope 0.47 samples/sample
    [11.0, 0.8, 0.1],            # 25 orbits: many table-cell and cycle boundaries
    [2.0, 0.05, 2.0],            # a table cell lasts ~480 samples (just beyond the phase span run detection accepts)
    [2.3, 0.3, 1.0],             # just inside that span: 2.9e-3 rad per group, amplitude 4 600 samples
    [3.0, 0.2, 0.5],             # 2.2e-3 rad per group, amplitude 3 000 samples, slope 0.42 samples/sample
    [0.3, 0.01, 4.0],            # a table cell lasts ~70 samples
    [900.0, -0.3, 3.0],          # negative amplitude
    [1500.0, 0.33, -40.0],       # negative phase, several cycles
    [1e6, 0.335, 1.0],           # a fraction of one table cell in the whole work-unit
])


@pytest.mark.parametrize("log2n", [16, 20, 22])
def test_run_detection_equals_per_sample_evaluation(oracle, log2n):
    """The production resampler evaluates the timing model at every fourth sample and proves the indices in between
    (csrc/resamp.cu, "Run detection").  Against the same kernel evaluating every sample, on the device, whole banks at a
    time: not one index may differ.  The per-sample path is itself pinned to the oracle above and, for the adversarial
    orbits, here."""
    nu = 1 << log2n
    series, _ = synth.make_series(nu, seed=7, pulsar=None)
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, 1.0, 400.0, 1000)
    s = brp_b200.Search(g, max_batch=1)
    s.set_series(series)
    n_bank = 6662 if log2n == 22 else 1500
    assert s.stage_resample_compare(synth.make_bank(n_bank)) == 0
    assert s.stage_resample_compare(small_bank(300, seed=3)) == 0
    assert s.stage_resample_compare(ADVERSARIAL) == 0
    if log2n <= 20:
        go = oracle.geometry(nu, synth.TSAMPLE_US, 1.0, 400.0, 1000)
        for row in ADVERSARIAL:
            t = brp_b200.derive_template(*row, g.step_inv)
            _, idx, ns, _ = oracle.resample(series, go, oracle.template(*row, go.step_inv))
            _, gidx, gns, _ = s.stage_resample(t, want_series=False)
            assert gns == ns and np.array_equal(gidx[:ns], idx[:ns]), row
    s.close()
