"""GPU parity, tap 2b: fused harmonic summing + per-bin table.  The float additions follow the reference's association
(hs_common.c:107,125-126,145-148), so above-threshold maxima must be BIT-EXACT, and the set of (harmonic, bin) identical."""
import numpy as np
import pytest

import brp_b200
from brp_b200 import synth
from helpers import bits

pytestmark = pytest.mark.gpu


def _spectrum(rng, n, lines=8):
    ps = rng.exponential(1.0, n).astype(np.float32)     # chi^2_2 / 2: the null distribution of a normalised periodogram
    for k in rng.integers(n // 60, n // 17, lines):           # a few strong lines with harmonics
        for h in range(1, 17):
            ps[min(k * h, n - 1)] += np.float32(rng.uniform(3, 15))
    ps[0] = 0.0
    return ps


def _expected(oracle, ps, go, thr):
    mx = oracle.harmonic_maxima(ps, go)
    out = {}
    for k in range(5):
        bins = np.nonzero(mx[k] > thr[k])[0]
        for b in bins[bins >= go.window_2]:
            out[(k, int(b))] = mx[k, b]
    return out


@pytest.mark.parametrize("log2nu,padding,f0,window", [(14, 1.0, 400.0, 200), (16, 3.0, 400.0, 100), (18, 1.0, 250.0, 1000),
                                                      (16, 1.0, 6000.0, 50), (22, 1.0, 400.0, 1000), (22, 3.0, 400.0, 1000)])
def test_hits_match_oracle(oracle, log2nu, padding, f0, window):
    nu = 1 << log2nu
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, padding, f0, window)
    go = oracle.geometry(nu, synth.TSAMPLE_US, padding, f0, window)
    rng = np.random.default_rng(log2nu * 7 + int(padding))
    ps = _spectrum(rng, g.fft_size)
    thr = brp_b200.base_thresholds(0.08, g.fft_size) * np.float32(0.7)     # lowered: a few hundred crossings
    s = brp_b200.Search(g, max_batch=1)
    hits = s.stage_harmonic_summing(ps, thr, tmpl_index=41)
    want = _expected(oracle, ps, go, thr)
    got = {(int(h["harm_idx"]), int(h["bin"])): h["power"] for h in hits}
    assert set(got) == set(want)
    assert len(want) > 20
    for key, p in want.items():
        assert np.float32(got[key]).tobytes() == np.float32(p).tobytes(), key
    assert np.all(hits["tmpl"] == 41)
    assert len(hits) == len(got)                                             # one entry per (harmonic, bin)
    s.close()


def test_reference_sumspec_above_threshold_equals_hits(oracle, reference):
    """Against the reference's own harmonic_summing(): every entry of its sumspec arrays that exceeds the threshold (the only
    ones the candidate scan reads, demod_binary.c:1350) is in the device table with the same bits, and nothing else is."""
    nu = 1 << 16
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, 1.0, 400.0, 100)
    go = oracle.geometry(nu, synth.TSAMPLE_US, 1.0, 400.0, 100)
    ps = _spectrum(np.random.default_rng(1), g.fft_size)
    thr = brp_b200.base_thresholds(0.08, g.fft_size) * np.float32(0.6)
    ss, dirty = reference.harmonic_summing(ps, go, thr)
    s = brp_b200.Search(g, max_batch=1)
    hits = s.stage_harmonic_summing(ps, thr)
    got = {(int(h["harm_idx"]), int(h["bin"])): h["power"] for h in hits}
    fu, w2 = g.fundamental_idx_hi, g.window_2
    want = {}
    for b in np.nonzero(ps[:fu] > thr[0])[0]:
        if b >= w2:
            want[(0, int(b))] = ps[b]
    for k in range(1, 5):
        for b in np.nonzero(ss[k - 1] > thr[k])[0]:
            if b >= w2:
                want[(k, int(b))] = ss[k - 1][b]
    assert set(got) == set(want) and len(want) > 20
    assert all(bits(np.float32(got[k]))[0] == bits(np.float32(v))[0] for k, v in want.items())
    for (k, b) in want:
        assert dirty[k, b >> 10] == 1
    s.close()
