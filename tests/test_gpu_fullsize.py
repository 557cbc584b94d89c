"""GPU parity at BASELINE.json's full sizes: the 2^22-sample work-unit with -P 1 (configs 1-3), -P 3 (the reference's own
bench command, FFT length 3*2^22) and -P 2 (config 4: 2^23-point padded series), a few templates each against the oracle
(bit-exact n_steps / gathered samples, spectrum within 1e-4, identical toplist), plus config 5 (several short 2^20-sample
work-units in flight at once on separate handles/streams)."""
import numpy as np
import pytest

import brp_b200
from brp_b200 import synth
from helpers import power_close
from test_gpu_search import _compare_toplists

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("padding", [1.0, 3.0, 2.0])
def test_full_size_workunit_matches_oracle(oracle, padding):
    nu = 1 << 22
    # the benchmark's work-unit; mean padding dilutes the normalised power by 1/padding, so the padded cases get a
    # proportionally stronger pulsar to keep the toplists populated
    series, _ = synth.make_series(nu, pulsar=synth.Pulsar(summed_power=16.0 + 104.0 * padding))
    bank = synth.make_bank(3)                                   # null template, the injected orbit, one random template
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, padding, 400.0, 1000)
    go = oracle.geometry(nu, synth.TSAMPLE_US, padding, 400.0, 1000)
    s = brp_b200.Search(g, fA=0.08, max_batch=2)
    s.set_series(series)
    # stage taps for the injected orbit
    t, to = brp_b200.derive_template(*bank[1], g.step_inv), oracle.template(*bank[1], go.step_inv)
    res, idx, ns, _ = oracle.resample(series, go, to)
    out, gidx, gns, mean = s.stage_resample(t)
    assert gns == ns and np.array_equal(gidx[:ns], idx[:ns]) and np.array_equal(out[:ns], res[:ns])
    ps = oracle.power_spectrum(res, go.nsamples)
    got = s.stage_spectrum(g.harmonic_idx_hi)
    err, _ = power_close(got, ps[: g.harmonic_idx_hi])
    assert err <= 1e-4, err
    # whole loop
    s.run(bank)
    cands, tidx = brp_b200.toplist_from_hits(s.collect(), bank)
    s.close()
    want, want_t = oracle.search(series, go, 0.08, bank)
    _compare_toplists(cands, tidx, want, want_t)
    assert np.count_nonzero(want["power"]) > 0
    k16 = slice(400, 500)
    assert tidx[k16][np.argmax(cands["power"][k16])] == 1      # the injected pulsar is found by its own template
    assert cands["power"][k16].max() > 60.0


def test_many_short_workunits_in_flight():
    """Config 5: independent 2^20-sample work-units, one handle (= one stream) each, all enqueued before any is awaited."""
    nu = 1 << 20
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, 1.0, 400.0, 1000)
    bank = synth.make_bank(12)
    wus = [synth.make_series(nu, seed=500 + k)[0] for k in range(4)]
    alone = []
    for x in wus:
        cands, tidx = brp_b200.search_bank(x, g, bank, max_batch=4)
        alone.append((cands.tobytes(), tidx.tobytes()))
    handles = [brp_b200.Search(g, max_batch=4) for _ in wus]
    for h, x in zip(handles, wus):
        h.set_series(x)
        h.run(bank)                                            # asynchronous: returns once the work is enqueued
    for h, ref in zip(handles, alone):
        cands, tidx = brp_b200.toplist_from_hits(h.collect(), bank)
        assert (cands.tobytes(), tidx.tobytes()) == ref
        h.close()


def test_results_do_not_depend_on_batching_or_repetition():
    """The device table is filled by 64-bit atomic maxima over (power bits, earliest template) and the mean by index-ordered
    sums, so neither the number of templates per launch nor the scheduling of a repeated run may change a single bit of
    the hit list (power, bin, template, harmonic) of a full-size work-unit."""
    nu = 1 << 22
    series, _ = synth.make_series(nu, seed=21)
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, 1.0, 400.0, 1000)
    bank = synth.make_bank(150, seed=22)
    lists = []
    for max_batch in (64, 64, 7, 1):
        s = brp_b200.Search(g, 0.08, max_batch)
        s.set_series(series)
        s.run(bank)
        hits = s.collect()
        s.close()
        order = np.lexsort((hits["tmpl"], hits["bin"], hits["harm_idx"]))
        lists.append(hits[order])
    assert len(lists[0]) > 20
    for other in lists[1:]:
        assert other.tobytes() == lists[0].tobytes()
