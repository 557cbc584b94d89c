"""brp_resample_noalloc: the resampler for callers that cannot allocate device memory (fixed memory budget, own arena,
or no tolerance for cudaMalloc's device-wide synchronisation).  Its scratch is a fixed array inside the library, so this
test checks through the C ABI that it gives what the allocating entry point (brp_stage_resample) gives -- on two calls
in a row, because the second call finds the first call's leftovers in the shared scratch array -- and, directly, what the
CPU oracle gives (indices are not an output of this entry point, so the gathered samples are compared bit for bit).

Collected last on purpose (file name): a run that stops at the first failure has then already run every other test.
It passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

import brp_b200
from brp_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("log2nu,padding", [(18, 3.0), (23, 1.0)])        # 2^23: beyond 2^22 samples the kernel indexes in double
def test_noalloc_resampler_matches_allocating_entry_point_twice(oracle, log2nu, padding):
    import torch

    nu = 1 << log2nu
    series, _ = synth.make_series(nu, seed=2718, pulsar=None)
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, padding, 400.0, 1000)   # padded: exercises the mean fill too
    go = oracle.geometry(nu, synth.TSAMPLE_US, padding, 400.0, 1000)
    ref = brp_b200.Search(g, max_batch=1)
    ref.set_series(series)

    # the caller owns every buffer the new entry point touches
    d_series = torch.from_numpy(series).cuda()
    d_out = torch.empty(g.nsamples, dtype=torch.float32, device="cuda")
    d_ns = torch.zeros(1, dtype=torch.int32, device="cuda")
    d_mean = torch.zeros(1, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()

    # two different templates back to back: different n_steps, different number of padded samples
    for row in ([35.0, 0.012, 0.7], [1462.994097917309, 0.192481315985, 1.753485476554]):
        t = brp_b200.derive_template(*row, g.step_inv)
        want, _, want_ns, want_mean = ref.stage_resample(t, want_index=False)
        d_out.fill_(float("nan"))
        torch.cuda.synchronize()
        brp_b200.resample_noalloc(d_series.data_ptr(), g, t, d_out.data_ptr(), d_ns.data_ptr(), d_mean.data_ptr(), 0)
        torch.cuda.synchronize()
        got, got_ns, got_mean = d_out.cpu().numpy(), int(d_ns.item()), float(d_mean.item())
        assert got_ns == want_ns                                                        # integer work: exact
        assert np.array_equal(got[:got_ns].view(np.uint32), want[:want_ns].view(np.uint32))   # gathered samples: bit-exact
        assert abs(got_mean - want_mean) <= 1e-6 * max(1.0, abs(want_mean))             # mean: the tolerance of test_gpu_resample.py
        assert np.all(got[got_ns:] == np.float32(got_mean))                             # padding [n_steps, N)
        # and against the CPU oracle directly (demod_binary_resamp_cpu.c:91-133 restated)
        res, idx, ns, _ = oracle.resample(series, go, oracle.template(*row, go.step_inv))
        assert got_ns == ns and np.array_equal(got[:ns].view(np.uint32), res[:ns].view(np.uint32))
        exact = float(np.mean(series[idx[:ns]].astype(np.float64)))
        assert abs(got_mean - exact) <= 2e-7 * max(1.0, abs(exact))
    ref.close()


def test_noalloc_resampler_rejects_series_beyond_its_scratch():
    import torch

    small = torch.zeros(16, dtype=torch.float32, device="cuda")
    ns = torch.zeros(1, dtype=torch.int32, device="cuda")
    g = brp_b200.Geometry()
    g.nsamples_unpadded = brp_b200.NOALLOC_MAX_SAMPLES + 4096          # rejected before any buffer is touched
    g.nsamples = g.nsamples_unpadded
    g.dt, g.step_inv = 6.5e-5, 15272.7
    t = brp_b200.derive_template(1000.0, 0.1, 0.5, g.step_inv)
    with pytest.raises(brp_b200.BrpError) as e:
        brp_b200.resample_noalloc(small.data_ptr(), g, t, small.data_ptr(), ns.data_ptr(), small.data_ptr(), 0)
    assert e.value.code == brp_b200.RADPUL_EVAL
