"""Script (not a test): error of the GPU power spectrum of one resampled 2^22-sample series against the oracle's
double-precision transform, in the metric of helpers.power_close.  Used for A/B runs of FFT variants:
    python tests/fft_error_probe.py"""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent))
import brp_b200  # noqa: E402
from brp_b200 import synth  # noqa: E402
from helpers import power_close  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402

orc = Oracle()
for log2n, padding in ((22, 1.0), (22, 3.0), (20, 1.0)):
    nu = 1 << log2n
    series, _ = synth.make_series(nu, seed=11)
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, padding, 400.0, 1000)
    s = brp_b200.Search(g, max_batch=1)
    s.set_series(series)
    t = brp_b200.derive_template(1462.994097917309, 0.192481315985, 1.753485476554, g.step_inv)
    res, _, ns, mean = s.stage_resample(t, want_index=False)
    ps = s.stage_spectrum(g.harmonic_idx_hi)
    ref = orc.power_spectrum(res, g.nsamples)[: g.harmonic_idx_hi]
    worst, err = power_close(ps, ref)
    print(f"2^{log2n} P={padding}: max err {worst:.3e}  rms err {float(np.sqrt(np.mean(err ** 2))):.3e}  (tolerance 1e-4)")
    s.close()
