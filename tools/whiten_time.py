"""Device time of the whitening pre-step (brp_whiten_series) for a 2^22-sample work-unit at -P 1 and -P 3."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, brp_b200
from brp_b200 import synth
series, _ = synth.make_series(1 << 22, seed=3)
for pad in (1.0, 3.0):
    for _ in range(2):
        w, ms = brp_b200.whiten_series(series, pad, 1000)
    print("whiten 2^22 padding", pad, "device ms", ms, float(np.std(w)))
