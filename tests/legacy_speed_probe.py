"""Script (not a test): wall-clock rate of the UNMODIFIED reference driver (oracle/_ref/demod_binary_b200 = demod_binary.c
-DUSE_CUDA linked against libbrp_b200.so) on a synthetic 2^22-sample work-unit, i.e. the drop-in path with one template
per call and a synchronous read-back after each stage.      python tests/legacy_speed_probe.py [n_templates]"""
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import brp_b200  # noqa: E402,F401
from brp_b200 import synth  # noqa: E402

n_t = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
exe = ROOT / "oracle" / "_ref" / "demod_binary_b200"
with tempfile.TemporaryDirectory() as d:
    d = Path(d)
    series, q = synth.make_series(1 << 22, seed=synth.SEED)
    synth.write_binary_wu(d / "wu.binary", q)
    for n in (8, n_t + 8):          # the difference of two runs removes start-up, ingest and output
        synth.write_bank(d / f"bank{n}.txt", synth.make_bank(n))
    times = {}
    for n in (8, 8, n_t + 8):       # the first run pays for loading the CUDA driver and is discarded
        t0 = time.perf_counter()
        r = subprocess.run([str(exe), "-i", str(d / "wu.binary"), "-o", str(d / "o.cand"), "-t", str(d / f"bank{n}.txt"), "-c", str(d / f"c{n}.cpt"),
                            "-f", "400", "-A", "0.08", "-P", "1.0", "-B", "1000", "-D", "0"], capture_output=True, text=True)
        times[n] = time.perf_counter() - t0
        assert r.returncode == 0, r.stderr[-1000:]
    dt = times[n_t + 8] - times[8]
    print(f"unmodified reference driver on this backend: {n_t} templates in {dt:.2f} s = {n_t / dt:.0f} templates/s (whole runs: {times}); "
          "includes the driver's own per-template work: the standalone (non-BOINC) driver writes its checkpoint file after EVERY "
          "template (demod_binary.c:1423-1427), so this figure depends on the file system under the temporary directory")

# the backend alone: the eleven legacy functions called in the driver's order from this process (no driver, no checkpoint file)
g = brp_b200.derive_geometry(1 << 22, synth.TSAMPLE_US, 1.0, 400.0, 1000)
be = brp_b200.LegacyBackend()
be.set_up(series, g)
thr = brp_b200.base_thresholds(0.08, g.fft_size)
bank = synth.make_bank(n_t + 20)
for row in bank[:20]:
    be.run_template(brp_b200.derive_template(*row, g.step_inv), thr)
t0 = time.perf_counter()
for row in bank[20:]:
    be.run_template(brp_b200.derive_template(*row, g.step_inv), thr)
dt = time.perf_counter() - t0
be.shutdown()
print(f"legacy backend functions alone (run_resampling + run_fft + run_harmonic_summing per template, ctypes caller): "
      f"{n_t} templates in {dt:.3f} s = {n_t / dt:.0f} templates/s")
