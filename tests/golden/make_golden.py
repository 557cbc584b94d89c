"""Generate tests/golden/golden_v1.npz from the REFERENCE's own code (oracle/_ref, i.e. the reference
translation units compiled unmodified from /root/reference) on seeded synthetic inputs.

Run in the build container (where /root/reference exists):   python tests/golden/make_golden.py
The inputs are regenerated from the seed by the tests, so only outputs are stored.

Cases
  A  Nu = 2^14, padding 1.0  : index array, n_steps, resampled series, spectrum, summed spectra, dirty pages
  B  Nu = 2^14, padding 3.0  : same taps for the 3*2^n FFT length
  C  Nu = 2^16, 12-template bank through the unmodified reference driver (demod_binary.c MAIN): candidate file,
     without and with the whitening pre-step (-W, empty zap list)
"""
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import brp_b200  # noqa: E402,F401
from brp_b200 import synth  # noqa: E402
from oracle.oracle import Oracle, Reference  # noqa: E402

CASE_TEMPLATES = np.array([
    [1000.0, 0.0, 0.0],
    [1462.994097917309, 0.192481315985, 1.753485476554],
    [31.25, 0.0123, 4.5],
    [12.5, 0.004, 0.3],
])


def small_bank(n, seed):
    bank = synth.make_bank(n, seed=seed)
    bank[2:, 0] = np.round(np.linspace(8.0, 90.0, n - 2), 12)   # short orbits: the series is only seconds long
    bank[2:, 1] = np.round(bank[2:, 1] * 0.05, 12)
    return bank


def main():
    O, R = Oracle(), Reference()
    out = {}
    for tag, padding in (("A", 1.0), ("B", 3.0)):
        nu = 1 << 14
        series, _ = synth.make_series(nu, seed=synth.SEED + 7)
        g = O.geometry(nu, synth.TSAMPLE_US, padding, 400.0, 200)
        thr = O.base_thresholds(0.08, g.fft_size) * np.float32(0.5)  # lowered so that dirty pages exist
        out[f"{tag}_thr"] = thr
        for it, row in enumerate(CASE_TEMPLATES):
            t = O.template(*row, g.step_inv)
            idx, ns = R.resamp_indices(g, t)
            res = R.resample(series, g, t)
            ps = R.fft_power(res, g.nsamples)
            ss, dirty = R.harmonic_summing(ps, g, thr)
            out[f"{tag}{it}_idx"] = idx
            out[f"{tag}{it}_nsteps"] = np.uint32(ns)
            out[f"{tag}{it}_resampled"] = res
            out[f"{tag}{it}_ps"] = ps
            out[f"{tag}{it}_sumspec"] = ss
            out[f"{tag}{it}_dirty"] = dirty
    # case C: whole driver
    nu = 1 << 16
    series, q = synth.make_series(nu, seed=synth.SEED + 11, pulsar=synth.Pulsar(f_spin=173.7, P_orb=40.0, tau=0.01, Psi0=1.0, summed_power=200.0))
    bank = small_bank(12, synth.SEED + 12)
    bank[1] = (40.0, 0.01, 1.0)
    with tempfile.TemporaryDirectory() as d:
        synth.write_binary_wu(d + "/wu.binary", q)
        synth.write_bank(d + "/bank.txt", bank)
        for tag, padding in (("C1", "1.0"), ("C3", "3.0")):
            r = Reference.run_cli(["-i", d + "/wu.binary", "-o", d + f"/{tag}.cand", "-t", d + "/bank.txt", "-c", d + f"/{tag}.cpt",
                                   "-f", "400", "-A", "0.08", "-P", padding, "-B", "200"])
            assert r.returncode == 0, r.stderr
            out[f"{tag}_candfile"] = np.frombuffer(open(d + f"/{tag}.cand", "rb").read(), np.uint8)
            # the same run with the whitening pre-step (-W) and an EMPTY zap list (no GSL noise draws)
            open(d + "/zap.txt", "w").close()
            r = Reference.run_cli(["-i", d + "/wu.binary", "-o", d + f"/{tag}W.cand", "-t", d + "/bank.txt", "-c", d + f"/{tag}W.cpt",
                                   "-l", d + "/zap.txt", "-f", "400", "-A", "0.08", "-P", padding, "-B", "200", "-W"])
            assert r.returncode == 0, r.stderr
            out[f"{tag}W_candfile"] = np.frombuffer(open(d + f"/{tag}W.cand", "rb").read(), np.uint8)
    out["C_bank"] = bank
    np.savez_compressed(Path(__file__).parent / "golden_v1.npz", **out)
    print("wrote golden_v1.npz with", len(out), "arrays")


if __name__ == "__main__":
    main()
