"""Golden vectors for the zap-list noise of the whitening pre-step (demod_binary.c:982-1033), from the REFERENCE driver
(oracle/_ref/demod_binary_ref, demod_binary.c compiled in place) and the GSL stand-in it is linked against
(oracle/shim_gsl.c: taus2 + ziggurat restated, tables from tools/gen_ziggurat.py; libgsl itself is not installable, so
parity with it is unpinned).  Run in the build container:   python tests/golden/make_golden_zap.py

  draws        first 20 000 gsl_ran_gaussian_ziggurat(r, 1.0) values after gsl_rng_set(r, SEED)  (float64)
  CZ1 / CZ3    case C of make_golden.py (2^16-sample work-unit, 12 templates) through the reference's own command line with
               -W and the NON-EMPTY zap list below, -P 1.0 / -P 3.0: candidate files
"""
import ctypes as C
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
from brp_b200_synth import synth  # noqa: E402
from oracle.oracle import REF_SO, Reference  # noqa: E402

ZAP_TEXT = "59.9 60.1\n119.9 120.3\n120.2 120.6\n300.0 300.05\n1.0 1.2\n"   # overlapping ranges, one inside the zeroed edge
SEED = 123456789


def small_bank(n, seed):
    bank = synth.make_bank(n, seed=seed)
    bank[2:, 0] = np.round(np.linspace(8.0, 90.0, n - 2), 12)
    bank[2:, 1] = np.round(bank[2:, 1] * 0.05, 12)
    return bank


def main():
    out = {"zap_text": np.frombuffer(ZAP_TEXT.encode(), np.uint8), "seed": np.int64(SEED)}
    L = C.CDLL(str(REF_SO))
    L.gsl_rng_alloc.restype = C.c_void_p
    L.gsl_rng_alloc.argtypes = [C.c_void_p]
    L.gsl_rng_set.argtypes = [C.c_void_p, C.c_ulong]
    L.gsl_ran_gaussian_ziggurat.restype = C.c_double
    L.gsl_ran_gaussian_ziggurat.argtypes = [C.c_void_p, C.c_double]
    r = L.gsl_rng_alloc(None)
    L.gsl_rng_set(r, SEED)
    out["draws"] = np.array([L.gsl_ran_gaussian_ziggurat(r, 1.0) for _ in range(20000)], np.float64)

    series, q = synth.make_series(1 << 16, seed=synth.SEED + 11, pulsar=synth.Pulsar(f_spin=173.7, P_orb=40.0, tau=0.01, Psi0=1.0, summed_power=200.0))
    bank = small_bank(12, synth.SEED + 12)
    bank[1] = (40.0, 0.01, 1.0)
    with tempfile.TemporaryDirectory() as d:
        synth.write_binary_wu(d + "/wu.binary", q)
        synth.write_bank(d + "/bank.txt", bank)
        Path(d, "zap.txt").write_text(ZAP_TEXT)
        for tag, padding in (("CZ1", "1.0"), ("CZ3", "3.0")):
            rr = Reference.run_cli(["-i", d + "/wu.binary", "-o", d + f"/{tag}.cand", "-t", d + "/bank.txt", "-c", d + f"/{tag}.cpt",
                                    "-l", d + "/zap.txt", "-f", "400", "-A", "0.08", "-P", padding, "-B", "200", "-W"])
            assert rr.returncode == 0, rr.stderr
            out[f"{tag}_candfile"] = np.frombuffer(open(d + f"/{tag}.cand", "rb").read(), np.uint8)
    np.savez_compressed(Path(__file__).parent / "golden_zap.npz", **out)
    print("wrote golden_zap.npz", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
