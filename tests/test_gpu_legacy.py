"""GPU parity through the reference's own backend interface (the drop-in boundary): the eleven legacy
functions driven with the reference's call sequence and fed-back thresholds (demod_binary.c:1268-1293), the driver's
candidate scan (:1310-1397) supplied by the oracle's restatement of it."""
import ctypes

import numpy as np
import pytest

import brp_b200
from brp_b200 import synth
from helpers import golden_case_C, power_close

pytestmark = pytest.mark.gpu


def _drive(oracle, series, g, go, bank, fA):
    be = brp_b200.LegacyBackend()
    be.set_up(series, g)
    from oracle.oracle import CAND_DTYPE
    cands = np.zeros(500, CAND_DTYPE)
    tidx = np.full(500, -1, np.int32)
    base = brp_b200.base_thresholds(fA, g.fft_size)
    taps = []
    for it, row in enumerate(bank):
        t = brp_b200.derive_template(*row, g.step_inv)
        thrA = np.maximum(base, cands["power"][99::100].astype(np.float32))      # demod_binary.c:1276-1281
        ss, dd = be.run_template(t, thrA)
        taps.append(([a.copy() for a in ss], [d.copy() for d in dd], thrA.copy()))
        sp = (ctypes.POINTER(ctypes.c_float) * 5)(*[ss[k].ctypes.data_as(ctypes.POINTER(ctypes.c_float)) for k in range(5)])
        dp = (ctypes.POINTER(ctypes.c_int32) * 5)(*[dd[k].ctypes.data_as(ctypes.POINTER(ctypes.c_int32)) for k in range(5)])
        to = oracle.template(*row, go.step_inv)
        oracle.L.brp_oracle_toplist_update(cands.ctypes.data_as(ctypes.c_void_p), tidx.ctypes.data_as(ctypes.c_void_p), sp, dp,
                                           ctypes.c_uint32(g.nr_pages), ctypes.byref(go), thrA.ctypes.data_as(ctypes.c_void_p),
                                           ctypes.byref(to), ctypes.c_int32(it))
    be.shutdown()
    return cands, tidx, taps


@pytest.mark.parametrize("tag,padding", [("C1", 1.0), ("C3", 3.0)])
def test_legacy_interface_reproduces_reference_candidates(oracle, golden, tag, padding):
    series, _ = golden_case_C()
    bank = golden["C_bank"]
    g = brp_b200.derive_geometry(1 << 16, synth.TSAMPLE_US, padding, 400.0, 200)
    go = oracle.geometry(1 << 16, synth.TSAMPLE_US, padding, 400.0, 200)
    cands, tidx, taps = _drive(oracle, series, g, go, bank, 0.08)
    want, want_t = oracle.search(series, go, 0.08, bank)
    for k in range(5):
        sl = slice(100 * k, 100 * k + 100)
        a = sorted(zip(tidx[sl], cands["f0"][sl], cands["n_harm"][sl]))
        b = sorted(zip(want_t[sl], want["f0"][sl], want["n_harm"][sl]))
        assert a == b
        pa = dict(zip(zip(tidx[sl], cands["f0"][sl]), cands["power"][sl]))
        pb = dict(zip(zip(want_t[sl], want["f0"][sl]), want["power"][sl]))
        for key, p in pb.items():
            assert abs(pa[key] - p) <= 1e-4 * max(p, 1e-30)
    # same batched result
    bc, bt = brp_b200.search_bank(series, g, bank, fA=0.08, max_batch=4)
    assert bc.tobytes() == cands.view(brp_b200.CAND_DTYPE).tobytes() and np.array_equal(bt, tidx)
    # what the driver's candidate scan reads (demod_binary.c:1310-1350): sumspec[0..4] above threshold and the page flags, for
    # several templates in a row (the arrays are cleared entry by entry between templates, not wholesale)
    for it in (0, len(bank) // 2, len(bank) - 1):
        ss, dd, thrA = taps[it]
        t = oracle.template(*bank[it], go.step_inv)
        res, _, _, _ = oracle.resample(series, go, t)
        ps = oracle.power_spectrum(res, go.nsamples)
        rss, rdirty = oracle.harmonic_summing(ps, go, thrA)
        for k in range(5):
            ref = ps[: g.fundamental_idx_hi].copy() if k == 0 else rss[k - 1]
            if k == 0:
                ref[: g.window_2] = 0                                   # the summing loop starts at window_2 (hs_common.c:74,79)
            above = np.nonzero(ref > thrA[k] * np.float32(1.001))[0]
            assert np.all(dd[k][above >> 10] == 1)
            e, _ = power_close(ss[k][above], ref[above])
            assert e <= 1e-4
            # nothing else may look like a candidate: every other entry is at or below the threshold
            rest = np.ones(g.fundamental_idx_hi, bool)
            rest[np.nonzero(ref > thrA[k] * np.float32(0.999))[0]] = False
            assert np.all(ss[k][rest] <= thrA[k])


def test_legacy_error_codes():
    be = brp_b200.LegacyBackend()
    series, _ = synth.make_series(5 << 10, seed=3)
    g = brp_b200.derive_geometry(5 << 10, synth.TSAMPLE_US, 1.0, 400.0, 20)
    with pytest.raises(brp_b200.BrpError) as e:       # FFT length 5*2^10: only 2^a and 3*2^a are supported
        be.set_up(series, g)
    assert e.value.code == brp_b200.RADPUL_CUDA_FFT_PLAN
    be.shutdown()


@pytest.mark.parametrize("tag,padding", [("C1", "1.0"), ("C3", "3.0")])
def test_unmodified_reference_driver_linked_against_this_backend(golden, tmp_path, tag, padding):
    """The real drop-in proof: oracle/_ref/demod_binary_b200 is the reference's own demod_binary.c,
    compiled unmodified with -DUSE_CUDA and linked against libbrp_b200.so instead of the reference's CUDA backend and
    cuFFT (oracle/Makefile).  Same command line, same work-unit and bank files as the golden CPU run of the reference:
    the candidate file must agree line by line (identical frequency, template and n_harm; power to the tolerance)."""
    import subprocess
    from pathlib import Path
    exe = Path(__file__).resolve().parent.parent / "oracle" / "_ref" / "demod_binary_b200"
    if not exe.exists():
        pytest.skip("oracle/_ref/demod_binary_b200 not built (needs /root/reference at build time)")
    _, q = golden_case_C()
    synth.write_binary_wu(tmp_path / "wu.binary", q)
    synth.write_bank(tmp_path / "bank.txt", golden["C_bank"])
    r = subprocess.run([str(exe), "-i", str(tmp_path / "wu.binary"), "-o", str(tmp_path / "out.cand"), "-t", str(tmp_path / "bank.txt"),
                        "-c", str(tmp_path / "cp.cpt"), "-f", "400", "-A", "0.08", "-P", padding, "-B", "200", "-D", "0"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    got = (tmp_path / "out.cand").read_text().splitlines()
    ref = bytes(golden[f"{tag}_candfile"]).decode().splitlines()
    assert len(got) == len(ref) and got[-1] == "%DONE%" and len(ref) > 5
    for a, b in zip(got[:-1], ref[:-1]):
        fa, fb = a.split(), b.split()
        assert fa[:4] == fb[:4] and fa[6] == fb[6], (a, b)
        assert abs(float(fa[4]) - float(fb[4])) <= 2e-4 * float(fb[4])
        assert abs(float(fa[5]) - float(fb[5])) <= 2e-3 * max(float(fb[5]), 1.0)
    # the reference writes its checkpoint through the same driver code (demod_binary.c:1743-1783); our reader parses it
    n, name, cands = brp_b200.checkpoint_read(tmp_path / "cp.cpt")
    assert n == len(golden["C_bank"]) and name.endswith("wu.binary") and np.count_nonzero(cands["power"]) > 5
