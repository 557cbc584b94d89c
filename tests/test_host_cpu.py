"""CPU-only: the C-ABI library loads and exports what include/brp_b200.h declares, and its host-side logic
(driver arithmetic around the loop) matches the oracle / the reference-generated golden files."""
import ctypes as C
import re
import tempfile
from pathlib import Path

import numpy as np
import pytest

import brp_b200
from brp_b200 import synth
from helpers import golden_case_C

ROOT = Path(__file__).resolve().parent.parent


def test_library_exports_every_declared_symbol():
    header = (ROOT / "include" / "brp_b200.h").read_text()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"^\s*(?:int|double|void|void \*|uint64_t)\s*\*?\s*([a-z_A-Z0-9]+)\s*\(", header, flags=re.M))
    assert declared, "header parse failed"
    assert declared == set(brp_b200.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(brp_b200.lib, name), f"libbrp_b200.so does not export {name}"


def test_no_cpu_fallback_without_gpu():
    """On a box without a CUDA device every compute entry point must fail loudly, not compute on the host."""
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    g = brp_b200.derive_geometry(1 << 14, synth.TSAMPLE_US, 1.0, 400.0, 200)
    with pytest.raises(brp_b200.BrpError) as e:
        brp_b200.Search(g)
    assert e.value.code == brp_b200.RADPUL_CUDA_DEVICE_FIND
    with pytest.raises(brp_b200.BrpError):
        brp_b200.LegacyBackend()
    with pytest.raises(brp_b200.BrpError):
        brp_b200.stage_sin_lut(np.zeros(4, np.float32))


def test_command_line_driver_argument_handling_and_no_cpu_fallback(tmp_path):
    """brp_b200_search: option checks of demod_binary.c:221-439 and of the wrapper's -i/-o pairing (erp_boinc_wrapper.cpp:412-416)
    happen before any device is touched; with a valid work-unit and no GPU the program must fail with the library's code."""
    import subprocess
    exe = Path(__file__).resolve().parent.parent / "boinc-app-eah-brp_b200" / "brp_b200_search"
    assert exe.exists(), "brp_b200_search not built (python boinc-app-eah-brp_b200/build.py)"

    def run(*args):
        return subprocess.run([str(exe)] + [str(a) for a in args], capture_output=True, text=True, timeout=120)

    assert run("-P", "0.5").returncode == brp_b200.RADPUL_EVAL
    assert run("-f", "-1").returncode == brp_b200.RADPUL_EVAL
    assert run("-A", "2").returncode == brp_b200.RADPUL_EVAL
    assert run("-B", "300000").returncode == brp_b200.RADPUL_EVAL
    assert run("--devices", "1-0").returncode == brp_b200.RADPUL_EVAL
    assert run("-i", "x.dat", "-o", "o", "-t", "t").returncode == brp_b200.RADPUL_EFILE     # extension, demod_binary.c:318
    assert run("-o", "o", "-t", "t").returncode == brp_b200.RADPUL_EMISC                     # no input
    assert run("-q").returncode == brp_b200.RADPUL_EMISC and run("-i").returncode == brp_b200.RADPUL_EMISC
    r = run("-i", "a.binary", "-i", "b.binary", "-o", "x", "-t", "t")
    assert r.returncode == brp_b200.RADPUL_EMISC and "differ" in r.stderr
    r = run("-i", tmp_path / "missing.binary", "-o", tmp_path / "o", "-t", tmp_path / "t")
    assert r.returncode == brp_b200.RADPUL_EIO                                                # demod_binary.c:657-661
    try:
        import torch
        if torch.cuda.is_available():
            return
    except ImportError:
        pass
    _, q = golden_case_C()
    synth.write_binary_wu(tmp_path / "wu.binary", q)
    synth.write_bank(tmp_path / "bank.txt", synth.make_bank(4))
    r = run("-i", tmp_path / "wu.binary", "-o", tmp_path / "out.cand", "-t", tmp_path / "bank.txt", "-f", "400", "-B", "200")
    # (a process exit status keeps the low eight bits of the RADPUL code; the code itself is in the message)
    assert r.returncode == brp_b200.RADPUL_CUDA_DEVICE_FIND % 256 and f"code {brp_b200.RADPUL_CUDA_DEVICE_FIND}" in r.stderr, r.stderr
    assert "no CPU fallback" in r.stderr and not (tmp_path / "out.cand").exists()


@pytest.mark.parametrize("nu,pad,f0,win", [(1 << 22, 1.0, 400.0, 1000), (1 << 22, 3.0, 400.0, 1000), (1 << 22, 2.0, 400.0, 1000),
                                           (1 << 20, 1.0, 400.0, 1000), (1 << 14, 1.0, 400.0, 200), (1 << 16, 3.0, 250.0, 1000),
                                           (1 << 15, 1.0, 5000.0, 100)])
def test_geometry_matches_oracle(oracle, nu, pad, f0, win):
    a = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, pad, f0, win)
    b = oracle.geometry(nu, synth.TSAMPLE_US, pad, f0, win)
    for (name, _) in a._fields_:
        va, vb = getattr(a, name), getattr(b, name)
        assert np.float32(va).tobytes() == np.float32(vb).tobytes() if isinstance(va, float) else va == vb, name


def test_geometry_rejects_nonsense():
    for pad in (0.5, 11.0):
        with pytest.raises(brp_b200.BrpError) as e:
            brp_b200.derive_geometry(1 << 14, synth.TSAMPLE_US, pad, 400.0, 200)
        assert e.value.code == brp_b200.RADPUL_EVAL
    with pytest.raises(brp_b200.BrpError):   # running-median window wider than the spectrum (demod_binary.c:1102)
        brp_b200.derive_geometry(1 << 10, synth.TSAMPLE_US, 1.0, 400.0, 5000)


def test_template_derivation_matches_oracle(oracle):
    bank = synth.make_bank(500)
    g = oracle.geometry(1 << 22, synth.TSAMPLE_US, 1.0, 400.0, 1000)
    for row in bank:
        a = brp_b200.derive_template(*row, g.step_inv)
        b = oracle.template(*row, g.step_inv)
        for (name, _) in a._fields_:
            assert np.float32(getattr(a, name)).tobytes() == np.float32(getattr(b, name)).tobytes(), name


def test_thresholds_match_oracle_and_scipy(oracle):
    from scipy.stats import chi2
    for fft_size in (8193, 2097153, 6291457):
        a, b = brp_b200.base_thresholds(0.08, fft_size), oracle.base_thresholds(0.08, fft_size)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    for n in (1, 2, 4, 8, 16):
        x = brp_b200.lib.brp_chisq_Qinv(3e-8, n)
        assert abs(x - chi2.isf(3e-8, 2 * n)) < 1e-9 * x


def test_bank_reader_follows_reference_eof_rule(tmp_path):
    bank = synth.make_bank(7)
    p = tmp_path / "bank.txt"
    synth.write_bank(p, bank)
    assert np.array_equal(brp_b200.bank_read(p), bank)
    # a last line without '\n' is dropped by the reference's fgets/feof loop (demod_binary.c:1180)
    p.write_text(p.read_text().rstrip("\n"))
    assert np.array_equal(brp_b200.bank_read(p), bank[:-1])
    p.write_text("1000.0 0.0\n")
    with pytest.raises(brp_b200.BrpError) as e:
        brp_b200.bank_read(p)
    assert e.value.code == brp_b200.RADPUL_EVAL
    with pytest.raises(brp_b200.BrpError):
        brp_b200.bank_read(tmp_path / "missing")
    p.write_text("")
    assert brp_b200.bank_read(p).shape == (0, 3)


def test_wu_readers(tmp_path):
    series, q = synth.make_series(1 << 12, seed=5)
    synth.write_binary_wu(tmp_path / "a.binary", q)
    got, ts = brp_b200.wu_read(tmp_path / "a.binary")
    assert ts == synth.TSAMPLE_US and np.array_equal(got.view(np.uint32), series.view(np.uint32))
    nib = np.random.default_rng(1).integers(0, 16, 1 << 12).astype(np.uint8)
    synth.write_bin4_wu(tmp_path / "b.bin4", nib, scale=1.4645e-4)
    got, _ = brp_b200.wu_read(tmp_path / "b.bin4")
    want = (nib.astype(np.float32).astype(np.float64) / 1.4645e-4).astype(np.float32)   # (float)(nibble)/scale, demod_binary.c:835
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    with pytest.raises(brp_b200.BrpError) as e:
        brp_b200.wu_read(tmp_path / "c.dat")
    assert e.value.code == brp_b200.RADPUL_EFILE
    (tmp_path / "short.binary").write_bytes(b"\0" * 100)
    with pytest.raises(brp_b200.BrpError):
        brp_b200.wu_read(tmp_path / "short.binary")


def _random_template_hits(rng, n_templates, fu, w2, thr):
    """Sparse above-threshold per-bin maxima per template, with collisions between templates on a few hot bins."""
    hot = rng.integers(w2, fu, 40)
    per_template = []
    for t in range(n_templates):
        hits = []
        for k in range(5):
            nh = rng.integers(0, 30)
            bins = np.unique(np.concatenate([rng.choice(hot, min(nh, 12)), rng.integers(w2, fu, nh)]))
            power = (thr[k] * (1.0 + rng.random(bins.size) * 2.0)).astype(np.float32)
            if t % 7 == 3 and bins.size:      # exact ties across templates: the earliest template must win
                power[0] = np.float32(thr[k] * 2.5)
            hits.append((bins.astype(np.uint32), power))
        per_template.append(hits)
    return per_template


def test_toplist_from_hits_equals_sequential_reference_rule(oracle):
    """the driver's sequential toplist with threshold feedback (demod_binary.c:1268-1397, restated in the
    oracle and pinned against the reference) equals 'top-100 of per-bin maxima, earliest template on ties'."""
    rng = np.random.default_rng(7)
    nu = 1 << 16
    g_o = oracle.geometry(nu, synth.TSAMPLE_US, 1.0, 400.0, 200)
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, 1.0, 400.0, 200)
    fu, w2 = g.fundamental_idx_hi, g.window_2
    thr = oracle.base_thresholds(0.08, g.fft_size)
    n_templates = 60
    bank = synth.make_bank(n_templates)
    data = _random_template_hits(rng, n_templates, fu, w2, thr)
    # sequential rule through the oracle
    import ctypes
    cands = np.zeros(500, brp_b200.CAND_DTYPE)
    tidx = np.full(500, -1, np.int32)
    nr_pages = g.nr_pages
    for t in range(n_templates):
        ss = np.zeros((5, fu), np.float32)
        dirty = np.zeros((5, nr_pages), np.int32)
        for k in range(5):
            b, p = data[t][k]
            ss[k, b] = p
            dirty[k, b >> 10] = 1
        thrA = np.maximum(thr, cands["power"][99::100].astype(np.float32))
        tp = oracle.template(*bank[t], g.step_inv)
        sp = (ctypes.POINTER(ctypes.c_float) * 5)(*[ss[k].ctypes.data_as(ctypes.POINTER(ctypes.c_float)) for k in range(5)])
        dp = (ctypes.POINTER(ctypes.c_int32) * 5)(*[dirty[k].ctypes.data_as(ctypes.POINTER(ctypes.c_int32)) for k in range(5)])
        oracle.L.brp_oracle_toplist_update(cands.ctypes.data_as(ctypes.c_void_p), tidx.ctypes.data_as(ctypes.c_void_p), sp, dp,
                                           ctypes.c_uint32(nr_pages), ctypes.byref(g_o), thrA.ctypes.data_as(ctypes.c_void_p),
                                           ctypes.byref(tp), ctypes.c_int32(t))
    # per-bin maxima (what the device table holds), earliest template on ties, then the product's host merge
    best = {}
    for t in range(n_templates):
        for k in range(5):
            for b, p in zip(*data[t][k]):
                key = (k, int(b))
                if key not in best or p > best[key][0]:
                    best[key] = (p, t)
    hits = np.zeros(len(best), brp_b200.HIT_DTYPE)
    for i, ((k, b), (p, t)) in enumerate(best.items()):
        hits[i] = (p, b, t, k)
    rng.shuffle(hits)
    # split into two "shards" with duplicate bins to exercise the multi-GPU concatenation rule
    got, got_t = brp_b200.toplist_from_hits(np.concatenate([hits, hits[: len(hits) // 3]]), bank)
    for k in range(5):
        sl = slice(100 * k, 100 * k + 100)
        a = sorted(zip(cands["f0"][sl], cands["power"][sl], tidx[sl], cands["n_harm"][sl]))
        b = sorted(zip(got["f0"][sl], got["power"][sl], got_t[sl], got["n_harm"][sl]))
        assert a == b
        assert np.all(np.diff(got["power"][sl]) <= 0)
    assert np.count_nonzero(cands["power"]) > 300
    for name in ("P_b", "tau", "Psi"):
        m = got_t >= 0
        order_a = np.lexsort((cands["f0"], cands["n_harm"]))
        order_b = np.lexsort((got["f0"], got["n_harm"]))
        assert np.array_equal(cands[name][order_a], got[name][order_b])
    # final statistics and candidate file
    fa, ta = oracle.finalize(cands.view(np.dtype(oracle_dtype())), tidx)
    fb, tb = brp_b200.finalize_candidates(got, got_t)
    txt_a = oracle.format_candidates(fa, g_o, synth.TSAMPLE_US)
    txt_b = brp_b200.format_candidates(fb, g, synth.TSAMPLE_US)
    assert txt_a == txt_b and txt_a.count("\n") == 101


def oracle_dtype():
    from oracle.oracle import CAND_DTYPE
    return CAND_DTYPE


def test_candidate_writer_and_checkpoint_roundtrip(tmp_path, oracle):
    rng = np.random.default_rng(3)
    g = brp_b200.derive_geometry(1 << 16, synth.TSAMPLE_US, 1.0, 400.0, 200)
    bank = synth.make_bank(10)
    hits = np.zeros(50, brp_b200.HIT_DTYPE)
    hits["power"] = rng.uniform(20, 90, 50).astype(np.float32)
    hits["bin"] = rng.choice(np.arange(g.window_2, g.fundamental_idx_hi), 50, replace=False)
    hits["tmpl"] = rng.integers(0, 10, 50)
    hits["harm_idx"] = rng.integers(0, 5, 50)
    cands, tidx = brp_b200.toplist_from_hits(hits, bank)
    brp_b200.checkpoint_write(tmp_path / "cp.cpt", 123, "wu.binary", cands)
    assert (tmp_path / "cp.cpt").stat().st_size == 260 + 500 * 48        # CP_Header + 500 CP_cand (structs.h:111-132)
    n, name, back = brp_b200.checkpoint_read(tmp_path / "cp.cpt")
    assert n == 123 and name == "wu.binary" and back.tobytes() == cands.tobytes()
    fin, _ = brp_b200.finalize_candidates(cands, tidx)
    brp_b200.write_candidates(tmp_path / "out.cand", fin, g, synth.TSAMPLE_US)
    text = (tmp_path / "out.cand").read_text()
    assert text == brp_b200.format_candidates(fin, g, synth.TSAMPLE_US)
    assert text.endswith("%DONE%\n") and not (tmp_path / "out.cand.tmp").exists()
    # at most one line per frequency bin (maximisation over n_harm, demod_binary.c:1654-1661)
    freqs = [line.split()[0] for line in text.splitlines()[:-1]]
    assert len(freqs) == len(set(freqs)) == len(np.unique(hits["bin"]))


def test_synthetic_series_is_reference_compatible(reference, tmp_path):
    """The 8-bit container written by synth is read back identically by the product reader (the reference reads the
    same bytes through demod_binary.c:657-842; covered end to end in test_oracle.py)."""
    series, q = golden_case_C()
    synth.write_binary_wu(tmp_path / "wu.binary", q)
    got, _ = brp_b200.wu_read(tmp_path / "wu.binary")
    assert np.array_equal(got, series)
    assert abs(float(series.mean())) < 5e-3 and abs(float(series.std()) - 1.0) < 0.02


def test_toplist_merge_resumes_exactly():
    """Checkpoint resume in the batched driver (brp_b200_search): candidates of the first part of the bank merged with
    the per-bin maxima of the rest must equal the candidates of the whole bank searched in one go (the reference keeps
    one running toplist through demod_binary.c:1310-1397, checkpointed at :1423-1427)."""
    rng = np.random.default_rng(5)
    n_bins, n_t = 4000, 60
    bank = np.column_stack([rng.uniform(600, 9000, n_t), rng.uniform(0, 1, n_t), rng.uniform(0, 6.28, n_t)])
    for split in (1, 17, 59):
        parts = []
        for lo, hi in ((0, split), (split, n_t)):
            hs = []
            for k in range(5):
                nb = 700 if k < 4 else 40                     # fewer than 100 hits for the last harmonic
                b = rng.choice(np.arange(100, n_bins), nb, replace=False)
                h = np.zeros(nb, brp_b200.HIT_DTYPE)
                h["bin"], h["harm_idx"] = b, k
                h["power"] = np.round(rng.gamma(3.0, 4.0, nb), 1).astype(np.float32) + 1   # coarse values: plenty of ties
                h["tmpl"] = rng.integers(lo, hi, nb)
                hs.append(h)
            parts.append(np.concatenate(hs))
        ca, _ = brp_b200.toplist_from_hits(parts[0], bank)
        cb, _ = brp_b200.toplist_from_hits(parts[1], bank)
        whole, _ = brp_b200.toplist_from_hits(np.concatenate(parts), bank)
        merged = brp_b200.toplist_merge(ca, cb)
        for k in range(5):
            m, w = merged[100 * k:100 * k + 100], whole[100 * k:100 * k + 100]
            assert np.array_equal(m["power"], w["power"])
            # entries strictly above the rank-100 power are the same candidates (ties AT the boundary may pick another bin)
            strict = w["power"] > w["power"][-1]
            key = lambda c: sorted(zip(c["power"].tolist(), c["f0"].tolist(), c["P_b"].tolist(), c["tau"].tolist(), c["Psi"].tolist()))
            assert key(m[strict]) == key(w[strict])
