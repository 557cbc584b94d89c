"""brp_b200_search: the command-line driver over the batched API (row N3 of SURVEY.md 8f: work-unit ingest, chunked
template loop, checkpoint/resume, candidate file) against the reference's own candidate files (tests/golden/)."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

import brp_b200
from brp_b200 import synth
from helpers import golden_case_C

pytestmark = pytest.mark.gpu

EXE = Path(brp_b200.__file__).resolve().parent / "brp_b200_search"
if not EXE.exists():  # brp_b200.py is an alias module at the repo root
    EXE = Path(__file__).resolve().parent.parent / "boinc-app-eah-brp_b200" / "brp_b200_search"


def run_cli(tmp, *extra, out="out.cand", cpt="cp.cpt", padding="1.0", check=True):
    cmd = [str(EXE), "-i", str(tmp / "wu.binary"), "-o", str(tmp / out), "-t", str(tmp / "bank.txt"), "-f", "400", "-A", "0.08",
           "-P", padding, "-B", "200", "-D", "0", "-z"]
    if cpt:
        cmd += ["-c", str(tmp / cpt)]
    r = subprocess.run(cmd + list(extra), capture_output=True, text=True, timeout=600)
    if check:
        assert r.returncode == 0, r.stderr[-2000:]
    return r


@pytest.fixture()
def wu(tmp_path, golden):
    assert EXE.exists(), "brp_b200_search not built (python boinc-app-eah-brp_b200/build.py)"
    _, q = golden_case_C()
    synth.write_binary_wu(tmp_path / "wu.binary", q)
    synth.write_bank(tmp_path / "bank.txt", golden["C_bank"])
    return tmp_path


def assert_same_candidates(got, ref):
    assert len(got) == len(ref) and got[-1] == "%DONE%" and len(ref) > 5
    for a, b in zip(got[:-1], ref[:-1]):
        fa, fb = a.split(), b.split()
        assert fa[:4] == fb[:4] and fa[6] == fb[6], (a, b)
        assert abs(float(fa[4]) - float(fb[4])) <= 2e-4 * float(fb[4])           # power: the FFT tolerance of SURVEY.md 8c
        assert abs(float(fa[5]) - float(fb[5])) <= 2e-3 * max(float(fb[5]), 1.0)  # -log10(fA)


@pytest.mark.parametrize("tag,padding", [("C1", "1.0"), ("C3", "3.0")])
def test_cli_reproduces_reference_candidate_file(wu, golden, tag, padding):
    run_cli(wu, padding=padding)
    got = (wu / "out.cand").read_text().splitlines()
    assert_same_candidates(got, bytes(golden[f"{tag}_candfile"]).decode().splitlines())
    n, name, cands = brp_b200.checkpoint_read(wu / "cp.cpt")          # final checkpoint, demod_binary.c:1498
    assert n == len(golden["C_bank"]) and name.endswith("wu.binary") and np.count_nonzero(cands["power"]) > 5


@pytest.mark.parametrize("stops", [(5,), (1, 4, 3)])
def test_cli_interrupted_and_resumed_gives_the_same_file(wu, golden, stops):
    run_cli(wu, out="whole.cand", cpt=None)
    done = 0
    for n in stops:                                # each invocation searches n more templates, then stops with a checkpoint
        r = run_cli(wu, "--max_templates", str(n), "--chunk", "2", "--checkpoint_period", "0")
        done += n
        assert not (wu / "out.cand").exists()
        assert brp_b200.checkpoint_read(wu / "cp.cpt")[0] == done, r.stderr
    r = run_cli(wu, "--chunk", "3", "--checkpoint_period", "0")
    assert f"at template no. {done}" in r.stderr
    assert (wu / "out.cand").read_bytes() == (wu / "whole.cand").read_bytes()


def test_cli_with_whitening(wu, golden):
    (wu / "zap.txt").write_text("")
    run_cli(wu, "-W", "-l", str(wu / "zap.txt"))
    got = (wu / "out.cand").read_text().splitlines()
    ref = bytes(golden["C1W_candfile"]).decode().splitlines()
    key = lambda line: tuple(line.split()[:4] + [line.split()[6]])
    assert got[-1] == "%DONE%" and [key(l) for l in got[:10]] == [key(l) for l in ref[:10]]
    assert len({key(l) for l in got[:-1]} & {key(l) for l in ref[:-1]}) >= 0.9 * (len(ref) - 1)


@pytest.mark.parametrize("tag,padding", [("CZ1", "1.0"), ("CZ3", "3.0")])
def test_cli_with_whitening_and_zap_list(wu, tag, padding):
    """-W -l <non-empty zap list>: the reference's production form of the command (bench_single.sh:28)."""
    gz = np.load(Path(__file__).resolve().parent / "golden" / "golden_zap.npz", allow_pickle=False)
    (wu / "zap.txt").write_bytes(bytes(gz["zap_text"]))
    run_cli(wu, "-W", "-l", str(wu / "zap.txt"), padding=padding)
    got = (wu / "out.cand").read_text().splitlines()
    ref = bytes(gz[f"{tag}_candfile"]).decode().splitlines()
    key = lambda line: tuple(line.split()[:4] + [line.split()[6]])
    assert got[-1] == "%DONE%" and [key(l) for l in got[:10]] == [key(l) for l in ref[:10]]
    assert len({key(l) for l in got[:-1]} & {key(l) for l in ref[:-1]}) >= 0.9 * (len(ref) - 1)


def test_cli_bank_split_over_several_devices_gives_the_same_file(wu):
    """--devices: one handle per listed device, contiguous pieces of every round of the bank, host merge of the hit lists.
    The candidate file and the final checkpoint must be byte-identical to the one-device run (an ordinal may repeat, so the
    split is exercised on a one-GPU box too; with more GPUs visible they are all used)."""
    import torch
    run_cli(wu, out="one.cand", cpt="one.cpt")
    lists = ["0,0,0"]
    n = torch.cuda.device_count()
    if n >= 2:
        lists.append(f"0-{n - 1}")
    for k, devs in enumerate(lists):
        r = run_cli(wu, "--devices", devs, "--chunk", "2", out=f"many{k}.cand", cpt=f"many{k}.cpt")
        assert f"on {len(devs.split(',')) if ',' in devs else n} device(s)" in r.stderr
        assert (wu / f"many{k}.cand").read_bytes() == (wu / "one.cand").read_bytes()
        a, b = brp_b200.checkpoint_read(wu / f"many{k}.cpt"), brp_b200.checkpoint_read(wu / "one.cpt")
        assert a[0] == b[0] and a[2].tobytes() == b[2].tobytes()
    r = run_cli(wu, "--devices", "0,x", cpt=None, check=False)
    assert r.returncode == brp_b200.RADPUL_EVAL


def test_cli_several_work_units_in_one_invocation(wu):
    """The passes of the reference's BOINC wrapper (erp_boinc_wrapper.cpp:430-474): -i/-o pairs are searched in turn, a pass
    whose output exists is skipped, the checkpoint file is removed after each completed pass; with --devices the work-units
    are dealt out over the devices (one each at a time, own checkpoint files) and give the same bytes."""
    rng = np.random.default_rng(5)
    base = np.frombuffer((wu / "wu.binary").read_bytes(), np.uint8)
    names = ["wu.binary"]
    for k in (1, 2):                                # two more work-units: the same header, other noise in the samples
        other = base.copy()
        samples = other[1168:].view(np.int8).astype(np.int16) + rng.integers(-3, 4, len(other) - 1168)
        other[1168:] = np.clip(samples, -128, 127).astype(np.int8).view(np.uint8)
        names.append(f"wu{k}.binary")
        (wu / names[-1]).write_bytes(other.tobytes())
    (wu / "zap.txt").write_text("50.0 50.5\n")       # whitening with a zap range too: concurrent passes each whiten their own series
    common = ["-t", str(wu / "bank.txt"), "-f", "400", "-A", "0.08", "-P", "1.0", "-B", "200", "-z", "-W", "-l", str(wu / "zap.txt")]
    for k, name in enumerate(names):                # one invocation each: the expected files
        r = subprocess.run([str(EXE), "-i", str(wu / name), "-o", str(wu / f"single{k}.cand")] + common, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
    assert (wu / "single0.cand").read_bytes() != (wu / "single1.cand").read_bytes()

    def pairs(prefix):
        out = []
        for k, name in enumerate(names):
            out += ["-i", str(wu / name), "-o", str(wu / f"{prefix}{k}.cand")]
        return out

    (wu / "seq1.cand").write_text("already there\n")                                   # pass 1 must be skipped
    r = subprocess.run([str(EXE)] + pairs("seq") + common + ["-c", str(wu / "seq.cpt")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "seq1.cand' already exists - skipping pass" in r.stderr
    assert (wu / "seq1.cand").read_text() == "already there\n"
    for k in (0, 2):
        assert (wu / f"seq{k}.cand").read_bytes() == (wu / f"single{k}.cand").read_bytes()
    assert not (wu / "seq.cpt").exists()                                               # removed after the last pass too
    import torch
    n = torch.cuda.device_count()
    r = subprocess.run([str(EXE)] + pairs("par") + common + ["-c", str(wu / "par.cpt"), "--devices", f"0-{n - 1}" if n > 1 else "0,0"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    for k in range(3):
        assert (wu / f"par{k}.cand").read_bytes() == (wu / f"single{k}.cand").read_bytes()
        assert not (wu / f"par.cpt.{k}").exists()
    r = subprocess.run([str(EXE), "-i", str(wu / "wu.binary")] + pairs("bad") + common, capture_output=True, text=True, timeout=60)
    assert r.returncode == brp_b200.RADPUL_EMISC and "differ" in r.stderr                # erp_boinc_wrapper.cpp:412-416


def test_cli_error_paths(wu):
    run_cli(wu, "--max_templates", "2", "--checkpoint_period", "0")
    (wu / "other.binary").write_bytes((wu / "wu.binary").read_bytes())
    cmd = [str(EXE), "-i", str(wu / "other.binary"), "-o", str(wu / "o"), "-t", str(wu / "bank.txt"), "-c", str(wu / "cp.cpt")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
    assert r.returncode == brp_b200.RADPUL_EVAL and "doesn't agree" in r.stderr       # demod_binary.c:589-593
    r = run_cli(wu, "-W", cpt=None, check=False)                                        # -W opens the zap list unconditionally
    assert r.returncode == brp_b200.RADPUL_EFILE and "zaplist" in r.stderr              # demod_binary.c:982-989
    (wu / "zap.txt").write_text("60.0\n")
    r = run_cli(wu, "-W", "-l", str(wu / "zap.txt"), cpt=None, check=False)            # incomplete line, :1003-1007
    assert r.returncode == brp_b200.RADPUL_EIO
    r = subprocess.run([str(EXE), "-i", "x.dat", "-o", "o", "-t", "t"], capture_output=True, text=True)
    assert r.returncode == brp_b200.RADPUL_EFILE
    r = subprocess.run([str(EXE), "-P", "0.5"], capture_output=True, text=True)
    assert r.returncode == brp_b200.RADPUL_EVAL


def test_cli_full_size_production_options_equal_the_api_path(tmp_path):
    """A 2^22-sample work-unit with the production options (-P 3.0 -W, tensor-copy FFT kernels, sub-transform pass 1): the
    driver's candidate file must be the one the same library calls give when made one by one from Python."""
    nu = 1 << 22
    series, q = synth.make_series(nu, seed=31, pulsar=synth.Pulsar(summed_power=400.0))
    bank = synth.make_bank(96, seed=32)
    synth.write_binary_wu(tmp_path / "wu.binary", q)
    synth.write_bank(tmp_path / "bank.txt", bank)
    (tmp_path / "zap.txt").write_text("")
    r = subprocess.run([str(EXE), "-i", str(tmp_path / "wu.binary"), "-o", str(tmp_path / "out.cand"), "-t", str(tmp_path / "bank.txt"),
                        "-l", str(tmp_path / "zap.txt"), "-f", "400", "-A", "0.08", "-P", "3.0", "-W", "-B", "1000", "-D", "0"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    got = (tmp_path / "out.cand").read_text()
    wu_series, ts = brp_b200.wu_read(tmp_path / "wu.binary")
    assert wu_series.size == nu
    g = brp_b200.derive_geometry(nu, ts, 3.0, 400.0, 1000)
    white, _ = brp_b200.whiten_series(wu_series, 3.0, 1000)
    cands, tidx = brp_b200.search_bank(white, g, synth.read_bank(tmp_path / "bank.txt"), fA=0.08, max_batch=64)
    fin, _ = brp_b200.finalize_candidates(cands, tidx)
    want = brp_b200.format_candidates(fin, g, ts)
    assert got == want and got.endswith("%DONE%\n") and got.count("\n") > 5
