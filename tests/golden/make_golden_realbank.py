"""Golden candidate files from the UNMODIFIED reference driver (oracle/_ref/demod_binary_ref = demod_binary.c MAIN compiled in
place) on the reference's own fixtures: the bundled template bank `stochastic_full.bank` (6 662 lines), the bundled real
work-unit `..._1099.bin4` and its zap list (debian/extra/einstein_bench/testwu/).

Run in the build container (where /root/reference exists):

    python tests/golden/make_golden_realbank.py [--full]        # --full adds the two 6 662-template runs (about an hour)

It copies the three DATA fixtures into tests/golden/realwu/ (the GPU box has no /root/reference) and writes
tests/golden/golden_realbank.npz:

  S1_64 / S3_64   synthetic 2^22-sample work-unit (the benchmark's series; for -P 3 with a three times stronger pulsar, as
                  mean padding dilutes the normalised power), first 64 bank lines, -P 1 / -P 3, -f 400 -A 0.08
                  (BASELINE.json config 1 as worded)
  R3W_64          the real work-unit with the reference's own bench options (bench_single.sh:28: -A 0.08 -P 3.0 -f 400.0 -W
                  -l <zap>), first 64 bank lines
  S1_full, R3W_full   the same two with all 6 662 lines (BASELINE.json config 2; only with --full)

Per case the candidate file and the final checkpoint (CP_Header + candidates_all[500] right after the loop) are stored.
The zap-list noise behind R3W_* is drawn by oracle/shim_gsl.c (GSL itself is not installable here; parity with libgsl is
unpinned).
"""
import shutil
import subprocess
import sys
import tempfile
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from brp_b200_synth import synth  # noqa: E402  (loaded by path: no CUDA library needed here)
from oracle.oracle import REF_BIN  # noqa: E402

REFWU = Path("/root/reference/debian/extra/einstein_bench/testwu")
BANK = REFWU / "stochastic_full.bank"
WU = REFWU / "p2030.20151015.G187.41-00.88.N.b2s0g0.00000_1099.bin4"
ZAP = REFWU / "p2030.20151015.G187.41-00.88.N.b2s0g0.00000.zap"
HERE = Path(__file__).resolve().parent
DEST = HERE / "realwu"


def run_ref(tag, wu, bank, extra, workdir):
    out = Path(workdir) / f"{tag}.cand"
    cmd = [str(REF_BIN), "-i", str(wu), "-o", str(out), "-t", str(bank), "-c", str(Path(workdir) / f"{tag}.cpt"), "-A", "0.08", "-f", "400.0"] + extra
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, (tag, r.stderr[-2000:])
    # the final checkpoint (demod_binary.c:1498) holds candidates_all[500] as they are right after the template loop
    return tag, np.frombuffer(out.read_bytes(), np.uint8), np.frombuffer((Path(workdir) / f"{tag}.cpt").read_bytes(), np.uint8)


def main():
    full = "--full" in sys.argv
    DEST.mkdir(exist_ok=True)
    for src, name in ((BANK, "stochastic_full.bank"), (WU, "real_wu_1099.bin4"), (ZAP, "real_wu.zap")):
        shutil.copyfile(src, DEST / name)
    with tempfile.TemporaryDirectory() as d:
        _, q = synth.make_series(1 << 22)
        synth.write_binary_wu(d + "/synth.binary", q)
        # mean padding dilutes the normalised power by 1/padding: the -P 3 case gets a proportionally stronger pulsar
        _, q3 = synth.make_series(1 << 22, pulsar=synth.Pulsar(summed_power=16.0 + 104.0 * 3.0))
        synth.write_binary_wu(d + "/synth3.binary", q3)
        lines = BANK.read_text().splitlines(keepends=True)
        (Path(d) / "bank64.txt").write_text("".join(lines[:64]))
        jobs = [
            ("S1_64", d + "/synth.binary", d + "/bank64.txt", ["-P", "1.0"]),
            ("S3_64", d + "/synth3.binary", d + "/bank64.txt", ["-P", "3.0"]),
            ("R3W_64", WU, d + "/bank64.txt", ["-P", "3.0", "-W", "-l", str(ZAP)]),
        ]
        if full:
            jobs += [
                ("S1_full", d + "/synth.binary", BANK, ["-P", "1.0"]),
                ("R3W_full", WU, BANK, ["-P", "3.0", "-W", "-l", str(ZAP)]),
            ]
        out = {}
        path = HERE / "golden_realbank.npz"
        if path.exists():  # keep what an earlier (--full) run produced
            out.update(np.load(path))
        with ThreadPoolExecutor(len(jobs)) as pool:
            for tag, cand, cpt in pool.map(lambda j: run_ref(j[0], j[1], j[2], j[3], d), jobs):
                out[f"{tag}_candfile"] = cand
                out[f"{tag}_checkpoint"] = cpt
                print(tag, len(cand), "bytes", flush=True)
                np.savez_compressed(path, **out)
    print("wrote", path, sorted(out))


if __name__ == "__main__":
    main()
