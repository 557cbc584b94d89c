#!/usr/bin/env python
"""Split the SASS of one kernel of an ncu report (--page source --csv) into phases at its block barriers and print, per
phase, the share of warp-stall samples, the instructions executed and the dominant stall reasons.

    ncu -i prof.ncu-rep --page source --csv --kernel-name regex:fft_pass1 > /tmp/k.csv; python tools/ncu_phases.py /tmp/k.csv
"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
body = [r for r in rows[2:] if len(r) == len(hdr) and r[0].startswith("0x")]
total = sum(int(r[ix["# Samples"]] or 0) for r in body)


def new():
    return {"samples": 0, "inst": 0, "stall": {s: 0 for s in stalls}, "wf": 0, "wf_ideal": 0}


acc, phases = new(), []
for r in body:
    sass = r[ix["Source"]].strip()
    acc["samples"] += int(r[ix["# Samples"]] or 0)
    acc["inst"] += int(r[ix["Instructions Executed"]] or 0)
    acc["wf"] += int(r[ix["L1 Wavefronts Shared"]] or 0)
    acc["wf_ideal"] += int(r[ix["L1 Wavefronts Shared Ideal"]] or 0)
    for s in stalls:
        acc["stall"][s] += int(r[ix[s]] or 0)
    if "BAR.SYNC" in sass or "SYNCS.PHASECHK" in sass:
        acc["end"] = sass
        phases.append(acc)
        acc = new()
acc["end"] = "(exit)"
phases.append(acc)
print(f"total samples {total}")
for i, p in enumerate(phases):
    top = sorted(p["stall"].items(), key=lambda kv: -kv[1])[:4]
    print(f"phase {i:2d}: {100.0 * p['samples'] / max(total, 1):5.1f}% of samples, {p['inst']:>10d} warp-instr, smem wavefronts {p['wf']} (ideal {p['wf_ideal']}), "
          f"ends at [{p['end'][:40]}]  " + ", ".join(f"{k[6:]} {100.0 * v / max(p['samples'], 1):.0f}%" for k, v in top))
