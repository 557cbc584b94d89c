#!/usr/bin/env python
"""Static SASS instruction counts per kernel of libbrp_b200.so (cuobjdump -sass): which kernels carry TMA tensor copies
(UTMALDG / UTMASTG / UTMAPF), bulk copies (UBLKCP), mbarrier operations (SYNCS), cp.async (LDGSTS), packed FP32
(FFMA2 / FADD2 / FMUL2) -- and that none carries tensor-core instructions (UTCMMA / HMMA): no stage is a contraction.

    python tools/sass_summary.py > profiles/r2_sass_summary.txt
"""
import collections
import re
import subprocess
import sys
from pathlib import Path

lib = Path(__file__).resolve().parent.parent / "boinc-app-eah-brp_b200" / "libbrp_b200.so"
out = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
pats = ["UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "SYNCS", "LDGSTS", "FFMA2", "FADD2", "FMUL2", "UTCMMA", "HMMA", "ATOMG", "LDS", "STS", "LDG", "STG",
        "BAR", "MUFU", "SHFL"]
counts, cur = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m:
        op = m.group(1)
        counts[cur]["total"] += 1
        for p in pats:
            if op.startswith(p):
                counts[cur][p] += 1
names = subprocess.run(["c++filt"], input="\n".join(counts), capture_output=True, text=True).stdout.splitlines()
print("Static SASS instruction counts per kernel of libbrp_b200.so (cuobjdump -sass, sm_100a); tools/sass_summary.py")
print("kernel | total | " + " | ".join(pats))
for (k, c), name in zip(counts.items(), names):
    name = re.sub(r"\(anonymous namespace\)::", "", name)
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*", "", name)
    print(f"{name} | {c['total']} | " + " | ".join(str(c[p]) for p in pats))
