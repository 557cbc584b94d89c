#!/usr/bin/env python
"""Run bench.py (extra arguments are passed on) and print one short line: rates, candidates, per-stage microseconds
per launch.  For A/B runs on the GPU box, e.g.  BRP_FFT_TMA=0 python tools/stage_times.py --steps 2 --warmup 2"""
import json
import subprocess
import sys
from pathlib import Path

root = Path(__file__).resolve().parent.parent
out = subprocess.run([sys.executable, str(root / "bench.py"), "--no-cpu-baseline", "--no-other-configs"] + sys.argv[1:], capture_output=True, text=True)
lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
if not lines:
    print("bench failed:", out.stderr[-1500:])
    sys.exit(1)
d = json.loads(lines[-1])
st = d["roofline"]["stages"]
front = st["resample_fft_power"]
print(f"value {d['value']:.0f}  e2e {d['e2e']['value']:.0f}  batch {d['config']['batch']}  cands {d['config']['candidates_found']}  "
      f"resample {front['resample_ms'] * 1e3:.1f}  pass1 {front['pass1_ms'] * 1e3:.1f}  "
      f"pass2 {front['pass2_ms'] * 1e3:.1f}  hs {st['harmonic_sum']['ms_per_batch'] * 1e3:.1f} us")
