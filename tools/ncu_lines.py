#!/usr/bin/env python
"""Top source lines by warp-stall samples from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`.

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [N]
"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur_file, hdr, lines = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0].strip().isdigit() and len(r) >= len(hdr):
        nm = len(hdr) - 2  # unescaped quotes/commas in source text can split it: metrics are the LAST nm columns
        d = dict(zip(hdr[2:], r[-nm:]))
        try:
            samples = int(d.get("# Samples", "0") or 0)
        except ValueError:
            continue
        lines.append((samples, cur_file, int(r[0]), ",".join(r[1:len(r) - nm]).strip(), d))
total = sum(l[0] for l in lines) or 1
lines.sort(key=lambda l: -l[0])
stall_cols = [c for c in (hdr or []) if c.startswith("stall_")]
print(f"total samples {total}")
for s, f, ln, src, d in lines[:top]:
    def _i(x):
        try:
            return int(x or 0)
        except ValueError:
            return 0
    stalls = sorted(((_i(d.get(c, "0")), c[6:]) for c in stall_cols), reverse=True)[:3]
    st = " ".join(f"{n}:{v}" for v, n in stalls if v)
    conf = d.get("L1 Wavefronts Shared Excessive", "0")
    print(f"{100.0 * s / total:5.1f}% {f}:{ln:<4} inst={d.get('Instructions Executed','?'):>9} shExc={conf:>8} | {st} | {src[:90]}")
