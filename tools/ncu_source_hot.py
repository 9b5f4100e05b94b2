#!/usr/bin/env python
"""Hottest CUDA source lines (warp stall samples) of one captured launch:
    python tools/ncu_source_hot.py report.ncu-rep 'regex:flash_attn_hd72' <launch-skip> [top]"""
import csv
import io
import subprocess
import sys

rep, kname, skip = sys.argv[1], sys.argv[2], sys.argv[3]
top_n = int(sys.argv[4]) if len(sys.argv) > 4 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda", "--csv", "--kernel-name", kname,
                      "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
out = []
fname, hdr = None, None
for r in rows:
    if not r:
        continue
    if r[0] == "File Name":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        ix = {h: i for i, h in enumerate(hdr)}
        continue
    if hdr and len(r) == len(hdr):
        out.append((fname, r))
tot = sum(int(r[ix["# Samples"]]) for _, r in out)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("total samples", tot)
for f, r in sorted(out, key=lambda fr: -int(fr[1][ix["# Samples"]]))[:top_n]:
    s = int(r[ix["# Samples"]])
    st = sorted(((int(r[ix[h]]), h[6:]) for h in stalls), reverse=True)[:2]
    print(f"{s:6d} {100 * s / tot:5.1f}% {f}:{r[0]:>4s} {r[ix['Source']].strip()[:100]:100s} {st}")
