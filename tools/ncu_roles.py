#!/usr/bin/env python
"""Stall samples of one captured launch grouped by code region (between marker opcodes) and by opcode:
    python tools/ncu_roles.py report.ncu-rep <launch-skip> idx0,idx1,... [names]"""
import collections
import csv
import io
import subprocess
import sys

rep, skip = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
print(rows[0][1][:100])
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
first = {}
for i, r in enumerate(data):
    for k in ("UTMALDG", "LDTM", "UTCHMMA", "UTMASTG", "MUFU.EX2", "EXIT", "STTM"):
        if k in r[ix["Source"]] and k not in first:
            first[k] = i
print("first occurrence:", first, "instructions:", len(data))
if len(sys.argv) > 3:
    cuts = [int(x) for x in sys.argv[3].split(",")]
    names = sys.argv[4].split(",") if len(sys.argv) > 4 else [str(i) for i in range(len(cuts))]
    tot = sum(int(r[ix["# Samples"]]) for r in data)
    for (a, b), name in zip(zip(cuts[:-1], cuts[1:]), names):
        t = sum(int(r[ix["# Samples"]]) for r in data[a:b])
        agg = {h[6:]: sum(int(r[ix[h]]) for r in data[a:b]) for h in stalls}
        ops = collections.Counter()
        waits = collections.Counter()
        for r in data[a:b]:
            s = int(r[ix["# Samples"]])
            src = r[ix["Source"]].strip()
            op = src.split()[1] if src.startswith("@") else src.split()[0]
            ops[op.split(".")[0]] += s
        print(f"{name:14s} [{a},{b}) {t:6d} samples ({100 * t / tot:4.1f}%)", sorted(agg.items(), key=lambda kv: -kv[1])[:4])
        print("       opcodes:", ops.most_common(7))
        hot = sorted(data[a:b], key=lambda r: -int(r[ix["# Samples"]]))[:5]
        for r in hot:
            print(f"         {int(r[ix['# Samples']]):6d}  {r[ix['Source']].strip()[:90]}")
