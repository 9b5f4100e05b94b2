#!/usr/bin/env python
"""Summarise an `ncu --set full` report of tools/ncu_top_kernels.py: per captured launch its duration, DRAM bytes, tensor
pipe / DRAM / issue / XU utilisation; writes profiles/<tag>_top_kernels_summary.json and profiles/<tag>_traffic.json (the
per-launch dram__bytes_read.sum + dram__bytes_write.sum that bench.py reports as roofline.traffic).

    python tools/ncu_summary.py gpurun_out/r2b_top.ncu-rep r2b "<commit the capture was taken on>"
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, tag = sys.argv[1], sys.argv[2]
commit = sys.argv[3] if len(sys.argv) > 3 else "unknown"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}


def num(r, key):
    try:
        return float(r[ix[key]].replace(",", ""))
    except Exception:
        return None


def to_bytes(r, key):
    v, u = num(r, key), units[ix[key]].lower()
    if v is None:
        return None
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)


def to_us(r, key):
    v, u = num(r, key), units[ix[key]].lower()
    return None if v is None else v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1)


out = []
for r in rows[2:]:
    out.append({
        "kernel": r[ix["Kernel Name"]][:96], "grid": r[ix["Grid Size"]], "block": r[ix["Block Size"]],
        "time_us": to_us(r, "gpu__time_duration.sum"),
        "dram_read_bytes": to_bytes(r, "dram__bytes_read.sum"), "dram_write_bytes": to_bytes(r, "dram__bytes_write.sum"),
        "tensor_pipe_pct": num(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")
        if "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active" in ix else num(r, "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"),
        "dram_pct": num(r, "dram__throughput.avg.pct_of_peak_sustained_elapsed"),
        "issue_pct": num(r, "sm__issue_active.avg.pct_of_peak_sustained_elapsed"),
        "xu_pct": num(r, "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
        "regs": r[ix["launch__registers_per_thread"]] if "launch__registers_per_thread" in ix else None,
    })
for o in out:
    if o["dram_read_bytes"] is not None and o["time_us"]:
        o["dram_GBps"] = (o["dram_read_bytes"] + o["dram_write_bytes"]) / (o["time_us"] * 1e-6) / 1e9
json.dump(out, open(os.path.join(ROOT, "profiles", f"{tag}_top_kernels_summary.json"), "w"), indent=1)
# the warm (second) launch of each of the four stage-3 GEMM forms: launches 1, 3, 5, 7 of the capture
gemm = [o for o in out if "gemm_bf16" in o["kernel"]]
warm = gemm[1::2][:4]
traffic = {
    "dram_bytes_per_launch_avg": sum(o["dram_read_bytes"] + o["dram_write_bytes"] for o in warm) / max(1, len(warm)),
    "per_form": [{"kernel": o["kernel"], "dram_bytes": o["dram_read_bytes"] + o["dram_write_bytes"], "time_us": o["time_us"], "tensor_pipe_pct": o["tensor_pipe_pct"]} for o in warm],
    "note": (f"bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum), mean over the warm launch of the four stage-3 GEMM forms "
             f"(attn.proj, fc2, qkv, fc1+GELU) at BATCH 32, one `ncu --set full --clock-control none` capture of tools/ncu_top_kernels.py "
             f"on the build of commit {commit}: profiles/{tag}_top_kernels_summary.json"),
}
json.dump(traffic, open(os.path.join(ROOT, "profiles", f"{tag}_traffic.json"), "w"), indent=1)
for o in out:
    print(f"{o['kernel'][:70]:70s} {o['time_us']:9.1f} us  tensor {o['tensor_pipe_pct']}  dram {o['dram_pct']}  issue {o['issue_pct']}  xu {o['xu_pct']}  {o.get('dram_GBps', 0):.0f} GB/s")
