#!/usr/bin/env python
"""Summarise an ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none --csv`) of
`python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-torch-gpu-baseline --video-frames 0`: the encoder kernels repeat
once per 32-image step, every step starts with `patch_im2col_kernel`; the fourth step (warm-up 3 + 1 timed) is summarised
per kernel into a markdown table.

    python tools/launch_summary.py profiles/r2_launches.csv "<what was captured>" > profiles/r2_launches_summary.md
"""
import csv
import re
import sys

path, what = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
rows = []
with open(path, newline="") as f:
    lines = [ln for ln in f if ln.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    rows.append((r["Kernel Name"], float(r["Metric Value"].replace(",", "")) * (1e-6 if r["Metric Unit"] in ("ns", "nsecond") else 1e-3)))
OURS = re.compile(r"sam2b::|gemm_bf16_tcgen05|flash_attn|window_attn|mlp_fused|patch_im2col|maxpool2x2|upsample2x_add")
starts = [i for i, (k, _) in enumerate(rows) if "patch_im2col_kernel" in k]
if len(starts) < 4:
    sys.exit(f"only {len(starts)} encoder steps in {path}")


def step_rows(i0):
    out = []
    for k, ms in rows[i0:]:
        if out and "patch_im2col_kernel" in k:
            break
        if not OURS.search(k):
            if len(out) > 50:
                break                    # the step is over: torch kernels (gather copies, checksum) follow
            continue
        out.append((k, ms))
    return out


step = step_rows(starts[3])
agg = {}
for k, ms in step:
    m = re.match(r"(?:void )?((?:sam2b::)?[\w:]+(?:<[^>]*>)?)", k)
    name = m.group(1) if m else k[:60]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += ms
total = sum(v[1] for v in agg.values())
print(f"# ncu launch list of {what}\n")
print(f"`ncu --metrics gpu__time_duration.sum --clock-control none --csv` (`{path}`): {len(rows)} launches captured, {len(starts)} encoder "
      "steps (each starts with `patch_im2col_kernel`); the fourth is the timed 32-image step, summarised here.  Per-launch times are "
      "cold-cache and serialised under ncu: compare SHARES with `roofline.share_of_step_kernel_time` / `kernel_time_ms_by_family` "
      "of the bench line of the same build, not absolute times.\n")
print("| kernel | launches | total ms | share |\n|---|---|---|---|")
for name, (c, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{name}` | {c} | {ms:.3f} | {ms / total:.1%} |")
fam = lambda pred: sum(v[1] for k, v in agg.items() if pred(k))
print(f"\nTotal {total:.2f} ms over {len(step)} launches; tcgen05 GEMM {fam(lambda k: 'gemm_bf16' in k) / total:.1%}, "
      f"fused MLP {fam(lambda k: 'mlp_fused' in k) / total:.1%}, attention kernels {fam(lambda k: 'attn' in k) / total:.1%}, "
      f"element-wise {fam(lambda k: not any(s in k for s in ('gemm_bf16', 'mlp_fused', 'attn'))) / total:.1%}.")
