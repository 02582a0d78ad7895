#!/usr/bin/env python
"""Collects the step-parity reports that `pytest -m gpu` writes (gpurun_out/parity_*.json, tests/test_step_gpu.py) into one markdown
file.  usage: tools/parity_report.py [title] > profiles/rNN_parity.md"""
import glob
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
title = sys.argv[1] if len(sys.argv) > 1 else "step parity reports"
print(f"# {title}\n")
print("Written by `pytest -m gpu` on 1xB200 (tests/test_step_gpu.py): max |cuda - oracle| / max |oracle| per tensor; metrics relative to "
      "max(1, |oracle|).\n")
for f in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "parity_*.json"))):
    d = json.load(open(f))
    head = ", ".join(f"{k} {d[k]}" for k in ("config", "mode", "scan_kernels", "rollout_kernel") if k in d)
    print(f"## {os.path.basename(f)}  ({head})\n")
    print("| quantity | value |\n|---|---|")
    for k, v in d.items():
        if k in ("config", "mode", "scan_kernels", "rollout_kernel"):
            continue
        if isinstance(v, float):
            print(f"| {k} | {v:.3e} |")
        elif isinstance(v, (int, str)):
            print(f"| {k} | {v} |")
    print()
