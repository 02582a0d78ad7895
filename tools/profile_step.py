#!/usr/bin/env python
"""Runs a few warm-up gradient steps of a bench workload, then N steps between cudaProfilerStart / Stop (for
`ncu --profile-from-start off ...`).  usage: profile_step.py [size] [steps]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

size = sys.argv[1] if len(sys.argv) > 1 else "S"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
run = bench.Run(size, "bf16", dev, 1, 0, 20000)
for _ in range(4):
    run.step_device()
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(steps):
    run.step_device()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled", steps, "steps of", size, "scan kernels", run.eng.ops.scan_kernel, run.eng.ops.bwd_scan_kernel, "rollout", run.eng.ops.rollout_kernel)
