#!/usr/bin/env python
"""One bf16 TMA contraction, timed with CUDA events (and a convenient single-kernel target for `ncu -k regex:gemm_bf16_tma`).
usage: gemm_one.py [mode NT|NN|TN] [M] [N] [K]      defaults: NT 16384 512 1536 (the 16 384-row first layers of the behaviour phase)"""
import sys

import torch

sys.path.insert(0, ".")
from sheeprl_prey_b200._lib import CudaOps  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "NT"
M, N, K = (int(v) for v in sys.argv[2:5]) if len(sys.argv) > 4 else (16384, 512, 1536)
ops = CudaOps("cuda:0", precision="bf16")
A = torch.randn(*((M, K) if mode[0] == "N" else (K, M)), device="cuda")
B = torch.randn(*((N, K) if mode[1] == "T" else (K, N)), device="cuda")
C = torch.zeros(M, N, device="cuda")
for t in (A, B):
    ops.mirror(t)
    ops.refresh(t)
for _ in range(3):
    ops.gemm(mode, A, B, C)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    ops.gemm(mode, A, B, C)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"{mode} M{M} N{N} K{K}: {ms * 1e3:.1f} us, {2.0 * M * N * K / ms / 1e9:.1f} TFLOP/s")
