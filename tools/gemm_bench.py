#!/usr/bin/env python
"""Per-shape timing of the contraction kernels: 20 launches captured in a CUDA graph (no CPU launch gaps), warm L2."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sheeprl_prey_b200._lib import CudaOps  # noqa: E402

dev = "cuda:0"
shapes = [("NT", 128, 128, 64), ("NT", 128, 128, 512), ("NT", 1024, 512, 64), ("NT", 1024, 512, 256), ("NT", 1024, 512, 512), ("NT", 1024, 512, 1536), ("NT", 1024, 1536, 1024), ("NT", 16384, 512, 512), ("NT", 16384, 512, 1536),
          ("NN", 15360, 512, 512), ("TN", 512, 1536, 15360), ("TN", 512, 512, 15360), ("NT", 16, 1536, 1024), ("NT", 16, 512, 1024),
          ("NT", 16, 1024, 512), ("NT", 1024, 640, 6144), ("NT", 262144, 96, 768), ("NT", 8192, 8192, 8192)]
import ctypes  # noqa: E402


def stamps(ops):
    buf = (ctypes.c_ulonglong * 16)()
    ops.lib.dv3_gemm_tc_debug_stamps.argtypes = [ctypes.c_void_p]
    ops.lib.dv3_gemm_tc_debug_stamps(ctypes.cast(buf, ctypes.c_void_p))
    v = list(buf)
    names = ["start", "alloc+init done", "first stage filled", "last stage filled", "accumulator ready", "epilogue done", "after final sync"]
    extra = f" || chunk0: ld {(v[7] - v[4]) / 1e3:.2f} sts {(v[8] - v[7]) / 1e3:.2f} store {(v[9] - v[8]) / 1e3:.2f} us"
    return " | ".join(f"{names[i]} +{(v[i] - v[0]) / 1e3:.2f}us" for i in range(1, 7)) + extra


print("precision | shape | us / launch | TFLOP/s")
for prec in ("bf16", "fp32"):
    ops = CudaOps(dev, precision=prec)
    for mode, M, N, K in shapes:
        if prec == "fp32" and M * N * K > 2e11:
            continue
        A = torch.randn((M, K) if mode != "TN" else (K, M), device=dev)
        B = torch.randn((N, K) if mode == "NT" else (K, N), device=dev)
        C = torch.zeros(M, N, device=dev)
        ops.gemm(mode, A, B, C)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        reps = 20
        with torch.cuda.graph(g):
            for _ in range(reps):
                ops.gemm(mode, A, B, C)
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / reps
        print(f"{prec} | {mode} M{M} N{N} K{K} | {us:.1f} | {2.0 * M * N * K / us / 1e6:.1f}" + (("   [" + stamps(ops) + "]") if prec == "bf16" and M >= 128 else ""))


# TMA-fed bf16-resident variant (NT only)
lib = CudaOps(dev).lib
print("tma bf16 | shape | us / launch | TFLOP/s")
for mode, M, N, K in shapes:
    if mode != "NT" or M < 128 or K % 8:
        continue
    A = torch.randn(M, K, device=dev).bfloat16()
    B = torch.randn(N, K, device=dev).bfloat16()
    C = torch.zeros(M, N, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    assert lib.dv3_gemm_bf16_tma(0, M, N, K, A.data_ptr(), K, B.data_ptr(), K, C.data_ptr(), N, None, 0, st) == 0
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20):
            lib.dv3_gemm_bf16_tma(0, M, N, K, A.data_ptr(), K, B.data_ptr(), K, C.data_ptr(), N, None, 0, torch.cuda.current_stream().cuda_stream)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20
    print(f"tma | NT M{M} N{N} K{K} | {us:.1f} | {2.0 * M * N * K / us / 1e6:.1f}")
