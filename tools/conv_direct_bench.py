#!/usr/bin/env python
"""Times the direct narrow-channel convolution kernels (csrc/conv_direct.cu) at config M shapes against the im2col / col2im + GEMM
composition they replace.  usage: conv_direct_bench.py [N] [Cout]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sheeprl_prey_b200._lib import CudaOps  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
O = int(sys.argv[2]) if len(sys.argv) > 2 else 48
dev = "cuda:0"
ops = CudaOps(dev, precision="bf16")
H, C = 64, 3
rows = N * 32 * 32
g = torch.Generator().manual_seed(0)
r = lambda *s: torch.randn(*s, generator=g).to(dev)  # noqa: E731
x, w, gam, bet = r(N * H * H, C), r(O, 16 * C) * 0.2, r(O) * 0.1 + 1, r(O) * 0.1
pre, act, mean, rstd = torch.zeros(rows, O, device=dev), torch.zeros(rows, O, device=dev), torch.zeros(rows, device=dev), torch.zeros(rows, device=dev)
col = torch.zeros(rows, 16 * C, device=dev)
ops.mirror(col); ops.mirror(w); ops.mirror(act); ops.refresh(w)
gy, dw = r(rows, O), torch.zeros(O, 16 * C, device=dev)
ops.mirror(gy); ops.refresh(gy)
xin, wr, bias, out = r(rows, O), r(O, 16 * C) * 0.2, r(C), torch.zeros(N * H * H, C, device=dev)
ops.mirror(xin); ops.mirror(wr); ops.refresh(xin); ops.refresh(wr)
colT = torch.zeros(rows, 16 * C, device=dev)
ops.mirror(colT)
dx = torch.zeros(rows, O, device=dev)


def timed(fn, n=10):
    fn(); fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def old_fwd():
    ops.im2col(x, N, H, H, C, col)
    ops.gemm("NT", col, w, pre)
    ops.ln_act_fwd(pre, gam, bet, 1e-3, 1, act, mean, rstd)


def old_dec():
    ops.gemm("NN", xin, wr, colT)
    ops.col2im(colT, N, 32, 32, C, bias, 0.5, out)


def old_dec_bwd():
    ops.im2col(out, N, H, H, C, colT)
    ops.gemm("TN", xin, colT, dw)
    ops.gemm("NT", colT, wr, dx)


res = {"N": N, "O": O,
       "enc_fwd_direct_ms": timed(lambda: ops.conv_narrow_fwd(x, N, H, H, C, w, pre, gam, bet, 1e-3, act, mean, rstd)),
       "enc_fwd_im2col_gemm_ln_ms": timed(old_fwd),
       "wgrad_direct_ms": timed(lambda: ops.conv_narrow_wgrad(x, N, H, H, C, gy, dw)),
       "wgrad_gemm_tn_ms": timed(lambda: ops.gemm("TN", gy, col, dw)),
       "dec_fwd_direct_ms": timed(lambda: ops.deconv_narrow_fwd(xin, N, 32, 32, wr, C, bias, 0.5, out)),
       "dec_fwd_gemm_col2im_ms": timed(old_dec),
       "dec_bwd_direct_ms": timed(lambda: (ops.conv_narrow_wgrad(out, N, H, H, C, xin, dw), ops.conv_narrow_fwd(out, N, H, H, C, wr, dx))),
       "dec_bwd_im2col_gemms_ms": timed(old_dec_bwd)}
print(json.dumps(res, indent=1))
