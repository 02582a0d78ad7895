"""Pixel encoder / decoder of DreamerV3 as sequences of C-ABI calls (no arithmetic here).

Reference: CNNEncoder (sheeprl/algos/dreamer_v3/agent.py:28-80): stages x [Conv2d(k4,s2,p1,no bias) -> LayerNormChannelLast
(eps 1e-3) -> SiLU], Flatten(C,H,W); CNNDecoder (agent.py:136-206): Linear(lat -> 8m*4*4) -> Unflatten(8m,4,4) ->
(stages-1) x [ConvTranspose2d(k4,s2,p1,no bias) -> LN-CL -> SiLU] -> ConvTranspose2d(k4,s2,p1,+bias) -> +0.5.

Layout: activations are channel-last [N*H*W, C] so the channel LayerNorm is the row LayerNorm kernel; a convolution is
im2col + GEMM, a transposed convolution GEMM + col2im (and the other way round in the backward pass); the contraction index
is tap-major (k = (kh*4+kw)*C + c), so torch-layout weights are re-ordered with dv3_permute_last2 once per step and their
gradients are re-ordered back into the flat gradient arena.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence

import torch

Tensor = torch.Tensor


def _direct(ops) -> bool:
    """The narrow-channel layers (image <-> first / last feature map) as direct kernels (csrc/conv_direct.cu) instead of
    im2col / col2im + contraction.  Opt-in ($DV3_CONV_DIRECT=1): at config M the CUDA-core kernels are correct but not faster than the
    contractions they replace (tools/conv_direct_bench.py: first layer forward 0.34 vs 0.58 ms, but weight gradient 0.40 vs 0.11 ms and
    last layer forward 0.53 vs 0.39 ms), so the step rate does not move (64.8 vs 64.9 steps/s)."""
    return hasattr(ops, "conv_narrow_fwd") and os.environ.get("DV3_CONV_DIRECT", "0") == "1"


def _mirror(ops, *tensors) -> None:
    """bf16 mirrors for the contraction operands (CudaOps in bf16 precision; see engine._mirror)."""
    fn = getattr(ops, "mirror", None)
    if fn is not None:
        for t in tensors:
            if t is not None:
                fn(t)


class ConvEncoder:
    def __init__(self, ops, device, N: int, in_channels: Sequence[int], image_hw: int, layers: List[dict], out_view: Tensor, ld_out: int):
        """layers[l]: dict(w [Cout,Cin,4,4], g, beta, dw, dg, dbeta).  out_view: [N, 16*C_last] view (row stride ld_out) that
        receives the flattened (C,H,W) features."""
        self.ops, self.N, self.layers = ops, N, layers
        f32 = dict(dtype=torch.float32, device=device)
        self.in_channels = list(in_channels)
        self.C0 = sum(in_channels)
        self.hw = image_hw
        self.x0 = torch.zeros(N * image_hw * image_hw, self.C0, **f32)      # NHWC float image in [0,1]
        self.col, self.pre, self.act, self.mean, self.rstd, self.wt, self.dwt = [], [], [], [], [], [], []
        H, Cin = image_hw, self.C0
        self.dims = []
        for L in layers:
            Cout = L["w"].shape[0]
            rows = N * (H // 2) * (H // 2)
            self.dims.append((H, Cin, Cout, rows))
            self.col.append(torch.empty(rows, 16 * Cin, **f32))
            self.pre.append(torch.empty(rows, Cout, **f32))
            self.act.append(torch.empty(rows, Cout, **f32))
            self.mean.append(torch.empty(rows, **f32))
            self.rstd.append(torch.empty(rows, **f32))
            self.wt.append(torch.empty(Cout, 16 * Cin, **f32))
            self.dwt.append(torch.empty(Cout, 16 * Cin, **f32))
            H, Cin = H // 2, Cout
        self.H_last, self.C_last = H, Cin
        self.out_view, self.ld_out = out_view, ld_out
        self.out_dim = self.C_last * H * H
        rmax = max(d[3] * d[2] for d in self.dims)
        self._dpre = torch.empty(rmax, **f32)
        self._dln = torch.empty(rmax, **f32)
        self._dact = torch.empty(max(max(d[3] * d[2] for d in self.dims), N * image_hw * image_hw * self.C0), **f32)
        _mirror(ops, *self.col, *self.wt, *self.act, self._dpre)

    def load_images(self, images: Sequence[Tensor]) -> None:
        """images[k]: uint8 [T,B,C_k,H,W] (or [N,C_k,H,W]); -> x0 (channel-last, /255; dreamer_v3.py:102)."""
        off = 0
        for img, Ck in zip(images, self.in_channels):
            self.ops.u8_to_nhwc(img.reshape(self.N, Ck, self.hw * self.hw), self.N, Ck, self.hw * self.hw, 1.0 / 255.0, self.x0,
                                self.C0, off)
            off += Ck

    def forward(self) -> None:
        ops = self.ops
        x = self.x0
        for l, L in enumerate(self.layers):
            H, Cin, Cout, rows = self.dims[l]
            ops.permute_last2(L["w"], Cout, Cin, 16, self.wt[l])                       # [Cout][Cin][16] -> [Cout][16][Cin]
            self.direct0 = l == 0 and _direct(ops) and ops.conv_narrow_fwd(x, self.N, H, H, Cin, self.wt[l], self.pre[l], L["g"], L["beta"], 1e-3,
                                                                           self.act[l], self.mean[l], self.rstd[l])
            if not (l == 0 and self.direct0):
                ops.im2col(x, self.N, H, H, Cin, self.col[l])
                ops.gemm("NT", self.col[l], self.wt[l], self.pre[l])
                ops.ln_act_fwd(self.pre[l], L["g"], L["beta"], 1e-3, 1, self.act[l], self.mean[l], self.rstd[l])
            x = self.act[l]
        # [N][HW][C] -> [N][C][HW]: the reference flattens NCHW (nn.Flatten(-3,-1), agent.py:73)
        HW = self.H_last * self.H_last
        ops.permute_last2(x, self.N, HW, self.C_last, self.out_view, lds=HW * self.C_last, ldd=self.ld_out)

    def backward(self, dout_view: Tensor, ld_dout: int) -> None:
        """dout_view: [N, 16*C_last] gradient wrt the flattened features (row stride ld_dout)."""
        ops = self.ops
        HW = self.H_last * self.H_last
        n_last = self.N * HW * self.C_last
        dact = self._dact[:n_last].view(self.N * HW, self.C_last)
        ops.permute_last2(dout_view, self.N, self.C_last, HW, dact, lds=ld_dout, ldd=HW * self.C_last)   # -> channel-last
        for l in range(len(self.layers) - 1, -1, -1):
            L = self.layers[l]
            H, Cin, Cout, rows = self.dims[l]
            dpre = self._dpre[: rows * Cout].view(rows, Cout)
            dln = self._dln[: rows * Cout].view(rows, Cout)
            ops.ln_act_bwd_fused(dact, self.pre[l], L["g"], L["beta"], self.mean[l], self.rstd[l], 1, dpre, dln, L["dg"], L["dbeta"], None)
            if not (l == 0 and getattr(self, "direct0", False) and ops.conv_narrow_wgrad(self.x0, self.N, H, H, Cin, dpre, self.dwt[l])):
                if l == 0 and getattr(self, "direct0", False):
                    ops.im2col(self.x0, self.N, H, H, Cin, self.col[l])                 # (the direct forward did not build it)
                ops.gemm("TN", dpre, self.col[l], self.dwt[l])
            ops.permute_last2(self.dwt[l], Cout, 16, Cin, L["dw"])                     # back to [Cout][Cin][16]
            if l > 0:
                dcol = self.col[l]                                                     # its forward content is no longer needed
                ops.gemm("NN", dpre, self.wt[l], dcol)
                dact = self._dact[: self.N * H * H * Cin].view(self.N * H * H, Cin)
                ops.col2im(dcol, self.N, H // 2, H // 2, Cin, None, 0.0, dact)


class ConvDecoder:
    def __init__(self, ops, device, N: int, lat: int, out_channels: Sequence[int], lin: dict, layers: List[dict]):
        """lin: dict(w [16*C0, lat], b, dw, db).  layers[l]: dict(w [Cin,Cout,4,4], b|None, g|None, beta|None, dw, db, dg, dbeta);
        the last layer has a bias and no LN / activation."""
        self.ops, self.N, self.lin, self.layers = ops, N, lin, layers
        f32 = dict(dtype=torch.float32, device=device)
        self.C0 = lin["w"].shape[0] // 16
        self.lin_out = torch.empty(N, 16 * self.C0, **f32)       # [N][C0][16] (NCHW flatten, as nn.Unflatten expects)
        self.x0 = torch.empty(N * 16, self.C0, **f32)             # channel-last
        self.dlin_out = torch.empty(N, 16 * self.C0, **f32)
        self.colT, self.pre, self.act, self.mean, self.rstd, self.wr, self.dwr, self.dims = [], [], [], [], [], [], [], []
        H, Cin = 4, self.C0
        for L in layers:
            Cout = L["w"].shape[1]
            rows_in, rows_out = N * H * H, N * 4 * H * H
            self.dims.append((H, Cin, Cout, rows_in, rows_out))
            self.colT.append(torch.empty(rows_in, 16 * Cout, **f32))
            self.pre.append(torch.empty(rows_out, Cout, **f32))
            has_ln = L.get("g") is not None
            self.act.append(torch.empty(rows_out, Cout, **f32) if has_ln else None)
            self.mean.append(torch.empty(rows_out, **f32) if has_ln else None)
            self.rstd.append(torch.empty(rows_out, **f32) if has_ln else None)
            self.wr.append(torch.empty(Cin, 16 * Cout, **f32))
            self.dwr.append(torch.empty(Cin, 16 * Cout, **f32))
            H, Cin = 2 * H, Cout
        self.out = self.pre[-1]                                   # [N*H*W, C_img] channel-last reconstruction (+0.5 applied)
        self.out_hw, self.out_c = H, Cin
        rmax = max(d[4] * d[2] for d in self.dims)
        self._dpre = torch.empty(rmax, **f32)
        self._dln = torch.empty(rmax, **f32)
        self._dx = torch.empty(max(d[3] * d[1] for d in self.dims), **f32)
        self.d_rec = torch.empty(N, self.out_hw * self.out_hw * self.out_c, **f32)
        self.loss_rows = torch.empty(N, **f32)
        _mirror(ops, self.x0, *self.colT, *self.wr, *[a for a in self.act if a is not None], self._dpre, self.d_rec, self.dlin_out)

    def forward(self, latent: Tensor) -> None:
        ops = self.ops
        ops.gemm("NT", latent, self.lin["w"], self.lin_out, bias=self.lin["b"])
        ops.permute_last2(self.lin_out, self.N, self.C0, 16, self.x0)              # [N][C][16] -> [N][16][C]
        x = self.x0
        last = len(self.layers) - 1
        for l, L in enumerate(self.layers):
            H, Cin, Cout, rows_in, rows_out = self.dims[l]
            ops.permute_last2(L["w"], Cin, Cout, 16, self.wr[l])                    # [Cin][Cout][16] -> [Cin][16][Cout]
            if not (l == last and _direct(ops) and ops.deconv_narrow_fwd(x, self.N, H, H, self.wr[l], Cout, L.get("b"), 0.5, self.pre[l])):
                ops.gemm("NN", x, self.wr[l], self.colT[l])
                ops.col2im(self.colT[l], self.N, H, H, Cout, L.get("b"), 0.5 if l == last else 0.0, self.pre[l])
            if L.get("g") is not None:
                ops.ln_act_fwd(self.pre[l], L["g"], L["beta"], 1e-3, 1, self.act[l], self.mean[l], self.rstd[l])
                x = self.act[l]
            else:
                x = self.pre[l]

    def loss(self, target_nhwc: Tensor, scale: float) -> None:
        """MSEDistribution(dims=3).log_prob: sum over C,H,W of (mode - x)^2 per (t,b) (distribution.py:212-221)."""
        D = self.out_hw * self.out_hw * self.out_c
        self.ops.mse_rows(self.out.view(self.N, D), target_nhwc.view(self.N, D), scale, self.loss_rows, self.d_rec)

    def backward(self, latent: Tensor, dlatent: Tensor, accumulate: bool) -> None:
        ops = self.ops
        last = len(self.layers) - 1
        dout = self.d_rec.view(self.N * self.out_hw * self.out_hw, self.out_c)
        for l in range(last, -1, -1):
            L = self.layers[l]
            H, Cin, Cout, rows_in, rows_out = self.dims[l]
            if L.get("g") is not None:
                dpre = self._dpre[: rows_out * Cout].view(rows_out, Cout)
                dln = self._dln[: rows_out * Cout].view(rows_out, Cout)
                ops.ln_act_bwd_fused(dout, self.pre[l], L["g"], L["beta"], self.mean[l], self.rstd[l], 1, dpre, dln, L["dg"], L["dbeta"], None)
            else:
                dpre = dout
            if L.get("b") is not None:
                ops.colsum(dpre, L["db"])
            x_in = self.x0 if l == 0 else self.act[l - 1]
            dx = self._dx[: rows_in * Cin].view(rows_in, Cin)
            if (l == last and _direct(ops) and dpre.is_contiguous() and ops.conv_narrow_wgrad(dpre, self.N, 2 * H, 2 * H, Cout, x_in, self.dwr[l])
                    and ops.conv_narrow_fwd(dpre, self.N, 2 * H, 2 * H, Cout, self.wr[l], dx)):
                pass                                                                 # weight and data gradient straight from the image-shaped dpre
            else:
                dcolT = self.colT[l]                                                 # forward content no longer needed
                ops.im2col(dpre, self.N, 2 * H, 2 * H, Cout, dcolT)
                ops.gemm("TN", x_in, dcolT, self.dwr[l])
                ops.gemm("NT", dcolT, self.wr[l], dx)
            ops.permute_last2(self.dwr[l], Cin, 16, Cout, L["dw"])                  # back to [Cin][Cout][16]
            dout = dx
        ops.permute_last2(dout, self.N, 16, self.C0, self.dlin_out)                 # -> [N][C][16]
        ops.gemm("TN", self.dlin_out, latent, self.lin["dw"])
        ops.colsum(self.dlin_out, self.lin["db"])
        ops.gemm("NN", self.dlin_out, self.lin["w"], dlatent, accumulate=accumulate)
