"""TEST INFRASTRUCTURE ONLY -- plain-PyTorch statement of every primitive the CUDA library exports.

The product's step engine (sheeprl_prey_b200/engine.py) is written against an `ops` object whose
methods map 1:1 onto the C-ABI entry points of libdv3b200.so (include/dv3_b200.h).  This file
implements the same methods with torch tensor ops so that

  * tests/test_engine_cpu.py can run the engine's hand-written backward / scan orchestration on
    CPU and compare it with the oracle's autograd (host-logic test, `-m "not gpu"`), and
  * tests/test_kernels_gpu.py can compare every CUDA kernel with its torch statement on the same
    inputs (`-m gpu`).

It is never imported by the product package: `engine.DV3Engine` requires the CUDA ops unless a
test injects this class explicitly.

All matrices are 2-D views whose last dimension is contiguous (row stride = leading dimension).
Outputs are written in place into caller-provided tensors, exactly like the C-ABI.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
_EPS32 = float(torch.finfo(torch.float32).eps)


def _inv_cdf(p: Tensor, u: Tensor) -> Tensor:
    c = torch.zeros_like(u)
    idx = torch.zeros(u.shape, dtype=torch.long, device=u.device)
    for k in range(p.shape[-1]):
        c = c + p[..., k]
        idx = idx + (c <= u).long()
    return idx.clamp_(max=p.shape[-1] - 1)


class TorchOps:
    name = "torch"

    # ------------------------------------------------------------------ dense contractions
    def gemm(self, mode: str, A: Tensor, B: Tensor, C: Tensor, bias: Optional[Tensor] = None,
             accumulate: bool = False) -> None:
        """NT: C[M,N] = A[M,K] B[N,K]^T;  NN: C = A[M,K] B[K,N];  TN: C = A[K,M]^T B[K,N];
        (+ bias[N]) ; `accumulate` adds into C instead of overwriting."""
        if mode == "NT":
            R = A @ B.t()
        elif mode == "NN":
            R = A @ B
        elif mode == "TN":
            R = A.t() @ B
        else:
            raise ValueError(mode)
        if bias is not None:
            R = R + bias
        if accumulate:
            C += R
        else:
            C.copy_(R)

    def transpose(self, src: Tensor, dst: Tensor) -> None:
        dst.copy_(src.t())

    def copy2d(self, src: Tensor, dst: Tensor) -> None:
        dst.copy_(src)

    # ------------------------------------------------------------------ LayerNorm (+SiLU)
    def ln_act_fwd(self, x: Tensor, gamma: Tensor, beta: Tensor, eps: float, act: int, y: Tensor,
                   mean: Tensor, rstd: Tensor) -> None:
        """y = act(LN(x)); act 0 = identity, 1 = SiLU. Biased variance; saves mean and 1/sqrt(var+eps)."""
        mu = x.mean(-1)
        var = ((x - mu[:, None]) ** 2).mean(-1)
        r = torch.rsqrt(var + eps)
        ln = (x - mu[:, None]) * r[:, None] * gamma + beta
        y.copy_(F.silu(ln) if act else ln)
        mean.copy_(mu)
        rstd.copy_(r)

    def ln_act_bwd(self, dy: Tensor, x: Tensor, gamma: Tensor, beta: Tensor, mean: Tensor, rstd: Tensor, act: int,
                   dx: Tensor, dln: Optional[Tensor]) -> None:
        """Row-wise backward of y = act(LN(x)).  dln (optional) receives dL/d(LN output); dx receives dL/dx.
        dx may alias dy."""
        xhat = (x - mean[:, None]) * rstd[:, None]
        if act:
            ln = xhat * gamma + beta
            s = torch.sigmoid(ln)
            g = dy * (s * (1 + ln * (1 - s)))
        else:
            g = dy.clone()
        if dln is not None:
            dln.copy_(g)
        gg = g * gamma
        dx.copy_(rstd[:, None] * (gg - gg.mean(-1, keepdim=True) - xhat * (gg * xhat).mean(-1, keepdim=True)))

    def ln_act_fwd_gather(self, x, gamma, beta, eps, act, y, mean, rstd, table, idx, head_offsets) -> bool:
        """ln_act_fwd on x + sum_h table[head_offsets[h] + idx[:, h]]."""
        xx = x.clone()
        for h, off in enumerate(head_offsets):
            xx += table[idx[:, h].long() + int(off)]
        self.ln_act_fwd(xx, gamma, beta, eps, act, y, mean, rstd)
        return True

    def ln_act_bwd_fused(self, dy, x, gamma, beta, mean, rstd, act, dx, dln, dgamma, dbeta, dbias) -> None:
        """ln_act_bwd, ln_param_grad and (dbias given) colsum(dx) of a dense layer's backward."""
        self.ln_act_bwd(dy, x, gamma, beta, mean, rstd, act, dx, dln)
        self.ln_param_grad(dln, x, mean, rstd, dgamma, dbeta)
        if dbias is not None:
            self.colsum(dx, dbias)

    def ln_param_grad(self, dln: Tensor, x: Tensor, mean: Tensor, rstd: Tensor, dgamma: Tensor, dbeta: Tensor) -> None:
        """dgamma[u] = sum_m dln[m,u] * xhat[m,u];  dbeta[u] = sum_m dln[m,u]  (overwrite)."""
        xhat = (x - mean[:, None]) * rstd[:, None]
        dgamma.copy_((dln * xhat).sum(0))
        dbeta.copy_(dln.sum(0))

    def colsum(self, dy: Tensor, db: Tensor) -> None:
        db.copy_(dy.sum(0))

    def symlog(self, x: Tensor, y: Tensor) -> None:
        y.copy_(torch.sign(x) * torch.log(1 + torch.abs(x)))

    # ------------------------------------------------------------------ pixel path (channel-last, k = tap*C + c)
    def im2col(self, x: Tensor, N: int, H: int, W: int, C: int, col: Tensor) -> None:
        """col[(n,oh,ow)][(kh*4+kw)*C + c] = x[n][2oh-1+kh][2ow-1+kw][c] (zero padded); x NHWC."""
        xp = F.pad(x.reshape(N, H, W, C), (0, 0, 1, 1, 1, 1))
        Ho, Wo = H // 2, W // 2
        taps = [xp[:, kh:kh + 2 * Ho:2, kw:kw + 2 * Wo:2, :] for kh in range(4) for kw in range(4)]
        col.copy_(torch.stack(taps, 3).reshape(N * Ho * Wo, 16 * C))

    def col2im(self, col: Tensor, N: int, Hi: int, Wi: int, C: int, bias: Optional[Tensor], add: float, y: Tensor) -> None:
        """y[n][2ih-1+kh][2iw-1+kw][c] += col[(n,ih,iw)][(kh*4+kw)*C + c]  (+ bias + add); y NHWC [N,2Hi,2Wi,C]."""
        c = col.reshape(N, Hi, Wi, 4, 4, C)
        out = torch.zeros(N, 2 * Hi + 2, 2 * Wi + 2, C, dtype=col.dtype, device=col.device)
        for kh in range(4):
            for kw in range(4):
                out[:, kh:kh + 2 * Hi:2, kw:kw + 2 * Wi:2, :] += c[:, :, :, kh, kw, :]
        out = out[:, 1:-1, 1:-1, :] + add
        if bias is not None:
            out = out + bias
        y.copy_(out.reshape(y.shape))

    def permute_last2(self, src: Tensor, A: int, Bn: int, Cn: int, dst: Tensor, lds=None, ldd=None) -> None:
        lds, ldd = lds or Bn * Cn, ldd or Bn * Cn
        s = torch.as_strided(src, (A, Bn, Cn), (lds, Cn, 1), src.storage_offset())
        d = torch.as_strided(dst, (A, Cn, Bn), (ldd, Bn, 1), dst.storage_offset())
        d.copy_(s.transpose(1, 2))

    def u8_to_nhwc(self, src: Tensor, N: int, C: int, HW: int, scale: float, dst: Tensor, Ctot: int, c_off: int) -> None:
        dst.view(N, HW, Ctot)[:, :, c_off:c_off + C] = src.reshape(N, C, HW).permute(0, 2, 1).float() * scale

    def mse_rows(self, mode: Tensor, target: Tensor, scale: float, loss_rows: Tensor, dmode: Tensor) -> None:
        diff = mode - target
        loss_rows.copy_((diff * diff).sum(-1))
        dmode.copy_(scale * 2 * diff)

    # ------------------------------------------------------------------ GRU gates
    def gru_gate_fwd(self, gates_pre: Tensor, lnw: Tensor, lnb: Tensor, h_in: Tensor, h_out: Tensor, mean: Tensor,
                     rstd: Tensor) -> None:
        """x = LN_1e-5(gates_pre [R,3D]); r,c,u = chunk(x); r=sig(r); c=tanh(r*c); u=sig(u-1); h_out = u*c+(1-u)*h_in."""
        mu = gates_pre.mean(-1)
        var = ((gates_pre - mu[:, None]) ** 2).mean(-1)
        rs = torch.rsqrt(var + 1e-5)
        x = (gates_pre - mu[:, None]) * rs[:, None] * lnw + lnb
        r, c, u = torch.chunk(x, 3, -1)
        r = torch.sigmoid(r)
        c = torch.tanh(r * c)
        u = torch.sigmoid(u - 1)
        h_out.copy_(u * c + (1 - u) * h_in)
        mean.copy_(mu)
        rstd.copy_(rs)

    def gru_gate_bwd(self, dh_out: Tensor, gates_pre: Tensor, lnw: Tensor, lnb: Tensor, mean: Tensor, rstd: Tensor,
                     h_in: Tensor, dgates_pre: Tensor, dln: Tensor, dh_in: Tensor) -> None:
        """Backward of gru_gate_fwd: dgates_pre [R,3D] (through the LN), dln [R,3D] (grad wrt LN output, for the LN
        parameter grads), dh_in [R,D] = dh_out*(1-u) (the direct path only)."""
        xhat = (gates_pre - mean[:, None]) * rstd[:, None]
        x = xhat * lnw + lnb
        rp, cp, up = torch.chunk(x, 3, -1)
        r = torch.sigmoid(rp)
        c = torch.tanh(r * cp)
        u = torch.sigmoid(up - 1)
        du = dh_out * (c - h_in)
        dc = dh_out * u
        dh_in.copy_(dh_out * (1 - u))
        dpre_c = dc * (1 - c * c)           # grad wrt (r*cp)
        dr = dpre_c * cp
        dcp = dpre_c * r
        drp = dr * r * (1 - r)
        dup = du * u * (1 - u)
        g = torch.cat((drp, dcp, dup), -1)
        dln.copy_(g)
        gg = g * lnw
        dgates_pre.copy_(rstd[:, None] * (gg - gg.mean(-1, keepdim=True) - xhat * (gg * xhat).mean(-1, keepdim=True)))

    # ------------------------------------------------------------------ categorical helpers
    def unimix_fwd(self, raw: Tensor, groups: int, K: int, unimix: float, logits: Tensor) -> None:
        """Per group of K classes (each row holds `groups` groups): p = softmax(raw); p = (1-mix)p + mix/K;
        logits = log(clamp(p, eps32, 1-eps32))."""
        R = raw.shape[0]
        p = raw.reshape(R, groups, K).softmax(-1)
        p = (1 - unimix) * p + unimix / K
        logits.copy_(torch.log(p.clamp(min=_EPS32, max=1 - _EPS32)).reshape(R, groups * K))

    def unimix_sample(self, raw: Tensor, groups: int, K: int, unimix: float, u: Optional[Tensor], forced: Optional[Tensor],
                      idx: Tensor, onehot: Tensor, logits: Optional[Tensor] = None) -> None:
        """unimix_fwd followed by sample_onehot(mode=0) on its output."""
        lg = torch.empty_like(raw) if logits is None else logits
        self.unimix_fwd(raw, groups, K, unimix, lg)
        self.sample_onehot(lg, groups, K, u, forced, 0, idx, onehot)

    def sample_onehot(self, logits: Tensor, groups: int, K: int, u: Optional[Tensor], forced: Optional[Tensor],
                      mode: int, idx: Tensor, onehot: Tensor) -> None:
        """probs = softmax(logits - logsumexp(logits)) per group; idx = forced, or argmax (mode=1), or the
        inverse-CDF of u (sequential fp32 cumsum; idx = #{k: c_k <= u} clamped).  idx [R,groups] int32;
        onehot [R, groups*K]."""
        R = logits.shape[0]
        l = logits.reshape(R, groups, K)
        p = (l - l.logsumexp(-1, keepdim=True)).softmax(-1)
        if forced is not None:
            ix = forced.long().reshape(R, groups)
        elif mode:
            ix = p.argmax(-1)
        else:
            ix = _inv_cdf(p, u.reshape(R, groups))
        idx.copy_(ix.to(idx.dtype))
        onehot.copy_(F.one_hot(ix, K).to(logits.dtype).reshape(R, groups * K))

    def unimix_bwd(self, raw: Tensor, groups: int, K: int, unimix: float, dlogits: Optional[Tensor],
                   dz: Optional[Tensor], draw: Tensor) -> None:
        """Backward of  raw -> unimix logits lu -> {consumers of lu (grad `dlogits`), straight-through sample
        z = onehot + p - sg(p), p = softmax(lu) (grad `dz`)}  to draw = dL/draw.  draw may alias dlogits."""
        R = raw.shape[0]
        ps = raw.reshape(R, groups, K).softmax(-1)
        pm = (1 - unimix) * ps + unimix / K
        inside = (pm >= _EPS32) & (pm <= 1 - _EPS32)
        lu = torch.log(pm.clamp(min=_EPS32, max=1 - _EPS32))
        g = torch.zeros_like(ps)
        if dlogits is not None:
            g = g + dlogits.reshape(R, groups, K)
        if dz is not None:
            p = (lu - lu.logsumexp(-1, keepdim=True)).softmax(-1)
            d = dz.reshape(R, groups, K)
            g = g + p * (d - (p * d).sum(-1, keepdim=True))
        gpm = torch.where(inside, g / pm, torch.zeros_like(g))
        gps = (1 - unimix) * gpm
        draw.copy_((ps * (gps - (ps * gps).sum(-1, keepdim=True))).reshape(R, groups * K))

    # ------------------------------------------------------------------ losses (forward + gradient in one pass)
    def kl_balance(self, post: Tensor, prior: Tensor, groups: int, K: int, free_nats: float, w_dyn: float,
                   w_rep: float, scale: float, rows_out: Tensor, dpost: Tensor, dprior: Tensor) -> None:
        """rows_out [R,4] = (kl, w_dyn*max(kl,free)+w_rep*max(kl,free), post_entropy, prior_entropy) per row, where
        kl = sum over groups of KL(post || prior) on the normalised logits.  dpost / dprior = scale * d(kl_loss)/d(logits)
        with the reference's stop-gradients (dyn -> prior only, rep -> post only)."""
        R = post.shape[0]
        lp = post.reshape(R, groups, K)
        lq = prior.reshape(R, groups, K)
        lp = lp - lp.logsumexp(-1, keepdim=True)
        lq = lq - lq.logsumexp(-1, keepdim=True)
        p, q = lp.softmax(-1), lq.softmax(-1)
        t = p * (lp - lq)
        t = torch.where(q == 0, torch.full_like(t, math.inf), t)
        t = torch.where(p == 0, torch.zeros_like(t), t)
        klg = t.sum(-1)
        kl = klg.sum(-1)
        active = (kl > free_nats).to(kl.dtype) + 0.5 * (kl == free_nats).to(kl.dtype)
        rows_out[:, 0] = kl
        rows_out[:, 1] = (w_dyn + w_rep) * torch.clamp(kl, min=free_nats)
        rows_out[:, 2] = -(lp.clamp(min=torch.finfo(lp.dtype).min) * p).sum(-1).sum(-1)
        rows_out[:, 3] = -(lq.clamp(min=torch.finfo(lq.dtype).min) * q).sum(-1).sum(-1)
        a = active[:, None, None]
        dprior.copy_((scale * w_dyn * a * (q - p)).reshape(R, groups * K))
        dpost.copy_((scale * w_rep * a * p * ((lp - lq) - klg[..., None])).reshape(R, groups * K))

    def twohot_loss(self, logits: Tensor, x: Tensor, w: Optional[Tensor], scale: float, loss_rows: Tensor,
                    dlogits: Tensor, accumulate: bool) -> None:
        """nll[m] = -sum_k target_k(x[m]) * log_softmax(logits[m])_k with the two-hot target over
        linspace(-20,20,K) of symlog(x).  loss_rows (+)= w[m]*nll[m];  dlogits (+)= scale*w[m]*(softmax - target)."""
        K = logits.shape[-1]
        bins = torch.linspace(-20.0, 20.0, K, device=logits.device)
        y = (torch.sign(x) * torch.log(1 + torch.abs(x))).reshape(-1, 1)
        below = ((bins <= y).sum(-1, keepdim=True) - 1).clamp(0, K - 1)
        above = (K - (bins > y).sum(-1, keepdim=True)).clamp(0, K - 1)
        eq = below == above
        db = torch.where(eq, torch.ones_like(y), (bins[below] - y).abs())
        da = torch.where(eq, torch.ones_like(y), (bins[above] - y).abs())
        tot = db + da
        target = F.one_hot(below.squeeze(-1), K) * (da / tot) + F.one_hot(above.squeeze(-1), K) * (db / tot)
        lsm = logits - logits.logsumexp(-1, keepdim=True)
        nll = -(target * lsm).sum(-1)
        ww = torch.ones_like(nll) if w is None else w.reshape(-1)
        g = scale * ww[:, None] * (lsm.exp() * target.sum(-1, keepdim=True) - target)
        if accumulate:
            loss_rows += ww * nll
            dlogits += g
        else:
            loss_rows.copy_(ww * nll)
            dlogits.copy_(g)

    def twohot_mean(self, logits: Tensor, out: Tensor) -> None:
        """out[m] = symexp(sum_k softmax(logits[m])_k * bins_k)"""
        bins = torch.linspace(-20.0, 20.0, logits.shape[-1], device=logits.device)
        v = (logits.softmax(-1) * bins).sum(-1)
        out.copy_(torch.sign(v) * (torch.exp(torch.abs(v)) - 1))

    def symlog_mse(self, mode: Tensor, target: Tensor, scale: float, loss_rows: Tensor, dmode: Tensor) -> None:
        """dist = (mode-target)^2, entries < 1e-8 zeroed; loss_rows[m] = sum_d dist; dmode = scale*2*(mode-target) (0 where zeroed)."""
        diff = mode - target
        d = diff * diff
        keep = d >= 1e-8
        loss_rows.copy_(torch.where(keep, d, torch.zeros_like(d)).sum(-1))
        dmode.copy_(torch.where(keep, scale * 2 * diff, torch.zeros_like(d)))

    def bernoulli_loss(self, logit: Tensor, dones: Tensor, scale: float, loss_rows: Tensor, dlogit: Tensor) -> None:
        """target t = 1 - dones (dreamer_v3.py:155); loss = -Bernoulli(logits).log_prob(t) = max(l,0) - l*t + log1p(exp(-|l|));
        dlogit = scale*(sigmoid(l)-t)."""
        l = logit.reshape(-1)
        t = 1 - dones.reshape(-1)
        loss_rows.copy_(F.binary_cross_entropy_with_logits(l, t, reduction="none"))
        dlogit.copy_((scale * (torch.sigmoid(l) - t)).reshape(dlogit.shape))

    def reduce_mean(self, x: Tensor, col: int, out: Tensor, slot: int, scale: float, accumulate: bool) -> None:
        """out[slot] (+)= scale * sum_m x[m, col]   (x is [M, C])."""
        v = scale * x[:, col].sum()
        out[slot] = out[slot] + v if accumulate else v

    # ------------------------------------------------------------------ behaviour-learning kernels
    def lambda_returns(self, values: Tensor, rewards: Tensor, cont_logit: Tensor, dones: Tensor, H: int, N: int,
                       gamma: float, lmbda: float, lam: Tensor, discount: Tensor) -> None:
        """values/rewards/cont_logit [(H+1)*N]; dones [N].  continues = Bernoulli.mode (NaN at p==0.5) with row 0
        replaced by 1-dones; lam [H*N] = lambda-returns on steps 1..H; discount [(H+1)*N] = cumprod(c*gamma)/gamma."""
        v = values.reshape(H + 1, N)
        r = rewards.reshape(H + 1, N)
        p = torch.sigmoid(cont_logit.reshape(H + 1, N))
        c = (p >= 0.5).to(v.dtype)
        c = torch.where(p == 0.5, torch.full_like(c, math.nan), c)
        c = torch.cat(((1 - dones.reshape(1, N)), c[1:]), 0)
        cg = c[1:] * gamma
        interm = r[1:] + cg * v[1:] * (1 - lmbda)
        nxt = v[H]
        out = torch.empty(H, N, dtype=v.dtype, device=v.device)
        for t in range(H - 1, -1, -1):
            nxt = interm[t] + cg[t] * lmbda * nxt
            out[t] = nxt
        lam.copy_(out.reshape(-1))
        discount.copy_((torch.cumprod(c * gamma, 0) / gamma).reshape(-1))

    def moments(self, x: Tensor, state: Tensor, out: Tensor, decay: float, max_: float, plow: float, phigh: float) -> None:
        """state = (low, high) EMA buffers; out = (offset, invscale).  Quantiles with linear interpolation
        (torch.quantile semantics)."""
        lo = torch.quantile(x.reshape(-1), plow)
        hi = torch.quantile(x.reshape(-1), phigh)
        state[0] = decay * state[0] + (1 - decay) * lo
        state[1] = decay * state[1] + (1 - decay) * hi
        out[0] = state[0]
        out[1] = torch.maximum(torch.tensor(1.0 / max_, device=x.device), state[1] - state[0])

    def actor_loss(self, raw: Tensor, head_dims, act_idx: Tensor, lam: Tensor, values: Tensor, discount: Tensor,
                   mom: Tensor, unimix: float, ent_coef: float, scale: float, loss_rows: Tensor, draw: Tensor) -> None:
        """REINFORCE + entropy objective on M rows (discrete heads), forward value and gradient wrt the raw head logits.
        loss_rows[m] = -discount[m]*(sum_heads logp(a) * adv + ent_coef * sum_heads H);  adv = (lam-off)/inv - (v-off)/inv.
        draw = scale * d(sum_m loss_rows)/d(raw)."""
        off, inv = mom[0], mom[1]
        adv = (lam - off) / inv - (values - off) / inv
        col = 0
        total = torch.zeros_like(lam)
        for i, K in enumerate(head_dims):
            r = raw[:, col:col + K]
            ps = r.softmax(-1)
            pm = (1 - unimix) * ps + unimix / K if unimix > 0 else ps
            inside = (pm >= _EPS32) & (pm <= 1 - _EPS32)
            lu = torch.log(pm.clamp(min=_EPS32, max=1 - _EPS32)) if unimix > 0 else r
            ln = lu - lu.logsumexp(-1, keepdim=True)
            p = ln.softmax(-1)
            a = act_idx[:, i].long()
            logp = ln.gather(-1, a[:, None]).squeeze(-1)
            ent = -(ln * p).sum(-1)
            total = total + logp * adv + ent_coef * ent
            onehot = F.one_hot(a, K).to(raw.dtype)
            g_lu = -(scale * discount)[:, None] * (adv[:, None] * (onehot - p) + ent_coef * (-p * (ln + ent[:, None])))
            if unimix > 0:
                gpm = torch.where(inside, g_lu / pm, torch.zeros_like(g_lu))
                gps = (1 - unimix) * gpm
                draw[:, col:col + K] = ps * (gps - (ps * gps).sum(-1, keepdim=True))
            else:
                draw[:, col:col + K] = g_lu
            col += K
        loss_rows.copy_(-discount * total)

    # ------------------------------------------------------------------ optimizer
    def adam_clip(self, p: Tensor, g: Tensor, m: Tensor, v: Tensor, step: Tensor, lr: float, beta1: float, beta2: float,
                  eps: float, max_norm: float, norm_out: Tensor, slot: int) -> None:
        """Global L2 norm over g -> norm_out[slot]; clip coef = min(1, max_norm/(norm+1e-6)); torch.optim.Adam update
        (no weight decay, bias-corrected) on the flat arena; step (device scalar) += 1."""
        norm = torch.sqrt((g.double() ** 2).sum()).float()
        norm_out[slot] = norm
        coef = torch.clamp(max_norm / (norm + 1e-6), max=1.0)
        gg = g * coef
        step += 1
        t = float(step)
        m.mul_(beta1).add_(gg, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(gg, gg, value=1 - beta2)
        bc1 = 1 - beta1 ** t
        bc2 = 1 - beta2 ** t
        denom = v.sqrt() / math.sqrt(bc2) + eps
        p.addcdiv_(m, denom, value=-lr / bc1)

    def ema(self, target: Tensor, src: Tensor, tau: float) -> None:
        """target = tau*src + (1-tau)*target  (sheeprl/algos/dreamer_v3/dreamer_v3.py:713-716)"""
        target.mul_(1 - tau).add_(src, alpha=tau)

    def scan_prep(self, first: Tensor, h_prev: Optional[Tensor], z_prev: Optional[Tensor], zreset: Tensor,
                  h_in: Tensor, z_in: Tensor) -> None:
        """Reset masking of RSSM.dynamic (agent.py:379-384): rows with first==1 get h=0, z=zreset; others copy.
        h_prev / z_prev None means t == 0 (all rows reset)."""
        f = first.reshape(-1, 1)
        if h_prev is None:
            h_in.zero_()
            z_in.copy_(zreset.expand_as(z_in))
        else:
            h_in.copy_((1 - f) * h_prev)
            z_in.copy_((1 - f) * z_prev + f * zreset)

    def scan_carry(self, first: Tensor, dz_prev: Tensor, dh_a: Tensor, dh_b: Tensor, dz_dst: Tensor, dh_dst: Tensor) -> None:
        """dz_dst += (1-first)*dz_prev ; dh_dst += (1-first)*(dh_a + dh_b)   (gradient of scan_prep into step t-1)."""
        f = first.reshape(-1, 1)
        dz_dst += (1 - f) * dz_prev
        dh_dst += (1 - f) * (dh_a + dh_b)

    def mask_shift_actions(self, actions: Tensor, first: Tensor, out: Tensor) -> None:
        """out[t] = (1-first[t]) * actions[t-1] (zeros for t=0): dreamer_v3.py:108 + agent.py:379.
        actions [T,B,A]; first [T,B]; out [T,B,A] view."""
        T = actions.shape[0]
        out[0] = 0
        if T > 1:
            out[1:] = (1 - first[1:, :, None]) * actions[:-1]

    def uniform_fill(self, out: Tensor, seed: int, counter: Tensor) -> None:
        """Fill with U[0,1) from a counter-based generator; counter (device scalar) is advanced.  (The torch
        statement uses torch's generator: streams differ from the CUDA kernel by design; tests inject noise.)"""
        g = torch.Generator(device=out.device).manual_seed(int(seed) + int(counter))
        out.copy_(torch.rand(out.shape, generator=g, device=out.device))
        counter += 1
