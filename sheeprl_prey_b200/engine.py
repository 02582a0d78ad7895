"""Step engine: the DreamerV3 training step (sheeprl/algos/dreamer_v3/dreamer_v3.py:51-351 of the
reference) restructured for the GPU and expressed as a fixed sequence of calls into the CUDA library.

Nothing here does arithmetic: every tensor op is a call on `ops`, whose methods are the C-ABI entry
points of libdv3b200.so (sheeprl_prey_b200/_lib.py; include/dv3_b200.h).  torch is used for device
memory only.  There is no autograd: the backward pass (BPTT through the RSSM scan included) is written
out by hand, which is what lets the whole step be captured as one CUDA graph (no host syncs, static
buffers, device-side scalars).

Restructurings relative to the reference (each keeps results within fp32 round-off):
  * the reset branch of RSSM.dynamic (agent.py:380-384) is the weight-only constant
    z_reset = mode(unimix(transition(0))), computed once per step;
  * the prior (transition) head does not feed the recurrence in the observe loop, so it is evaluated
    for all T*B rows after the scan as one batched stack (the reference evaluates it per step, twice);
  * weight gradients of everything inside the scan are accumulated after the reverse scan as single
    K = T*B contractions over saved per-step activations; only dh / dz propagation is sequential;
  * discrete actions: the imagination rollout keeps no backward state (dreamer_v3.py:303-312 detaches
    everything that would use it); the actor / critic activations computed during the rollout and the
    value pass are reused for the policy / value losses instead of re-running the networks
    (the reference's second actor pass, dreamer_v3.py:293, and third critic pass, :328, recompute the
    same values with the same weights).
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Dict, List, Optional, Sequence

import torch

Tensor = torch.Tensor

METRIC_KEYS = [
    "Loss/world_model_loss", "Loss/observation_loss", "Loss/reward_loss", "Loss/state_loss", "Loss/continue_loss",
    "State/kl", "State/post_entropy", "State/prior_entropy", "Grads/world_model",
    "Loss/policy_loss", "Grads/actor", "Loss/value_loss", "Grads/critic",
]
_SLOT = {k: i for i, k in enumerate(METRIC_KEYS)}


def _round_up(n: int, m: int) -> int:
    return (n + m - 1) // m * m


def _mirror(ops, *tensors: Tensor) -> None:
    """Asks the ops object for a bf16 mirror of each buffer (CudaOps in bf16 precision; see CudaOps.mirror).  Contractions whose
    operands are both mirrored read bf16 through TMA; every writer of a mirrored buffer either has a fused bf16 side output or
    is followed by `_refresh`."""
    fn = getattr(ops, "mirror", None)
    if fn is not None:
        for t in tensors:
            fn(t)


def _refresh(ops, *views: Tensor) -> None:
    fn = getattr(ops, "refresh", None)
    if fn is not None:
        for v in views:
            fn(v)


class Arena:
    """One flat fp32 buffer holding every parameter of an optimizer group (+ grad, Adam m / v, step).
    Each named parameter is a view at a 256-byte-aligned offset."""

    def __init__(self, shapes: "OrderedDict[str, Sequence[int]]", device, with_opt: bool = True):
        self.offsets: Dict[str, tuple] = OrderedDict()
        off = 0
        for name, shape in shapes.items():
            n = 1
            for s in shape:
                n *= int(s)
            self.offsets[name] = (off, n, tuple(shape))
            off += _round_up(n, 64)
        self.size = off
        self.numel = sum(n for _, n, _ in self.offsets.values())
        self.flat = torch.zeros(off, dtype=torch.float32, device=device)
        self.p = OrderedDict((k, self.flat[o:o + n].view(s)) for k, (o, n, s) in self.offsets.items())
        if with_opt:
            self.grad = torch.zeros_like(self.flat)
            self.m = torch.zeros_like(self.flat)
            self.v = torch.zeros_like(self.flat)
            self.step = torch.zeros(1, dtype=torch.float32, device=device)
            self.g = OrderedDict((k, self.grad[o:o + n].view(s)) for k, (o, n, s) in self.offsets.items())

    def load(self, sd: Dict[str, Tensor]) -> None:
        for k, v in self.p.items():
            v.copy_(sd[k].to(v.device, v.dtype))


class DenseStack:
    """L x [Linear(+bias) -> LayerNorm -> SiLU] followed by zero or more Linear heads written side by side into one
    output matrix (reference: sheeprl/models/models.py:15-118 `MLP`, plus the per-key / per-action heads of
    MLPDecoder agent.py:253-257 and Actor agent.py:672-697).  Saves pre-LN activations and LN statistics for the
    hand-written backward."""

    def __init__(self, ops, device, rows: int, in_dim: int, layers: List[dict], heads: List[dict],
                 last_act_out: Optional[Tensor] = None, need_bwd: bool = True):
        self.ops, self.rows, self.in_dim = ops, rows, in_dim
        self.layers, self.heads = layers, heads
        f32 = dict(dtype=torch.float32, device=device)
        self.pre, self.act, self.mean, self.rstd = [], [], [], []
        umax = 1
        for i, L in enumerate(layers):
            U = L["w"].shape[0]
            umax = max(umax, U)
            self.pre.append(torch.empty(rows, U, **f32))
            if i == len(layers) - 1 and last_act_out is not None:
                self.act.append(last_act_out)
            else:
                self.act.append(torch.empty(rows, U, **f32))
            self.mean.append(torch.empty(rows, **f32))
            self.rstd.append(torch.empty(rows, **f32))
        self.out_dim = sum(h["w"].shape[0] for h in heads)
        self.out = torch.empty(rows, self.out_dim, **f32) if heads else None
        _mirror(ops, *[a for a in self.act if a is not last_act_out])
        if need_bwd:
            self._dact = torch.empty(rows * umax, **f32)
            self._dpre = torch.empty(rows * umax, **f32)
            self._dln = torch.empty(rows * umax, **f32)
            _mirror(ops, self._dpre)

    def forward(self, x: Tensor, r0: int = 0, r1: Optional[int] = None) -> None:
        """Rows [r0, r1) of the stack's buffers <- stack(x); x is [r1-r0, in_dim]."""
        ops = self.ops
        r1 = self.rows if r1 is None else r1
        inp = x
        fuse = getattr(ops, "gemm_ln", None)
        for i, L in enumerate(self.layers):
            if fuse is None or not fuse(inp, L["w"], L.get("b"), L["g"], L["beta"], L["eps"], 1, self.pre[i][r0:r1], self.act[i][r0:r1],
                                        self.mean[i][r0:r1], self.rstd[i][r0:r1]):
                ops.gemm("NT", inp, L["w"], self.pre[i][r0:r1], bias=L.get("b"))
                ops.ln_act_fwd(self.pre[i][r0:r1], L["g"], L["beta"], L["eps"], 1, self.act[i][r0:r1],
                               self.mean[i][r0:r1], self.rstd[i][r0:r1])
            inp = self.act[i][r0:r1]
        col = 0
        for h in self.heads:
            k = h["w"].shape[0]
            ops.gemm("NT", inp, h["w"], self.out[r0:r1, col:col + k], bias=h["b"])
            col += k

    def backward(self, x: Tensor, rows: int, dout: Tensor, dx: Optional[Tensor] = None, dx_accumulate: bool = False) -> None:
        """Gradients of rows [0, rows).  dout = dL/d(out) if the stack has heads, else dL/d(last activation).
        Parameter gradients overwrite L['dw'] etc.; dx (optional) receives / accumulates dL/dx."""
        ops = self.ops
        nl = len(self.layers)
        if self.heads:
            U = self.layers[-1]["w"].shape[0]
            dact = self._dact[: rows * U].view(rows, U)
            col = 0
            for j, h in enumerate(self.heads):
                k = h["w"].shape[0]
                d = dout[:rows, col:col + k]
                ops.gemm("TN", d, self.act[-1][:rows], h["dw"])
                ops.colsum(d, h["db"])
                ops.gemm("NN", d, h["w"], dact, accumulate=j > 0)
                col += k
        else:
            dact = dout[:rows]
        for i in range(nl - 1, -1, -1):
            L = self.layers[i]
            U = L["w"].shape[0]
            dpre = self._dpre[: rows * U].view(rows, U)
            dln = self._dln[: rows * U].view(rows, U)
            ops.ln_act_bwd_fused(dact, self.pre[i][:rows], L["g"], L["beta"], self.mean[i][:rows], self.rstd[i][:rows], 1, dpre, dln,
                                 L["dg"], L["dbeta"], L["db"] if L.get("b") is not None else None)
            inp = self.act[i - 1][:rows] if i > 0 else x[:rows]
            ops.gemm("TN", dpre, inp, L["dw"])
            if i > 0:
                Up = self.layers[i - 1]["w"].shape[0]
                dact = self._dact[: rows * Up].view(rows, Up)
                ops.gemm("NN", dpre, L["w"], dact)
            elif dx is not None:
                ops.gemm("NN", dpre, L["w"], dx, accumulate=dx_accumulate)


def _stack_spec(P: Dict[str, Tensor], G: Optional[Dict[str, Tensor]], prefix: str, n_layers: int, eps: float,
                head_keys: Sequence[str]):
    """Builds the layer / head dicts of a DenseStack from state_dict-style names."""
    layers, heads = [], []
    for i in range(n_layers):
        wk, bk = f"{prefix}{3 * i}.weight", f"{prefix}{3 * i}.bias"
        gk, bek = f"{prefix}{3 * i + 1}.weight", f"{prefix}{3 * i + 1}.bias"
        L = dict(w=P[wk], b=P.get(bk), g=P[gk], beta=P[bek], eps=eps)
        if G is not None:
            L.update(dw=G[wk], db=G.get(bk), dg=G[gk], dbeta=G[bek])
        layers.append(L)
    for hk in head_keys:
        h = dict(w=P[hk + ".weight"], b=P[hk + ".bias"])
        if G is not None:
            h.update(dw=G[hk + ".weight"], db=G[hk + ".bias"])
        heads.append(h)
    return layers, heads


class DV3Engine:
    def __init__(self, cfg, obs_dims: Dict[str, int], actions_dim: Sequence[int], wm: Arena, actor: Arena,
                 critic: Arena, target: Arena, ops, device, dist=None, seed: int = 0):
        a, w = cfg.algo, cfg.algo.world_model
        self.cfg, self.ops, self.device, self.dist = cfg, ops, device, dist
        self.wm, self.actor, self.critic, self.target = wm, actor, critic, target
        # first offset of the parameters whose gradients are complete before the reverse scan (heads at the tail of the arena)
        tail = [o for k, (o, n, shp) in wm.offsets.items() if k.startswith(("observation_model.", "reward_model.", "continue_model."))]
        head_max = max([o + n for k, (o, n, shp) in wm.offsets.items()
                        if not k.startswith(("observation_model.", "reward_model.", "continue_model."))] or [0])
        self.wm_tail_from = min(tail) if tail and min(tail) >= head_max else 0
        # first offset of the RSSM parameters (recurrent / representation / transition model): their gradients are complete after the
        # scan's weight-gradient contractions, before the encoder's backward -- the second early reduction (everything in front of
        # it must be encoder parameters)
        rssm = [o for k, (o, n, shp) in wm.offsets.items() if k.startswith("rssm.")]
        enc_max = max([o + n for k, (o, n, shp) in wm.offsets.items() if k.startswith("encoder.")] or [0])
        self.wm_mid_from = min(rssm) if rssm and self.wm_tail_from > 0 and min(rssm) >= enc_max and min(rssm) > 0 else 0
        self.keys = list(cfg.mlp_keys.encoder)
        self.cnn_keys = list(cfg.cnn_keys.encoder)
        if list(cfg.cnn_keys.decoder) != self.cnn_keys or list(cfg.mlp_keys.decoder) != self.keys:
            raise NotImplementedError("decoder keys must equal encoder keys")
        shp = {k: ((v,) if isinstance(v, int) else tuple(v)) for k, v in obs_dims.items()}
        self.obs_dims = [shp[k][0] for k in self.keys]
        self.Od = sum(self.obs_dims)
        self.img_channels = [shp[k][0] for k in self.cnn_keys]
        self.img_hw = shp[self.cnn_keys[0]][-1] if self.cnn_keys else 0
        if self.cnn_keys and any(shp[k][-1] != 64 or shp[k][-2] != 64 for k in self.cnn_keys):
            raise NotImplementedError("pixel observations must be 64x64 (4 conv stages), as in every DreamerV3 config of the reference")
        self.actions_dim = list(actions_dim)
        self.A = sum(actions_dim)
        self.T, self.B = cfg.per_rank_sequence_length, cfg.per_rank_batch_size
        self.N = self.T * self.B
        self.H = a.horizon
        self.M1, self.M0 = (self.H + 1) * self.N, self.H * self.N
        self.G, self.K = w.stochastic_size, w.discrete_size
        self.S = self.G * self.K
        self.D = w.recurrent_model.recurrent_state_size
        self.U = w.recurrent_model.dense_units
        self.Hh = w.representation_model.hidden_size
        self.cnn_out = 8 * w.encoder.cnn_channels_multiplier * 16 if self.cnn_keys else 0
        self.E = self.cnn_out + (w.encoder.dense_units if self.keys else 0)
        self.lat = self.S + self.D
        self.bins = w.reward_model.bins
        self.seed = seed
        self.opt_hparams = {
            "world": dict(lr=w.optimizer.lr, eps=w.optimizer.eps, beta1=0.9, beta2=0.999),
            "actor": dict(lr=a.actor.optimizer.lr, eps=a.actor.optimizer.eps, beta1=0.9, beta2=0.999),
            "critic": dict(lr=a.critic.optimizer.lr, eps=a.critic.optimizer.eps, beta1=0.9, beta2=0.999),
        }
        T, B, N, S, D, U, Hh, E, A, lat, H, M1, M0 = (self.T, self.B, self.N, self.S, self.D, self.U, self.Hh, self.E,
                                                       self.A, self.lat, self.H, self.M1, self.M0)
        f32 = dict(dtype=torch.float32, device=device)
        i32 = dict(dtype=torch.int32, device=device)
        z = lambda *s: torch.zeros(*s, **f32)  # noqa: E731
        zp = lambda r, c: torch.zeros(r, _round_up(c, 8), **f32)[:, :c]  # noqa: E731  (row stride padded to 8: TMA-loadable mirror)
        P, Gd = wm.p, wm.g
        _mirror(ops, wm.flat, actor.flat, critic.flat, target.flat)

        # ---------------- world-model buffers
        self.obs_cat, self.obs_sym = z(N, max(self.Od, 1)), z(N, max(self.Od, 1))
        self.actions, self.rewards, self.dones, self.first = z(T, B, A), z(N), z(N), z(T, B)
        self.x_rep, self.x_gru, self.x_mlp, self.latent = z(N, D + E), z(N, D + U), zp(N, S + A), z(N, lat)
        self.feat_pre, self.feat_mean, self.feat_rstd = z(N, U), z(N), z(N)
        self.gates_pre, self.gru_mean, self.gru_rstd = z(N, 3 * D), z(N), z(N)
        self.hidr_pre, self.hidr, self.hidr_mean, self.hidr_rstd = z(N, Hh), z(N, Hh), z(N), z(N)
        self.post_raw, self.post_logits, self.prior_logits = z(N, S), z(N, S), z(N, S)
        self.post_idx = torch.zeros(N, self.G, **i32)
        self.zreset, self.zreset_logits, self.zreset_idx = z(1, S), z(1, S), torch.zeros(1, self.G, **i32)
        self.zero_h = z(1, D)
        self.kl_rows, self.obs_rows, self.rew_rows, self.cont_rows = z(N, 4), z(N, 1), z(N, 1), z(N, 1)
        self.d_rec, self.d_rew, self.d_cont = zp(N, max(self.Od, 1)), zp(N, self.bins), z(N, 1)
        self.dpost_logits, self.dprior_logits, self.dlatent = z(N, S), z(N, S), z(N, lat)
        self.dpost_raw, self.dhidr_act, self.dhidr_pre, self.dln_r = z(N, S), z(B, Hh), z(N, Hh), z(N, Hh)
        self.dgates_pre, self.dln_gru, self.dh_direct, self.dx_gru = z(N, 3 * D), z(N, 3 * D), z(B, D), z(B, D + U)
        self.dfeat_pre, self.dln_feat, self.dz_prev, self.de = z(N, U), z(N, U), z(B, S), z(N, E)
        self.W_r1T, self.W_r0hT, self.W_gruT, self.W_mlpzT = z(Hh, S), z(D, Hh), z(D + U, 3 * D), z(S, U)
        self.eW = z(N, Hh)                       # e_t . Wr0[:, D:]^T + b_r0 for all t (hoisted out of the scan)
        self.ub_ps, self.ub_coef, self.ub_pn = z(N, S), z(N, S), z(N, S)   # unimix Jacobian factors (cluster reverse scan)
        self.use_fused_bwd = os.environ.get("DV3_FUSED_BWD_SCAN", "1") != "0"
        self.aW = z(N, U)                        # a_t . Wmlp[:, S:]^T for all t (hoisted out of the scan; cluster kernel)
        self.W_mlpaT = z(A, _round_up(U, 4))[:, :U]   # Wmlp[:, S:]^T: discrete actions enter the rollout's first layer as a row gather
        self.act_head_off = [int(sum(self.actions_dim[:j])) for j in range(len(self.actions_dim))]
        self.Wmlp_z = z(U, S)                    # compact copy of Wmlp[:, :S] for the rollout (row stride S+A is not TMA-loadable)
        _mirror(ops, self.x_rep, self.x_gru, self.x_mlp._base, self.latent, self.hidr, self.Wmlp_z, self.dpost_raw, self.dhidr_pre,
                self.dgates_pre, self.dfeat_pre, self.dprior_logits, self.d_rec._base, self.d_rew._base)
        self.use_fused_scan = hasattr(ops, "rssm_observe_fwd") and os.environ.get("DV3_FUSED_SCAN", "1") != "0"
        self.metrics = z(16)
        self.u_post = z(T, B, self.G)
        nh = len(self.actions_dim)
        self.u_act0, self.u_z, self.u_act = z(N, nh), z(H, N, self.G), z(H, N, nh)
        self.rng_counter = torch.zeros(1, dtype=torch.int64, device=device)

        # ---------------- world-model stacks
        self.enc = self.dec = self.cnn_enc = self.cnn_dec = None
        if self.keys:
            enc_l, _ = _stack_spec(P, Gd, "encoder.mlp_encoder.model._model.", w.encoder.mlp_layers, 1e-3, [])
            self.enc = DenseStack(ops, device, N, self.Od, enc_l, [], last_act_out=self.x_rep[:, D + self.cnn_out:])
        if self.cnn_keys:
            from .conv_stack import ConvDecoder, ConvEncoder

            ep, dp = "encoder.cnn_encoder.model.0._model.", "observation_model.cnn_decoder.model."
            el = [dict(w=P[f"{ep}{3 * i}.weight"], g=P[f"{ep}{3 * i + 1}.weight"], beta=P[f"{ep}{3 * i + 1}.bias"],
                       dw=Gd[f"{ep}{3 * i}.weight"], dg=Gd[f"{ep}{3 * i + 1}.weight"], dbeta=Gd[f"{ep}{3 * i + 1}.bias"]) for i in range(4)]
            self.cnn_enc = ConvEncoder(ops, device, N, self.img_channels, self.img_hw, el, self.x_rep[:, D:D + self.cnn_out], D + E)
            dl = []
            for i in range(4):
                wk = f"{dp}2._model.{3 * i}"
                Ld = dict(w=P[wk + ".weight"], dw=Gd[wk + ".weight"], b=P.get(wk + ".bias"), db=Gd.get(wk + ".bias"))
                if i < 3:
                    Ld.update(g=P[f"{dp}2._model.{3 * i + 1}.weight"], beta=P[f"{dp}2._model.{3 * i + 1}.bias"],
                              dg=Gd[f"{dp}2._model.{3 * i + 1}.weight"], dbeta=Gd[f"{dp}2._model.{3 * i + 1}.bias"])
                dl.append(Ld)
            lin = dict(w=P[dp + "0.weight"], b=P[dp + "0.bias"], dw=Gd[dp + "0.weight"], db=Gd[dp + "0.bias"])
            self.cnn_dec = ConvDecoder(ops, device, N, lat, self.img_channels, lin, dl)
        tr_l, tr_h = _stack_spec(P, Gd, "rssm.transition_model._model.", 1, 1e-5, ["rssm.transition_model._model.3"])
        self.trans = DenseStack(ops, device, N, D, tr_l, tr_h)
        tr0_l, tr0_h = _stack_spec(P, None, "rssm.transition_model._model.", 1, 1e-5, ["rssm.transition_model._model.3"])
        self.trans0 = DenseStack(ops, device, 1, D, tr0_l, tr0_h, need_bwd=False)
        if self.keys:
            dec_l, dec_h = _stack_spec(P, Gd, "observation_model.mlp_decoder.model._model.", w.observation_model.mlp_layers, 1e-3,
                                       [f"observation_model.mlp_decoder.heads.{i}" for i in range(len(self.keys))])
            self.dec = DenseStack(ops, device, N, lat, dec_l, dec_h)
        Lr, Lc = w.reward_model.mlp_layers, w.discount_model.mlp_layers
        rw_l, rw_h = _stack_spec(P, Gd, "reward_model._model.", Lr, 1e-5, [f"reward_model._model.{3 * Lr}"])
        self.rew = DenseStack(ops, device, N, lat, rw_l, rw_h)
        ct_l, ct_h = _stack_spec(P, Gd, "continue_model._model.", Lc, 1e-5, [f"continue_model._model.{3 * Lc}"])
        self.cont = DenseStack(ops, device, N, lat, ct_l, ct_h)
        self.rew_img = DenseStack(ops, device, M1, lat, rw_l, rw_h, need_bwd=False)
        self.cont_img = DenseStack(ops, device, M1, lat, ct_l, ct_h, need_bwd=False)
        # RSSM pieces that live inside the scan
        self.Wmlp, self.g_mlp, self.b_mlp = (P["rssm.recurrent_model.mlp._model.0.weight"],
                                             P["rssm.recurrent_model.mlp._model.1.weight"],
                                             P["rssm.recurrent_model.mlp._model.1.bias"])
        self.Wgru, self.g_gru, self.b_gru = (P["rssm.recurrent_model.rnn.linear.weight"],
                                             P["rssm.recurrent_model.rnn.layer_norm.weight"],
                                             P["rssm.recurrent_model.rnn.layer_norm.bias"])
        rp = "rssm.representation_model._model."
        self.Wr0, self.br0, self.g_r, self.b_r, self.Wr1, self.br1 = (P[rp + "0.weight"], P[rp + "0.bias"], P[rp + "1.weight"],
                                                                      P[rp + "1.bias"], P[rp + "3.weight"], P[rp + "3.bias"])

        # ---------------- behaviour buffers / stacks
        La, Lv = a.actor.mlp_layers, a.critic.mlp_layers
        ac_l, ac_h = _stack_spec(actor.p, actor.g, "model._model.", La, 1e-3, [f"mlp_heads.{i}" for i in range(nh)])
        self.act_stack = DenseStack(ops, device, M1, lat, ac_l, ac_h)
        cr_l, cr_h = _stack_spec(critic.p, critic.g, "_model.", Lv, 1e-5, [f"_model.{3 * Lv}"])
        self.crit_stack = DenseStack(ops, device, M1, lat, cr_l, cr_h)
        tc_l, tc_h = _stack_spec(target.p, None, "_model.", Lv, 1e-5, [f"_model.{3 * Lv}"])
        self.tcrit_stack = DenseStack(ops, device, M0, lat, tc_l, tc_h, need_bwd=False)
        self.traj, self.acts = z(M1, lat), z(M1, A)
        self.img_feat_pre, self.img_feat, self.img_gates = z(N, U), z(N, U), z(N, 3 * D)
        _mirror(ops, self.traj, self.img_feat)
        self.img_mean, self.img_rstd = z(N), z(N)
        self.img_logits, self.act_logits = z(N, S), z(N, A)
        self.z_idx = torch.zeros(H, N, self.G, **i32)
        self.act_idx = torch.zeros(M1, nh, **i32)
        self.values, self.rewards_img, self.tvalues = z(M1), z(M1), z(M0)
        self.lam, self.discount = z(M0), z(M1)
        self.mom_state, self.mom_out = z(2), z(2)
        self.lam_all = z(M0 * (dist.world_size if dist is not None else 1))
        self.pl_rows, self.d_actor_raw = z(M0, 1), zp(M0, A)
        self.vl_rows, self.d_crit = z(M0, 1), zp(M0, self.bins)
        _mirror(ops, self.d_actor_raw._base, self.d_crit._base)

    # ------------------------------------------------------------------ inputs
    def load_batch(self, data: Dict[str, Tensor]) -> None:
        """data: [T,B,...] tensors on the device (the reference's `data` TensorDict, dreamer_v3.py:102-108)."""
        T, B, N = self.T, self.B, self.N
        col = 0
        for k, d in zip(self.keys, self.obs_dims):
            self.ops.copy2d(data[k].reshape(N, d), self.obs_cat[:, col:col + d])
            col += d
        if self.cnn_enc is not None:
            self.cnn_enc.load_images([data[k] for k in self.cnn_keys])
        self.ops.copy2d(data["actions"].reshape(N, self.A), self.actions.view(N, self.A))
        self.ops.copy2d(data["rewards"].reshape(N, 1), self.rewards.view(N, 1))
        self.ops.copy2d(data["dones"].reshape(N, 1), self.dones.view(N, 1))
        self.ops.copy2d(data["is_first"].reshape(N, 1), self.first.view(N, 1))
        self.first[0].fill_(1.0)           # dreamer_v3.py:104 (the caller's tensor is patched by train())

    def load_noise(self, noise: Optional[Dict[str, Tensor]]) -> None:
        if noise is None:
            for buf in (self.u_post, self.u_act0, self.u_z, self.u_act):
                self.ops.uniform_fill(buf, self.seed, self.rng_counter)
        else:
            self.u_post.copy_(noise["obs_post"])
            self.u_act0.copy_(noise["act0"])
            self.u_z.copy_(noise["img_z"])
            self.u_act.copy_(noise["img_act"])

    # ------------------------------------------------------------------ dynamic learning
    def world_forward(self, forced_post_idx: Optional[Tensor] = None) -> None:
        ops = self.ops
        T, B, N, S, D, U, E, A, lat = self.T, self.B, self.N, self.S, self.D, self.U, self.E, self.A, self.lat
        G, K, mix = self.G, self.K, self.cfg.algo.unimix
        _refresh(ops, self.wm.flat)
        # a2/a3: symlog + encoder (embedding lands in x_rep[:, D:])
        if self.cnn_enc is not None:
            self.cnn_enc.forward()
            _refresh(ops, self.x_rep[:, D:D + self.cnn_out])
        if self.enc is not None:
            ops.symlog(self.obs_cat, self.obs_sym)
            self.enc.forward(self.obs_sym)
        ops.mask_shift_actions(self.actions, self.first, self.x_mlp.unflatten(0, (T, B))[:, :, S:])
        # z_reset = mode(unimix(transition(0)))
        self.trans0.forward(self.zero_h)
        ops.unimix_fwd(self.trans0.out, G, K, mix, self.zreset_logits)
        ops.sample_onehot(self.zreset_logits, G, K, None, None, 1, self.zreset_idx, self.zreset)
        ops.transpose(self.Wmlp[:, :S], self.W_mlpzT)         # one-hot z input = row gather (cluster scan); reverse scan operand
        # a4: observe scan -- one persistent cooperative kernel when the shape fits, else the per-step composition
        fused = False
        if self.use_fused_scan:
            ops.gemm("NT", self.x_rep[:, D:], self.Wr0[:, D:], self.eW, bias=self.br0)
            ops.gemm("NT", self.x_mlp[:, S:], self.Wmlp[:, S:], self.aW)
            fused = ops.rssm_observe_fwd(self, forced_post_idx)
            if not fused:
                self.use_fused_scan = False
        for t in range(0 if not fused else T, T):
            r = slice(t * B, (t + 1) * B)
            rp = slice((t - 1) * B, t * B)
            ops.scan_prep(self.first[t], None if t == 0 else self.x_rep[rp, :D], None if t == 0 else self.latent[rp, :S],
                          self.zreset, self.x_gru[r, :D], self.x_mlp[r, :S])
            ops.gemm("NT", self.x_mlp[r], self.Wmlp, self.feat_pre[r])
            ops.ln_act_fwd(self.feat_pre[r], self.g_mlp, self.b_mlp, 1e-3, 1, self.x_gru[r, D:], self.feat_mean[r], self.feat_rstd[r])
            ops.gemm("NT", self.x_gru[r], self.Wgru, self.gates_pre[r])
            ops.gru_gate_fwd(self.gates_pre[r], self.g_gru, self.b_gru, self.x_gru[r, :D], self.x_rep[r, :D],
                             self.gru_mean[r], self.gru_rstd[r])
            ops.gemm("NT", self.x_rep[r], self.Wr0, self.hidr_pre[r], bias=self.br0)
            ops.ln_act_fwd(self.hidr_pre[r], self.g_r, self.b_r, 1e-5, 1, self.hidr[r], self.hidr_mean[r], self.hidr_rstd[r])
            ops.gemm("NT", self.hidr[r], self.Wr1, self.post_raw[r], bias=self.br1)
            ops.unimix_fwd(self.post_raw[r], G, K, mix, self.post_logits[r])
            ops.sample_onehot(self.post_logits[r], G, K, self.u_post[t].reshape(B, G),
                              None if forced_post_idx is None else forced_post_idx[r], 0,
                              self.post_idx[r], self.latent[r, :S])
        _refresh(ops, self.x_rep[:, :D], self.x_gru, self.x_mlp, self.hidr, self.latent[:, :S])   # scan products
        ops.copy2d(self.x_rep[:, :D], self.latent[:, S:])
        # prior head, batched over all T*B rows
        self.trans.forward(self.x_rep[:, :D])
        ops.unimix_fwd(self.trans.out, G, K, mix, self.prior_logits)
        # a5: decoder / reward / continue heads
        if self.dec is not None:
            self.dec.forward(self.latent)
        if self.cnn_dec is not None:
            self.cnn_dec.forward(self.latent)
        self.rew.forward(self.latent)
        self.cont.forward(self.latent)

    def world_losses(self) -> None:
        """a6: per-row losses and their gradients (scale 1/(T*B) = the .mean() of loss.py:102)."""
        ops, w, N = self.ops, self.cfg.algo.world_model, self.N
        sc = 1.0 / N
        if self.dec is not None:
            ops.symlog_mse(self.dec.out, self.obs_sym, sc, self.obs_rows.view(N), self.d_rec)
        if self.cnn_dec is not None:
            self.cnn_dec.loss(self.cnn_enc.x0, sc)
        ops.twohot_loss(self.rew.out, self.rewards, None, sc, self.rew_rows.view(N), self.d_rew, False)
        ops.bernoulli_loss(self.cont.out, self.dones, sc * w.continue_scale_factor, self.cont_rows.view(N), self.d_cont)
        ops.kl_balance(self.post_logits, self.prior_logits, self.G, self.K, w.kl_free_nats, w.kl_dynamic, w.kl_representation,
                       sc * w.kl_regularizer, self.kl_rows, self.dpost_logits, self.dprior_logits)
        _refresh(ops, self.d_rec, self.d_rew)
        m = self.metrics
        first_obs = True
        if self.dec is not None:
            ops.reduce_mean(self.obs_rows, 0, m, _SLOT["Loss/observation_loss"], sc, False)
            first_obs = False
        if self.cnn_dec is not None:
            ops.reduce_mean(self.cnn_dec.loss_rows.view(N, 1), 0, m, _SLOT["Loss/observation_loss"], sc, not first_obs)
        ops.reduce_mean(self.rew_rows, 0, m, _SLOT["Loss/reward_loss"], sc, False)
        ops.reduce_mean(self.cont_rows, 0, m, _SLOT["Loss/continue_loss"], sc * w.continue_scale_factor, False)
        ops.reduce_mean(self.kl_rows, 0, m, _SLOT["State/kl"], sc, False)
        ops.reduce_mean(self.kl_rows, 1, m, _SLOT["Loss/state_loss"], sc, False)
        ops.reduce_mean(self.kl_rows, 2, m, _SLOT["State/post_entropy"], sc, False)
        ops.reduce_mean(self.kl_rows, 3, m, _SLOT["State/prior_entropy"], sc, False)
        # rec_loss = mean(kl_regularizer*kl_loss + obs + reward + continue)  (loss.py:102); the mean is linear
        tot = _SLOT["Loss/world_model_loss"]
        ops.reduce_mean(self.kl_rows, 1, m, tot, sc * w.kl_regularizer, False)
        if self.dec is not None:
            ops.reduce_mean(self.obs_rows, 0, m, tot, sc, True)
        if self.cnn_dec is not None:
            ops.reduce_mean(self.cnn_dec.loss_rows.view(N, 1), 0, m, tot, sc, True)
        ops.reduce_mean(self.rew_rows, 0, m, tot, sc, True)
        ops.reduce_mean(self.cont_rows, 0, m, tot, sc * w.continue_scale_factor, True)

    def world_backward(self) -> None:
        ops = self.ops
        T, B, N, S, D, U, E, A, lat, Hh = self.T, self.B, self.N, self.S, self.D, self.U, self.E, self.A, self.lat, self.Hh
        G, K, mix = self.G, self.K, self.cfg.algo.unimix
        Gd = self.wm.g
        # heads -> dlatent
        acc = False
        if self.dec is not None:
            self.dec.backward(self.latent, N, self.d_rec, dx=self.dlatent, dx_accumulate=False)
            acc = True
        if self.cnn_dec is not None:
            self.cnn_dec.backward(self.latent, self.dlatent, acc)
        self.rew.backward(self.latent, N, self.d_rew, dx=self.dlatent, dx_accumulate=True)
        self.cont.backward(self.latent, N, self.d_cont, dx=self.dlatent, dx_accumulate=True)
        # prior head (batched): d prior_logits -> d raw -> transition stack -> dh
        ops.unimix_bwd(self.trans.out, G, K, mix, self.dprior_logits, None, self.dprior_logits)
        _refresh(ops, self.dprior_logits)
        self.trans.backward(self.x_rep[:, :D], N, self.dprior_logits, dx=self.dlatent[:, S:], dx_accumulate=True)
        # a15: the gradients of the decoder / reward / continue heads (the tail of the flat arena) are final here: their all-reduce
        # runs on the communication stream while the reverse scan (16 SMs) and the remaining weight gradients are computed
        self._wm_pending, self._wm_done_from = None, -1
        if self.dist is not None and self.wm_tail_from > 0:
            h = self._start_reduce(self.wm.grad[self.wm_tail_from:])
            if h is not None:
                self._wm_pending, self._wm_done_from = [h], self.wm_tail_from
        # reverse scan: only dh / dz propagate sequentially -- one cluster kernel when available, else the per-step composition
        fused_bwd = False
        if self.use_fused_bwd and hasattr(ops, "rssm_observe_bwd"):
            fused_bwd = ops.rssm_observe_bwd(self)
            if not fused_bwd:
                self.use_fused_bwd = False
        if not fused_bwd:                                     # transposed copies of the weights the per-step reverse scan contracts against
            ops.transpose(self.Wr1, self.W_r1T)
            ops.transpose(self.Wr0[:, :D], self.W_r0hT)
            ops.transpose(self.Wgru, self.W_gruT)
        for t in range(T - 1, -1, -1) if not fused_bwd else ():
            r = slice(t * B, (t + 1) * B)
            rp = slice((t - 1) * B, t * B)
            ops.unimix_bwd(self.post_raw[r], G, K, mix, self.dpost_logits[r], self.dlatent[r, :S], self.dpost_raw[r])
            ops.gemm("NT", self.dpost_raw[r], self.W_r1T, self.dhidr_act)
            ops.ln_act_bwd(self.dhidr_act, self.hidr_pre[r], self.g_r, self.b_r, self.hidr_mean[r], self.hidr_rstd[r], 1,
                           self.dhidr_pre[r], self.dln_r[r])
            ops.gemm("NT", self.dhidr_pre[r], self.W_r0hT, self.dlatent[r, S:], accumulate=True)
            ops.gru_gate_bwd(self.dlatent[r, S:], self.gates_pre[r], self.g_gru, self.b_gru, self.gru_mean[r], self.gru_rstd[r],
                             self.x_gru[r, :D], self.dgates_pre[r], self.dln_gru[r], self.dh_direct)
            ops.gemm("NT", self.dgates_pre[r], self.W_gruT, self.dx_gru)
            ops.ln_act_bwd(self.dx_gru[:, D:], self.feat_pre[r], self.g_mlp, self.b_mlp, self.feat_mean[r], self.feat_rstd[r], 1,
                           self.dfeat_pre[r], self.dln_feat[r])
            if t > 0:
                ops.gemm("NT", self.dfeat_pre[r], self.W_mlpzT, self.dz_prev)
                ops.scan_carry(self.first[t], self.dz_prev, self.dx_gru[:, :D], self.dh_direct,
                               self.dlatent[rp, :S], self.dlatent[rp, S:])
        # parameter gradients of the scan, one K = T*B contraction each
        if not fused_bwd:
            _refresh(ops, self.dpost_raw, self.dgates_pre)      # (dhidr_pre / dfeat_pre mirrors are written by ln_act_bwd)
        rpfx = "rssm.representation_model._model."
        ops.gemm("TN", self.dpost_raw, self.hidr, Gd[rpfx + "3.weight"])
        ops.colsum(self.dpost_raw, Gd[rpfx + "3.bias"])
        ops.gemm("TN", self.dhidr_pre, self.x_rep, Gd[rpfx + "0.weight"])
        ops.colsum(self.dhidr_pre, Gd[rpfx + "0.bias"])
        ops.ln_param_grad(self.dln_r, self.hidr_pre, self.hidr_mean, self.hidr_rstd, Gd[rpfx + "1.weight"], Gd[rpfx + "1.bias"])
        ops.gemm("TN", self.dgates_pre, self.x_gru, Gd["rssm.recurrent_model.rnn.linear.weight"])
        ops.ln_param_grad(self.dln_gru, self.gates_pre, self.gru_mean, self.gru_rstd,
                          Gd["rssm.recurrent_model.rnn.layer_norm.weight"], Gd["rssm.recurrent_model.rnn.layer_norm.bias"])
        ops.gemm("TN", self.dfeat_pre, self.x_mlp, Gd["rssm.recurrent_model.mlp._model.0.weight"])
        ops.ln_param_grad(self.dln_feat, self.feat_pre, self.feat_mean, self.feat_rstd,
                          Gd["rssm.recurrent_model.mlp._model.1.weight"], Gd["rssm.recurrent_model.mlp._model.1.bias"])
        # a15: the RSSM gradients are final: reduce them while the encoder's backward runs; only the encoder's bucket stays exposed
        if self._wm_pending is not None and 0 < self.wm_mid_from < self.wm_tail_from and os.environ.get("DV3_DP_MID_BUCKET", "1") != "0":
            h = self._start_reduce(self.wm.grad[self.wm_mid_from:self.wm_tail_from])
            if h is not None:
                self._wm_pending.append(h)
                self._wm_done_from = self.wm_mid_from
        # encoder
        ops.gemm("NN", self.dhidr_pre, self.Wr0[:, D:], self.de)
        if self.cnn_enc is not None:
            self.cnn_enc.backward(self.de[:, :self.cnn_out], E)
        if self.enc is not None:
            self.enc.backward(self.obs_sym, N, self.de[:, self.cnn_out:])

    def _start_reduce(self, grad: Tensor):
        """a15: launch the data-parallel mean of a (slice of a) flat gradient arena; None when there is nothing to do."""
        if self.dist is None or not hasattr(self.dist, "start_reduce_mean"):
            return None
        return self.dist.start_reduce_mean(grad)

    def _optimize(self, arena: Arena, tag: str, max_norm: float, slot: str, pending=None, done_from: int = -1) -> None:
        """a7 / a15: gradient all-reduce (data parallel), global-norm clip, Adam.  `pending`: handles of reductions started earlier
        (overlapped with independent kernels); `done_from` >= 0: arena.grad[done_from:] is covered by them, only the head is
        reduced here."""
        if self.dist is not None:
            if pending is None:
                self.dist.all_reduce_mean(arena.grad)
            else:
                if done_from > 0:
                    self.dist.all_reduce_mean(arena.grad[:done_from])
                for h in pending:
                    self.dist.finish_reduce(h)
        hp = self.opt_hparams[tag]
        if max_norm is None or max_norm <= 0:      # the reference skips clipping then (dreamer_v3.py:180-186); inf makes the coefficient clamp to 1
            max_norm = float("inf")
        self.ops.adam_clip(arena.flat, arena.grad, arena.m, arena.v, arena.step, hp["lr"], hp["beta1"], hp["beta2"], hp["eps"],
                           max_norm, self.metrics, _SLOT[slot])

    # ------------------------------------------------------------------ behaviour learning
    def imagine(self, forced: Optional[Dict[str, Tensor]] = None) -> None:
        """a8: H-step rollout from every posterior state (dreamer_v3.py:219-257), forward only."""
        ops = self.ops
        N, S, D, U, A, lat, H = self.N, self.S, self.D, self.U, self.A, self.lat, self.H
        G, K, mix = self.G, self.K, self.cfg.algo.unimix
        _refresh(ops, self.wm.flat, self.actor.flat)               # the world model was just updated (dreamer_v3.py:187)
        ops.copy2d(self.Wmlp[:, :S], self.Wmlp_z)
        ops.transpose(self.Wmlp[:, S:], self.W_mlpaT)
        gather_ok = hasattr(ops, "ln_act_fwd_gather") and len(self.actions_dim) <= 4
        ops.copy2d(self.latent, self.traj[:N])
        fused = getattr(ops, "rollout_tc", None)
        if fused is not None and fused(self, forced):       # one persistent kernel for all H steps (rollout_tc.cu)
            return

        def policy(i, u, fidx):
            r = slice(i * N, (i + 1) * N)
            self.act_stack.forward(self.traj[r], i * N, (i + 1) * N)
            col = 0
            for j, k in enumerate(self.actions_dim):
                ops.unimix_sample(self.act_stack.out[r, col:col + k], 1, k, mix, u[:, j:j + 1], None if fidx is None else fidx[r, j:j + 1],
                                  self.act_idx[r, j:j + 1], self.acts[r, col:col + k])
                col += k

        fa = None if forced is None else forced["act_idx"].reshape(self.M1, -1)
        policy(0, self.u_act0, fa)
        for i in range(1, H + 1):
            r, rp = slice(i * N, (i + 1) * N), slice((i - 1) * N, i * N)
            ops.gemm("NT", self.traj[rp, :S], self.Wmlp_z, self.img_feat_pre)
            if gather_ok:      # one-hot actions: a . Wmlp[:, S:]^T = rows of the transposed weight, added inside the LayerNorm kernel
                gather_ok = ops.ln_act_fwd_gather(self.img_feat_pre, self.g_mlp, self.b_mlp, 1e-3, 1, self.img_feat, self.img_mean,
                                                  self.img_rstd, self.W_mlpaT, self.act_idx[rp], self.act_head_off)
            if not gather_ok:
                ops.gemm("NT", self.acts[rp], self.Wmlp[:, S:], self.img_feat_pre, accumulate=True)
                ops.ln_act_fwd(self.img_feat_pre, self.g_mlp, self.b_mlp, 1e-3, 1, self.img_feat, self.img_mean, self.img_rstd)
            ops.gemm("NT", self.traj[rp, S:], self.Wgru[:, :D], self.img_gates)
            ops.gemm("NT", self.img_feat, self.Wgru[:, D:], self.img_gates, accumulate=True)
            ops.gru_gate_fwd(self.img_gates, self.g_gru, self.b_gru, self.traj[rp, S:], self.traj[r, S:], self.img_mean, self.img_rstd)
            self.trans.forward(self.traj[r, S:])
            ops.unimix_sample(self.trans.out, G, K, mix, self.u_z[i - 1], None if forced is None else forced["z_idx"][i - 1],
                              self.z_idx[i - 1], self.traj[r, :S])
            policy(i, self.u_act[i - 1], fa)

    def behaviour(self) -> None:
        ops, a = self.ops, self.cfg.algo
        N, H, M0, M1 = self.N, self.H, self.M0, self.M1
        _refresh(ops, self.critic.flat, self.target.flat)
        # a9: values / rewards / continues on the imagined states
        self.crit_stack.forward(self.traj)
        self.rew_img.forward(self.traj)
        self.cont_img.forward(self.traj)
        ops.twohot_mean(self.crit_stack.out, self.values)
        ops.twohot_mean(self.rew_img.out, self.rewards_img)
        # a10: lambda-returns + discount
        ops.lambda_returns(self.values, self.rewards_img, self.cont_img.out.view(M1), self.dones, H, N, a.gamma, a.lmbda,
                           self.lam, self.discount)
        # a11 / a15: the other ranks' lambda-values (Moments) travel while the critic's loss and backward run: neither needs them
        mo = a.actor.moments
        h_lam = None
        if self.dist is not None:
            if hasattr(self.dist, "start_all_gather") and hasattr(self.dist, "start_reduce_mean"):
                h_lam = self.dist.start_all_gather(self.lam, self.lam_all)
            else:
                self.dist.all_gather(self.lam, self.lam_all)
        overlap = h_lam is not None

        def critic_part():
            # a13: critic loss (activations of the value pass above; target critic forward)
            self.tcrit_stack.forward(self.traj[:M0])
            ops.twohot_mean(self.tcrit_stack.out, self.tvalues)
            ops.twohot_loss(self.crit_stack.out[:M0], self.lam, self.discount[:M0], 1.0 / M0, self.vl_rows.view(M0), self.d_crit, False)
            ops.twohot_loss(self.crit_stack.out[:M0], self.tvalues, self.discount[:M0], 1.0 / M0, self.vl_rows.view(M0), self.d_crit, True)
            ops.reduce_mean(self.vl_rows, 0, self.metrics, _SLOT["Loss/value_loss"], 1.0 / M0, False)
            _refresh(ops, self.d_crit)
            self.crit_stack.backward(self.traj, M0, self.d_crit)

        h_critic = None
        if overlap:        # data parallel: critic first, its reduction then overlaps the actor's loss and backward
            critic_part()
            h_critic = self._start_reduce(self.critic.grad)
            self.dist.finish_reduce(h_lam)
        src = self.lam_all if self.dist is not None else self.lam
        ops.moments(src, self.mom_state, self.mom_out, mo.decay, mo.max, mo.percentile.low, mo.percentile.high)
        # a12: actor loss on steps 0..H-1 (activations come from the rollout)
        ops.actor_loss(self.act_stack.out[:M0], self.actions_dim, self.act_idx[:M0], self.lam, self.values[:M0], self.discount[:M0],
                       self.mom_out, a.unimix, a.actor.ent_coef, 1.0 / M0, self.pl_rows.view(M0), self.d_actor_raw)
        ops.reduce_mean(self.pl_rows, 0, self.metrics, _SLOT["Loss/policy_loss"], 1.0 / M0, False)
        _refresh(ops, self.d_actor_raw)
        self.act_stack.backward(self.traj, M0, self.d_actor_raw)
        if overlap:
            h_actor = self._start_reduce(self.actor.grad)
            self._optimize(self.critic, "critic", a.critic.clip_gradients, "Grads/critic", pending=[h_critic])   # (runs while the actor's bucket travels)
            self._optimize(self.actor, "actor", a.actor.clip_gradients, "Grads/actor", pending=[h_actor])
        else:
            self._optimize(self.actor, "actor", a.actor.clip_gradients, "Grads/actor")
            critic_part()
            self._optimize(self.critic, "critic", a.critic.clip_gradients, "Grads/critic")

    # ------------------------------------------------------------------ whole step
    def step(self, forced: Optional[dict] = None) -> None:
        """One gradient step on the batch / noise currently loaded (dreamer_v3.py:102-346 order: world-model update
        BEFORE imagination; rollout with the pre-update actor; values with the pre-update critic)."""
        w = self.cfg.algo.world_model
        self.world_forward(None if forced is None else forced.get("post_idx"))
        self.world_losses()
        self.world_backward()
        self._optimize(self.wm, "world", w.clip_gradients, "Grads/world_model", pending=getattr(self, "_wm_pending", None),
                       done_from=getattr(self, "_wm_done_from", -1))
        self.imagine(None if forced is None else forced.get("imagine"))
        self.behaviour()

    def timed_step(self) -> Dict[str, float]:
        """One eager step with CUDA events between the phases (diagnostics for bench.py / DESIGN.md; milliseconds)."""
        w = self.cfg.algo.world_model
        marks = []

        def mark(name):
            e = torch.cuda.Event(enable_timing=True)
            e.record(torch.cuda.current_stream(self.device))
            marks.append((name, e))

        mark("start")
        self.world_forward()
        mark("world_forward (encoder, observe scan, prior head, decoder/reward/continue heads)")
        self.world_losses()
        mark("world_losses")
        self.world_backward()
        mark("world_backward (heads, prior head, reverse scan, weight grads, encoder)")
        self._optimize(self.wm, "world", w.clip_gradients, "Grads/world_model", pending=getattr(self, "_wm_pending", None),
                       done_from=getattr(self, "_wm_done_from", -1))
        mark("world clip+Adam")
        self.imagine()
        mark("imagination rollout (H steps x [GRU, prior, sample, actor])")
        self.behaviour()
        mark("behaviour (value/reward/continue heads, lambda, moments, actor+critic loss/backward/Adam)")
        torch.cuda.synchronize(self.device)
        return {marks[i][0]: marks[i - 1][1].elapsed_time(marks[i][1]) for i in range(1, len(marks))}

    def update_target(self, tau: float) -> None:
        self.ops.ema(self.target.flat, self.critic.flat, tau)

    def metrics_dict(self) -> Dict[str, float]:
        vals = self.metrics.detach().cpu().tolist()
        return {k: vals[i] for k, i in _SLOT.items()}
