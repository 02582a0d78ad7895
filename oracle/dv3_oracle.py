"""TEST INFRASTRUCTURE ONLY -- the checker, never the product.

CPU restatement (plain eager PyTorch fp32, autograd, torch.optim.Adam) of the reference's
DreamerV3 training step: vector ("mlp_keys") and 64x64 pixel ("cnn_keys") observations, discrete actions.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
import this file.  The product package (sheeprl_prey_b200/) never does; it raises if its CUDA
extension is missing.

Parity status: PINNED.  tests/test_oracle.py runs the unmodified reference
(`/root/reference/sheeprl/algos/dreamer_v3/*.py`, imported through oracle/ref_shim.py) and this
restatement on the same weights, batch and injected sampling noise and requires agreement to
fp32 round-off; oracle/make_golden.py stores reference outputs in tests/golden/ so that the
same check runs where /root/reference is absent (the GPU box).

Every function cites the reference lines it restates (paths relative to /root/reference/).

Sampling contract: the reference draws categoricals with torch.multinomial,
whose CPU and CUDA algorithms differ, so "identical RNG draws" is defined as *injected uniforms*:
index = #{k : cumsum_k(probs) <= u} (sequential fp32 left-to-right cumsum), clamped to K-1.
The reference is patched the same way when goldens are made (ref_shim.NoiseTape).  Draw order
(SURVEY Appendix D): per observe step prior then posterior; actor on the start states; per
imagination step transition then actor; the reference's discarded second-actor draw needs no
uniforms here because its result is unused.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# ----------------------------------------------------------------------------- small math
def symlog(x: Tensor) -> Tensor:
    """sheeprl/utils/utils.py:120-121"""
    return torch.sign(x) * torch.log(1 + torch.abs(x))


def symexp(x: Tensor) -> Tensor:
    """sheeprl/utils/utils.py:124-125"""
    return torch.sign(x) * (torch.exp(torch.abs(x)) - 1)


def inverse_cdf(probs: Tensor, u: Tensor) -> Tensor:
    """Sampling contract above (replaces torch.multinomial at sheeprl/utils/distribution.py:355-360)."""
    K = probs.shape[-1]
    c = torch.zeros_like(u)
    idx = torch.zeros(u.shape, dtype=torch.long, device=u.device)
    for k in range(K):
        c = c + probs[..., k]
        idx = idx + (c <= u).long()
    return idx.clamp_(max=K - 1)


def unimix_logits(raw: Tensor, unimix: float) -> Tensor:
    """sheeprl/algos/dreamer_v3/agent.py:390-402 (RSSM) and :741-747 (Actor); last dim = classes.
    probs_to_logits = log(clamp(p, eps, 1-eps)) (torch.distributions.utils)."""
    if unimix <= 0.0:
        return raw
    p = raw.softmax(-1)
    p = (1 - unimix) * p + unimix / p.shape[-1]
    eps = torch.finfo(p.dtype).eps
    return torch.log(p.clamp(min=eps, max=1 - eps))


def categorical_probs(logits: Tensor) -> Tensor:
    """torch Categorical(logits=l): l <- l - logsumexp(l); probs = softmax(l)
    (used through sheeprl/utils/distribution.py:308-331)."""
    ln = logits - logits.logsumexp(-1, keepdim=True)
    return ln.softmax(-1)


def st_sample(logits: Tensor, u: Optional[Tensor], forced_idx: Optional[Tensor] = None):
    """Straight-through one-hot sample: sheeprl/utils/distribution.py:390-393 and
    sheeprl/algos/dreamer_v2/utils.py:36-38.  Returns (sample with ST gradient, indices)."""
    p = categorical_probs(logits)
    idx = forced_idx if forced_idx is not None else inverse_cdf(p.detach(), u)
    onehot = F.one_hot(idx, p.shape[-1]).to(p)
    return onehot + (p - p.detach()), idx


def categorical_entropy(logits: Tensor) -> Tensor:
    """torch Categorical.entropy: -sum p * clamp(logits_normalised, min=finfo.min)."""
    ln = logits - logits.logsumexp(-1, keepdim=True)
    p = ln.softmax(-1)
    return -(ln.clamp(min=torch.finfo(ln.dtype).min) * p).sum(-1)


def kl_categorical(p_logits: Tensor, q_logits: Tensor) -> Tensor:
    """torch _kl_categorical_categorical (sheeprl/utils/distribution.py:396-398):
    t = p*(log p - log q); t[q==0]=inf; t[p==0]=0; sum over classes."""
    pl = p_logits - p_logits.logsumexp(-1, keepdim=True)
    ql = q_logits - q_logits.logsumexp(-1, keepdim=True)
    pp, qp = pl.softmax(-1), ql.softmax(-1)
    t = pp * (pl - ql)
    t = torch.where(qp == 0, torch.full_like(t, math.inf), t)
    t = torch.where(pp == 0, torch.zeros_like(t), t)
    return t.sum(-1)


def twohot_log_prob(logits: Tensor, x: Tensor, low: float = -20.0, high: float = 20.0) -> Tensor:
    """sheeprl/utils/distribution.py:253-270 with dims=1; logits [...,K], x [...,1] -> [...]."""
    K = logits.shape[-1]
    bins = torch.linspace(low, high, K, device=logits.device)
    y = symlog(x)
    below = (bins <= y).to(torch.int32).sum(-1, keepdim=True) - 1
    above = K - (bins > y).to(torch.int32).sum(-1, keepdim=True)
    below = below.clamp(0, K - 1).long()
    above = above.clamp(0, K - 1).long()
    equal = below == above
    d_below = torch.where(equal, torch.ones_like(y), (bins[below] - y).abs())
    d_above = torch.where(equal, torch.ones_like(y), (bins[above] - y).abs())
    total = d_below + d_above
    w_below, w_above = d_above / total, d_below / total
    target = F.one_hot(below.squeeze(-1), K) * w_below + F.one_hot(above.squeeze(-1), K) * w_above
    log_pred = logits - logits.logsumexp(-1, keepdim=True)
    return (target * log_pred).sum(-1)


def twohot_mean(logits: Tensor, low: float = -20.0, high: float = 20.0) -> Tensor:
    """sheeprl/utils/distribution.py:245-247 -> [...,1]"""
    bins = torch.linspace(low, high, logits.shape[-1], device=logits.device)
    return symexp((logits.softmax(-1) * bins).sum(-1, keepdim=True))


def lambda_values(rewards: Tensor, values: Tensor, continues: Tensor, lmbda: float) -> Tensor:
    """sheeprl/algos/dreamer_v3/utils.py:44-55 (reverse scan over the horizon)."""
    H = rewards.shape[0]
    interm = rewards + continues * values * (1 - lmbda)
    out: List[Tensor] = [None] * H  # type: ignore
    nxt = values[-1]
    for t in range(H - 1, -1, -1):
        nxt = interm[t] + continues[t] * lmbda * nxt
        out[t] = nxt
    return torch.stack(out, 0)


# ----------------------------------------------------------------------------- network blocks
def dense_stack(x: Tensor, P: Dict[str, Tensor], prefix: str, n_layers: int, eps: float) -> Tensor:
    """n x [Linear(+bias if present) -> LayerNorm(eps) -> SiLU]: sheeprl/models/models.py:83-99 with
    the miniblock order of sheeprl/utils/model.py:71-87."""
    for i in range(n_layers):
        w = P[f"{prefix}{3 * i}.weight"]
        b = P.get(f"{prefix}{3 * i}.bias")
        x = F.linear(x, w, b)
        x = F.layer_norm(x, (w.shape[0],), P[f"{prefix}{3 * i + 1}.weight"], P[f"{prefix}{3 * i + 1}.bias"], eps)
        x = F.silu(x)
    return x


def mlp_with_head(x: Tensor, P: Dict[str, Tensor], prefix: str, n_layers: int, eps: float) -> Tensor:
    """MLP with a final Linear (reward / continue / critic / transition / representation)."""
    x = dense_stack(x, P, prefix, n_layers, eps)
    return F.linear(x, P[f"{prefix}{3 * n_layers}.weight"], P[f"{prefix}{3 * n_layers}.bias"])


class DV3Oracle:
    """Functional restatement; parameters are plain dicts keyed by the reference's state_dict names
    (without the Fabric wrapper segment)."""

    def __init__(self, cfg, obs_dims: Dict[str, int], actions_dim: Sequence[int],
                 wm: Dict[str, Tensor], actor: Dict[str, Tensor], critic: Dict[str, Tensor],
                 target_critic: Dict[str, Tensor]):
        self.cfg = cfg
        a = cfg.algo
        w = a.world_model
        self.keys = list(cfg.mlp_keys.encoder)
        self.cnn_keys = list(cfg.cnn_keys.encoder)
        self.obs_dims = obs_dims
        self.actions_dim = list(actions_dim)
        self.A = sum(actions_dim)
        self.stoch, self.disc = w.stochastic_size, w.discrete_size
        self.S = self.stoch * self.disc
        self.D = w.recurrent_model.recurrent_state_size
        self.unimix = a.unimix
        self.L_enc = w.encoder.mlp_layers
        self.L_dec = w.observation_model.mlp_layers
        self.L_rew = w.reward_model.mlp_layers
        self.L_cont = w.discount_model.mlp_layers
        self.L_actor = a.actor.mlp_layers
        self.L_critic = a.critic.mlp_layers
        self.wm, self.actor, self.critic, self.target_critic = wm, actor, critic, target_critic

    # ---- world model pieces -------------------------------------------------------------
    @staticmethod
    def _ln_channel_last(x: Tensor, w: Tensor, b: Tensor) -> Tensor:
        """LayerNormChannelLast, sheeprl/utils/model.py:225-235 (eps 1e-3 at agent.py:67, :194)."""
        return F.layer_norm(x.permute(0, 2, 3, 1), (x.shape[1],), w, b, 1e-3).permute(0, 3, 1, 2)

    def cnn_encoder(self, obs: Dict[str, Tensor]) -> Tensor:
        """sheeprl/algos/dreamer_v3/agent.py:61-80: 4 x [Conv2d k4 s2 p1 no bias, LN channel-last, SiLU], Flatten(C,H,W)."""
        x = torch.cat([obs[k] for k in self.cnn_keys], -3)
        lead = x.shape[:-3]
        x = x.reshape(-1, *x.shape[-3:])
        p = "encoder.cnn_encoder.model.0._model."
        for i in range(4):
            x = F.conv2d(x, self.wm[f"{p}{3 * i}.weight"], None, stride=2, padding=1)
            x = F.silu(self._ln_channel_last(x, self.wm[f"{p}{3 * i + 1}.weight"], self.wm[f"{p}{3 * i + 1}.bias"]))
        return x.flatten(-3).reshape(*lead, -1)

    def cnn_decoder(self, latent: Tensor) -> Tensor:
        """sheeprl/algos/dreamer_v3/agent.py:178-206: Linear -> Unflatten(C,4,4) -> 3 x [ConvT k4 s2 p1, LN-CL, SiLU] -> ConvT(+bias) -> +0.5."""
        p = "observation_model.cnn_decoder.model."
        lead = latent.shape[:-1]
        x = F.linear(latent.reshape(-1, latent.shape[-1]), self.wm[p + "0.weight"], self.wm[p + "0.bias"])
        x = x.reshape(x.shape[0], -1, 4, 4)
        for i in range(3):
            x = F.conv_transpose2d(x, self.wm[f"{p}2._model.{3 * i}.weight"], None, stride=2, padding=1)
            x = F.silu(self._ln_channel_last(x, self.wm[f"{p}2._model.{3 * i + 1}.weight"], self.wm[f"{p}2._model.{3 * i + 1}.bias"]))
        x = F.conv_transpose2d(x, self.wm[p + "2._model.9.weight"], self.wm[p + "2._model.9.bias"], stride=2, padding=1)
        return x.reshape(*lead, *x.shape[1:]) + 0.5

    def encoder(self, obs: Dict[str, Tensor]) -> Tensor:
        """MultiEncoder sheeprl/models/models.py:452-460: cat(cnn features, mlp features).
        MLP part: agent.py:131-133 (symlog, concat, MLP no-bias, LN eps 1e-3)."""
        outs = []
        if self.cnn_keys:
            outs.append(self.cnn_encoder(obs))
        if self.keys:
            x = torch.cat([symlog(obs[k]) for k in self.keys], -1)
            outs.append(dense_stack(x, self.wm, "encoder.mlp_encoder.model._model.", self.L_enc, 1e-3))
        return torch.cat(outs, -1)

    def recurrent(self, z: Tensor, a: Tensor, h: Tensor) -> Tensor:
        """sheeprl/algos/dreamer_v3/agent.py:296-309 + LayerNormGRUCell sheeprl/models/models.py:388-395."""
        P = self.wm
        feat = dense_stack(torch.cat((z, a), -1), P, "rssm.recurrent_model.mlp._model.", 1, 1e-3)
        x = F.linear(torch.cat((h, feat), -1), P["rssm.recurrent_model.rnn.linear.weight"])
        x = F.layer_norm(x, (x.shape[-1],), P["rssm.recurrent_model.rnn.layer_norm.weight"],
                         P["rssm.recurrent_model.rnn.layer_norm.bias"], 1e-5)
        r, c, u = torch.chunk(x, 3, -1)
        r = torch.sigmoid(r)
        c = torch.tanh(r * c)
        u = torch.sigmoid(u - 1)
        return u * c + (1 - u) * h

    def transition_logits(self, h: Tensor) -> Tensor:
        """sheeprl/algos/dreamer_v3/agent.py:422-437 -> unimix'd logits [..., stoch, disc]."""
        raw = mlp_with_head(h, self.wm, "rssm.transition_model._model.", 1, 1e-5)
        return unimix_logits(raw.view(*raw.shape[:-1], self.stoch, self.disc), self.unimix)

    def representation_logits(self, h: Tensor, e: Tensor) -> Tensor:
        """sheeprl/algos/dreamer_v3/agent.py:404-420"""
        raw = mlp_with_head(torch.cat((h, e), -1), self.wm, "rssm.representation_model._model.", 1, 1e-5)
        return unimix_logits(raw.view(*raw.shape[:-1], self.stoch, self.disc), self.unimix)

    def z_reset(self) -> Tensor:
        """mode(transition(0)): the reset branch of RSSM.dynamic (agent.py:380-384) evaluated at the
        masked (zero) recurrent state; a weight-only constant.  Returns indices [stoch]."""
        with torch.no_grad():
            l = self.transition_logits(torch.zeros(1, self.D))
            return categorical_probs(l).argmax(-1)[0]

    def observe(self, embed: Tensor, actions: Tensor, is_first: Tensor, u_post: Optional[Tensor],
                forced_post_idx: Optional[Tensor] = None):
        """Loop of sheeprl/algos/dreamer_v3/dreamer_v3.py:108-130 over RSSM.dynamic (agent.py:350-388).
        embed [T,B,E]; actions [T,B,A] (un-shifted); is_first [T,B,1] (row 0 already forced to 1).
        The prior *sample* of each step is unused by train() (dreamer_v3.py:123) and is skipped."""
        T, B = embed.shape[:2]
        a_shift = torch.cat((torch.zeros_like(actions[:1]), actions[:-1]), 0)
        h = torch.zeros(B, self.D)
        z = torch.zeros(B, self.S)
        zr = F.one_hot(self.z_reset(), self.disc).float().reshape(1, self.S)
        hs, zs, idxs, post_l, prior_l = [], [], [], [], []
        for t in range(T):
            f = is_first[t]
            a = (1 - f) * a_shift[t]
            h = (1 - f) * h
            z = (1 - f) * z + f * zr
            h = self.recurrent(z, a, h)
            pl = self.transition_logits(h)
            ql = self.representation_logits(h, embed[t])
            zt, idx = st_sample(ql, None if u_post is None else u_post[t],
                                None if forced_post_idx is None else forced_post_idx[t])
            z = zt.reshape(B, self.S)
            hs.append(h); zs.append(z); idxs.append(idx); post_l.append(ql); prior_l.append(pl)
        return (torch.stack(hs), torch.stack(zs), torch.stack(idxs), torch.stack(post_l), torch.stack(prior_l))

    def decoder(self, latent: Tensor) -> Dict[str, Tensor]:
        """MultiDecoder models.py:483-489; MLPDecoder sheeprl/algos/dreamer_v3/agent.py:255-257; CNNDecoder :204-206"""
        out: Dict[str, Tensor] = {}
        if self.cnn_keys:
            rec = self.cnn_decoder(latent)
            chans = [self.obs_dims[k][0] for k in self.cnn_keys]
            out.update({k: v for k, v in zip(self.cnn_keys, torch.split(rec, chans, -3))})
        if self.keys:
            x = dense_stack(latent, self.wm, "observation_model.mlp_decoder.model._model.", self.L_dec, 1e-3)
            out.update({k: F.linear(x, self.wm[f"observation_model.mlp_decoder.heads.{i}.weight"],
                                    self.wm[f"observation_model.mlp_decoder.heads.{i}.bias"]) for i, k in enumerate(self.keys)})
        return out

    def reward_logits(self, latent: Tensor) -> Tensor:
        return mlp_with_head(latent, self.wm, "reward_model._model.", self.L_rew, 1e-5)

    def continue_logits(self, latent: Tensor) -> Tensor:
        return mlp_with_head(latent, self.wm, "continue_model._model.", self.L_cont, 1e-5)

    def critic_logits(self, latent: Tensor, P: Optional[Dict[str, Tensor]] = None) -> Tensor:
        return mlp_with_head(latent, self.critic if P is None else P, "_model.", self.L_critic, 1e-5)

    def actor_logits(self, latent: Tensor) -> List[Tensor]:
        """sheeprl/algos/dreamer_v3/agent.py:696-697,729-733: trunk (no bias, LN 1e-3) + per-head Linear + unimix."""
        x = dense_stack(latent, self.actor, "model._model.", self.L_actor, 1e-3)
        return [unimix_logits(F.linear(x, self.actor[f"mlp_heads.{i}.weight"], self.actor[f"mlp_heads.{i}.bias"]),
                              self.unimix) for i in range(len(self.actions_dim))]

    # ---- losses ---------------------------------------------------------------------------
    def world_loss(self, data: Dict[str, Tensor], u_post: Optional[Tensor], forced_post_idx=None):
        """dreamer_v3.py:102-178 + loss.py:65-110.  Returns dict with the scalar loss and intermediates."""
        w = self.cfg.algo.world_model
        obs = {k: data[k] for k in self.keys}
        obs.update({k: data[k].float() / 255.0 for k in self.cnn_keys})        # dreamer_v3.py:102
        is_first = data["is_first"].clone()
        is_first[0] = 1.0
        embed = self.encoder(obs)
        hs, zs, idx, post_l, prior_l = self.observe(embed, data["actions"], is_first, u_post, forced_post_idx)
        latent = torch.cat((zs, hs), -1)
        rec = self.decoder(latent)
        # SymlogDistribution.log_prob: sheeprl/utils/distribution.py:177-193
        obs_loss = 0
        for k in self.cnn_keys:                 # MSEDistribution(dims=3).log_prob: sheeprl/utils/distribution.py:212-221
            obs_loss = obs_loss + ((rec[k] - obs[k]) ** 2).sum((-3, -2, -1))
        for k in self.keys:
            d = (rec[k] - symlog(obs[k])) ** 2
            d = torch.where(d < 1e-8, torch.zeros_like(d), d)
            obs_loss = obs_loss + d.sum(-1)
        rew_logits = self.reward_logits(latent)
        reward_loss = -twohot_log_prob(rew_logits, data["rewards"])
        cont_logit = self.continue_logits(latent)
        cont_target = 1 - data["dones"]
        # -Bernoulli(logits).log_prob(target) summed over the last dim (Independent(.,1)); dreamer_v3.py:150-155
        continue_loss = w.continue_scale_factor * F.binary_cross_entropy_with_logits(
            cont_logit, cont_target, reduction="none").sum(-1)
        kl = kl_categorical(post_l.detach(), prior_l).sum(-1)          # loss.py:70-83
        dyn = w.kl_dynamic * torch.maximum(kl, torch.tensor(w.kl_free_nats))
        rep = kl_categorical(post_l, prior_l.detach()).sum(-1)         # loss.py:85-97
        rep = w.kl_representation * torch.maximum(rep, torch.tensor(w.kl_free_nats))
        kl_loss = dyn + rep
        rec_loss = (w.kl_regularizer * kl_loss + obs_loss + reward_loss + continue_loss).mean()
        return dict(rec_loss=rec_loss, kl=kl.mean(), state_loss=kl_loss.mean(), reward_loss=reward_loss.mean(),
                    observation_loss=obs_loss.mean(), continue_loss=continue_loss.mean(),
                    post_entropy=categorical_entropy(post_l.detach()).sum(-1).mean(),
                    prior_entropy=categorical_entropy(prior_l.detach()).sum(-1).mean(),
                    embed=embed, recurrent_states=hs, posteriors=zs, post_idx=idx,
                    post_logits=post_l, prior_logits=prior_l, latent=latent,
                    rec_obs=rec, reward_logits=rew_logits, continue_logits=cont_logit,
                    obs_loss_tb=obs_loss, reward_loss_tb=reward_loss, continue_loss_tb=continue_loss, kl_tb=kl)

    def imagine(self, z0: Tensor, h0: Tensor, u_act0: Optional[Tensor], u_z: Optional[Tensor], u_act: Optional[Tensor],
                forced=None):
        """dreamer_v3.py:219-257 (rollout over the horizon; discrete actions => no autograd state needed).
        z0 [N,S], h0 [N,D]; u_act0 [N,heads]; u_z [H,N,stoch]; u_act [H,N,heads].
        `forced` = dict(z_idx [H,N,stoch], act_idx [H+1,N,heads]) teacher-forces the draws."""
        H = self.cfg.algo.horizon
        N = z0.shape[0]
        with torch.no_grad():
            def act(latent, u, fidx):
                outs, idxs = [], []
                for i, l in enumerate(self.actor_logits(latent)):
                    s, ix = st_sample(l, None if u is None else u[:, i], None if fidx is None else fidx[:, i])
                    outs.append(s); idxs.append(ix)
                return torch.cat(outs, -1), torch.stack(idxs, -1)

            z, h = z0, h0
            traj = [torch.cat((z, h), -1)]
            a, ai = act(traj[0], u_act0, None if forced is None else forced["act_idx"][0])
            acts, act_idx, z_idx = [a], [ai], []
            for i in range(H):
                h = self.recurrent(z, a, h)
                zt, zi = st_sample(self.transition_logits(h), None if u_z is None else u_z[i],
                                   None if forced is None else forced["z_idx"][i])
                z = zt.reshape(N, self.S)
                traj.append(torch.cat((z, h), -1))
                a, ai = act(traj[-1], None if u_act is None else u_act[i],
                            None if forced is None else forced["act_idx"][i + 1])
                acts.append(a); act_idx.append(ai); z_idx.append(zi)
        return torch.stack(traj), torch.stack(acts), torch.stack(act_idx), torch.stack(z_idx)

    def behaviour_targets(self, traj: Tensor, dones: Tensor):
        """dreamer_v3.py:260-280: values / rewards / continues on the imagined states, lambda-returns, discount."""
        a = self.cfg.algo
        with torch.no_grad():
            values = twohot_mean(self.critic_logits(traj))
            rewards = twohot_mean(self.reward_logits(traj))
            cont_logit = self.continue_logits(traj)
            self.last_continue_logits_img = cont_logit    # pre-threshold value, kept for the parity tests
            cont_p = torch.sigmoid(cont_logit)
            continues = (cont_p >= 0.5).to(traj.dtype)  # Bernoulli.mode (NaN at exactly 0.5 is not reproduced)
            true_done = (1 - dones).flatten().reshape(1, -1, 1)
            continues = torch.cat((true_done, continues[1:]))
            lam = lambda_values(rewards[1:], values[1:], continues[1:] * a.gamma, a.lmbda)
            discount = torch.cumprod(continues * a.gamma, 0) / a.gamma
        return values, rewards, continues, lam, discount

    @staticmethod
    def moments_update(lam: Tensor, low: Tensor, high: Tensor, decay: float, max_: float, plow: float, phigh: float):
        """sheeprl/algos/dreamer_v3/utils.py:34-41 (world_size 1: all_gather is the identity)."""
        x = lam.detach()
        lo = torch.quantile(x, plow)
        hi = torch.quantile(x, phigh)
        new_low = decay * low + (1 - decay) * lo
        new_high = decay * high + (1 - decay) * hi
        invscale = torch.maximum(torch.tensor(1 / max_), new_high - new_low)
        return new_low, new_high, invscale

    def actor_loss(self, traj: Tensor, act_idx: Tensor, lam: Tensor, values: Tensor, discount: Tensor,
                   offset: Tensor, invscale: Tensor):
        """dreamer_v3.py:292-317 (discrete branch)."""
        a = self.cfg.algo
        logits = self.actor_logits(traj.detach())
        adv = (lam - offset) / invscale - (values[:-1] - offset) / invscale
        logp, ent = 0, 0
        for i, l in enumerate(logits):
            ln = l - l.logsumexp(-1, keepdim=True)
            logp = logp + ln.gather(-1, act_idx[..., i:i + 1])     # OneHotCategorical.log_prob (distribution.py:362-366)
            ent = ent + categorical_entropy(l)
        objective = logp[:-1] * adv.detach()
        entropy = a.actor.ent_coef * ent
        return -torch.mean(discount[:-1].detach() * (objective + entropy.unsqueeze(-1)[:-1]))

    def critic_loss(self, traj: Tensor, lam: Tensor, discount: Tensor):
        """dreamer_v3.py:328-337"""
        qv = self.critic_logits(traj.detach()[:-1])
        with torch.no_grad():
            tv = twohot_mean(self.critic_logits(traj.detach()[:-1], self.target_critic))
        loss = -twohot_log_prob(qv, lam.detach()) - twohot_log_prob(qv, tv.detach())
        return torch.mean(loss * discount[:-1].squeeze(-1))


# ----------------------------------------------------------------------------- full step
def _adam(params: Dict[str, Tensor], lr: float, eps: float):
    return torch.optim.Adam(list(params.values()), lr=lr, eps=eps, betas=(0.9, 0.999), weight_decay=0)


class OracleTrainer:
    """Holds parameters (leaf tensors), three torch Adam optimizers and the Moments buffers, and runs
    the whole reference training step (dreamer_v3.py:51-351) on CPU.  `step()` returns metrics with the
    reference's aggregator keys plus intermediates for parity checks."""

    def __init__(self, cfg, obs_dims, actions_dim, wm, actor, critic, target_critic=None):
        def leaf(d):
            return {k: v.detach().clone().float().requires_grad_(True) for k, v in d.items()}

        self.wm, self.actor, self.critic = leaf(wm), leaf(actor), leaf(critic)
        tc = critic if target_critic is None else target_critic
        self.target_critic = {k: v.detach().clone().float() for k, v in tc.items()}
        self.cfg = cfg
        self.model = DV3Oracle(cfg, obs_dims, actions_dim, self.wm, self.actor, self.critic, self.target_critic)
        a = cfg.algo
        self.world_opt = _adam(self.wm, a.world_model.optimizer.lr, a.world_model.optimizer.eps)
        self.actor_opt = _adam(self.actor, a.actor.optimizer.lr, a.actor.optimizer.eps)
        self.critic_opt = _adam(self.critic, a.critic.optimizer.lr, a.critic.optimizer.eps)
        self.low = torch.zeros(())
        self.high = torch.zeros(())

    def update_target(self, tau: float):
        """dreamer_v3.py:713-716"""
        with torch.no_grad():
            for k, v in self.critic.items():
                self.target_critic[k].mul_(1 - tau).add_(tau * v)

    @staticmethod
    def _clip(params: Dict[str, Tensor], max_norm: float) -> Tensor:
        return torch.nn.utils.clip_grad_norm_(list(params.values()), max_norm, error_if_nonfinite=False)

    def step(self, data: Dict[str, Tensor], noise: Optional[Dict[str, Tensor]] = None, forced=None,
             keep: bool = True):
        a = self.cfg.algo
        m = self.model
        noise = noise or {}
        out: Dict[str, object] = {}
        # ---- dynamic learning (dreamer_v3.py:102-187)
        self.world_opt.zero_grad(set_to_none=True)
        W = m.world_loss(data, noise.get("obs_post"), None if forced is None else forced.get("post_idx"))
        W["rec_loss"].backward()
        out["Grads/world_model"] = self._clip(self.wm, a.world_model.clip_gradients)
        if keep:
            out["world_grads"] = {k: v.grad.detach().clone() for k, v in self.wm.items()}
        self.world_opt.step()
        out.update({"Loss/world_model_loss": W["rec_loss"].detach(), "Loss/observation_loss": W["observation_loss"].detach(),
                    "Loss/reward_loss": W["reward_loss"].detach(), "Loss/state_loss": W["state_loss"].detach(),
                    "Loss/continue_loss": W["continue_loss"].detach(), "State/kl": W["kl"].detach(),
                    "State/post_entropy": W["post_entropy"], "State/prior_entropy": W["prior_entropy"]})
        # ---- behaviour learning (dreamer_v3.py:219-346); imagination uses the UPDATED world model
        N = W["posteriors"].shape[0] * W["posteriors"].shape[1]
        z0 = W["posteriors"].detach().reshape(N, -1)
        h0 = W["recurrent_states"].detach().reshape(N, -1)
        traj, acts, act_idx, z_idx = m.imagine(z0, h0, noise.get("act0"), noise.get("img_z"), noise.get("img_act"),
                                               None if forced is None else forced.get("imagine"))
        values, rewards, continues, lam, discount = m.behaviour_targets(traj, data["dones"])
        mo = a.actor.moments
        self.low, self.high, invscale = m.moments_update(lam, self.low, self.high, mo.decay, mo.max,
                                                         mo.percentile.low, mo.percentile.high)
        self.actor_opt.zero_grad(set_to_none=True)
        pl = m.actor_loss(traj, act_idx, lam, values, discount, self.low, invscale)
        pl.backward()
        out["Grads/actor"] = self._clip(self.actor, a.actor.clip_gradients)
        if keep:
            out["actor_grads"] = {k: v.grad.detach().clone() for k, v in self.actor.items()}
        self.actor_opt.step()
        out["Loss/policy_loss"] = pl.detach()
        self.critic_opt.zero_grad(set_to_none=True)
        vl = m.critic_loss(traj, lam, discount)
        vl.backward()
        out["Grads/critic"] = self._clip(self.critic, a.critic.clip_gradients)
        if keep:
            out["critic_grads"] = {k: v.grad.detach().clone() for k, v in self.critic.items()}
        self.critic_opt.step()
        out["Loss/value_loss"] = vl.detach()
        if keep:
            for k in ("embed", "recurrent_states", "posteriors", "post_idx", "post_logits", "prior_logits",
                      "obs_loss_tb", "reward_loss_tb", "continue_loss_tb", "kl_tb", "reward_logits", "continue_logits"):
                out[k] = W[k].detach()
            out.update(traj=traj, imagined_actions=acts, act_idx=act_idx, z_idx=z_idx, values=values, rewards=rewards,
                       continues=continues, continue_logits_img=m.last_continue_logits_img, lambda_values=lam, discount=discount,
                       moments_low=self.low.clone(), moments_high=self.high.clone(), invscale=invscale)
        return out
