"""Environment-interaction step of the agent on the training step's C-ABI kernels.

Reference: `PlayerDV3` (sheeprl/algos/dreamer_v3/agent.py:458-599): per env step, for `num_envs` rows,
    embedded = encoder(obs); h = recurrent_model([z, a], h); z ~ representation(h, embedded); a ~ actor([z, h])
(`get_greedy_action`, agent.py:556-599) and `init_states` (agent.py:512-535: h = 0, a = 0, z = mode of transition(0)).

The weights are the training engine's arenas (no copies: a gradient step is visible to the next env step), the state lives on the
device, and the ~25 launches of a step are captured once as a CUDA graph.  Rows are few (num_envs), so the contractions run on the
fp32 CUDA-core kernels (exact parity with the eager module; tensor cores would idle at 4 rows).  Draws follow the library's contract:
uniforms from `dv3_uniform_fill`, idx = #{k : cumsum_k(probs) <= u}.
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Sequence, Tuple

import torch

from . import _lib
from .engine import DenseStack, DV3Engine, _stack_spec

Tensor = torch.Tensor


class PlayerStep:
    def __init__(self, eng: DV3Engine, num_envs: int, seed: int = 0):
        if eng.device.type != "cuda":
            raise RuntimeError("PlayerStep needs the CUDA engine; there is no CPU path")
        self.eng, self.n = eng, num_envs
        dev = eng.device
        self.ops = ops = _lib.CudaOps(dev, scratch_bytes=4 << 20, precision="fp32")
        n, S, D, U, Hh, E, A, lat = num_envs, eng.S, eng.D, eng.U, eng.Hh, eng.E, eng.A, eng.lat
        f32 = dict(dtype=torch.float32, device=dev)
        z = lambda *s: torch.zeros(*s, **f32)  # noqa: E731
        P = eng.wm.p
        w, a = eng.cfg.algo.world_model, eng.cfg.algo
        self.obs_cat, self.obs_sym = z(n, max(eng.Od, 1)), z(n, max(eng.Od, 1))
        self.x_mlp, self.x_gru, self.x_rep, self.latent = z(n, S + A), z(n, D + U), z(n, D + E), z(n, lat)
        self.feat_pre, self.gates_pre, self.hidr_pre, self.hidr, self.post_raw = z(n, U), z(n, 3 * D), z(n, Hh), z(n, Hh), z(n, S)
        self.stat = z(6, n)
        self.z_idx = torch.zeros(n, eng.G, dtype=torch.int32, device=dev)
        nh = len(eng.actions_dim)
        self.act_idx = torch.zeros(n, nh, dtype=torch.int32, device=dev)
        self.acts = z(n, A)
        self.u_z, self.u_a = z(n, eng.G), z(n, nh)
        self.seed = seed
        self.rng_counter = torch.zeros(1, dtype=torch.int64, device=dev)
        self.zero_h, self.zreset, self.zreset_logits = z(1, D), z(1, S), z(1, S)
        self.zreset_idx = torch.zeros(1, eng.G, dtype=torch.int32, device=dev)
        self.enc = self.cnn_enc = None
        if eng.keys:
            enc_l, _ = _stack_spec(P, None, "encoder.mlp_encoder.model._model.", w.encoder.mlp_layers, 1e-3, [])
            self.enc = DenseStack(ops, dev, n, eng.Od, enc_l, [], last_act_out=self.x_rep[:, D + eng.cnn_out:], need_bwd=False)
        if eng.cnn_keys:
            from .conv_stack import ConvEncoder

            ep = "encoder.cnn_encoder.model.0._model."
            el = [dict(w=P[f"{ep}{3 * i}.weight"], g=P[f"{ep}{3 * i + 1}.weight"], beta=P[f"{ep}{3 * i + 1}.bias"]) for i in range(4)]
            self.cnn_enc = ConvEncoder(ops, dev, n, eng.img_channels, eng.img_hw, el, self.x_rep[:, D:D + eng.cnn_out], D + E)
        tr_l, tr_h = _stack_spec(P, None, "rssm.transition_model._model.", 1, 1e-5, ["rssm.transition_model._model.3"])
        self.trans0 = DenseStack(ops, dev, 1, D, tr_l, tr_h, need_bwd=False)
        ac_l, ac_h = _stack_spec(eng.actor.p, None, "model._model.", a.actor.mlp_layers, 1e-3, [f"mlp_heads.{i}" for i in range(nh)])
        self.actor = DenseStack(ops, dev, n, lat, ac_l, ac_h, need_bwd=False)
        self._graphs: Dict[bool, "torch.cuda.CUDAGraph"] = {}
        self.use_graph = os.environ.get("DV3_PLAYER_GRAPH", "1") != "0"
        self.init_states()

    # state views in the reference's shapes ([1, num_envs, .], agent.py:527-535)
    @property
    def recurrent_state(self) -> Tensor:
        return self.x_rep[:, :self.eng.D].unsqueeze(0)

    @property
    def stochastic_state(self) -> Tensor:
        return self.latent[:, :self.eng.S].unsqueeze(0)

    @property
    def actions(self) -> Tensor:
        return self.acts.unsqueeze(0)

    def init_states(self, reset_envs: Optional[Sequence[int]] = None) -> None:
        eng, ops = self.eng, self.ops
        self.trans0.forward(self.zero_h)                      # z_reset = mode(unimix(transition(0))) with the current weights
        ops.unimix_fwd(self.trans0.out, eng.G, eng.K, eng.cfg.algo.unimix, self.zreset_logits)
        ops.sample_onehot(self.zreset_logits, eng.G, eng.K, None, None, 1, self.zreset_idx, self.zreset)
        if reset_envs is None or len(reset_envs) == 0:
            self.x_rep[:, :eng.D].zero_()
            self.acts.zero_()
            self.latent[:, :eng.S] = self.zreset
        else:
            idx = torch.as_tensor(list(reset_envs), dtype=torch.long, device=eng.device)
            self.x_rep[idx, :eng.D] = 0.0
            self.acts[idx] = 0.0
            self.latent[idx, :eng.S] = self.zreset

    def _load_obs(self, obs: Dict[str, Tensor]) -> None:
        eng, n = self.eng, self.n
        col = 0
        for k, d in zip(eng.keys, eng.obs_dims):
            self.ops.copy2d(obs[k].reshape(n, d).to(torch.float32), self.obs_cat[:, col:col + d])
            col += d
        if self.cnn_enc is not None:
            imgs = [obs[k] for k in eng.cnn_keys]
            if all(i.dtype == torch.uint8 for i in imgs):
                self.cnn_enc.load_images(imgs)
            else:       # already scaled to [0, 1] by the caller (dreamer_v3.py:577): NCHW float -> channel-last
                off = 0
                hw = eng.img_hw * eng.img_hw
                for img, ck in zip(imgs, eng.img_channels):
                    self.cnn_enc.x0.view(n, hw, self.cnn_enc.C0)[:, :, off:off + ck] = img.reshape(n, ck, hw).transpose(1, 2).to(torch.float32)
                    off += ck

    def _body(self, is_training: bool) -> None:
        eng, ops = self.eng, self.ops
        S, D, G, K, mix = eng.S, eng.D, eng.G, eng.K, eng.cfg.algo.unimix
        ops.uniform_fill(self.u_z, self.seed, self.rng_counter)
        ops.uniform_fill(self.u_a, self.seed, self.rng_counter)
        if self.cnn_enc is not None:
            self.cnn_enc.forward()
        if self.enc is not None:
            ops.symlog(self.obs_cat, self.obs_sym)
            self.enc.forward(self.obs_sym)
        ops.copy2d(self.latent[:, :S], self.x_mlp[:, :S])
        ops.copy2d(self.acts, self.x_mlp[:, S:])
        ops.copy2d(self.x_rep[:, :D], self.x_gru[:, :D])
        ops.gemm("NT", self.x_mlp, eng.Wmlp, self.feat_pre)
        ops.ln_act_fwd(self.feat_pre, eng.g_mlp, eng.b_mlp, 1e-3, 1, self.x_gru[:, D:], self.stat[0], self.stat[1])
        ops.gemm("NT", self.x_gru, eng.Wgru, self.gates_pre)
        ops.gru_gate_fwd(self.gates_pre, eng.g_gru, eng.b_gru, self.x_gru[:, :D], self.x_rep[:, :D], self.stat[2], self.stat[3])
        ops.gemm("NT", self.x_rep, eng.Wr0, self.hidr_pre, bias=eng.br0)
        ops.ln_act_fwd(self.hidr_pre, eng.g_r, eng.b_r, 1e-5, 1, self.hidr, self.stat[4], self.stat[5])
        ops.gemm("NT", self.hidr, eng.Wr1, self.post_raw, bias=eng.br1)
        ops.unimix_sample(self.post_raw, G, K, mix, self.u_z, None, self.z_idx, self.latent[:, :S])
        ops.copy2d(self.x_rep[:, :D], self.latent[:, S:])
        self.actor.forward(self.latent)
        col = 0
        for j, k in enumerate(eng.actions_dim):
            raw = self.actor.out[:, col:col + k]
            if is_training:
                ops.unimix_sample(raw, 1, k, mix, self.u_a[:, j:j + 1], None, self.act_idx[:, j:j + 1], self.acts[:, col:col + k])
            else:       # mode of the unimix distribution = argmax of the raw logits
                ops.sample_onehot(raw, 1, k, None, None, 1, self.act_idx[:, j:j + 1], self.acts[:, col:col + k])
            col += k

    def step(self, obs: Dict[str, Tensor], is_training: bool = True, mask=None) -> Tuple[Tensor, ...]:
        if mask is not None:
            raise NotImplementedError("action masks (MineDojo) are outside the accelerated path")
        self._load_obs(obs)
        g = self._graphs.get(is_training) if self.use_graph else None
        if g is not None:
            g.replay()
        else:
            self._body(is_training)                          # the first step of each mode runs eagerly ...
            if self.use_graph:                               # ... and is then captured (capture enqueues nothing) for the later ones
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._body(is_training)
                self._graphs[is_training] = g
        return self._actions()

    def _actions(self) -> Tuple[Tensor, ...]:
        out, col = [], 0
        for k in self.eng.actions_dim:
            out.append(self.acts[:, col:col + k].unsqueeze(0))
            col += k
        return tuple(out)
