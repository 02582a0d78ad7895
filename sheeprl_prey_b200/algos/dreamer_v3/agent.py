"""Module tree of the DreamerV3 agent with the reference's names and `build_models` signature
(sheeprl/algos/dreamer_v3/agent.py:831-1068), so that state_dict keys, checkpoints and the caller-side code that
reaches into the modules (PlayerDV3 agent.py:494-501, target-critic EMA dreamer_v3.py:713-716) keep working.

The modules are parameter containers: every nn.Parameter is a view into one of three flat fp32 arenas
(world model / actor / critic, plus a fourth for the target critic) that the CUDA step engine reads and updates
in place.  `train()` never calls these modules' forward; their plain-torch `forward` exists only for the
caller-side player / evaluation code, which is outside the accelerated path.
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn
from torch import Tensor

from ...engine import Arena, DV3Engine


# ----------------------------------------------------------------------------- init (algos/dreamer_v3/utils.py:116-159)
def init_weights(m: nn.Module) -> None:
    if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
        space = m.kernel_size[0] * m.kernel_size[1]
        denom = (space * m.in_channels + space * m.out_channels) / 2.0
        std = math.sqrt(1.0 / denom) / 0.87962566103423978
        nn.init.trunc_normal_(m.weight.data, mean=0.0, std=std, a=-2.0, b=2.0)
        if m.bias is not None:
            m.bias.data.fill_(0.0)
    elif isinstance(m, nn.Linear):
        denom = (m.in_features + m.out_features) / 2.0
        std = math.sqrt(1.0 / denom) / 0.87962566103423978
        nn.init.trunc_normal_(m.weight.data, mean=0.0, std=std, a=-2.0 * std, b=2.0 * std)
        if m.bias is not None:
            m.bias.data.fill_(0.0)
    elif isinstance(m, nn.LayerNorm):
        m.weight.data.fill_(1.0)
        m.bias.data.fill_(0.0)


def uniform_init_weights(given_scale: float):
    def f(m: nn.Module) -> None:
        if isinstance(m, nn.Linear):
            denom = (m.in_features + m.out_features) / 2.0
            limit = math.sqrt(3 * given_scale / denom)
            nn.init.uniform_(m.weight.data, a=-limit, b=limit)
            if m.bias is not None:
                m.bias.data.fill_(0.0)
        elif isinstance(m, nn.LayerNorm):
            m.weight.data.fill_(1.0)
            m.bias.data.fill_(0.0)

    return f


# ----------------------------------------------------------------------------- blocks (sheeprl/models/models.py)
class MLP(nn.Module):
    """hidden x [Linear, LayerNorm, SiLU] (+ final Linear): models.py:15-118 with the DV3 arguments."""

    def __init__(self, input_dims: int, output_dim: Optional[int], hidden_sizes: Sequence[int], bias: bool, eps: float):
        super().__init__()
        dims = [input_dims] + list(hidden_sizes)
        layers: List[nn.Module] = []
        for i, o in zip(dims[:-1], dims[1:]):
            layers += [nn.Linear(i, o, bias=bias), nn.LayerNorm(o, eps=eps), nn.SiLU()]
        if output_dim is not None:
            layers.append(nn.Linear(dims[-1], output_dim))
        self._model = nn.Sequential(*layers)
        self._output_dim = output_dim or dims[-1]

    @property
    def model(self) -> nn.Module:
        return self._model

    @property
    def output_dim(self) -> int:
        return self._output_dim

    def forward(self, x: Tensor) -> Tensor:
        return self._model(x)


class LayerNormGRUCell(nn.Module):
    """models.py:330-402"""

    def __init__(self, input_size: int, hidden_size: int):
        super().__init__()
        self.input_size, self.hidden_size = input_size, hidden_size
        self.linear = nn.Linear(input_size + hidden_size, 3 * hidden_size, bias=False)
        self.layer_norm = nn.LayerNorm(3 * hidden_size)

    def forward(self, input: Tensor, hx: Tensor) -> Tensor:
        three_d = input.dim() == 3
        if three_d:
            input, hx = input.squeeze(0), hx.squeeze(0)
        x = self.layer_norm(self.linear(torch.cat((hx, input), -1)))
        r, c, u = torch.chunk(x, 3, -1)
        r = torch.sigmoid(r)
        c = torch.tanh(r * c)
        u = torch.sigmoid(u - 1)
        hx = u * c + (1 - u) * hx
        return hx.unsqueeze(0) if three_d else hx


class LayerNormChannelLast(nn.LayerNorm):
    """sheeprl/utils/model.py:225-235"""

    def forward(self, x: Tensor) -> Tensor:
        return super().forward(x.permute(0, 2, 3, 1)).permute(0, 3, 1, 2)


class _ConvSeq(nn.Module):
    """Container with the `_model` Sequential of sheeprl.models.models.CNN / DeCNN (models.py:121-284)."""

    def __init__(self, layers: List[nn.Module]):
        super().__init__()
        self._model = nn.Sequential(*layers)

    def forward(self, x: Tensor) -> Tensor:
        return self._model(x)


class CNNEncoder(nn.Module):
    """agent.py:28-80: stages x [Conv2d(k4,s2,p1,no bias), LayerNormChannelLast(eps 1e-3), SiLU], Flatten."""

    def __init__(self, keys: Sequence[str], input_channels: Sequence[int], image_size, channels_multiplier: int, stages: int = 4):
        super().__init__()
        self.keys = list(keys)
        self.input_dim = (sum(input_channels), *image_size)
        chans = [sum(input_channels)] + [(2 ** i) * channels_multiplier for i in range(stages)]
        layers: List[nn.Module] = []
        for i, o in zip(chans[:-1], chans[1:]):
            layers += [nn.Conv2d(i, o, kernel_size=4, stride=2, padding=1, bias=False), LayerNormChannelLast(o, eps=1e-3), nn.SiLU()]
        self.model = nn.Sequential(_ConvSeq(layers), nn.Flatten(-3, -1))
        self.output_dim = chans[-1] * (image_size[0] // 2 ** stages) * (image_size[1] // 2 ** stages)

    def forward(self, obs: Dict[str, Tensor]) -> Tensor:
        x = torch.cat([obs[k] for k in self.keys], -3)
        lead = x.shape[:-3]
        return self.model(x.reshape(-1, *x.shape[-3:])).reshape(*lead, -1)


class CNNDecoder(nn.Module):
    """agent.py:136-206"""

    def __init__(self, keys: Sequence[str], output_channels: Sequence[int], channels_multiplier: int, latent_state_size: int,
                 cnn_encoder_output_dim: int, image_size, stages: int = 4):
        super().__init__()
        self.keys, self.output_channels = list(keys), list(output_channels)
        self.output_dim = (sum(output_channels), *image_size)
        chans = [(2 ** (stages - 1)) * channels_multiplier] + [(2 ** i) * channels_multiplier for i in reversed(range(stages - 1))]
        layers: List[nn.Module] = []
        for i, o in zip(chans[:-1], chans[1:]):
            layers += [nn.ConvTranspose2d(i, o, kernel_size=4, stride=2, padding=1, bias=False), LayerNormChannelLast(o, eps=1e-3), nn.SiLU()]
        layers.append(nn.ConvTranspose2d(chans[-1], self.output_dim[0], kernel_size=4, stride=2, padding=1))
        self.model = nn.Sequential(nn.Linear(latent_state_size, cnn_encoder_output_dim), nn.Unflatten(1, (-1, 4, 4)), _ConvSeq(layers))

    def forward(self, latent: Tensor) -> Dict[str, Tensor]:
        lead = latent.shape[:-1]
        out = self.model(latent.reshape(-1, latent.shape[-1])).reshape(*lead, *self.output_dim) + 0.5
        return {k: v for k, v in zip(self.keys, torch.split(out, self.output_channels, -3))}


class MLPEncoder(nn.Module):
    """agent.py:83-133"""

    def __init__(self, keys: Sequence[str], input_dims: Sequence[int], mlp_layers: int, dense_units: int):
        super().__init__()
        self.keys = list(keys)
        self.input_dim = sum(input_dims)
        self.model = MLP(self.input_dim, None, [dense_units] * mlp_layers, bias=False, eps=1e-3)
        self.output_dim = dense_units

    def forward(self, obs: Dict[str, Tensor]) -> Tensor:
        x = torch.cat([torch.sign(obs[k]) * torch.log(1 + torch.abs(obs[k])) for k in self.keys], -1)
        return self.model(x)


class MLPDecoder(nn.Module):
    """agent.py:209-257"""

    def __init__(self, keys: Sequence[str], output_dims: Sequence[int], latent_state_size: int, mlp_layers: int, dense_units: int):
        super().__init__()
        self.keys, self.output_dims = list(keys), list(output_dims)
        self.model = MLP(latent_state_size, None, [dense_units] * mlp_layers, bias=False, eps=1e-3)
        self.heads = nn.ModuleList([nn.Linear(dense_units, d) for d in self.output_dims])

    def forward(self, latent: Tensor) -> Dict[str, Tensor]:
        x = self.model(latent)
        return {k: h(x) for k, h in zip(self.keys, self.heads)}


class MultiEncoder(nn.Module):
    """models.py:405-460 (vector branch)"""

    def __init__(self, cnn_encoder, mlp_encoder):
        super().__init__()
        if cnn_encoder is None and mlp_encoder is None:
            raise ValueError("There must be at least one encoder, both cnn and mlp encoders are None")
        self.cnn_encoder, self.mlp_encoder = cnn_encoder, mlp_encoder
        self.cnn_output_dim = cnn_encoder.output_dim if cnn_encoder is not None else 0
        self.mlp_output_dim = mlp_encoder.output_dim if mlp_encoder is not None else 0
        self.output_dim = self.cnn_output_dim + self.mlp_output_dim

    def forward(self, obs):
        outs = [e(obs) for e in (self.cnn_encoder, self.mlp_encoder) if e is not None]
        return torch.cat(outs, -1)


class MultiDecoder(nn.Module):
    """models.py:463-489 (vector branch)"""

    def __init__(self, cnn_decoder, mlp_decoder):
        super().__init__()
        self.cnn_decoder, self.mlp_decoder = cnn_decoder, mlp_decoder

    def forward(self, x):
        out = {}
        if self.cnn_decoder is not None:
            out.update(self.cnn_decoder(x))
        if self.mlp_decoder is not None:
            out.update(self.mlp_decoder(x))
        return out


class RecurrentModel(nn.Module):
    """agent.py:260-309"""

    def __init__(self, input_size: int, recurrent_state_size: int, dense_units: int):
        super().__init__()
        self.mlp = MLP(input_size, None, [dense_units], bias=False, eps=1e-3)
        self.rnn = LayerNormGRUCell(dense_units, recurrent_state_size)

    def forward(self, input: Tensor, recurrent_state: Tensor) -> Tensor:
        return self.rnn(self.mlp(input), recurrent_state)


class RSSM(nn.Module):
    """agent.py:312-455 (container; the training-time scan runs in the CUDA engine)"""

    def __init__(self, recurrent_model, representation_model, transition_model, distribution_cfg, discrete=32, unimix=0.01):
        super().__init__()
        self.recurrent_model, self.representation_model, self.transition_model = recurrent_model, representation_model, transition_model
        self.discrete, self.unimix, self.distribution_cfg = discrete, unimix, distribution_cfg

    # Caller-side (player / evaluation) single-step methods in plain torch, agent.py:390-437.  Not used by train().
    def _uniform_mix(self, logits: Tensor) -> Tensor:
        logits = logits.view(*logits.shape[:-1], -1, self.discrete)
        if self.unimix > 0.0:
            probs = (1 - self.unimix) * logits.softmax(-1) + self.unimix / self.discrete
            logits = torch.distributions.utils.probs_to_logits(probs)
        return logits.view(*logits.shape[:-2], -1)

    def _stoch(self, logits: Tensor, sample: bool) -> Tensor:
        l = logits.view(*logits.shape[:-1], -1, self.discrete)
        d = torch.distributions.OneHotCategoricalStraightThrough(logits=l)
        return d.rsample() if sample else torch.nn.functional.one_hot(d.probs.argmax(-1), self.discrete).to(l)

    def _representation(self, recurrent_state: Tensor, embedded_obs: Tensor) -> Tuple[Tensor, Tensor]:
        logits = self._uniform_mix(self.representation_model(torch.cat((recurrent_state, embedded_obs), -1)))
        return logits, self._stoch(logits, True)

    def _transition(self, recurrent_out: Tensor, sample_state: bool = True) -> Tuple[Tensor, Tensor]:
        logits = self._uniform_mix(self.transition_model(recurrent_out))
        return logits, self._stoch(logits, sample_state)


class WorldModel(nn.Module):
    """sheeprl/algos/dreamer_v2/agent.py:630-655"""

    def __init__(self, encoder, rssm, observation_model, reward_model, continue_model):
        super().__init__()
        self.encoder, self.rssm, self.observation_model = encoder, rssm, observation_model
        self.reward_model, self.continue_model = reward_model, continue_model


class Actor(nn.Module):
    """agent.py:602-747 (discrete heads)"""

    def __init__(self, latent_state_size: int, actions_dim: Sequence[int], is_continuous: bool, dense_units: int, mlp_layers: int,
                 unimix: float = 0.01, distribution_cfg=None):
        super().__init__()
        if is_continuous:
            raise NotImplementedError("continuous actions are SURVEY.md section 8(f) item 1 (not built yet)")
        self.model = MLP(latent_state_size, None, [dense_units] * mlp_layers, bias=False, eps=1e-3)
        self.mlp_heads = nn.ModuleList([nn.Linear(dense_units, a) for a in actions_dim])
        self.actions_dim, self.is_continuous, self._unimix = list(actions_dim), is_continuous, unimix
        self.distribution_cfg = distribution_cfg if distribution_cfg is not None else {"validate_args": False}

    def forward(self, state: Tensor, is_training: bool = True, mask=None):
        """Caller-side policy evaluation in plain torch (agent.py:682-739, discrete branch).  Not used by train()."""
        out = self.model(state)
        actions, dists = [], []
        for head in self.mlp_heads:
            logits = head(out)
            if self._unimix > 0.0:
                probs = (1 - self._unimix) * logits.softmax(-1) + self._unimix / logits.shape[-1]
                logits = torch.distributions.utils.probs_to_logits(probs)
            d = torch.distributions.OneHotCategoricalStraightThrough(logits=logits)
            dists.append(d)
            actions.append(d.rsample() if is_training else torch.nn.functional.one_hot(d.probs.argmax(-1), logits.shape[-1]).to(logits))
        return tuple(actions), tuple(dists)


class PlayerDV3(nn.Module):
    """Environment-interaction side of the agent (agent.py:458-599): shares the trained modules' weights and steps
    encoder -> GRU -> posterior -> actor once per env step.  With modules from `build_models` on a CUDA device the step runs on the
    library's kernels (`sheeprl_prey_b200/player.py`, state on the device, one CUDA graph per step); otherwise (CPU modules) it is the
    same arithmetic in plain torch.  It is the caller of the hot path, not part of it."""

    def __init__(self, encoder, rssm: RSSM, actor, actions_dim: Sequence[int], num_envs: int, stochastic_size: int,
                 recurrent_state_size: int, device="cpu", discrete_size: int = 32):
        super().__init__()
        unwrap = lambda m: getattr(m, "module", m)  # noqa: E731
        self.encoder, self.actor = unwrap(encoder), unwrap(actor)
        self.rssm = RSSM(unwrap(rssm.recurrent_model), unwrap(rssm.representation_model), unwrap(rssm.transition_model),
                         rssm.distribution_cfg, rssm.discrete, rssm.unimix)
        self.actions_dim, self.num_envs = list(actions_dim), num_envs
        self.stochastic_size, self.discrete_size, self.recurrent_state_size = stochastic_size, discrete_size, recurrent_state_size
        self.device = torch.device(device)
        eng = getattr(unwrap(rssm), "_dv3_engine", None)
        self._cuda = None
        if eng is not None and eng.device.type == "cuda" and self.device.type == "cuda":
            from ...player import PlayerStep

            object.__setattr__(self, "_cuda", PlayerStep(eng, num_envs, seed=eng.seed + 0x5DEECE66D))

    def _sync_views(self) -> None:
        self.actions, self.recurrent_state, self.stochastic_state = self._cuda.actions, self._cuda.recurrent_state, self._cuda.stochastic_state

    @torch.no_grad()
    def init_states(self, reset_envs: Optional[Sequence[int]] = None) -> None:
        if self._cuda is not None:
            self._cuda.init_states(reset_envs)
            self._sync_views()
            return
        if reset_envs is None or len(reset_envs) == 0:
            self.actions = torch.zeros(1, self.num_envs, sum(self.actions_dim), device=self.device)
            self.recurrent_state = torch.zeros(1, self.num_envs, self.recurrent_state_size, device=self.device)
            self.stochastic_state = self.rssm._transition(self.recurrent_state, sample_state=False)[1].reshape(1, self.num_envs, -1)
        else:
            self.actions[:, reset_envs] = 0
            self.recurrent_state[:, reset_envs] = 0
            self.stochastic_state[:, reset_envs] = self.rssm._transition(
                self.recurrent_state[:, reset_envs], sample_state=False)[1].reshape(1, len(reset_envs), -1)

    @torch.no_grad()
    def get_greedy_action(self, obs: Dict[str, Tensor], is_training: bool = True, mask=None) -> Sequence[Tensor]:
        if self._cuda is not None:
            actions = self._cuda.step(obs, is_training, mask)
            self._sync_views()
            return actions
        embedded = self.encoder(obs)
        self.recurrent_state = self.rssm.recurrent_model(torch.cat((self.stochastic_state, self.actions), -1), self.recurrent_state)
        _, stoch = self.rssm._representation(self.recurrent_state, embedded)
        self.stochastic_state = stoch.view(*stoch.shape[:-2], self.stochastic_size * self.discrete_size)
        actions, _ = self.actor(torch.cat((self.stochastic_state, self.recurrent_state), -1), is_training, mask)
        self.actions = torch.cat(actions, -1)
        return actions

    def get_exploration_action(self, obs: Dict[str, Tensor], mask=None) -> Sequence[Tensor]:
        return self.get_greedy_action(obs, True, mask)      # expl_amount is 0 in every DreamerV3 config (algo/dreamer_v3.yaml:141-146)


def _arena_from(module: nn.Module, device) -> Arena:
    shapes = OrderedDict((k, tuple(p.shape)) for k, p in module.named_parameters())
    arena = Arena(shapes, device)
    for k, p in module.named_parameters():
        arena.p[k].copy_(p.data)
        p.data = arena.p[k]          # the nn.Parameter now aliases the arena
    return arena


def _mix_seed(seed: int, rank: int) -> int:
    """splitmix64(seed) ^ rank: decorrelates nearby seeds, keeps ranks of one run on distinct streams."""
    z = (seed + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return (z ^ (z >> 31)) ^ rank


def build_models(
    fabric: Any,
    actions_dim: Sequence[int],
    is_continuous: bool,
    cfg: Dict[str, Any],
    obs_space: Dict[str, Any],
    world_model_state: Optional[Dict[str, Tensor]] = None,
    actor_state: Optional[Dict[str, Tensor]] = None,
    critic_state: Optional[Dict[str, Tensor]] = None,
    target_critic_state: Optional[Dict[str, Tensor]] = None,
) -> Tuple[WorldModel, nn.Module, nn.Module, nn.Module]:
    """Same positional contract as the reference's build_models.  The CUDA library is required (there is no CPU path)."""
    w, a = cfg.algo.world_model, cfg.algo
    device = torch.device(fabric.device)
    keys = list(cfg.mlp_keys.encoder or [])
    ckeys = list(cfg.cnn_keys.encoder or [])
    obs_dims = {k: int(obs_space[k].shape[0]) for k in keys}
    obs_shapes = dict(obs_dims)
    obs_shapes.update({k: tuple(int(x) for x in obs_space[k].shape[-3:]) for k in ckeys})
    S = w.stochastic_size * w.discrete_size
    D = w.recurrent_model.recurrent_state_size
    lat = S + D
    cnn_enc = (CNNEncoder(ckeys, [obs_shapes[k][0] for k in ckeys], obs_shapes[ckeys[0]][-2:], w.encoder.cnn_channels_multiplier)
               if ckeys else None)
    mlp_enc = MLPEncoder(keys, [obs_dims[k] for k in keys], w.encoder.mlp_layers, w.encoder.dense_units) if keys else None
    enc = MultiEncoder(cnn_enc, mlp_enc)
    rec = RecurrentModel(int(sum(actions_dim) + S), D, w.recurrent_model.dense_units)
    rep = MLP(D + enc.output_dim, S, [w.representation_model.hidden_size], bias=True, eps=1e-5)
    tra = MLP(D, S, [w.transition_model.hidden_size], bias=True, eps=1e-5)
    rssm = RSSM(rec.apply(init_weights), rep.apply(init_weights), tra.apply(init_weights), cfg.distribution, w.discrete_size, a.unimix)
    cnn_dec = (CNNDecoder(list(cfg.cnn_keys.decoder), [obs_shapes[k][0] for k in cfg.cnn_keys.decoder],
                          w.observation_model.cnn_channels_multiplier, lat, cnn_enc.output_dim, obs_shapes[ckeys[0]][-2:])
               if ckeys else None)
    mlp_dec = (MLPDecoder(list(cfg.mlp_keys.decoder), [obs_dims[k] for k in cfg.mlp_keys.decoder], lat,
                          w.observation_model.mlp_layers, w.observation_model.dense_units) if keys else None)
    dec = MultiDecoder(cnn_dec, mlp_dec)
    rew = MLP(lat, w.reward_model.bins, [w.reward_model.dense_units] * w.reward_model.mlp_layers, bias=True, eps=1e-5)
    con = MLP(lat, 1, [w.discount_model.dense_units] * w.discount_model.mlp_layers, bias=True, eps=1e-5)
    world_model = WorldModel(enc.apply(init_weights), rssm, dec.apply(init_weights), rew.apply(init_weights), con.apply(init_weights))
    actor = Actor(lat, actions_dim, is_continuous, a.actor.dense_units, a.actor.mlp_layers, a.unimix, cfg.distribution)
    critic = MLP(lat, a.critic.bins, [a.critic.dense_units] * a.critic.mlp_layers, bias=True, eps=1e-5)
    actor.apply(init_weights)
    critic.apply(init_weights)
    if a.hafner_initialization:   # agent.py:1033-1043
        actor.mlp_heads.apply(uniform_init_weights(1.0))
        critic.model[-1].apply(uniform_init_weights(0.0))
        tra.model[-1].apply(uniform_init_weights(1.0))
        rep.model[-1].apply(uniform_init_weights(1.0))
        rew.model[-1].apply(uniform_init_weights(0.0))
        con.model[-1].apply(uniform_init_weights(1.0))
        if mlp_dec is not None:
            mlp_dec.heads.apply(uniform_init_weights(1.0))
    if world_model_state:
        world_model.load_state_dict(world_model_state)
    if actor_state:
        actor.load_state_dict(actor_state)
    if critic_state:
        critic.load_state_dict(critic_state)
    import copy

    target_critic = copy.deepcopy(critic)
    if target_critic_state:
        target_critic.load_state_dict(target_critic_state)

    # move everything into flat arenas on the device; parameters become views
    for m in (world_model, actor, critic, target_critic):
        m.to(device)
    arenas = [_arena_from(m, device) for m in (world_model, actor, critic, target_critic)]
    from ... import _lib

    ops = _lib.CudaOps(device)
    dist = None
    if getattr(fabric, "world_size", 1) > 1:
        from ...parallel import TorchDist

        dist = TorchDist(fabric.world_size)
    # sampling noise: counter-based generator keyed by (cfg.seed, rank) -- the reference seeds every rank through
    # fabric.seed_everything(cfg.seed) (sheeprl/cli.py); runs with different seeds must not share their training-step noise
    engine = DV3Engine(cfg, obs_shapes, list(actions_dim), *arenas, ops, device, dist=dist,
                       seed=_mix_seed(int(cfg.get("seed", 42) or 0), int(getattr(fabric, "global_rank", 0))))
    object.__setattr__(world_model, "_dv3_engine", engine)   # not a submodule / not in state_dict
    object.__setattr__(world_model.rssm, "_dv3_engine", engine)   # PlayerDV3 is handed (encoder, rssm, actor), not the world model

    # Fabric wrapping, same call sites as agent.py:1054-1063 (DDP hooks are inert: the engine all-reduces its own grads)
    wm = world_model
    wm.encoder = fabric.setup_module(wm.encoder)
    wm.observation_model = fabric.setup_module(wm.observation_model)
    wm.reward_model = fabric.setup_module(wm.reward_model)
    wm.rssm.recurrent_model = fabric.setup_module(wm.rssm.recurrent_model)
    wm.rssm.representation_model = fabric.setup_module(wm.rssm.representation_model)
    wm.rssm.transition_model = fabric.setup_module(wm.rssm.transition_model)
    wm.continue_model = fabric.setup_module(wm.continue_model)
    actor = fabric.setup_module(actor)
    critic = fabric.setup_module(critic)
    return world_model, actor, critic, target_critic
