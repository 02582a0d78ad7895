"""Plain-Python mirror of the reference's resolved Hydra config for DreamerV3.

The reference turns its resolved Hydra tree into a `dotdict` (sheeprl/utils/utils.py:13-32,
sheeprl/cli.py:28) before calling `main`/`train`; `train()` reads exactly the keys built here
(sheeprl/algos/dreamer_v3/dreamer_v3.py:95-103 and the `cfg.algo.*` uses further down).
Defaults follow sheeprl/configs/algo/dreamer_v3.yaml:1-147; the named sizes follow SURVEY.md
section 8 (the BASELINE.json configs).
"""
from __future__ import annotations

import copy
from typing import Any, Dict


class dotdict(dict):
    """dict with attribute access, nested (same behaviour as sheeprl/utils/utils.py:13-32)."""

    __getattr__ = dict.get
    __setattr__ = dict.__setitem__
    __delattr__ = dict.__delitem__

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        for k, v in list(self.items()):
            if isinstance(v, dict) and not isinstance(v, dotdict):
                self[k] = dotdict(v)

    def __getstate__(self):
        return dict(self)

    def __setstate__(self, d):
        self.update(d)

    def __deepcopy__(self, memo):
        return dotdict(copy.deepcopy(dict(self), memo))


# name -> (deter D, hidden Hh, dense units U, mlp layers L, cnn multiplier m)
SIZES = {
    "XS": (256, 256, 256, 1, 24),
    "S": (512, 512, 512, 2, 32),
    "prey": (256, 256, 1024, 2, 96),  # exp/dreamer_v3_prey.yaml:40-50 as shipped
    "M": (1024, 640, 640, 3, 48),
    "L": (2048, 768, 768, 4, 64),
    "XL": (4096, 1024, 1024, 5, 96),
}


def make_cfg(
    size: str = "S",
    batch_size: int = 16,
    sequence_length: int = 64,
    horizon: int = 15,
    mlp_keys=("state",),
    cnn_keys=(),
    deter: int | None = None,
    hidden: int | None = None,
    units: int | None = None,
    layers: int | None = None,
    cnn_mult: int | None = None,
    stochastic_size: int = 32,
    discrete_size: int = 32,
    bins: int = 255,
) -> dotdict:
    D, Hh, U, L, m = SIZES[size]
    D, Hh, U, L, m = deter or D, hidden or Hh, units or U, layers or L, cnn_mult or m
    act = "torch.nn.SiLU"

    def mlp_block():
        return {"dense_act": act, "mlp_layers": L, "layer_norm": True, "dense_units": U}

    def adam(lr, eps):
        return {"_target_": "torch.optim.Adam", "lr": lr, "eps": eps, "weight_decay": 0, "betas": [0.9, 0.999]}

    cfg: Dict[str, Any] = {
        "seed": 42,                                           # sheeprl/configs/config.yaml
        "per_rank_batch_size": batch_size,
        "per_rank_sequence_length": sequence_length,
        "distribution": {"validate_args": False, "type": "auto"},
        "cnn_keys": {"encoder": list(cnn_keys), "decoder": list(cnn_keys)},
        "mlp_keys": {"encoder": list(mlp_keys), "decoder": list(mlp_keys)},
        "env": {"screen_size": 64},
        "algo": {
            "name": "dreamer_v3",
            "gamma": 0.996996996996997,
            "lmbda": 0.95,
            "horizon": horizon,
            "layer_norm": True,
            "dense_units": U,
            "mlp_layers": L,
            "dense_act": act,
            "cnn_act": act,
            "unimix": 0.01,
            "hafner_initialization": True,
            "world_model": {
                "discrete_size": discrete_size,
                "stochastic_size": stochastic_size,
                "kl_dynamic": 0.5,
                "kl_representation": 0.1,
                "kl_free_nats": 1.0,
                "kl_regularizer": 1.0,
                "continue_scale_factor": 1.0,
                "clip_gradients": 1000.0,
                "encoder": {"cnn_channels_multiplier": m, "cnn_act": act, **mlp_block()},
                "recurrent_model": {"recurrent_state_size": D, "layer_norm": True, "dense_units": U},
                "transition_model": {"hidden_size": Hh, "dense_act": act, "layer_norm": True},
                "representation_model": {"hidden_size": Hh, "dense_act": act, "layer_norm": True},
                "observation_model": {"cnn_channels_multiplier": m, "cnn_act": act, **mlp_block()},
                "reward_model": {**mlp_block(), "bins": bins},
                "discount_model": {"learnable": True, **mlp_block()},
                "optimizer": adam(1e-4, 1e-8),
            },
            "actor": {
                "cls": "sheeprl.algos.dreamer_v3.agent.Actor",
                "ent_coef": 3e-4,
                "min_std": 0.1,
                "init_std": 0.0,
                "objective_mix": 1.0,
                **mlp_block(),
                "clip_gradients": 100.0,
                "moments": {"decay": 0.99, "max": 1.0, "percentile": {"low": 0.05, "high": 0.95}},
                "optimizer": adam(8e-5, 1e-5),
            },
            "critic": {
                **mlp_block(),
                "target_network_update_freq": 1,
                "tau": 0.02,
                "bins": bins,
                "clip_gradients": 100.0,
                "optimizer": adam(8e-5, 1e-5),
            },
            "player": {"expl_min": 0.0, "expl_amount": 0.0, "expl_decay": False, "max_step_expl_decay": 0,
                       "discrete_size": discrete_size},
        },
    }
    return dotdict(cfg)


class ObsSpace(dict):
    """Minimal stand-in for gymnasium.spaces.Dict: maps key -> object with `.shape`."""

    class _Box:
        def __init__(self, shape):
            self.shape = tuple(shape)

    def __init__(self, shapes: Dict[str, tuple]):
        super().__init__({k: ObsSpace._Box(v) for k, v in shapes.items()})
