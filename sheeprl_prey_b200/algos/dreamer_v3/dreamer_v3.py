"""`train()` with the reference's signature and side effects (sheeprl/algos/dreamer_v3/dreamer_v3.py:51-351),
executed by the CUDA step engine.  Callable from the reference's `main` loop unchanged (dreamer_v3.py:717-732)."""
from __future__ import annotations

import os
from typing import Any, Dict, Optional, Sequence

import torch

from ...engine import METRIC_KEYS, DV3Engine


def get_engine(world_model) -> DV3Engine:
    eng = getattr(world_model, "_dv3_engine", None)
    if eng is None:
        raise RuntimeError("world_model was not created by sheeprl_prey_b200.algos.dreamer_v3.build_models: "
                           "there is no PyTorch fallback for the training step")
    return eng


def _graphs_enabled(eng: DV3Engine) -> bool:
    """CUDA-graph replay of the step (the step has no host syncs, static buffers and device-side scalars).
    DV3_CUDA_GRAPH=0 disables it.  With data parallelism the NCCL all-reduces / all-gather are captured in the graph too
    (measured on 2 B200: 17.8 ms/step graphed vs 20.1 ms eager)."""
    flag = os.environ.get("DV3_CUDA_GRAPH")
    if flag is not None:
        return flag == "1"
    return eng.ops.name == "cuda"


def _graphed_step(eng: DV3Engine) -> None:
    """First call runs eagerly (warm-up: lazy module loading, NCCL setup), the second captures noise generation +
    the whole step into one CUDA graph, later calls replay it."""
    calls = getattr(eng, "_graph_calls", 0)
    eng._graph_calls = calls + 1
    if calls == 0:
        eng.load_noise(None)
        eng.step()
        return
    if getattr(eng, "_graph", None) is None:
        torch.cuda.synchronize(eng.device)
        l0 = eng.ops.launches
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            eng.load_noise(None)
            eng.step()
        eng._graph_launches = eng.ops.launches - l0     # kernels of ours inside one replay
        eng._graph = g
    eng._graph.replay()


def _adopt_optimizer(eng: DV3Engine, arena, optimizer, tag: str) -> None:
    """Make a caller-created torch.optim.Adam the owner of the engine's flat Adam state: its per-parameter
    exp_avg / exp_avg_sq become views into the arena buffers, so optimizer.state_dict() (checkpoints,
    dreamer_v3.py:781-795) stays valid while the fused kernel does the update."""
    if optimizer is None or getattr(optimizer, "_dv3_adopted", None) is arena:
        return
    group = optimizer.param_groups[0]
    b1, b2 = group.get("betas", (0.9, 0.999))
    eng.opt_hparams[tag] = dict(lr=float(group["lr"]), eps=float(group["eps"]), beta1=float(b1), beta2=float(b2))
    if group.get("weight_decay", 0) not in (0, 0.0):
        raise NotImplementedError("weight decay is not part of the reference's DreamerV3 config (optim/adam.yaml)")
    by_ptr = {v.data_ptr(): k for k, v in arena.p.items()}
    for p in group["params"]:
        name = by_ptr.get(p.data_ptr())
        if name is None:
            raise RuntimeError("optimizer parameter is not an arena view; build the models with sheeprl_prey_b200.build_models")
        off, n, shape = arena.offsets[name]
        st = optimizer.state[p]
        if "exp_avg" in st:   # resumed optimizer: import its moments
            arena.m[off:off + n].view(shape).copy_(st["exp_avg"])
            arena.v[off:off + n].view(shape).copy_(st["exp_avg_sq"])
            arena.step.fill_(float(st["step"]))
        st["exp_avg"] = arena.m[off:off + n].view(shape)
        st["exp_avg_sq"] = arena.v[off:off + n].view(shape)
        st["step"] = arena.step.view(())
    optimizer._dv3_adopted = arena


def train(
    fabric: Any,
    world_model,
    actor,
    critic,
    target_critic,
    world_optimizer,
    actor_optimizer,
    critic_optimizer,
    data: Dict[str, torch.Tensor],
    aggregator,
    cfg: Dict[str, Any],
    is_continuous: bool,
    actions_dim: Sequence[int],
    moments,
    noise: Optional[Dict[str, torch.Tensor]] = None,
) -> None:
    """One gradient step of world model, actor and critic.  Side effects as in the reference: parameters and optimizer
    states advance in place, `moments.low/high` update, `data["is_first"][0]` is overwritten with ones, the 13 metric keys
    are pushed to `aggregator`.  `noise` (optional) injects the uniforms behind every categorical draw (parity tests)."""
    if is_continuous:
        raise NotImplementedError("continuous actions: SURVEY.md section 8(f) item 1")
    eng = get_engine(world_model)
    data["is_first"][0, :] = 1.0                     # dreamer_v3.py:104
    _adopt_optimizer(eng, eng.wm, world_optimizer, "world")
    _adopt_optimizer(eng, eng.actor, actor_optimizer, "actor")
    _adopt_optimizer(eng, eng.critic, critic_optimizer, "critic")
    if moments is not None and not getattr(moments, "_dv3_bound", False):
        eng.mom_state[0:1].copy_(moments.low.reshape(1))
        eng.mom_state[1:2].copy_(moments.high.reshape(1))
        moments._dv3_bound = True
    eng.load_batch(data)
    if noise is not None or not _graphs_enabled(eng):
        eng.load_noise(noise)
        eng.step()
    else:
        _graphed_step(eng)
    if moments is not None:
        moments.low = eng.mom_state[0].clone()
        moments.high = eng.mom_state[1].clone()
    if aggregator is not None:
        m = eng.metrics
        for i, k in enumerate(METRIC_KEYS):
            aggregator.update(k, m[i])


class Moments(torch.nn.Module):
    """State holder with the reference's buffers (algos/dreamer_v3/utils.py:16-41); the percentile / EMA arithmetic runs
    in dv3_moments inside the step."""

    def __init__(self, fabric=None, decay: float = 0.99, max_: float = 1e8, percentile_low: float = 0.05, percentile_high: float = 0.95):
        super().__init__()
        self._decay, self._max, self._percentile_low, self._percentile_high = decay, max_, percentile_low, percentile_high
        self.register_buffer("low", torch.zeros((), dtype=torch.float32))
        self.register_buffer("high", torch.zeros((), dtype=torch.float32))
