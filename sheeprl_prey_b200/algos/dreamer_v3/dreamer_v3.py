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
    dreamer_v3.py:781-795) stays valid while the fused kernel does the update.
    Checked on every call: `optimizer.load_state_dict()` replaces the state tensors, so when they no longer alias the arena (a
    resume after the first train() call) the loaded moments are imported again; a changed lr / betas / eps is picked up and the
    captured CUDA graph, which froze the old values, is dropped."""
    if optimizer is None:
        return
    group = optimizer.param_groups[0]
    b1, b2 = group.get("betas", (0.9, 0.999))
    hp = dict(lr=float(group["lr"]), eps=float(group["eps"]), beta1=float(b1), beta2=float(b2))
    if eng.opt_hparams.get(tag) != hp:
        eng.opt_hparams[tag] = hp
        eng._graph = None                    # hyper-parameters are baked into the captured launches
    probe = getattr(optimizer, "_dv3_probe", None)       # (first parameter, address its exp_avg must have while adopted)
    if getattr(optimizer, "_dv3_adopted", None) is arena and probe is not None:
        st0 = optimizer.state.get(probe[0])
        if st0 and "exp_avg" in st0 and st0["exp_avg"].data_ptr() == probe[1]:
            return                           # still aliased: nothing to do
    if group.get("weight_decay", 0) not in (0, 0.0):
        raise NotImplementedError("weight decay is not part of the reference's DreamerV3 config (optim/adam.yaml)")
    by_ptr = {v.data_ptr(): k for k, v in arena.p.items()}
    for p in group["params"]:
        name = by_ptr.get(p.data_ptr())
        if name is None:
            raise RuntimeError("optimizer parameter is not an arena view; build the models with sheeprl_prey_b200.build_models")
        off, n, shape = arena.offsets[name]
        st = optimizer.state[p]
        if "exp_avg" in st and st["exp_avg"].data_ptr() != arena.m[off:].data_ptr():   # resumed optimizer: import its moments
            arena.m[off:off + n].view(shape).copy_(st["exp_avg"])
            arena.v[off:off + n].view(shape).copy_(st["exp_avg_sq"])
            arena.step.fill_(float(st["step"]))
        st["exp_avg"] = arena.m[off:off + n].view(shape)
        st["exp_avg_sq"] = arena.v[off:off + n].view(shape)
        st["step"] = arena.step.view(())
    optimizer._dv3_adopted = arena
    p0 = group["params"][0]
    optimizer._dv3_probe = (p0, optimizer.state[p0]["exp_avg"].data_ptr())


def _bind_moments(eng: DV3Engine, moments) -> None:
    """`moments.low / high` (and our `rng_counter`) become VIEWS of the engine's device state: the step updates them in place, and
    `moments.load_state_dict()` -- which copies into the existing buffers -- writes straight into the engine, whenever it is called.
    Works for the reference's own Moments module too (same buffer names, algos/dreamer_v3/utils.py:28-29)."""
    if moments is None:
        return
    low = getattr(moments, "low", None)
    if low is not None and low.data_ptr() == eng.mom_state.data_ptr():
        return
    eng.mom_state[0:1].copy_(moments.low.reshape(1).to(eng.mom_state.device))
    eng.mom_state[1:2].copy_(moments.high.reshape(1).to(eng.mom_state.device))
    moments.low = eng.mom_state[0]
    moments.high = eng.mom_state[1]
    if hasattr(moments, "rng_counter"):      # resume continues the noise stream instead of replaying steps 0..k
        eng.rng_counter.copy_(moments.rng_counter.reshape(1).to(eng.rng_counter.device))
        moments.rng_counter = eng.rng_counter


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
    _bind_moments(eng, moments)
    eng.load_batch(data)
    if noise is not None or not _graphs_enabled(eng):
        eng.load_noise(noise)
        eng.step()
    else:
        _graphed_step(eng)
    if aggregator is not None:
        m = eng.metrics
        for i, k in enumerate(METRIC_KEYS):
            aggregator.update(k, m[i])


class Moments(torch.nn.Module):
    """The reference's return-normalisation state (algos/dreamer_v3/utils.py:16-41) with the same buffers, constructor and
    `forward`.  Inside train() the percentile / EMA arithmetic runs in dv3_moments on the engine's copy of the state, which
    `low` / `high` alias after the first call; `forward` is the same arithmetic for any other caller.  `rng_counter` (ours) is the
    position of the step's counter-based noise generator, saved with the module so that a resumed run continues the stream."""

    def __init__(self, fabric=None, decay: float = 0.99, max_: float = 1e8, percentile_low: float = 0.05, percentile_high: float = 0.95):
        super().__init__()
        self._fabric = fabric
        self._decay, self._max, self._percentile_low, self._percentile_high = decay, max_, percentile_low, percentile_high
        self.register_buffer("low", torch.zeros((), dtype=torch.float32))
        self.register_buffer("high", torch.zeros((), dtype=torch.float32))
        self.register_buffer("rng_counter", torch.zeros(1, dtype=torch.int64))

    def forward(self, x: torch.Tensor, fabric=None):
        """utils.py:34-41: EMA of the 5 % / 95 % quantiles over all ranks' values; returns (offset, invscale)."""
        fabric = fabric if fabric is not None else self._fabric
        gathered = fabric.all_gather(x).detach() if fabric is not None and hasattr(fabric, "all_gather") else x.detach()
        gathered = gathered.float().reshape(-1)
        n = gathered.numel()
        if gathered.is_cuda:
            # exact torch.quantile (linear interpolation) through the library's radix select, EMA included
            from ..._lib import load_library

            state = torch.stack((self.low.reshape(()), self.high.reshape(()))).contiguous()
            out = torch.empty(2, dtype=torch.float32, device=gathered.device)
            rc = load_library().dv3_moments(n, gathered.contiguous().data_ptr(), state.data_ptr(), out.data_ptr(), self._decay, self._max,
                                            self._percentile_low, self._percentile_high, torch.cuda.current_stream(gathered.device).cuda_stream)
            if rc != 0:
                raise RuntimeError(f"dv3_moments failed ({rc})")
            self.low.copy_(state[0])
            self.high.copy_(state[1])
            return out[0], out[1]
        raise RuntimeError("Moments.forward needs CUDA tensors: sheeprl_prey_b200 has no CPU path")
