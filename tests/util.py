"""Shared helpers for the test-suite (test infrastructure)."""
import os
from collections import OrderedDict

import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    return torch.load(os.path.join(GOLDEN, name), weights_only=False)


def load_baseline_golden(name="dv3_xs_b16t64.npz"):
    """The BASELINE-size golden (oracle/make_golden.py::make_baseline_golden): (meta dict, dict of index arrays)."""
    import json

    import numpy as np

    z = np.load(os.path.join(GOLDEN, name))
    meta = json.loads(bytes(z["meta"]).decode())
    return meta, {k: torch.from_numpy(z[k].astype("int64")) for k in z.files if k != "meta"}


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / (b.abs().max() + 1e-20))


def make_arena(Arena, d, device, opt=True):
    a = Arena(OrderedDict((k, tuple(v.shape)) for k, v in d.items()), device, with_opt=opt)
    a.load(d)
    return a


def engine_from_params(cfg, obs_dims, actions_dim, params, ops, device):
    from sheeprl_prey_b200.engine import Arena, DV3Engine

    return DV3Engine(cfg, obs_dims, actions_dim, make_arena(Arena, params["wm"], device), make_arena(Arena, params["actor"], device),
                     make_arena(Arena, params["critic"], device), make_arena(Arena, params["target"], device, False), ops, device)


class _CpuKernels:
    """torch statement of the two buffer kernels (dv3_buffer_add, dv3_seq_gather) for host-logic tests on CPU.
    The product itself has no CPU path: the seam lives here."""

    def _ring_write(self, src, ring, env, pos, L, row_bytes):
        idx = (pos + torch.arange(L)) % self._buffer_size
        ring[env, idx.to(ring.device)] = src

    def _gather_rows(self, keys, out, env_id, start, col, samp, batch_size, T, n_samples):
        t = torch.arange(T)
        for j in range(env_id.numel()):
            rows = (int(start[j]) + t) % self._buffer_size
            for k in keys:
                out[k][int(samp[j]), :, int(col[j])] = self._store[k][int(env_id[j]), rows]


def _make_cpu_buffers():
    from sheeprl_prey_b200.data import buffers as B

    class AsyncReplayBufferCpu(_CpuKernels, B.AsyncReplayBuffer):
        pass

    class SequentialReplayBufferCpu(_CpuKernels, B.SequentialReplayBuffer):
        pass

    for c in (AsyncReplayBufferCpu, SequentialReplayBufferCpu):      # module-level names: picklable by reference
        c.__qualname__ = c.__name__
        globals()[c.__name__] = c
    return {"AsyncReplayBuffer": AsyncReplayBufferCpu, "SequentialReplayBuffer": SequentialReplayBufferCpu}


_CPU_BUFFERS = None


def cpu_replay_buffer(cls_name="AsyncReplayBuffer"):
    global _CPU_BUFFERS
    if _CPU_BUFFERS is None:
        _CPU_BUFFERS = _make_cpu_buffers()
    return _CPU_BUFFERS[cls_name]


class cpu_ops:
    """`with cpu_ops(): build_models(...)` -- build_models always constructs `_lib.CudaOps(device)`; for host-logic tests on CPU this
    context swaps that class for the torch statement of the kernels (oracle/torch_ops.py).  The seam lives in tests/, not in the product."""

    def __enter__(self):
        from oracle.torch_ops import TorchOps
        from sheeprl_prey_b200 import _lib

        self._saved = _lib.CudaOps
        _lib.CudaOps = lambda device, *a, **k: TorchOps()
        return self

    def __exit__(self, *exc):
        from sheeprl_prey_b200 import _lib

        _lib.CudaOps = self._saved
        return False


class FakeFabric:
    """Single-process stand-in for lightning.Fabric (only what build_models / train touch)."""

    world_size = 1
    global_rank = 0

    def __init__(self, device="cpu"):
        self.device = torch.device(device)

    def setup_module(self, m):
        return m


class Aggregator:
    def __init__(self):
        self.values = {}

    def update(self, k, v):
        self.values[k] = float(v)


def random_params_like(cfg, obs_dims, actions_dim, seed=0):
    """Random, non-degenerate parameters with the reference's state_dict names, made by the product's build_models on CPU
    with the torch statement of the kernels (no reference needed -> usable on the GPU box)."""
    from sheeprl_prey_b200.algos.dreamer_v3.agent import build_models
    from sheeprl_prey_b200.configs import ObsSpace

    torch.manual_seed(seed)
    with cpu_ops():
        wm, actor, critic, target = build_models(FakeFabric("cpu"), actions_dim, False, cfg,
                                                 ObsSpace({k: ((d,) if isinstance(d, int) else tuple(d)) for k, d in obs_dims.items()}))
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for m in (wm, actor, critic, target):
            for n, p in m.named_parameters():
                if p.dim() == 1:
                    p.add_(0.1 * torch.randn(p.shape, generator=g))
        for m in (critic, target, wm.reward_model):
            for p in m.parameters():
                p.add_(0.05 * torch.randn(p.shape, generator=g))
    sd = lambda m: {k: v.detach().clone() for k, v in m.state_dict().items()}  # noqa: E731
    return dict(wm=sd(wm), actor=sd(actor), critic=sd(critic), target=sd(target))
