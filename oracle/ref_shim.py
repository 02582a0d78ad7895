"""TEST INFRASTRUCTURE ONLY -- never imported by the product package.

Import shim that lets the reference's own DreamerV3 source files
(/root/reference/sheeprl/...) be imported *unmodified* in the build container,
where lightning / tensordict / hydra / gymnasium / torchmetrics are not
installed.  It is used for two things only:

  * oracle/make_golden.py  -- runs the real reference `build_models()`,
    `train()` and `SequentialReplayBuffer` to produce tests/golden/*.pt
  * tests/test_oracle_vs_reference.py -- pins oracle/dv3_oracle.py (our CPU
    restatement) against the real reference when /root/reference exists.

/root/reference does not exist on the GPU box, so nothing that runs there may
import this module (it raises if the reference tree is missing).

None of the stubbed packages contain hot-path arithmetic: they provide process
launch, DDP wrapping, a dict-of-tensors container, config parsing and metric
averaging.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

import torch  # must be imported before the stubs are installed

REFERENCE_ROOT = os.environ.get("DV3_REFERENCE_ROOT", "/root/reference")

_STUBS = [
    "gymnasium", "gymnasium.core", "gymnasium.spaces", "gymnasium.wrappers", "gymnasium.vector",
    "hydra", "hydra.utils",
    "lightning", "lightning.fabric", "lightning.fabric.wrappers", "lightning.fabric.loggers",
    "lightning.fabric.loggers.tensorboard", "lightning.fabric.plugins",
    "lightning.fabric.plugins.collectives", "lightning.fabric.utilities",
    "lightning.fabric.utilities.distributed",
    "omegaconf", "tensordict", "tensordict.tensordict", "torchmetrics",
    "pytorch_lightning", "pytorch_lightning.utilities",
    "lightning_utilities", "lightning_utilities.core", "lightning_utilities.core.imports",
]


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "sheeprl", "algos", "dreamer_v3"))


class _Meta(type):
    def __getattr__(cls, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Any


class _Any(metaclass=_Meta):
    """Dummy class: any attribute of it is itself, any call returns an instance."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Any()


def _stub(name: str) -> None:
    if name in sys.modules and not getattr(sys.modules[name], "__dv3_stub__", False):
        return  # a real package is installed: leave it alone
    m = types.ModuleType(name)
    m.__path__ = []
    m.__dv3_stub__ = True

    def _ga(attr):
        if attr.startswith("__"):
            raise AttributeError(attr)
        return _Any

    m.__getattr__ = _ga
    sys.modules[name] = m
    if "." in name:
        parent, child = name.rsplit(".", 1)
        setattr(sys.modules[parent], child, m)


class TD:
    """~60-line stand-in for tensordict.TensorDict: a dict of tensors sharing leading
    `batch_size` dims.  Just enough for sheeprl/data/buffers.py to run unmodified."""

    def __init__(self, source=None, batch_size=(), device=None):
        self._d = dict(source or {})
        self.batch_size = torch.Size(batch_size)
        self.device = device

    # -- container protocol -------------------------------------------------
    @property
    def shape(self):
        return self.batch_size

    def keys(self):
        return self._d.keys()

    def items(self):
        return self._d.items()

    def get(self, k):
        return self._d[k]

    def set(self, k, v, inplace=False):
        if inplace and k in self._d:
            self._d[k].copy_(v)
        else:
            self._d[k] = v
        return self

    def _new(self, d, bs):
        return TD(d, bs, self.device)

    def _bs_after(self, fn):
        return fn(torch.empty(self.batch_size, device="meta")).shape

    def __getitem__(self, idx):
        if isinstance(idx, str):
            return self._d[idx]
        return self._new({k: v[idx] for k, v in self._d.items()}, self._bs_after(lambda t: t[idx]))

    def __setitem__(self, idx, val):
        if isinstance(idx, str):
            self._d[idx] = val
            return
        for k, v in (val.items() if isinstance(val, TD) else val.items()):
            if k not in self._d:
                self._d[k] = torch.zeros(*self.batch_size, *v.shape[len(val.batch_size):], dtype=v.dtype)
            self._d[k][idx] = v

    def _map(self, name, *a, nb=None):
        nd = len(self.batch_size)
        out = {}
        for k, v in self._d.items():
            tail = v.shape[nd:]
            if name in ("view", "reshape"):
                out[k] = getattr(v, name)(*a, *tail)
            elif name == "permute":
                dims = [d % nd for d in a] + list(range(nd, v.dim()))
                out[k] = v.permute(*dims)
            elif name == "unsqueeze":
                d = a[0] if a[0] >= 0 else nd + 1 + a[0]
                out[k] = v.unsqueeze(d)
        return self._new(out, self._bs_after(lambda t: getattr(t, name)(*a)))

    def view(self, *s):
        return self._map("view", *s)

    def reshape(self, *s):
        return self._map("reshape", *s)

    def permute(self, *d):
        return self._map("permute", *d)

    def unsqueeze(self, d):
        return self._map("unsqueeze", d)

    def clone(self):
        return self._new({k: v.clone() for k, v in self._d.items()}, self.batch_size)

    def to(self, device):
        return self._new({k: v.to(device) for k, v in self._d.items()}, self.batch_size)

    @classmethod
    def __torch_function__(cls, func, types_, args=(), kwargs=None):
        kwargs = kwargs or {}
        if func is torch.cat:
            tds = args[0]
            dim = kwargs.get("dim", args[1] if len(args) > 1 else 0)
            out = {k: torch.cat([t._d[k] for t in tds], dim=dim) for k in tds[0].keys()}
            bs = torch.cat([torch.empty(t.batch_size, device="meta") for t in tds], dim=dim).shape
            return TD(out, bs, tds[0].device)
        return NotImplemented


_installed = False


def install() -> None:
    """Install stubs + a synthetic `sheeprl` parent package that bypasses sheeprl/__init__.py
    (which imports every algorithm, dotenv, and a torch-version check)."""
    global _installed
    if _installed:
        return
    if not reference_available():
        raise RuntimeError(f"reference tree not found under {REFERENCE_ROOT} (expected in the build container only)")
    for n in _STUBS:
        _stub(n)
    sys.modules["pytorch_lightning.utilities"].rank_zero_only = lambda f: f
    sys.modules["lightning_utilities.core.imports"].RequirementCache = lambda *a, **k: False

    def _get_class(path):
        mod, cls = path.rsplit(".", 1)
        return getattr(importlib.import_module(mod), cls)

    sys.modules["hydra.utils"].get_class = _get_class
    td = sys.modules["tensordict"]
    td.TensorDict = TD
    td.MemmapTensor = object
    tdt = sys.modules["tensordict.tensordict"]
    tdt.TensorDictBase = TD
    tdt.TensorDict = TD
    pkg = types.ModuleType("sheeprl")
    pkg.__path__ = [os.path.join(REFERENCE_ROOT, "sheeprl")]
    sys.modules["sheeprl"] = pkg
    _installed = True


def ref_modules():
    """Returns (dreamer_v3, agent, loss, utils, buffers, distribution) reference modules."""
    install()
    names = [
        "sheeprl.algos.dreamer_v3.dreamer_v3", "sheeprl.algos.dreamer_v3.agent",
        "sheeprl.algos.dreamer_v3.loss", "sheeprl.algos.dreamer_v3.utils",
        "sheeprl.data.buffers", "sheeprl.utils.distribution",
    ]
    return tuple(importlib.import_module(n) for n in names)


class Wrapped(torch.nn.Module):
    """Stand-in for lightning's _FabricModule: exposes `.module`, forwards calls/attributes."""

    def __init__(self, m):
        super().__init__()
        self._m = m

    @property
    def module(self):
        return self._m

    def forward(self, *a, **k):
        return self._m(*a, **k)

    def __getattr__(self, n):
        try:
            return super().__getattr__(n)
        except AttributeError:
            return getattr(super().__getattr__("_m"), n)


class FakeFabric:
    """Single-process stand-in for lightning.Fabric (the four calls train() makes)."""

    world_size = 1
    global_rank = 0

    def __init__(self, device="cpu"):
        self.device = torch.device(device)

    def setup_module(self, m):
        return Wrapped(m)

    def backward(self, loss):
        loss.backward()

    def clip_gradients(self, module, optimizer, max_norm, error_if_nonfinite=False):
        return torch.nn.utils.clip_grad_norm_(module.parameters(), max_norm, error_if_nonfinite=error_if_nonfinite)

    def all_gather(self, x):
        return x


class Aggregator:
    """Records aggregator.update(key, value) calls (stand-in for MetricAggregator)."""

    def __init__(self):
        self.values = {}

    def update(self, k, v):
        self.values[k] = float(v.detach().float().mean()) if torch.is_tensor(v) else float(v)


class NoiseTape:
    """Injected-noise contract (SURVEY.md fact 9 / hard part H2): every categorical draw is the
    sequential fp32 inverse-CDF of a supplied uniform.  Patches the reference's
    OneHotCategoricalValidateArgs.sample to consume uniforms from a tape in call order
    (Appendix D order) and to record the drawn indices."""

    def __init__(self, uniforms):
        self.uniforms = list(uniforms)  # one tensor per sample() call, shape == batch_shape
        self.pos = 0
        self.indices = []

    def install(self, distribution_module):
        tape = self
        cls = distribution_module.OneHotCategoricalValidateArgs
        self._orig = cls.sample

        def sample(self_, sample_shape=torch.Size()):
            probs = self_._categorical.probs
            u = tape.uniforms[tape.pos]
            tape.pos += 1
            u = u.reshape(probs.shape[:-1])
            idx = inverse_cdf(probs, u)
            tape.indices.append(idx)
            return torch.nn.functional.one_hot(idx, probs.shape[-1]).to(probs)

        cls.sample = sample
        self._cls = cls

    def uninstall(self):
        self._cls.sample = self._orig


def inverse_cdf(probs: torch.Tensor, u: torch.Tensor) -> torch.Tensor:
    """idx = #{k : cumsum_k(p) <= u}, clamped to K-1; cumsum is a sequential fp32 left-to-right sum."""
    K = probs.shape[-1]
    c = torch.zeros_like(u)
    idx = torch.zeros(u.shape, dtype=torch.long, device=u.device)
    for k in range(K):
        c = c + probs[..., k]
        idx = idx + (c <= u).long()
    return idx.clamp_(max=K - 1)
