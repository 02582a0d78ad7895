"""TEST INFRASTRUCTURE ONLY.  Seeded synthetic batches / sampling noise / parameter sets shared by the
oracle, the golden-vector generator, the GPU parity tests and bench.py's CPU baseline
(shapes follow SURVEY.md section 8d)."""
from __future__ import annotations

from typing import Dict, Sequence

import torch


def synthetic_batch(T: int, B: int, obs_dims: Dict[str, int], actions_dim: Sequence[int], seed: int = 1234,
                    done_p: float = 1 / 200) -> Dict[str, torch.Tensor]:
    """vectors ~ N(0,1) clipped to +-5; actions one-hot uniform; rewards ~ N(0,1);
    dones ~ Bernoulli(done_p); is_first[t] = dones[t-1]."""
    g = torch.Generator().manual_seed(seed)
    data = {}
    for k, d in obs_dims.items():
        if isinstance(d, int):
            data[k] = torch.randn(T, B, d, generator=g).clamp_(-5, 5)
        else:       # image key: uint8 ~ U{0..255}, shape (C, H, W)
            data[k] = torch.randint(0, 256, (T, B, *d), generator=g, dtype=torch.uint8)
    acts = [torch.nn.functional.one_hot(torch.randint(0, a, (T, B), generator=g), a).float() for a in actions_dim]
    data["actions"] = torch.cat(acts, -1)
    data["rewards"] = torch.randn(T, B, 1, generator=g)
    dones = (torch.rand(T, B, 1, generator=g) < done_p).float()
    data["dones"] = dones
    data["is_first"] = torch.cat((torch.zeros(1, B, 1), dones[:-1]), 0)
    return data


def synthetic_noise(T: int, B: int, H: int, stoch: int, n_heads: int, seed: int = 99) -> Dict[str, torch.Tensor]:
    """Uniforms for every categorical draw of one training step (SURVEY Appendix D order)."""
    g = torch.Generator().manual_seed(seed)
    N = T * B
    return {
        "obs_prior": torch.rand(T, B, stoch, generator=g),   # consumed by the reference, result unused
        "obs_post": torch.rand(T, B, stoch, generator=g),
        "act0": torch.rand(N, n_heads, generator=g),
        "img_z": torch.rand(H, N, stoch, generator=g),
        "img_act": torch.rand(H, N, n_heads, generator=g),
    }


def strip_wrapper_keys(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """ref_shim.Wrapped adds a '_m.' segment where Fabric's _FabricModule adds nothing."""
    return {k.replace("._m.", ".").replace("_m.", "", 1) if k.startswith("_m.") else k.replace("._m.", "."): v.detach().clone()
            for k, v in sd.items()}


def noise_tape_order(noise: Dict[str, torch.Tensor], T: int, H: int, n_heads: int, N: int):
    """Flatten `noise` into the per-call list the reference consumes (ref_shim.NoiseTape)."""
    tape = []
    for t in range(T):
        tape.append(noise["obs_prior"][t])
        tape.append(noise["obs_post"][t])
    for i in range(n_heads):
        tape.append(noise["act0"][:, i].reshape(1, N))
    for h in range(H):
        tape.append(noise["img_z"][h].reshape(1, N, -1))
        for i in range(n_heads):
            tape.append(noise["img_act"][h][:, i].reshape(1, N))
    # discarded second-actor draw over [(H+1), N] (dreamer_v3.py:293): any uniforms do
    for i in range(n_heads):
        tape.append(torch.full((H + 1, N), 0.5))
    return tape
