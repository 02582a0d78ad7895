"""Episode-wise replay storage (the DreamerV2 / P2E sibling of the sequential buffers).

Reference: `EpisodeBuffer` (sheeprl/data/buffers.py:342-525): whole episodes are kept apart, the oldest are evicted when a new one does
not fit, and `sample` draws fixed-length windows that never cross an episode boundary, optionally favouring episode ends.

Host logic only (which episode, which start: same torch-CPU-RNG call order as the reference, buffers.py:505-516); the episodes live on
the buffer's device and a sample is one index-gather per key over the concatenated storage.  `memmap` / `memmap_dir` are accepted for
signature compatibility: storage is device memory.
"""
from __future__ import annotations

from typing import Dict, List, Mapping, Optional, Union

import torch

from .buffers import TensorBatch

Tensor = torch.Tensor


class EpisodeBuffer:
    def __init__(self, buffer_size: int, sequence_length: int, device: Union[torch.device, str] = "cuda", memmap: bool = False,
                 memmap_dir=None) -> None:
        if buffer_size <= 0:
            raise ValueError(f"The buffer size must be greater than zero, got: {buffer_size}")
        if sequence_length <= 0:
            raise ValueError(f"The sequence length must be greater than zero, got: {sequence_length}")
        if buffer_size < sequence_length:
            raise ValueError(f"The sequence length must be lower than the buffer size, got: bs = {buffer_size} and sl = {sequence_length}")
        self._buffer_size, self._sequence_length = buffer_size, sequence_length
        self._device = torch.device(device)
        self._memmap, self._memmap_dir = memmap, memmap_dir
        self._buf: List[TensorBatch] = []
        self._cum_lengths: List[int] = []
        self._chunk = torch.arange(sequence_length).reshape(1, -1)

    # ------------------------------------------------------------------ properties of the reference API
    @property
    def buffer(self) -> Optional[List[TensorBatch]]:
        return self._buf

    @property
    def buffer_size(self) -> int:
        return self._buffer_size

    @property
    def sequence_length(self) -> int:
        return self._sequence_length

    @property
    def device(self) -> torch.device:
        return self._device

    @property
    def is_memmap(self) -> bool:
        return self._memmap

    @property
    def full(self) -> bool:
        # "no room for another shortest episode" (buffers.py:413-415)
        return self._cum_lengths[-1] + self._sequence_length > self._buffer_size if self._buf else False

    def __getitem__(self, key: int) -> TensorBatch:
        if not isinstance(key, int):
            raise TypeError("`key` must be an integer")
        return self._buf[key]

    def __len__(self) -> int:
        return self._cum_lengths[-1] if self._buf else 0

    # ------------------------------------------------------------------ add (buffers.py:423-483)
    def add(self, episode: Mapping[str, Tensor]) -> None:
        """episode: tensors [L, ...] with exactly one `dones` flag, on the last step."""
        if "dones" not in episode:
            raise KeyError('key "dones" not found in the episode')
        dones = episode["dones"]
        n_done = int(torch.count_nonzero(dones))
        if n_done != 1:
            raise RuntimeError(f"The episode must contain exactly one done, got: {n_done}")
        if float(dones[-1].reshape(-1)[0]) != 1:
            raise RuntimeError(f"The last step must contain a done, got: {dones[-1]}")
        ep_len = int(dones.shape[0])
        if ep_len < self._sequence_length:
            raise RuntimeError(f"Episode too short (at least {self._sequence_length} steps), got: {ep_len} steps")
        if ep_len > self._buffer_size:
            raise RuntimeError(f"Episode too long (at most {self._buffer_size} steps), got: {ep_len} steps")
        if self.full or len(self) + ep_len > self._buffer_size:
            # evict the oldest episodes until the new one fits
            total, drop = len(self), 0
            while total - self._cum_lengths[drop] + ep_len > self._buffer_size:
                drop += 1
            base = self._cum_lengths[drop]
            self._buf = self._buf[drop + 1:]
            self._cum_lengths = [c - base for c in self._cum_lengths[drop + 1:]]
        self._cum_lengths.append(len(self) + ep_len)
        self._buf.append(TensorBatch({k: v.to(self._device) for k, v in episode.items()}, [ep_len]))

    # ------------------------------------------------------------------ sample (buffers.py:485-525)
    def sample(self, batch_size: int, n_samples: int = 1, prioritize_ends: bool = False, clone: bool = False) -> TensorBatch:
        """Windows of `sequence_length` steps, each inside one episode: batch dims [n_samples, sequence_length, batch_size]."""
        if batch_size <= 0:
            raise ValueError(f"Batch size must be greater than 0, got: {batch_size}")
        if n_samples <= 0:
            raise ValueError(f"The number of samples must be greater than 0, got: {n_samples}")
        if len(self) == 0:
            raise RuntimeError("No sample has been added to the buffer. Please add at least one sample calling `self.add()`")
        sl = self._sequence_length
        per_episode = torch.bincount(torch.randint(0, len(self._buf), (batch_size * n_samples,)))
        chunks: Dict[str, List[Tensor]] = {k: [] for k in self._buf[0]}
        for i, n in enumerate(per_episode.tolist()):
            ep = self._buf[i]
            ep_len = ep.batch_size[0]
            upper = ep_len - sl + 1
            if prioritize_ends:
                upper += sl                       # starts past the last full window are clamped onto it: ends are drawn more often
            starts = torch.min(torch.randint(0, upper, size=(n,)).reshape(-1, 1), torch.tensor(ep_len - sl))
            idx = (starts + self._chunk).to(self._device)
            for k in chunks:
                chunks[k].append(ep[k][idx])
        out = {}
        for k, parts in chunks.items():
            t = torch.cat(parts, 0)                                            # [n_samples * batch, sl, ...]
            t = t.reshape(n_samples, batch_size, sl, *t.shape[2:]).transpose(1, 2)
            out[k] = t.clone() if clone else t
        return TensorBatch(out, [n_samples, sl, batch_size])
