"""Device-resident replay buffers with the reference's API (sheeprl/data/buffers.py):
`ReplayBuffer` (:16-216), `SequentialReplayBuffer` (:219-339), `AsyncReplayBuffer` (:528-690; alias
`EnvIndependentReplayBuffer`, the name later SheepRL versions use).

Storage is one HBM allocation per key, laid out [n_envs][buffer_size][*feature] in the key's own dtype (uint8 pixels
stay uint8).  `add` is a ring write (dv3_buffer_add); `sample` picks (env, start) on the host with exactly the
reference's torch-CPU-RNG call order -- so the sampled indices are identical under the same seed -- and moves the bytes
with one dv3_seq_gather launch straight into the time-major [n_samples, T, B, ...] batch the training step consumes
(the reference fancy-indexes CPU memmaps, concatenates, then copies host->device: dreamer_v3.py:703-709).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Mapping, Optional, Sequence, Tuple, Union

import torch

Tensor = torch.Tensor


class TensorBatch(dict):
    """dict of tensors sharing leading batch dims (the small part of TensorDict the caller side uses)."""

    def __init__(self, source: Mapping[str, Tensor], batch_size: Sequence[int]):
        super().__init__(source)
        self.batch_size = torch.Size(batch_size)

    @property
    def shape(self):
        return self.batch_size

    def __getitem__(self, idx):
        if isinstance(idx, str):
            return dict.__getitem__(self, idx)
        probe = torch.empty(self.batch_size, device="meta")[idx]
        return TensorBatch({k: v[idx] for k, v in self.items()}, probe.shape)

    def view(self, *shape):
        nd = len(self.batch_size)
        return TensorBatch({k: v.view(*shape, *v.shape[nd:]) for k, v in self.items()}, shape)

    def to(self, device, non_blocking: bool = False):
        return TensorBatch({k: v.to(device, non_blocking=non_blocking) for k, v in self.items()}, self.batch_size)

    def clone(self):
        return TensorBatch({k: v.clone() for k, v in self.items()}, self.batch_size)


def plan_sequential_starts(pos: int, full: bool, size: int, count: int, sequence_length: int) -> Tensor:
    """Start indices of `count` sequences, drawn like SequentialReplayBuffer.sample (buffers.py:283-300): never
    straddling the write head `pos`.  Consumes the torch CPU global RNG exactly like the reference (one randint of
    `count` values over the same range), without materialising the reference's O(size) index list."""
    if full:
        first_end = pos - sequence_length + 1
        second_end = size if first_end >= 0 else size + first_end
        n_first = max(first_end, 0)
        n_valid = n_first + max(second_end - pos, 0)
        r = torch.randint(0, n_valid, size=(count,))
        return torch.where(r < n_first, r, pos + (r - n_first))
    return torch.randint(0, pos - sequence_length + 1, size=(count,))


class _EnvView:
    """What `AsyncReplayBuffer.buffer[i]` exposes to the caller (dreamer_v3.py:636-640, utils/callback.py:36-40):
    `_pos`, `buffer_size`, `full`, and `buf["dones"][idx] = v` through a [size, 1, *feat] view of env i's ring."""

    def __init__(self, parent: "AsyncReplayBuffer", env: int):
        self._parent, self._env = parent, env

    @property
    def _pos(self) -> int:
        return self._parent._pos[self._env]

    @property
    def full(self) -> bool:
        return self._parent._full[self._env]

    _full = full

    @property
    def buffer_size(self) -> int:
        return self._parent.buffer_size

    @property
    def n_envs(self) -> int:
        return 1

    def __len__(self) -> int:
        return self._parent.buffer_size

    def __getitem__(self, key: str) -> Tensor:
        if not isinstance(key, str):
            raise TypeError("`key` must be a string")
        if self._parent._store is None:
            raise RuntimeError("The buffer has not been initialized. Try to add some data first.")
        return self._parent._store[key][self._env].unsqueeze(1)

    def __setitem__(self, key: str, t: Tensor) -> None:
        self[key].copy_(t)


class AsyncReplayBuffer:
    def __init__(self, buffer_size: int, n_envs: int = 1, device: Union[torch.device, str] = "cuda", memmap: bool = False,
                 memmap_dir=None, sequential: bool = False):
        if buffer_size <= 0:
            raise ValueError(f"The buffer size must be greater than zero, got: {buffer_size}")
        if n_envs <= 0:
            raise ValueError(f"The number of environments must be greater than zero, got: {n_envs}")
        self._buffer_size, self._n_envs = buffer_size, n_envs
        self._device = torch.device(device)
        self._sequential = sequential
        self._memmap, self._memmap_dir = memmap, memmap_dir   # storage is HBM; the directory is used by flush_memmap / load_memmap
        self._store: Optional[Dict[str, Tensor]] = None
        self._pos: List[int] = [0] * n_envs
        self._full: List[bool] = [False] * n_envs
        self._lib = None
        self._draw_env_counts = True

    # ------------------------------------------------------------------ checkpoints (utils/callback.py:23-64 pickles `rb`)
    def __getstate__(self):
        """Picklable state: the rings as CPU tensors, no library handle (a ctypes.CDLL cannot be pickled)."""
        st = {k: v for k, v in self.__dict__.items() if k not in ("_lib", "_keepalive", "_store", "_stage", "_stage_ev", "_stage_i")}
        st["_store"] = None if self._store is None else {k: v.cpu() for k, v in self._store.items()}
        return st

    def __setstate__(self, st):
        self.__dict__.update(st)
        self._lib = None
        if self._store is not None:
            dev = self._device if (self._device.type != "cuda" or torch.cuda.is_available()) else torch.device("cpu")
            self._store = {k: v.to(dev) for k, v in self._store.items()}

    # ------------------------------------------------------------------ on-disk copy in the reference's layout
    # Reference (memmap=True): `memmap_dir/env_{i}/{key}.memmap`, one raw array [buffer_size, 1, *feature] per key and environment
    # (buffers.py:134-147, :637), plus tensordict's `meta.pt`.  Here the rings live in HBM and sampling never touches the host, so the
    # files are not the working storage: `flush_memmap()` writes them (a checkpoint of the replay data another process can mmap),
    # `load_memmap()` reads them back.  `meta.pt` holds what is needed to reopen the raw files: shapes, dtypes, write positions.
    def flush_memmap(self, memmap_dir=None) -> None:
        import os

        import numpy as np

        root = memmap_dir if memmap_dir is not None else self._memmap_dir
        if root is None:
            raise ValueError("no memmap_dir: pass one here or to the constructor")
        if self._store is None:
            raise RuntimeError("the buffer is empty")
        meta = {"buffer_size": self._buffer_size, "n_envs": self._n_envs, "pos": list(self._pos), "full": list(self._full), "keys": {}}
        for k, ring in self._store.items():
            host = ring.cpu()
            meta["keys"][k] = {"shape": tuple(host.shape[2:]), "dtype": str(host.dtype).replace("torch.", "")}
            for i in range(self._n_envs):
                d = os.path.join(str(root), f"env_{i}")
                os.makedirs(d, exist_ok=True)
                arr = host[i].unsqueeze(1).contiguous().numpy()                  # [buffer_size, 1, *feature]
                mm = np.memmap(os.path.join(d, f"{k}.memmap"), dtype=arr.dtype, mode="w+", shape=arr.shape)
                mm[...] = arr
                mm.flush()
                del mm
        torch.save(meta, os.path.join(str(root), "meta.pt"))

    def load_memmap(self, memmap_dir=None) -> None:
        import os

        import numpy as np

        root = memmap_dir if memmap_dir is not None else self._memmap_dir
        meta = torch.load(os.path.join(str(root), "meta.pt"), weights_only=False)
        if meta["buffer_size"] != self._buffer_size or meta["n_envs"] != self._n_envs:
            raise ValueError(f"files hold a {meta['buffer_size']} x {meta['n_envs']} buffer, this one is {self._buffer_size} x {self._n_envs}")
        store = {}
        for k, info in meta["keys"].items():
            dt = getattr(torch, info["dtype"])
            ring = torch.zeros(self._n_envs, self._buffer_size, *info["shape"], dtype=dt, device=self._device)
            for i in range(self._n_envs):
                shape = (self._buffer_size, 1, *info["shape"])
                mm = np.memmap(os.path.join(str(root), f"env_{i}", f"{k}.memmap"), dtype=np.dtype(info["dtype"]), mode="r", shape=shape)
                ring[i].copy_(torch.from_numpy(np.array(mm)).squeeze(1))
                del mm
            store[k] = ring
        self._store, self._pos, self._full = store, list(meta["pos"]), list(meta["full"])

    # ------------------------------------------------------------------ properties of the reference API
    @property
    def buffer(self) -> Optional[Tuple[_EnvView, ...]]:
        return None if self._store is None else tuple(_EnvView(self, i) for i in range(self._n_envs))

    @property
    def buffer_size(self) -> int:
        return self._buffer_size

    @property
    def full(self):
        return None if self._store is None else tuple(self._full)

    @property
    def n_envs(self) -> int:
        return self._n_envs

    @property
    def device(self):
        return self._device

    def __len__(self) -> int:
        return self._buffer_size

    # ------------------------------------------------------------------ C-ABI
    def _library(self):
        if self._lib is None:
            from .._lib import load_library

            self._lib = load_library()
        return self._lib

    def _stream(self):
        return torch.cuda.current_stream(self._device).cuda_stream

    # ------------------------------------------------------------------ add (buffers.py:607-644 -> :99-151 per env)
    def add(self, data: Mapping[str, Tensor], indices: Optional[Sequence[int]] = None) -> None:
        if not isinstance(data, Mapping):
            raise TypeError("`data` must be a mapping of tensors (TensorDict-like)")
        first = next(iter(data.values()))
        if first.dim() < 2:
            raise RuntimeError("`data` must have 2 batch dimensions: [sequence_length, n_envs]. Shape is: {}".format(tuple(first.shape)))
        L = first.shape[0]
        if indices is None:
            indices = tuple(range(self._n_envs))
        if self._store is None:
            self._store = {k: torch.zeros(self._n_envs, self._buffer_size, *v.shape[2:], dtype=v.dtype, device=self._device)
                           for k, v in data.items()}
        size = self._buffer_size
        skip = max(L - size, 0)          # oversize add: only the last `size` rows survive (buffers.py:129-131)
        for j, env in enumerate(indices):
            pos = self._pos[env]
            for k, v in data.items():
                ring = self._store[k]
                src = v[skip:, j].to(self._device, dtype=ring.dtype).contiguous()
                row_bytes = ring[0, 0].numel() * ring.element_size()
                self._ring_write(src, ring, env, (pos + skip) % size, L - skip, row_bytes)
            if pos + L >= size:
                self._full[env] = True
            self._pos[env] = (pos + L) % size

    def _ring_write(self, src: Tensor, ring: Tensor, env: int, pos: int, L: int, row_bytes: int) -> None:
        rc = self._library().dv3_buffer_add(src.data_ptr(), ring.data_ptr(), env, pos, L, self._buffer_size, row_bytes, self._stream())
        if rc != 0:
            raise RuntimeError(f"dv3_buffer_add failed ({rc})")

    # ------------------------------------------------------------------ sample (buffers.py:646-690)
    def plan(self, batch_size: int, sequence_length: int, n_samples: int):
        """Host-side index selection with the reference's RNG call order.  Returns int tensors
        (env_id, start, col, samp), one entry per sampled sequence."""
        if batch_size <= 0 or n_samples <= 0:
            raise ValueError(f"`batch_size` ({batch_size}) and `n_samples` ({n_samples}) must be both greater than 0")
        if self._store is None:
            if not self._draw_env_counts:      # SequentialReplayBuffer checks "nothing added" first (buffers.py:271-276)
                raise ValueError("No sample has been added to the buffer. Please add at least one sample calling `self.add()`")
            raise RuntimeError("The buffer has not been initialized. Try to add some data first.")
        if self._draw_env_counts:
            bs_per_buf = torch.bincount(torch.randint(0, self._n_envs, (batch_size,)))      # buffers.py:678
        else:
            bs_per_buf = torch.tensor([batch_size])
        env_id, start, col, samp = [], [], [], []
        col0 = 0
        for env, bs in enumerate(bs_per_buf.tolist()):
            if bs <= 0:
                continue
            pos, full, size = self._pos[env], self._full[env], self._buffer_size
            if not full and pos == 0:
                raise ValueError("No sample has been added to the buffer. Please add at least one sample calling `self.add()`")
            if not full and pos - sequence_length + 1 < 1:
                raise ValueError(f"too long sequence length ({sequence_length})")
            if full and sequence_length > size:
                raise ValueError(f"too long sequence length ({sequence_length})")
            cnt = bs * n_samples
            if self._sequential:
                st = plan_sequential_starts(pos, full, size, cnt, sequence_length)
            else:                                     # ReplayBuffer.sample, buffers.py:175-190 (sequence_length == 1)
                if full:
                    second_end = size if pos >= 0 else size + pos
                    n_valid = pos + (second_end - pos)
                    r = torch.randint(0, n_valid, size=(cnt,))
                    st = r
                else:
                    st = torch.randint(0, pos, size=(cnt,))
            torch.randint(0, 1, size=(cnt,))          # the per-env `env_idxes` draw of _get_samples (:197, :325-327)
            j = torch.arange(cnt)
            env_id.append(torch.full((cnt,), env, dtype=torch.int32))
            start.append(st.to(torch.int64))
            col.append((col0 + j % bs).to(torch.int32))  # reshape(n_samples, bs, T): j = s*bs + b   (:307)
            samp.append((j // bs).to(torch.int32))
            col0 += bs
        return torch.cat(env_id), torch.cat(start), torch.cat(col), torch.cat(samp)

    def sample(self, batch_size: int, sample_next_obs: bool = False, clone: bool = False, sequence_length: int = 1,
               n_samples: int = 1) -> TensorBatch:
        """-> TensorBatch with batch_size [n_samples, sequence_length, batch_size] (sequential), on the device."""
        if sample_next_obs:
            raise NotImplementedError("sample_next_obs is not used by DreamerV3 (dreamer_v3.py:703-708)")
        T = sequence_length if self._sequential else 1
        env_id, start, col, samp = self.plan(batch_size, T, n_samples)
        out = self.gather(env_id, start, col, samp, batch_size, T, n_samples)
        if not self._sequential:
            out = TensorBatch({k: v.reshape(n_samples * batch_size, 1, *v.shape[3:]) for k, v in out.items()}, (n_samples * batch_size, 1))
        return out

    MAX_KEYS_PER_LAUNCH = 8      # dv3_seq_gather takes up to 8 keys per launch (kernel argument block); more keys = more launches

    def gather(self, env_id: Tensor, start: Tensor, col: Tensor, samp: Tensor, batch_size: int, T: int, n_samples: int) -> TensorBatch:
        keys = list(self._store.keys())
        out = {k: torch.empty(n_samples, T, batch_size, *self._store[k].shape[2:], dtype=self._store[k].dtype, device=self._device)
               for k in keys}
        self._gather_rows(keys, out, env_id, start, col, samp, batch_size, T, n_samples)
        return TensorBatch(out, (n_samples, T, batch_size))

    def _gather_rows(self, keys, out, env_id, start, col, samp, batch_size, T, n_samples) -> None:
        dev = self._device
        # the four index vectors travel in ONE pinned staging buffer and one async copy (they are a few hundred bytes)
        n_seq = env_id.numel()
        host = self._index_staging(5 * n_seq)
        host[:2 * n_seq].view(torch.int64).copy_(start)         # (int64 first: keeps its 8-byte alignment)
        host[2 * n_seq:3 * n_seq] = env_id
        host[3 * n_seq:4 * n_seq] = col
        host[4 * n_seq:] = samp
        didx = host.to(dev, non_blocking=True)
        p_start, p_env = didx.data_ptr(), didx.data_ptr() + 8 * n_seq
        p_col, p_samp = didx.data_ptr() + 12 * n_seq, didx.data_ptr() + 16 * n_seq
        for k0 in range(0, len(keys), self.MAX_KEYS_PER_LAUNCH):
            kk = keys[k0:k0 + self.MAX_KEYS_PER_LAUNCH]
            n = len(kk)
            src = (C.c_void_p * n)(*[self._store[k].data_ptr() for k in kk])
            dst = (C.c_void_p * n)(*[out[k].data_ptr() for k in kk])
            rb = (C.c_longlong * n)(*[self._store[k][0, 0].numel() * self._store[k].element_size() for k in kk])
            rc = self._library().dv3_seq_gather(n, C.cast(src, C.c_void_p), C.cast(dst, C.c_void_p), C.cast(rb, C.c_void_p),
                                                self._buffer_size, T, batch_size, n_samples, p_env, p_start, p_col, p_samp, self._stream())
            if rc != 0:
                raise RuntimeError(f"dv3_seq_gather failed ({rc}) for keys {kk}")
        self._keepalive = didx
        if dev.type == "cuda":
            self._stage_ev[self._stage_i].record(torch.cuda.current_stream(dev))

    def _index_staging(self, n: int) -> Tensor:
        """Pinned host staging for the index upload: a small ring of reusable buffers (allocating pinned memory per call costs more
        than the gather itself); a buffer is reused only after the copy that read it has completed."""
        if self._device.type != "cuda":
            return torch.empty(n, dtype=torch.int32)
        ring = self.__dict__.setdefault("_stage", [])
        if not ring or ring[0].numel() < n:
            ring[:] = [torch.empty(max(n, 1024), dtype=torch.int32).pin_memory() for _ in range(8)]
            self._stage_ev = [torch.cuda.Event() for _ in range(8)]
            self._stage_i = -1
        self._stage_i = (self._stage_i + 1) % len(ring)
        self._stage_ev[self._stage_i].synchronize()
        return ring[self._stage_i][:n]


class SequentialReplayBuffer(AsyncReplayBuffer):
    """Single-ring sequential buffer (buffers.py:219-339).  DreamerV3 only ever instantiates it with n_envs=1 (one per
    environment inside AsyncReplayBuffer, buffers.py:630-640), which is what is supported here."""

    def __init__(self, buffer_size: int, n_envs: int = 1, device="cuda", memmap: bool = False, memmap_dir=None):
        if n_envs != 1:
            raise NotImplementedError("SequentialReplayBuffer with n_envs > 1 is not on the DreamerV3 path; use AsyncReplayBuffer")
        super().__init__(buffer_size, n_envs, device, memmap, memmap_dir, sequential=True)
        self._draw_env_counts = False

    @property
    def _pos1(self) -> int:
        return self._pos[0]

    @property
    def full(self) -> bool:                     # a single ring: a bool, as ReplayBuffer.full (buffers.py:71-73)
        return self._full[0]

    def __getitem__(self, key: str) -> Tensor:
        """rb[key]: the ring as [buffer_size, 1, *feature] (ReplayBuffer.__getitem__, buffers.py:205-211)."""
        if not isinstance(key, str):
            raise TypeError("`key` must be a string")
        if self._store is None:
            raise RuntimeError("The buffer has not been initialized. Try to add some data first.")
        return self._store[key][0].unsqueeze(1)


EnvIndependentReplayBuffer = AsyncReplayBuffer
