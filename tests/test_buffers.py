"""Replay buffers: host logic on CPU (index selection with the reference's RNG order, ring semantics) against golden
sessions recorded from the reference AsyncReplayBuffer, property tests in the style of the reference's
tests/test_data/test_sequential_buffer.py, and (-m gpu) the bit-exact gather / ring-write kernels."""
import pytest
import torch

from sheeprl_prey_b200.data.buffers import AsyncReplayBuffer, SequentialReplayBuffer, plan_sequential_starts
from util import cpu_replay_buffer, load_golden

CpuAsync, CpuSequential = cpu_replay_buffer("AsyncReplayBuffer"), cpu_replay_buffer("SequentialReplayBuffer")


def _replay_session(sess, device, cls):
    torch.manual_seed(sess["seed"])
    rb = cls(sess["size"], sess["n_envs"], device=device, sequential=True)
    it = iter(sess["samples"])
    for ev in sess["script"]:
        if ev[0] == "add":
            rb.add(ev[1], ev[2])
        else:
            s = rb.sample(ev[1], sequence_length=sess["T"], n_samples=ev[2])
            ref = next(it)
            for k, v in ref.items():
                assert s[k].dtype == v.dtype and tuple(s[k].shape) == tuple(v.shape), k
                assert torch.equal(s[k].cpu(), v), k          # bit-exact, uint8 and fp32 alike
    assert [b._pos for b in rb.buffer] == sess["pos"] and [b.full for b in rb.buffer] == sess["full"]


def test_golden_sessions_host_logic():
    for sess in load_golden("buffers.pt"):
        _replay_session(sess, "cpu", CpuAsync)


@pytest.mark.gpu
def test_golden_sessions_gpu():
    for sess in load_golden("buffers.pt"):
        _replay_session(sess, "cuda:0", AsyncReplayBuffer)


def test_constructor_guards_and_errors():
    with pytest.raises(ValueError):
        AsyncReplayBuffer(-1)
    with pytest.raises(ValueError):
        AsyncReplayBuffer(1, -1)
    rb = CpuSequential(10, 1, device="cpu")
    with pytest.raises(ValueError, match="No sample has been added"):        # reference: test_seq_replay_buffer_sample_no_add
        rb.sample(2, sequence_length=2)
    with pytest.raises(RuntimeError):                                         # the per-environment container reports it as uninitialised
        CpuAsync(10, 2, device="cpu", sequential=True).sample(2, sequence_length=2)
    rb.add({"t": torch.rand(5, 1, 1)})
    with pytest.raises(ValueError):
        rb.sample(0, sequence_length=2)
    with pytest.raises(ValueError):
        rb.sample(2, sequence_length=8)      # too long for a not-full buffer
    assert rb.sample(3, sequence_length=5, n_samples=2)["t"].shape == (2, 5, 3, 1)


def test_reference_sequential_buffer_cases():
    """The reference's tests/test_data/test_sequential_buffer.py cases, on the same class names and calls."""
    rb = CpuSequential(5, 1, device="cpu")
    td1, td2, td3 = {"t": torch.rand(2, 1, 1)}, {"t": torch.rand(2, 1, 1)}, {"t": torch.rand(3, 1, 1)}
    rb.add(td1)
    rb.add(td2)
    rb.add(td3)                                                    # test_seq_replay_buffer_add_tds: wraps around
    assert rb.full
    assert rb["t"][0] == td3["t"][-2] and rb["t"][1] == td3["t"][-1]
    torch.testing.assert_close(rb["t"][2:4], td2["t"])
    rb = CpuSequential(5, 1, device="cpu")
    td = {"t": torch.rand(6, 1, 1)}
    rb.add(td)                                                     # test_seq_replay_buffer_add_single_td: oversize add
    assert rb.full and rb["t"][0] == td["t"][-1]
    rb = CpuSequential(10, 1, device="cpu")
    rb.add({"t": torch.rand(11, 1, 1)})
    assert rb.sample(4, sequence_length=2)["t"].shape == (1, 2, 4, 1)     # test_seq_replay_buffer_sample
    rb = CpuSequential(1, 1, device="cpu")
    one = {"t": torch.rand(1, 1, 1)}
    rb.add(one)                                                    # test_seq_replay_buffer_sample_one_element
    assert rb.full and rb.sample(1, sequence_length=1)["t"].view(-1) == one["t"].view(-1)
    with pytest.raises(ValueError):
        rb.sample(1, sequence_length=2)
    for seq_len in (50, 1000):                                     # test_seq_replay_buffer_sample_full(_large_sl)
        size = 10000
        rb = CpuSequential(size, 1, device="cpu")
        rb.add({"t": (torch.arange(10500) % size).view(-1, 1, 1)})
        s = rb.sample(1000, sequence_length=seq_len, n_samples=5)["t"][..., 0]
        pos = rb._pos[0]
        assert not torch.logical_and(s[:, 0, :] < pos, s[:, -1, :] >= pos).any()
        assert not torch.logical_and(s[:, 0, :] >= size + pos - seq_len + 1, s[:, -1, :] < pos).any()
    rb = CpuSequential(10, 1, device="cpu")
    rb.add({"t": torch.arange(5).view(-1, 1, 1)})
    with pytest.raises(ValueError, match="too long sequence length"):     # test_seq_replay_buffer_sample_fail_not_full
        rb.sample(5, sequence_length=8, n_samples=1)
    rb = CpuSequential(10, 1, device="cpu")
    rb.add({"t": torch.arange(7).view(-1, 1, 1) * 1.0})
    assert (rb.sample(2, sequence_length=5, n_samples=2)["t"] < 7).all()  # test_seq_replay_buffer_sample_not_full


def test_sequences_never_straddle_the_write_head():
    size, T = 20, 6
    rb = CpuSequential(size, 1, device="cpu")
    rb.add({"t": torch.arange(27, dtype=torch.float32).view(27, 1, 1)})   # full, pos = 7
    assert rb.full is True and rb._pos[0] == 7
    s = rb.sample(256, sequence_length=T)["t"][0, :, :, 0]                 # [T, 256]
    assert bool((s[1:] - s[:-1] == 1).all())                               # consecutive in time => never crosses pos
    # not full: stays below pos
    rb2 = CpuSequential(size, 1, device="cpu")
    rb2.add({"t": torch.arange(9, dtype=torch.float32).view(9, 1, 1)})
    s2 = rb2.sample(128, sequence_length=4)["t"]
    assert float(s2.max()) <= 8


def test_start_distribution_matches_reference_list_semantics():
    # the analytic index map must equal indexing the reference's explicit list (buffers.py:292-297)
    for size, pos, T in ((23, 7, 5), (23, 2, 5), (23, 0, 5), (23, 22, 5), (8, 3, 8)):
        first_end = pos - T + 1
        second_end = size if first_end >= 0 else size + first_end
        valid = list(range(0, first_end)) + list(range(pos, second_end))
        torch.manual_seed(4)
        mine = plan_sequential_starts(pos, True, size, 64, T)
        torch.manual_seed(4)
        ref = torch.tensor(valid)[torch.randint(0, len(valid), size=(64,))]
        assert torch.equal(mine, ref)


def test_env_view_pokes_storage():
    rb = CpuAsync(6, 2, device="cpu", sequential=True)
    rb.add({"dones": torch.zeros(3, 2, 1)})
    rb.buffer[1]["dones"][(rb.buffer[1]._pos - 1) % rb.buffer_size] = 1.0     # dreamer_v3.py:636-640
    assert float(rb._store["dones"][1, 2, 0]) == 1.0 and float(rb._store["dones"][0].sum()) == 0.0


@pytest.mark.gpu
def test_gather_large_bit_exact():
    dev = "cuda:0"
    size, n_envs, T, B = 4096, 4, 64, 16
    rb = AsyncReplayBuffer(size, n_envs, device=dev, sequential=True)
    g = torch.Generator().manual_seed(0)
    chunk = {"rgb": torch.randint(0, 256, (size + 100, n_envs, 3, 64, 64), generator=g, dtype=torch.uint8),
             "state": torch.randn(size + 100, n_envs, 14, generator=g), "rewards": torch.randn(size + 100, n_envs, 1, generator=g)}
    for i in range(0, size + 100, 512):
        rb.add({k: v[i:i + 512] for k, v in chunk.items()})
    torch.manual_seed(9)
    env_id, start, col, samp = rb.plan(B, T, 2)
    out = rb.gather(env_id, start, col, samp, B, T, 2)
    torch.cuda.synchronize()
    t = torch.arange(T)
    for j in range(env_id.numel()):
        rows = (int(start[j]) + t) % size
        # ring content: element at ring position r is the last row written there
        for k in chunk:
            src = rb._store[k][int(env_id[j])].cpu()[rows]
            assert torch.equal(out[k][int(samp[j]), :, int(col[j])].cpu(), src), k
    # ring content itself equals the reference semantics: last `size` rows, wrapped
    L = size + 100
    expect = torch.empty(size, 14)
    idx = torch.arange(L - size, L)
    expect[idx % size] = chunk["state"][idx, 2]
    assert torch.equal(rb._store["state"][2].cpu(), expect)


def test_buffer_pickles_without_the_library_handle():
    """The checkpoint callback pickles the buffer (utils/callback.py:23-64): state survives, the ctypes handle does not travel."""
    import pickle

    rb = CpuAsync(9, 2, device="cpu", sequential=True)
    rb.add({"state": torch.randn(5, 2, 3), "dones": torch.zeros(5, 2, 1)})
    rb._lib = object()                       # stands for the loaded CDLL
    rb2 = pickle.loads(pickle.dumps(rb))
    assert rb2._lib is None and rb2._pos == rb._pos and rb2._full == rb._full
    assert all(torch.equal(rb2._store[k], rb._store[k]) for k in rb._store)
    torch.manual_seed(1)
    a = rb.sample(4, sequence_length=3)
    torch.manual_seed(1)
    b = rb2.sample(4, sequence_length=3)
    assert all(torch.equal(a[k], b[k]) for k in a)


def test_memmap_directory_round_trip(tmp_path):
    """flush_memmap / load_memmap: `memmap_dir/env_{i}/{key}.memmap` raw arrays [buffer_size, 1, *feature] as the reference lays them
    out (buffers.py:134-147, :637) + meta.pt; a fresh buffer loaded from the directory samples the same sequences."""
    import numpy as np

    rb = CpuAsync(7, 2, device="cpu", memmap=True, memmap_dir=tmp_path, sequential=True)
    rb.add({"state": torch.randn(10, 2, 3), "rgb": torch.randint(0, 256, (10, 2, 3, 4, 4), dtype=torch.uint8), "dones": torch.zeros(10, 2, 1)})
    rb.add({"state": torch.randn(2, 1, 3), "rgb": torch.zeros(2, 1, 3, 4, 4, dtype=torch.uint8), "dones": torch.ones(2, 1, 1)}, [1])
    rb.flush_memmap()
    raw = np.memmap(tmp_path / "env_1" / "state.memmap", dtype=np.float32, mode="r", shape=(7, 1, 3))
    assert np.array_equal(np.asarray(raw)[:, 0], rb._store["state"][1].numpy())
    assert (tmp_path / "env_0" / "rgb.memmap").stat().st_size == 7 * 3 * 4 * 4
    rb2 = CpuAsync(7, 2, device="cpu", memmap=True, memmap_dir=tmp_path, sequential=True)
    rb2.load_memmap()
    assert rb2._pos == rb._pos and rb2._full == rb._full
    assert all(torch.equal(rb2._store[k], rb._store[k]) and rb2._store[k].dtype == rb._store[k].dtype for k in rb._store)
    torch.manual_seed(3)
    a = rb.sample(4, sequence_length=3)
    torch.manual_seed(3)
    b = rb2.sample(4, sequence_length=3)
    assert all(torch.equal(a[k], b[k]) for k in a)
    with pytest.raises(ValueError):
        CpuAsync(8, 2, device="cpu").load_memmap(tmp_path)


@pytest.mark.gpu
def test_gather_with_more_keys_than_one_launch_takes():
    """More than 8 keys (4 bookkeeping keys + many observation keys): the gather is issued in chunks of 8 keys."""
    dev = "cuda:0"
    rb = AsyncReplayBuffer(64, 2, device=dev, sequential=True)
    g = torch.Generator().manual_seed(3)
    data = {f"obs{i}": torch.randn(80, 2, 3 + i, generator=g) for i in range(9)}
    data.update(actions=torch.randn(80, 2, 4, generator=g), rewards=torch.randn(80, 2, 1, generator=g), dones=torch.zeros(80, 2, 1),
                is_first=torch.zeros(80, 2, 1))
    rb.add(data)
    torch.manual_seed(5)
    env_id, start, col, samp = rb.plan(6, 7, 2)
    out = rb.gather(env_id, start, col, samp, 6, 7, 2)
    t = torch.arange(7)
    for j in range(env_id.numel()):
        rows = (int(start[j]) + t) % 64
        for k in data:
            assert torch.equal(out[k][int(samp[j]), :, int(col[j])].cpu(), rb._store[k][int(env_id[j])].cpu()[rows]), k
