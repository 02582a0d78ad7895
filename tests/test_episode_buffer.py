"""EpisodeBuffer (sheeprl/data/buffers.py:342-525, the DreamerV2 / P2E sibling of the sequential buffers): a scripted session against
the golden recorded from the reference class (oracle/make_golden.py::make_episode_golden), the live reference where it is present,
and the cases of the reference's tests/test_data/test_episode_buffer.py."""
import os

import pytest
import torch

from oracle.make_golden import episode_session
from sheeprl_prey_b200.data.episode_buffer import EpisodeBuffer
from util import load_golden


def _ep(L, extra=None):
    d = {"dones": torch.zeros(L, 1)}
    d["dones"][-1] = 1
    d.update(extra or {})
    return d


def _same(a, b):
    assert a["lengths"] == b["lengths"] and a["total"] == b["total"] and a["full"] == b["full"]
    assert len(a["samples"]) == len(b["samples"])
    for sa, sb in zip(a["samples"], b["samples"]):
        for k in sb:
            assert sa[k].dtype == sb[k].dtype and torch.equal(sa[k], sb[k]), k           # bit-exact, uint8 and fp32 alike


def test_session_matches_reference_golden():
    _same(episode_session(lambda size, sl: EpisodeBuffer(size, sl, device="cpu")), load_golden("episode_buffer.pt"))


@pytest.mark.skipif(not os.path.isdir("/root/reference/sheeprl"), reason="reference tree only exists in the build container")
def test_session_matches_live_reference():
    import warnings

    from oracle import ref_shim

    buffers = ref_shim.ref_modules()[4]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        ref = episode_session(lambda size, sl: buffers.EpisodeBuffer(size, sl, device="cpu", memmap=False),
                              lambda d, L: ref_shim.TD(d, batch_size=[L]))
    _same(episode_session(lambda size, sl: EpisodeBuffer(size, sl, device="cpu")), ref)


def test_constructor_guards():
    for args in ((-1, 10), (1, -1), (5, 10)):
        with pytest.raises(ValueError):
            EpisodeBuffer(*args, device="cpu")


def test_add_episodes_and_eviction():
    sl = 5
    rb = EpisodeBuffer(30, sl, device="cpu")
    eps = [_ep(sl), _ep(sl + 5), _ep(sl + 10), _ep(sl)]
    for e in eps:
        rb.add(e)
    assert rb.full                                              # 5 + 10 + 15 = 30, then the oldest went for the last one
    assert torch.equal(rb[-1]["dones"], eps[3]["dones"]) and torch.equal(rb[0]["dones"], eps[1]["dones"])
    rb = EpisodeBuffer(5, 5, device="cpu")
    rb.add(_ep(5))
    assert rb.full and len(rb) == 5


def test_add_errors():
    rb = EpisodeBuffer(10, 5, device="cpu")
    d = {"dones": torch.zeros(3, 1)}
    with pytest.raises(RuntimeError, match="The episode must contain exactly one done"):
        rb.add(d)
    d["dones"][-3:] = 1
    with pytest.raises(RuntimeError, match="The episode must contain exactly one done"):
        rb.add(d)
    d["dones"][-2:] = 0
    with pytest.raises(RuntimeError, match="The last step must contain a done"):
        rb.add(d)
    d["dones"][-3] = 0
    d["dones"][-1] = 1
    with pytest.raises(RuntimeError, match="Episode too short"):
        rb.add(d)
    with pytest.raises(RuntimeError, match="Episode too long"):
        rb.add(_ep(15))
    with pytest.raises(KeyError, match='key "dones" not found'):
        rb.add({"t": torch.zeros(15, 1)})


def test_sample_shapes_and_windows():
    sl = 15
    rb = EpisodeBuffer(100, sl, device="cpu")
    with pytest.raises(RuntimeError, match="No sample has been added"):
        rb.sample(2, 2)
    for L, tag in ((40, -1.0), (45, -2.0), (50, -3.0)):
        rb.add(_ep(L, {"t": torch.full((L, 1), tag)}))
    with pytest.raises(ValueError, match="Batch size must be greater than 0"):
        rb.sample(-1, n_samples=2)
    with pytest.raises(ValueError, match="The number of samples must be greater than 0"):
        rb.sample(2, -1)
    s = rb.sample(50, n_samples=5)
    assert s.shape == torch.Size([5, sl, 50]) and s["t"].shape == (5, sl, 50, 1)
    t = s["t"][..., 0].transpose(1, 2).reshape(-1, sl)          # one row per sampled window
    d = s["dones"][..., 0].transpose(1, 2).reshape(-1, sl)
    assert bool(((t == t[:, :1]).all(1)).all())                 # a window never mixes episodes
    assert bool(((d.sum(1) == 0) | (d[:, -1] == 1)).all())      # a done can only be a window's last step
    assert bool((rb.sample(50, n_samples=5, prioritize_ends=True)["dones"] == 1).any())
    one = EpisodeBuffer(5, 5, device="cpu")
    e = _ep(5, {"t": torch.rand(5, 1)})
    one.add(e)
    s = one.sample(1, 1)
    assert torch.equal(s["t"][0, :, 0], e["t"]) and torch.equal(s["dones"][0, :, 0], e["dones"])
