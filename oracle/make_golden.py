"""TEST INFRASTRUCTURE ONLY.  Generates tests/golden/*.pt by running the UNMODIFIED reference
(/root/reference, imported through oracle/ref_shim.py) -- only possible in the build container.

    python -m oracle.make_golden

dv3_step.pt : reference build_models() weights (perturbed so no head is degenerate), a synthetic batch, the injected
              sampling noise, and the reference train() outputs for two consecutive gradient steps (the 13 aggregator
              metrics, drawn indices, Moments buffers, parameters after step 2).
dv3_xs_b16t64.npz : BASELINE.json config 1 at its real shape (DreamerV3-XS, obs 4 / Discrete(2), B16 x T64, H15): the reference train()
              run for two gradient steps on seeded parameters / batch / noise that the tests regenerate from the seeds
              (tests/util.py::random_params_like, oracle/harness.py), so the file holds only results: the 13 metrics, Moments,
              every drawn index (uint8) and per-tensor parameter norms before and after.
episode_buffer.pt : a scripted add / evict / sample session on the reference EpisodeBuffer (buffers.py:342-525)
buffers.pt  : a scripted add/sample session on the reference AsyncReplayBuffer(sequential=True) under a fixed
              torch seed (everything added, every sampled batch).
"""
from __future__ import annotations

import os

import torch

from oracle import harness, ref_shim
from sheeprl_prey_b200.configs import ObsSpace, make_cfg

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

STEP_CFG = dict(size="XS", batch_size=3, sequence_length=6, horizon=4, deter=32, hidden=24, units=40, layers=2,
                stochastic_size=8, discrete_size=16)
STEP_OBS = {"state": 6}
STEP_ACTIONS = [3, 5]


PIX_CFG = dict(size="XS", batch_size=3, sequence_length=4, horizon=3, deter=24, hidden=16, units=24, layers=1, stochastic_size=4,
               discrete_size=8, cnn_keys=("rgb",), mlp_keys=("state",), cnn_mult=2)
PIX_OBS = {"rgb": (3, 64, 64), "state": 5}
PIX_ACTIONS = [4]


def make_step_golden(STEP_CFG=STEP_CFG, STEP_OBS=STEP_OBS, STEP_ACTIONS=STEP_ACTIONS, fname="dv3_step.pt"):
    dv3, agent, loss, utils, buffers, dist = ref_shim.ref_modules()
    from sheeprl.utils.utils import dotdict as ref_dotdict

    torch.manual_seed(0)
    cfg = ref_dotdict(make_cfg(**STEP_CFG))
    T, B, H = cfg.per_rank_sequence_length, cfg.per_rank_batch_size, cfg.algo.horizon
    G = cfg.algo.world_model.stochastic_size
    fab = ref_shim.FakeFabric()
    wm, actor, critic, tcritic = agent.build_models(fab, STEP_ACTIONS, False, cfg,
                                                    ObsSpace({k: ((d,) if isinstance(d, int) else d) for k, d in STEP_OBS.items()}))
    with torch.no_grad():
        for p in list(critic.parameters()) + list(wm.reward_model.parameters()):
            p.add_(0.05 * torch.randn_like(p))
        for p in tcritic.parameters():
            p.add_(0.03 * torch.randn_like(p))
        for m_ in (wm, actor, critic):
            for _, p in m_.named_parameters():
                if p.dim() == 1:
                    p.add_(0.1 * torch.randn_like(p))
    S = harness.strip_wrapper_keys
    params0 = dict(wm=S(wm.state_dict()), actor=S(actor.state_dict()), critic=S(critic.state_dict()), target=S(tcritic.state_dict()))
    data = harness.synthetic_batch(T, B, STEP_OBS, STEP_ACTIONS, done_p=0.2)
    noise = harness.synthetic_noise(T, B, H, G, len(STEP_ACTIONS))
    a = cfg.algo
    wo = torch.optim.Adam(wm.parameters(), lr=a.world_model.optimizer.lr, eps=a.world_model.optimizer.eps)
    ao = torch.optim.Adam(actor.parameters(), lr=a.actor.optimizer.lr, eps=a.actor.optimizer.eps)
    co = torch.optim.Adam(critic.parameters(), lr=a.critic.optimizer.lr, eps=a.critic.optimizer.eps)
    mom = utils.Moments(fab, a.actor.moments.decay, a.actor.moments.max, a.actor.moments.percentile.low, a.actor.moments.percentile.high)
    steps = []
    for it in range(2):
        agg = ref_shim.Aggregator()
        tape = ref_shim.NoiseTape(harness.noise_tape_order(noise, T, H, len(STEP_ACTIONS), T * B))
        tape.install(dist)
        try:
            dv3.train(fab, wm, actor, critic, tcritic, wo, ao, co, {k: v.clone() for k, v in data.items()}, agg, cfg, False,
                      STEP_ACTIONS, mom)
        finally:
            tape.uninstall()
        ix = tape.indices
        nh = len(STEP_ACTIONS)
        post_idx = torch.stack([ix[2 * t + 1] for t in range(T)]).reshape(T, B, G)
        o = 2 * T
        act_idx = [torch.stack([ix[o + i].reshape(-1) for i in range(nh)], -1)]
        o += nh
        z_idx = []
        for h in range(H):
            z_idx.append(ix[o].reshape(T * B, G))
            o += 1
            act_idx.append(torch.stack([ix[o + i].reshape(-1) for i in range(nh)], -1))
            o += nh
        steps.append(dict(metrics=dict(agg.values), post_idx=post_idx.int(), z_idx=torch.stack(z_idx).int(),
                          act_idx=torch.stack(act_idx).int(), moments_low=float(mom.low), moments_high=float(mom.high)))
    params2 = dict(wm=S(wm.state_dict()), actor=S(actor.state_dict()), critic=S(critic.state_dict()))
    torch.save(dict(cfg=STEP_CFG, obs=STEP_OBS, actions_dim=STEP_ACTIONS, params0=params0, data=data, noise=noise, steps=steps,
                    params2=params2, torch_version=torch.__version__), os.path.join(GOLDEN_DIR, fname))


BASE_KW = dict(size="XS", batch_size=16, sequence_length=64, horizon=15)
BASE_OBS = {"state": 4}
BASE_ACTIONS = [2]
BASE_SEEDS = dict(params=11, data=1234, noise=99)


def make_baseline_golden(fname="dv3_xs_b16t64.npz"):
    """BASELINE config 1 at full size through the UNMODIFIED reference; inputs are regenerated from seeds by the tests."""
    import json
    import sys

    import numpy as np

    sys.path.insert(0, os.path.join(os.path.dirname(GOLDEN_DIR)))
    from util import random_params_like

    dv3, agent, loss, utils, buffers, dist = ref_shim.ref_modules()
    from sheeprl.utils.utils import dotdict as ref_dotdict

    cfg = ref_dotdict(make_cfg(**BASE_KW))
    T, B, H = cfg.per_rank_sequence_length, cfg.per_rank_batch_size, cfg.algo.horizon
    G, nh = cfg.algo.world_model.stochastic_size, len(BASE_ACTIONS)
    params = random_params_like(make_cfg(**BASE_KW), BASE_OBS, BASE_ACTIONS, seed=BASE_SEEDS["params"])
    fab = ref_shim.FakeFabric()
    wm, actor, critic, tcritic = agent.build_models(fab, BASE_ACTIONS, False, cfg, ObsSpace({k: (d,) for k, d in BASE_OBS.items()}),
                                                    params["wm"], params["actor"], params["critic"], params["target"])
    data = harness.synthetic_batch(T, B, BASE_OBS, BASE_ACTIONS, seed=BASE_SEEDS["data"], done_p=1 / 50)
    noise = harness.synthetic_noise(T, B, H, G, nh, seed=BASE_SEEDS["noise"])
    a = cfg.algo
    wo = torch.optim.Adam(wm.parameters(), lr=a.world_model.optimizer.lr, eps=a.world_model.optimizer.eps)
    ao = torch.optim.Adam(actor.parameters(), lr=a.actor.optimizer.lr, eps=a.actor.optimizer.eps)
    co = torch.optim.Adam(critic.parameters(), lr=a.critic.optimizer.lr, eps=a.critic.optimizer.eps)
    mom = utils.Moments(fab, a.actor.moments.decay, a.actor.moments.max, a.actor.moments.percentile.low, a.actor.moments.percentile.high)
    S = harness.strip_wrapper_keys
    arrays, meta = {}, dict(kw=BASE_KW, obs=BASE_OBS, actions_dim=BASE_ACTIONS, seeds=BASE_SEEDS, torch_version=torch.__version__, steps=[])
    meta["norms0"] = {n: {k: float(v.double().norm()) for k, v in params[n].items()} for n in ("wm", "actor", "critic", "target")}
    for it in range(2):
        agg = ref_shim.Aggregator()
        tape = ref_shim.NoiseTape(harness.noise_tape_order(noise, T, H, nh, T * B))
        tape.install(dist)
        try:
            dv3.train(fab, wm, actor, critic, tcritic, wo, ao, co, {k: v.clone() for k, v in data.items()}, agg, cfg, False, BASE_ACTIONS, mom)
        finally:
            tape.uninstall()
        ix = tape.indices
        post_idx = torch.stack([ix[2 * t + 1] for t in range(T)]).reshape(T, B, G)
        o = 2 * T
        act_idx = [torch.stack([ix[o + i].reshape(-1) for i in range(nh)], -1)]
        o += nh
        z_idx = []
        for h in range(H):
            z_idx.append(ix[o].reshape(T * B, G))
            o += 1
            act_idx.append(torch.stack([ix[o + i].reshape(-1) for i in range(nh)], -1))
            o += nh
        arrays[f"post_idx{it}"] = post_idx.to(torch.uint8).numpy()
        arrays[f"z_idx{it}"] = torch.stack(z_idx).to(torch.uint8).numpy()
        arrays[f"act_idx{it}"] = torch.stack(act_idx).to(torch.uint8).numpy()
        meta["steps"].append(dict(metrics={k: float(v) for k, v in agg.values.items()}, moments_low=float(mom.low), moments_high=float(mom.high)))
    meta["norms2"] = {n: {k: float(v.double().norm()) for k, v in S(m.state_dict()).items()} for n, m in (("wm", wm), ("actor", actor), ("critic", critic))}
    arrays["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(GOLDEN_DIR, fname), **arrays)


def make_buffer_golden():
    dv3, agent, loss, utils, buffers, dist = ref_shim.ref_modules()
    torch.manual_seed(1234)
    size, n_envs, T = 23, 3, 5
    rb = buffers.AsyncReplayBuffer(size, n_envs, device="cpu", memmap=False, sequential=True)
    g = torch.Generator().manual_seed(5)
    script, samples = [], []

    def mk(L, n):
        return {"state": torch.randn(L, n, 4, generator=g), "rgb": torch.randint(0, 256, (L, n, 3, 4, 4), generator=g, dtype=torch.uint8),
                "dones": (torch.rand(L, n, 1, generator=g) < 0.2).float()}

    def add(L, idx=None):
        d = mk(L, rb.n_envs if idx is None else len(idx))
        rb.add(ref_shim.TD(d, batch_size=[L, d["state"].shape[1]]), idx)
        script.append(("add", d, idx))

    def sample(bs, n):
        s = rb.sample(bs, sequence_length=T, n_samples=n)
        script.append(("sample", bs, n))
        samples.append({k: s[k].clone() for k in s.keys()})

    for _ in range(9):
        add(1)
    sample(4, 1)
    add(1, [0, 2])
    add(3)
    sample(6, 2)
    for _ in range(12):
        add(1)            # wraps around: buffers become full
    sample(5, 1)
    add(1, [1])
    sample(7, 3)
    sessions = [dict(size=size, n_envs=n_envs, T=T, seed=1234, script=script, samples=samples,
                     pos=[b._pos for b in rb.buffer], full=[b.full for b in rb.buffer])]
    # session 2: oversize adds on a not-yet-full ring (buffers.py:123-131; reference test_buffers.py:50-66)
    torch.manual_seed(77)
    size, n_envs, T = 7, 2, 3
    rb = buffers.AsyncReplayBuffer(size, n_envs, device="cpu", memmap=False, sequential=True)
    script, samples = [], []
    add(2)
    add(1)
    add(9)
    sample(5, 2)
    add(3, [1])
    sample(3, 1)
    sessions.append(dict(size=size, n_envs=n_envs, T=T, seed=77, script=script, samples=samples,
                         pos=[b._pos for b in rb.buffer], full=[b.full for b in rb.buffer]))
    torch.save(sessions, os.path.join(GOLDEN_DIR, "buffers.pt"))


def episode_session(make_buffer, make_episode=lambda d, L: d):
    """The scripted EpisodeBuffer session behind tests/golden/episode_buffer.pt: adds with eviction, plain and end-prioritised samples.
    `make_buffer(size, sl)` builds the buffer under test, `make_episode(dict, L)` wraps an episode in the container it expects."""
    torch.manual_seed(4321)
    g = torch.Generator().manual_seed(8)
    size, sl = 60, 7
    rb = make_buffer(size, sl)
    samples = []

    def add(L):
        d = {"state": torch.randn(L, 3, generator=g), "rgb": torch.randint(0, 256, (L, 2, 3, 3), generator=g, dtype=torch.uint8),
             "dones": torch.zeros(L, 1)}
        d["dones"][-1] = 1
        rb.add(make_episode(d, L))

    def sample(bs, n, ends=False):
        s = rb.sample(bs, n_samples=n, prioritize_ends=ends)
        samples.append({k: s[k].clone() for k in s.keys()})

    for L in (9, 14, 7):
        add(L)
    sample(5, 2)
    add(22)
    sample(4, 1, ends=True)
    add(19)                  # 9 + 14 + 7 + 22 + 19 > 60: the two oldest episodes go
    sample(6, 3)
    add(60)                  # an episode as long as the buffer evicts everything else
    sample(3, 2, ends=True)
    add(8)
    sample(7, 1)
    return dict(size=size, sl=sl, samples=samples, lengths=[int(e.shape[0]) for e in rb.buffer], total=len(rb), full=bool(rb.full))


def make_episode_golden():
    dv3, agent, loss, utils, buffers, dist = ref_shim.ref_modules()
    import warnings

    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        sess = episode_session(lambda size, sl: buffers.EpisodeBuffer(size, sl, device="cpu", memmap=False),
                               lambda d, L: ref_shim.TD(d, batch_size=[L]))
    torch.save(sess, os.path.join(GOLDEN_DIR, "episode_buffer.pt"))


if __name__ == "__main__":
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    make_step_golden()
    make_step_golden(PIX_CFG, PIX_OBS, PIX_ACTIONS, "dv3_step_pixels.pt")
    make_buffer_golden()
    make_episode_golden()
    make_baseline_golden()
    for f in sorted(os.listdir(GOLDEN_DIR)):
        print(f, os.path.getsize(os.path.join(GOLDEN_DIR, f)))
