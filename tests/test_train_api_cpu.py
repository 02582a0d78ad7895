"""Reference-facing Python API on CPU (host logic; the kernels are replaced by their torch statement through the test seam):
module tree / state_dict keys, train() side effects, optimizer adoption, player step."""
import pytest
import torch

from oracle import harness
from sheeprl_prey_b200.algos.dreamer_v3.agent import PlayerDV3, build_models
from sheeprl_prey_b200.algos.dreamer_v3.dreamer_v3 import Moments, train
from sheeprl_prey_b200.configs import ObsSpace, make_cfg
from sheeprl_prey_b200.engine import METRIC_KEYS
from util import Aggregator, FakeFabric, cpu_ops, load_golden


@pytest.mark.parametrize("fixture", ["dv3_step.pt", "dv3_step_pixels.pt"])
def test_state_dict_keys_match_reference(fixture):
    g = load_golden(fixture)
    cfg = make_cfg(**g["cfg"])
    space = ObsSpace({k: ((d,) if isinstance(d, int) else tuple(d)) for k, d in g["obs"].items()})
    with cpu_ops():
        wm, actor, critic, target = build_models(FakeFabric("cpu"), g["actions_dim"], False, cfg, space)
    for mod, name in ((wm, "wm"), (actor, "actor"), (critic, "critic"), (target, "target")):
        ref = g["params0"][name]
        sd = mod.state_dict()
        assert list(sd.keys()) == list(ref.keys()), name          # same names, same order
        for k in ref:
            assert tuple(sd[k].shape) == tuple(ref[k].shape), k
    # loading the reference's weights through the positional state arguments (agent.py:1046-1051)
    with cpu_ops():
        wm2, *_ = build_models(FakeFabric("cpu"), g["actions_dim"], False, cfg, space, g["params0"]["wm"], g["params0"]["actor"],
                               g["params0"]["critic"], g["params0"]["target"])
    for k, v in g["params0"]["wm"].items():
        assert torch.equal(wm2.state_dict()[k], v)


def test_train_side_effects_and_reference_metrics():
    g = load_golden("dv3_step.pt")
    cfg = make_cfg(**g["cfg"])
    space = ObsSpace({k: (d,) for k, d in g["obs"].items()})
    fab = FakeFabric("cpu")
    with cpu_ops():
        wm, actor, critic, target = build_models(fab, g["actions_dim"], False, cfg, space, g["params0"]["wm"], g["params0"]["actor"],
                                                 g["params0"]["critic"], g["params0"]["target"])
    a = cfg.algo
    wo = torch.optim.Adam(wm.parameters(), lr=a.world_model.optimizer.lr, eps=a.world_model.optimizer.eps)
    ao = torch.optim.Adam(actor.parameters(), lr=a.actor.optimizer.lr, eps=a.actor.optimizer.eps)
    co = torch.optim.Adam(critic.parameters(), lr=a.critic.optimizer.lr, eps=a.critic.optimizer.eps)
    mom = Moments(fab, 0.99, 1.0, 0.05, 0.95)
    for it in range(2):
        data = {k: v.clone() for k, v in g["data"].items()}
        data["is_first"][0] = 0.0
        agg = Aggregator()
        train(fab, wm, actor, critic, target, wo, ao, co, data, agg, cfg, False, g["actions_dim"], mom, noise=g["noise"])
        ref = g["steps"][it]
        assert float(data["is_first"][0].min()) == 1.0                       # dreamer_v3.py:104
        for k in METRIC_KEYS:
            assert abs(agg.values[k] - ref["metrics"][k]) <= 3e-5 * max(1.0, abs(ref["metrics"][k])), (it, k)
        assert abs(float(mom.low) - ref["moments_low"]) < 1e-6 and abs(float(mom.high) - ref["moments_high"]) < 1e-6
    st = wo.state_dict()["state"]
    assert float(st[0]["step"]) == 2.0 and float(st[0]["exp_avg"].abs().sum()) > 0     # adopted optimizer carries the fused state
    assert all(p.grad is None for p in wm.parameters())                                  # dreamer_v3.py:349-351
    for k, v in g["params2"]["wm"].items():
        assert float((wm.state_dict()[k] - v).abs().max()) <= 5e-5 * float(v.abs().max() + 1e-12), k


def test_player_steps_with_trained_modules():
    cfg = make_cfg("XS", batch_size=2, sequence_length=4, horizon=2, deter=16, hidden=16, units=16, layers=1)
    with cpu_ops():
        wm, actor, critic, target = build_models(FakeFabric("cpu"), [3], False, cfg, ObsSpace({"state": (5,)}))
    w = cfg.algo.world_model
    player = PlayerDV3(wm.encoder, wm.rssm, actor, [3], 4, w.stochastic_size, w.recurrent_model.recurrent_state_size, "cpu", w.discrete_size)
    player.init_states()
    acts = player.get_greedy_action({"state": torch.randn(1, 4, 5)})
    assert acts[0].shape == (1, 4, 3) and float(acts[0].sum()) == 4.0
    player.init_states([1, 3])
    assert float(player.recurrent_state[:, [1, 3]].abs().sum()) == 0.0


def test_state_loaded_after_the_first_step_reaches_the_engine():
    """Resume path: `moments.load_state_dict` / `optimizer.load_state_dict` called AFTER train() has run once must not be ignored
    (the engine owns the device state; the module buffers / optimizer state tensors alias it)."""
    import copy

    from sheeprl_prey_b200.algos.dreamer_v3.dreamer_v3 import get_engine

    g = load_golden("dv3_step.pt")
    cfg = make_cfg(**g["cfg"])
    space = ObsSpace({k: (d,) for k, d in g["obs"].items()})
    fab = FakeFabric("cpu")
    with cpu_ops():
        wm, actor, critic, target = build_models(fab, g["actions_dim"], False, cfg, space, g["params0"]["wm"], g["params0"]["actor"],
                                                 g["params0"]["critic"], g["params0"]["target"])
    a = cfg.algo
    wo = torch.optim.Adam(wm.parameters(), lr=a.world_model.optimizer.lr, eps=a.world_model.optimizer.eps)
    ao = torch.optim.Adam(actor.parameters(), lr=a.actor.optimizer.lr, eps=a.actor.optimizer.eps)
    co = torch.optim.Adam(critic.parameters(), lr=a.critic.optimizer.lr, eps=a.critic.optimizer.eps)
    mom = Moments(fab, 0.99, 1.0, 0.05, 0.95)
    eng = get_engine(wm)

    def step():
        train(fab, wm, actor, critic, target, wo, ao, co, {k: v.clone() for k, v in g["data"].items()}, Aggregator(), cfg, False,
              g["actions_dim"], mom, noise=g["noise"])

    step()
    saved_opt, saved_mom = copy.deepcopy(wo.state_dict()), {k: v.clone() for k, v in mom.state_dict().items()}
    m_after_1 = eng.wm.m.clone()
    step()
    assert not torch.equal(eng.wm.m, m_after_1)
    # resume from the checkpoint taken after step 1
    wo.load_state_dict(saved_opt)
    mom.load_state_dict(saved_mom)
    assert abs(float(eng.mom_state[0]) - float(saved_mom["low"])) < 1e-12          # buffers alias the engine state
    step()                                                                          # re-imports the optimizer moments first
    st = wo.state_dict()["state"]
    assert float(st[0]["step"]) == 2.0                                              # step 1 (loaded) + this one
    assert wo.state[wo.param_groups[0]["params"][0]]["exp_avg"].data_ptr() == eng.wm.m.data_ptr()
    # lr changes are picked up (and would drop a captured graph)
    wo.param_groups[0]["lr"] = 3e-5
    step()
    assert eng.opt_hparams["world"]["lr"] == 3e-5


def test_checkpoint_round_trip_host_logic(tmp_path, monkeypatch):
    """The save -> file -> resume test of the GPU suite (tests/test_step_gpu.py) run through the CPU seam: state_dict layout, optimizer
    adoption and the noise counter carried by `moments.state_dict()` are host logic."""
    import test_step_gpu as gpu_tests

    monkeypatch.setattr(gpu_tests, "DEV", "cpu")
    with cpu_ops():
        gpu_tests.test_checkpoint_round_trip_on_gpu(tmp_path)
