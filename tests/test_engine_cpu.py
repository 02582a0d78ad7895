"""Host-logic test (no GPU): the step engine's hand-written backward / scan orchestration, run with the plain-PyTorch
statement of the kernels (oracle/torch_ops.py), must reproduce the reference's train() (golden) and the oracle's autograd."""
import pytest
import torch

from oracle import dv3_oracle as O
from oracle.torch_ops import TorchOps
from sheeprl_prey_b200.configs import make_cfg
from sheeprl_prey_b200.engine import METRIC_KEYS
from util import engine_from_params, load_golden, rel


@pytest.mark.parametrize("fixture", ["dv3_step.pt", "dv3_step_pixels.pt"])
def test_engine_reproduces_reference_golden(fixture):
    g = load_golden(fixture)
    cfg = make_cfg(**g["cfg"])
    eng = engine_from_params(cfg, g["obs"], g["actions_dim"], g["params0"], TorchOps(), "cpu")
    T, B = cfg.per_rank_sequence_length, cfg.per_rank_batch_size
    for it in range(2):
        eng.load_batch(g["data"])
        eng.load_noise(g["noise"])
        ref = g["steps"][it]
        forced = None
        if it == 1:   # second step: teacher-forced draws (both routes must agree)
            forced = dict(post_idx=ref["post_idx"].reshape(T * B, -1), imagine=dict(z_idx=ref["z_idx"], act_idx=ref["act_idx"]))
        eng.step(forced)
        md = eng.metrics_dict()
        for k in METRIC_KEYS:
            assert abs(md[k] - ref["metrics"][k]) <= 3e-5 * max(1.0, abs(ref["metrics"][k])), (it, k, md[k], ref["metrics"][k])
        assert torch.equal(eng.post_idx.reshape(T, B, -1), ref["post_idx"])
        assert torch.equal(eng.z_idx, ref["z_idx"])
        assert torch.equal(eng.act_idx.reshape(ref["act_idx"].shape), ref["act_idx"])
        assert abs(float(eng.mom_state[0]) - ref["moments_low"]) < 1e-6 and abs(float(eng.mom_state[1]) - ref["moments_high"]) < 1e-6
    for name, arena in (("wm", eng.wm), ("actor", eng.actor), ("critic", eng.critic)):
        for k, v in g["params2"][name].items():
            assert rel(arena.p[k], v) < 5e-5, (name, k)


@pytest.mark.parametrize("fixture", ["dv3_step.pt", "dv3_step_pixels.pt"])
def test_engine_gradients_match_oracle_autograd(fixture):
    g = load_golden(fixture)
    cfg = make_cfg(**g["cfg"])
    eng = engine_from_params(cfg, g["obs"], g["actions_dim"], g["params0"], TorchOps(), "cpu")
    tr = O.OracleTrainer(cfg, g["obs"], g["actions_dim"], g["params0"]["wm"], g["params0"]["actor"], g["params0"]["critic"],
                         g["params0"]["target"])
    out = tr.step({k: v.clone() for k, v in g["data"].items()}, g["noise"])
    eng.load_batch(g["data"])
    eng.load_noise(g["noise"])
    eng.step()
    for nm, arena, gk, mx in (("world_model", eng.wm, "world_grads", 1000.0), ("actor", eng.actor, "actor_grads", 100.0),
                              ("critic", eng.critic, "critic_grads", 100.0)):
        coef = min(1.0, mx / (float(out[f"Grads/{nm}"]) + 1e-6))   # oracle grads are post-clip, the arena holds pre-clip
        for k in arena.g:
            assert rel(arena.g[k] * coef, out[gk][k]) < 2e-5, (nm, k)
    assert rel(eng.traj.reshape(out["traj"].shape), out["traj"]) < 1e-5
    assert rel(eng.lam.reshape(out["lambda_values"].shape), out["lambda_values"]) < 1e-5
