"""Data-parallel path on CPU: world_size 2 over gloo (host logic of SURVEY.md 8e: per-rank batches, gradient
all-reduce(mean) per flat arena, all_gather of lambda-values inside Moments, weights identical across ranks)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    from oracle import harness
    from oracle.torch_ops import TorchOps
    from sheeprl_prey_b200.configs import make_cfg
    from sheeprl_prey_b200.engine import Arena, DV3Engine
    from sheeprl_prey_b200.parallel import TorchDist
    from util import load_golden, make_arena

    g = load_golden("dv3_step.pt")
    cfg = make_cfg(**g["cfg"])
    T, B, H = cfg.per_rank_sequence_length, cfg.per_rank_batch_size, cfg.algo.horizon
    P = g["params0"]

    def engine(d):
        return DV3Engine(cfg, g["obs"], g["actions_dim"], make_arena(Arena, P["wm"], "cpu"), make_arena(Arena, P["actor"], "cpu"),
                         make_arena(Arena, P["critic"], "cpu"), make_arena(Arena, P["target"], "cpu", False), TorchOps(), "cpu", dist=d)

    data = [harness.synthetic_batch(T, B, g["obs"], g["actions_dim"], seed=100 + r, done_p=0.2) for r in range(world)]
    noise = [harness.synthetic_noise(T, B, H, cfg.algo.world_model.stochastic_size, len(g["actions_dim"]), seed=200 + r) for r in range(world)]
    eng = engine(TorchDist(world))
    eng.load_batch(data[rank])
    eng.load_noise(noise[rank])
    eng.step()
    # single-process engines on every rank's data: their mean world-model gradient must equal the all-reduced one
    locals_ = []
    for r in range(world):
        e = engine(None)
        e.load_batch(data[r])
        e.load_noise(noise[r])
        e.world_forward()
        e.world_losses()
        e.world_backward()
        locals_.append(e.wm.grad.clone())
    mean_grad = sum(locals_) / world
    ok_grad = float((eng.wm.grad - mean_grad).abs().max()) <= 1e-6 * float(mean_grad.abs().max())
    ok_grad = ok_grad and 0 < eng.wm_mid_from < eng.wm_tail_from and len(eng._wm_pending) == 2     # both early buckets were used
    torch.save(dict(wm=eng.wm.flat.clone(), actor=eng.actor.flat.clone(), critic=eng.critic.flat.clone(), mom=eng.mom_state.clone(),
                    ok_grad=ok_grad, lam_all=eng.lam_all.clone(), lam=eng.lam.clone()), os.path.join(tmp, f"rank{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_world_size_2_gloo(tmp_path):
    world, port = 2, 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(os.path.join(tmp_path, f"rank{i}.pt")) for i in range(world)]
    assert r[0]["ok_grad"] and r[1]["ok_grad"]
    for k in ("wm", "actor", "critic", "mom"):
        assert torch.equal(r[0][k], r[1][k]), k                  # replicas stay bit-identical
    assert torch.equal(r[0]["lam_all"], torch.cat((r[0]["lam"], r[1]["lam"])))
    assert not torch.equal(r[0]["lam"], r[1]["lam"])             # ranks really saw different data
