#!/usr/bin/env python
"""Times the imagination rollout alone (persistent kernel vs per-op composition) and prints the clock64 stamps of one epilogue thread of
the persistent kernel during imagination step 2 (wait acc / epilogue / end-of-stage barrier per stage).  usage: rollout_bench.py [size] [B] [T]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("DV3_SCAN_DEBUG", "1")
from oracle import harness  # noqa: E402
from sheeprl_prey_b200._lib import CudaOps  # noqa: E402
from sheeprl_prey_b200.configs import make_cfg  # noqa: E402
from util import engine_from_params, random_params_like  # noqa: E402

size = sys.argv[1] if len(sys.argv) > 1 else "S"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
T = int(sys.argv[3]) if len(sys.argv) > 3 else 64
dev = "cuda:0"
cfg = make_cfg(size, batch_size=B, sequence_length=T, horizon=15)
od, acts = {"state": 14}, [100]
params = random_params_like(cfg, od, acts, seed=5)
ops = CudaOps(dev, precision="bf16")
eng = engine_from_params(cfg, od, acts, {n: {k: v.to(dev) for k, v in d.items()} for n, d in params.items()}, ops, dev)
eng.load_batch({k: v.to(dev) for k, v in harness.synthetic_batch(T, B, od, acts, seed=1).items()})
eng.load_noise(None)
eng.world_forward()
torch.cuda.synchronize()
out = {"size": size, "B": B, "T": T}
for fused in (True, False):
    ops.use_rollout_tc = fused
    eng.imagine()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        eng.imagine()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    out["rollout_ms_" + ("persistent_kernel" if fused else "per_op_graph")] = e0.elapsed_time(e1) / 5
    if fused:
        st = ops.scratch[:128].view(torch.int64)[:60].cpu().tolist()
        names = []
        for s_ in ("S1 feat", "S2 gru", "S3 thid", "S4 sample z", "A0", "A1", "head"):
            names += [s_ + ": wait acc", s_ + ": epilogue", s_ + ": fence + cluster barrier (to next stage start)"]
        out["stage_clk"] = {names[i]: st[i + 1] - st[i] for i in range(min(len(names), 59)) if st[i + 1] > 0}
        out["step_clk"] = st[21] - st[0] if st[21] > 0 else None
print(json.dumps(out, indent=1))
