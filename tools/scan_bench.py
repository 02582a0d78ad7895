#!/usr/bin/env python
"""Times dv3_rssm_observe_fwd alone (config S shapes) and prints CTA-0 clock stamps of one step (DV3_SCAN_DEBUG=1)."""
import os
import sys
import time

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
dev = "cuda:0"
cfg = make_cfg(size, batch_size=16, sequence_length=64, horizon=15)
od, acts = {"state": 14}, [100]
params = random_params_like(cfg, od, acts, seed=5)
ops = CudaOps(dev)
eng = engine_from_params(cfg, od, acts, {n: {k: v.to(dev) for k, v in d.items()} for n, d in params.items()}, ops, dev)
eng.load_batch({k: v.to(dev) for k, v in harness.synthetic_batch(64, 16, od, acts, seed=1).items()})
eng.load_noise(None)
eng.world_forward()
torch.cuda.synchronize()
for fused in (True, False):
    eng.use_fused_scan = fused
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    eng.world_forward()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        if fused:
            ops.rssm_observe_fwd(eng, None)
        else:
            eng.world_forward()
    e1.record()
    torch.cuda.synchronize()
    print("fused scan kernel" if fused else "whole world_forward per-op", e0.elapsed_time(e1) / 5, "ms")
NC = torch.cuda.get_device_properties(0).multi_processor_count
off = NC * 16 * 2
raw = ops.scratch[off:off + 64].view(torch.int64)[:19].cpu().tolist()
print("fine stamps (clk since step start): P0 loaded", raw[11] - raw[0], "| P0 saved", raw[12] - raw[0], "| P1 done", raw[1] - raw[0])
print("  P2: feat loaded->LN done", raw[13] - raw[2], "| writes", raw[14] - raw[13], "| gemv", raw[15] - raw[14], "| tail", raw[3] - raw[15])
print("  P5: load", raw[16] - raw[8], "| LN", raw[17] - raw[16], "| hid write+gemv", raw[18] - raw[17], "| sync+sample", raw[9] - raw[18])
stamps = raw[:11]
names = ["P0 start", "P1 done", "sync1+load feat", "P2 done (LN, gates gemv)", "sync2+load partials", "P3 done", "sync3+load h",
         "P4 done", "sync4", "P5 done", "sync5"]
for i in range(1, 11):
    print(f"{names[i]:28s} {stamps[i] - stamps[i - 1]:8d} clk")
print("step total", stamps[10] - stamps[0], "clk")


# ---- cluster kernel (bf16 precision)
ops16 = CudaOps(dev, precision="bf16")
eng16 = engine_from_params(cfg, od, acts, {n: {k: v.to(dev) for k, v in d.items()} for n, d in params.items()}, ops16, dev)
eng16.load_batch({k: v.to(dev) for k, v in harness.synthetic_batch(64, 16, od, acts, seed=1).items()})
eng16.load_noise(None)
eng16.world_forward()
torch.cuda.synchronize()
print("scan kernel in bf16 mode:", ops16.scan_kernel)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    ops16.rssm_observe_fwd(eng16, None)
e1.record()
torch.cuda.synchronize()
print("cluster scan kernel", e0.elapsed_time(e1) / 5, "ms")
st = ops16.scratch[:64].view(torch.int64)[:32].cpu().tolist()
nm = ["P0 masks/saves", "P1 gather+partials", "sync1+LN+allgather", "sync2+gates contraction", "gates partials..sync3..h'..allgather",
      "sync4+hidr contraction+partials", "sync5+LN+allgather", "sync6+logits contraction", "sampling", "sync7"]
for i in range(10):
    print(f"  {nm[i]:40s} {st[i + 1] - st[i]:8d} clk")
print("  step total", st[10] - st[0], "clk")
print("  cluster.sync waits (CTA 0, thread 0):", [st[10 + 2 * n] - st[10 + 2 * n - 1] for n in range(1, 8)])


# ---- reverse scan: per-step composition vs cluster kernel
eng16.world_losses()
for fused in (False, True):
    eng16.use_fused_bwd = fused
    eng16.world_backward()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        eng16.world_backward()
    e1.record()
    torch.cuda.synchronize()
    print("world_backward with", "cluster reverse scan" if fused else "per-step reverse scan (eager)", e0.elapsed_time(e1) / 3, "ms", ops16.bwd_scan_kernel)
e0.record()
for _ in range(5):
    ops16.rssm_observe_bwd(eng16)
e1.record()
torch.cuda.synchronize()
print("cluster reverse scan kernel (+prep)", e0.elapsed_time(e1) / 5, "ms")
sb = ops16.scratch[:64].view(torch.int64)[:16].cpu().tolist()
nb = ["B1 unimix bwd + allgather", "sync1", "store + contract Wr1T + LN bwd a + push", "sync2", "combine + LN bwd b + allgather", "sync3",
      "stores + contract Wr0hT + GRU bwd + push", "sync4", "combine + LN bwd b x3 + allgather x3", "sync5",
      "stores + contract WgruT + LN bwd a + push", "sync6", "combine + LN bwd b + allgather", "sync7", "stores + contract WmlpzT + carry"]
for i in range(15):
    print(f"  {nb[i]:46s} {sb[i + 1] - sb[i]:8d} clk")
print("  step total", sb[15] - sb[0], "clk")
