#!/usr/bin/env python
"""Times the tcgen05 scan kernels alone (CUDA events around repeated launches, weights re-packed outside the timed region) and prints the
clock64 stamps the epilogue of CTA 0 records during step 2 (DV3_SCAN_DEBUG=1).  usage: scan_bench_tc.py [size] [B] [T] [--old]"""
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

args = [a for a in sys.argv[1:] if not a.startswith("--")]
size = args[0] if len(args) > 0 else "S"
B = int(args[1]) if len(args) > 1 else 16
T = int(args[2]) if len(args) > 2 else 64
old = "--old" in sys.argv
dev = "cuda:0"
cfg = make_cfg(size, batch_size=B, sequence_length=T, horizon=15)
od, acts = {"state": 14}, [100]
params = random_params_like(cfg, od, acts, seed=5)
ops = CudaOps(dev, precision="bf16")
ops.use_tc_scan = not old
eng = engine_from_params(cfg, od, acts, {n: {k: v.to(dev) for k, v in d.items()} for n, d in params.items()}, ops, dev)
eng.load_batch({k: v.to(dev) for k, v in harness.synthetic_batch(T, B, od, acts, seed=1).items()})
eng.load_noise(None)
eng.world_forward()
eng.world_losses()
eng.world_backward()
torch.cuda.synchronize()
out = {"size": size, "B": B, "T": T, "fwd_kernel": ops.scan_kernel, "bwd_kernel": ops.bwd_scan_kernel}


def timed(fn, n=10):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


out["fwd_ms_incl_pack"] = timed(lambda: ops.rssm_observe_fwd(eng, None))
st = ops.scratch[:128].view(torch.int64)[:64].cpu().tolist()
names = ["P0 arm + wait idx + cp.async wait + bar", "P1 gather", "P1 stats exchange", "P1 LN/SiLU -> stage + fence + bar", "P1 push feat", "P1 deferred stores",
         "wait gates acc", "P2 tmem->outb + bar", "P2 stats exchange", "P2 gates -> stage + fence + bar", "P2 push h", "P2 deferred stores", "wait r0 acc",
         "P3 tmem->outb + bar", "P3 stats exchange", "P3 LN/SiLU -> stage + fence + bar", "P3 push hid", "P3 deferred stores", "wait r1 acc",
         "P4 tmem->outp + bar", "P4 sample + idx push", "P4 fetch + bar"]
if ops.scan_kernel == "tcgen05":
    out["fwd_phases_clk"] = {names[i]: st[i + 1] - st[i] for i in range(22)}
    out["fwd_step_clk"] = st[22] - st[0]
    out["fwd_gru_stats_exchange_clk"] = {"partials computed (since tmem->outb bar)": st[32] - st[8], "st.async issue": st[33] - st[32], "wait": st[34] - st[33], "combine": st[9] - st[34]}
    e = st[0]
    out["fwd_mma_warp_clk_since_step_start"] = {"arm feat": st[40] - e, "feat landed": st[41] - e, "gf issued": st[42] - e, "h landed": st[43] - e, "r0 issued": st[44] - e,
                                                "gh(t+1) issued": st[45] - e, "hid landed": st[46] - e, "r1 issued": st[47] - e}
    out["fwd_epilogue_clk_since_step_start"] = {"push feat done": st[5] - e, "gates acc ready": st[7] - e, "push h done": st[11] - e, "r0 acc ready": st[13] - e,
                                                "push hid done": st[17] - e, "r1 acc ready": st[19] - e}
out["bwd_ms_incl_prep_pack"] = timed(lambda: ops.rssm_observe_bwd(eng))
if ops.bwd_scan_kernel == "tcgen05":
    sb = ops.scratch[:64].view(torch.int64)[:16].cpu().tolist()
    nb = ["B1 wait inputs + unimix/ST Jacobian + stage + push", "B1 prefetch + wait r1T acc", "B2 tmem->outb + LN bwd a + row-sum exchange", "B2 LN bwd b + stage + push",
          "B2 prefetch + wait r0T acc", "B3 tmem->outb + GRU bwd + row-sum exchange", "B3 LN bwd b + stage + push x3", "B3 prefetch + wait gruT acc",
          "B4 tmem->outb + LN bwd a + row-sum exchange", "B4 LN bwd b + stage + push", "B4 prefetch + wait mlpT acc", "B5 tmem->outb + carry"]
    out["bwd_phases_clk"] = {nb[i]: sb[i + 1] - sb[i] for i in range(12)}
    out["bwd_step_clk"] = sb[12] - sb[0]
ops.reuse_packed_weights = True
out["fwd_ms_kernel_only"] = timed(lambda: ops.rssm_observe_fwd(eng, None))
out["bwd_ms_kernel_plus_prep"] = timed(lambda: ops.rssm_observe_bwd(eng))
print(json.dumps(out, indent=1))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", f"scan_bench_tc_{size}_B{B}{'_old' if old else ''}.json"), "w") as f:
    json.dump(out, f, indent=1)
