#!/usr/bin/env python
"""Loss curves overlaid (north_star): N free-running gradient steps of DreamerV3-XS (BASELINE config 1: obs 4 / Discrete(2), B16 x T64,
H15) on fixed synthetic data, run by (a) the CPU oracle (fp32, the pinned restatement of the reference's train()), (b) the CUDA library
in fp32 mode and (c) in bf16 tensor-core mode, all from the same parameters, batches and sampling uniforms.  Free-running: every
implementation samples its own categoricals from the shared uniforms, so trajectories separate wherever a uniform sits next to a CDF
step; what must agree is the curve, not the draw.  Writes the 13 metrics per step and the gap between the smoothed curves.

    python tools/loss_curves.py [--steps 200] [--out profiles/r02_loss_curves.json]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

SIZE, B, T, H = "XS", 16, 64, 15
OBS, ACTS = {"state": 4}, [2]
N_BATCHES = 4          # the fixed data set: four batches, visited round-robin


def _inputs(step):
    from oracle import harness

    data = harness.synthetic_batch(T, B, OBS, ACTS, seed=1234 + (step % N_BATCHES), done_p=1 / 50)
    noise = harness.synthetic_noise(T, B, H, 32, len(ACTS), seed=50_000 + step)
    return data, noise


def run_curves(n_steps: int, impls=("oracle", "cuda-fp32", "cuda-bf16"), device="cuda:0", log=None):
    from oracle import dv3_oracle as O
    from sheeprl_prey_b200.configs import make_cfg
    from sheeprl_prey_b200.engine import METRIC_KEYS
    from util import engine_from_params, random_params_like

    cfg = make_cfg(SIZE, batch_size=B, sequence_length=T, horizon=H)
    params = random_params_like(cfg, OBS, ACTS, seed=11)
    tau = cfg.algo.critic.tau
    curves = {}
    for impl in impls:
        t0 = time.time()
        rows = {k: [] for k in METRIC_KEYS}
        if impl == "oracle":
            tr = O.OracleTrainer(cfg, OBS, ACTS, params["wm"], params["actor"], params["critic"], params["target"])
            for s in range(n_steps):
                data, noise = _inputs(s)
                tr.update_target(1.0 if s == 0 else tau)          # dreamer_v3.py:711-716 (before the train() call)
                out = tr.step(data, noise, keep=False)
                for k in METRIC_KEYS:
                    rows[k].append(float(out[k]))
        else:
            from sheeprl_prey_b200._lib import CudaOps

            ops = CudaOps(device, precision="fp32" if impl == "cuda-fp32" else "bf16")
            eng = engine_from_params(cfg, OBS, ACTS, {n: {k: v.to(device) for k, v in d.items()} for n, d in params.items()}, ops, device)
            for s in range(n_steps):
                data, noise = _inputs(s)
                eng.update_target(1.0 if s == 0 else tau)
                eng.load_batch({k: v.to(device) for k, v in data.items()})
                eng.load_noise({k: v.to(device) for k, v in noise.items()})
                eng.step()
                md = eng.metrics_dict()
                for k in METRIC_KEYS:
                    rows[k].append(md[k])
        curves[impl] = rows
        if log:
            log(f"{impl}: {n_steps} steps in {time.time() - t0:.1f} s")
    return curves


def smooth(xs, w):
    out, acc = [], 0.0
    for i, x in enumerate(xs):
        acc += x
        if i >= w:
            acc -= xs[i - w]
        out.append(acc / min(i + 1, w))
    return out


def gaps(curves, ref="oracle", window=20, skip=0):
    """max over steps >= skip of |smooth(impl) - smooth(ref)| / max(1, |smooth(ref)|), per metric."""
    res = {}
    for impl, rows in curves.items():
        if impl == ref:
            continue
        res[impl] = {}
        for k, xs in rows.items():
            a, b = smooth(xs, window), smooth(curves[ref][k], window)
            res[impl][k] = max(abs(x - y) / max(1.0, abs(y)) for x, y in list(zip(a, b))[skip:])
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "loss_curves.json"))
    args = ap.parse_args()
    curves = run_curves(args.steps, log=lambda m: print(m, file=sys.stderr, flush=True))
    g = gaps(curves, window=20, skip=20)
    out = {"config": f"DreamerV3-{SIZE} obs {OBS} actions {ACTS} B{B}xT{T} H{H}, {N_BATCHES} fixed synthetic batches round-robin, free-running draws from shared uniforms, "
                     "target critic EMA every step", "steps": args.steps, "smoothing_window": 20,
           "gap_of_smoothed_curves_vs_oracle (max over steps >= 20 of |a-b|/max(1,|b|))": g, "curves": curves, "torch": torch.__version__}
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(out, f)
    print(json.dumps({"steps": args.steps, "gaps": g}, indent=1))


if __name__ == "__main__":
    main()
