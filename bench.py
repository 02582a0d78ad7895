#!/usr/bin/env python
"""bench.py -- DreamerV3 gradient steps/sec (B16 x T64, H15) on N B200s, one JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--size S]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" = one replay-sample (device gather) + one full train() (world model, actor, critic updates).
`value`  : steps/s with the replay shard resident in HBM, CUDA-event timed, max over ranks.
`e2e`    : the same metric through the reference-facing train() API with the batch in pinned HOST memory:
           H2D copy of the batch + train() + D2H read of the 13 metrics inside the timed region, every step.
`--impl reference`: the reference's algorithm on the host cores (oracle/dv3_oracle.py: the CPU restatement pinned
           against the reference's own train(); the reference itself is Python and is not present on the GPU box).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[1]: DreamerV3 S, prey_env vector obs (Box(14,)), Discrete(100)
    "S": dict(size="S", obs={"state": 14}, actions=[100], desc="DreamerV3-S prey_env vector obs 14 / Discrete(100)"),
    # BASELINE.json configs[0]: DreamerV3 XS, CartPole-v1 (Box(4,), Discrete(2)) -- the reference's CPU-runnable case
    "XS": dict(size="XS", obs={"state": 4}, actions=[2], desc="DreamerV3-XS CartPole-v1 vector obs 4 / Discrete(2)"),
}
WORKLOADS["M"] = dict(size="M", obs={"rgb": (3, 64, 64)}, actions=[18], desc="DreamerV3-M synthetic 64x64x3 pixel obs (Atari-shaped) / Discrete(18)")
GFLOP_PER_STEP = {"S": 542.0, "XS": 174.9, "M": 1913.4}   # SURVEY.md section 8(d), reference op count (discrete actions)


def cfg_for(wl, B, T, H):
    from sheeprl_prey_b200.configs import make_cfg

    cnn = tuple(k for k, d in wl["obs"].items() if not isinstance(d, int))
    mlp = tuple(k for k, d in wl["obs"].items() if isinstance(d, int))
    return make_cfg(wl["size"], batch_size=B, sequence_length=T, horizon=H, cnn_keys=cnn, mlp_keys=mlp)


class Fab:
    """What build_models()/train() need from lightning.Fabric."""

    def __init__(self, device, world_size, rank):
        import torch

        self.device, self.world_size, self.global_rank = torch.device(device), world_size, rank

    def setup_module(self, m):
        return m


class Agg:
    def __init__(self):
        self.v = {}

    def update(self, k, v):
        self.v[k] = v


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tc_burst=d["bf16_tflops"], tc_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    src="MEASURED_PEAKS.json")
    return dict(hbm=6650.0, tc_burst=1590.0, tc_sustained=1400.0, src="fallback (B200_PROFILING.md)")


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.p, self.index = [], None, index

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i",
                                       str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=2)
        except Exception:
            self.p.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 8:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------- CPU reference arm
def cpu_reference(workload: str, steps: int, warmup: int):
    """Times the oracle (CPU restatement of the reference train()) on the host cores. Returns (steps/s, cores, sample)."""
    import torch

    from oracle import dv3_oracle as O
    from oracle import harness
    from sheeprl_prey_b200.configs import make_cfg
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from util import random_params_like

    wl = WORKLOADS[workload]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = cfg_for(wl, 16, 64, 15)
    params = random_params_like(cfg, wl["obs"], wl["actions"], seed=0)
    tr = O.OracleTrainer(cfg, wl["obs"], wl["actions"], params["wm"], params["actor"], params["critic"], params["target"])
    data = harness.synthetic_batch(64, 16, wl["obs"], wl["actions"], seed=1234)
    noise = harness.synthetic_noise(64, 16, 15, 32, len(wl["actions"]), seed=99)
    for _ in range(warmup):
        tr.step({k: v.clone() for k, v in data.items()}, noise, keep=False)
    t0 = time.perf_counter()
    for _ in range(steps):
        tr.step({k: v.clone() for k, v in data.items()}, noise, keep=False)
    dt = time.perf_counter() - t0
    return steps / dt, cores, f"{steps} full train() steps of the {wl['size']} workload after {warmup} warm-up, torch {torch.__version__} CPU, {torch.get_num_threads()} threads"


def run_reference(args, rank):
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 3)), max(1, min(args.warmup, 1))
    v, cores, sample = cpu_reference(args.size, steps, warmup)
    line = {"impl": "reference", "metric": "DreamerV3 grad steps/sec (B16xT64)", "value": v, "unit": "steps/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": 1000.0 / v, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOADS[args.size]["desc"] + ", B16xT64, H15", "global_batch": 16, "seq_len": 64},
            "cpu_baseline": {"value": v, "unit": "steps/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------- our arm
def synthetic_steps(L, n_envs, obs, actions, seed):
    """Synthetic replay content: vectors ~ N(0,1) clipped to +-5, one-hot uniform actions, rewards ~ N(0,1),
    dones ~ Bernoulli(1/200), is_first[t] = dones[t-1]."""
    import torch

    g = torch.Generator().manual_seed(seed)
    d = {}
    for k, n in obs.items():
        if isinstance(n, int):
            d[k] = torch.randn(L, n_envs, n, generator=g).clamp_(-5, 5)
        else:
            d[k] = torch.randint(0, 256, (L, n_envs, *n), generator=g, dtype=torch.uint8)
    d["actions"] = torch.cat([torch.nn.functional.one_hot(torch.randint(0, a, (L, n_envs), generator=g), a).float() for a in actions], -1)
    d["rewards"] = torch.randn(L, n_envs, 1, generator=g)
    dones = (torch.rand(L, n_envs, 1, generator=g) < 1 / 200).float()
    d["dones"] = dones
    d["is_first"] = torch.cat((torch.zeros(1, n_envs, 1), dones[:-1]), 0)
    return d


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    from sheeprl_prey_b200.algos.dreamer_v3 import build_models, train
    from sheeprl_prey_b200.algos.dreamer_v3.dreamer_v3 import Moments, get_engine
    from sheeprl_prey_b200.configs import ObsSpace, make_cfg
    from sheeprl_prey_b200.data.buffers import AsyncReplayBuffer

    dev = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    wl = WORKLOADS[args.size]
    B, T, H = 16, 64, 15
    cfg = cfg_for(wl, B, T, H)
    fab = Fab(dev, world, rank)
    torch.manual_seed(0)                                    # identical initial weights on every rank
    wm, actor, critic, target = build_models(fab, wl["actions"], False, cfg, ObsSpace({k: ((d,) if isinstance(d, int) else d) for k, d in wl["obs"].items()}))
    a = cfg.algo
    wo = torch.optim.Adam(wm.parameters(), lr=a.world_model.optimizer.lr, eps=a.world_model.optimizer.eps)
    ao = torch.optim.Adam(actor.parameters(), lr=a.actor.optimizer.lr, eps=a.actor.optimizer.eps)
    co = torch.optim.Adam(critic.parameters(), lr=a.critic.optimizer.lr, eps=a.critic.optimizer.eps)
    mom = Moments(fab, a.actor.moments.decay, a.actor.moments.max, a.actor.moments.percentile.low, a.actor.moments.percentile.high).to(dev)
    eng = get_engine(wm)

    # per-rank replay shard, device resident, ring full with the write head mid-ring
    n_envs, size = 4, args.buffer
    if any(not isinstance(d, int) for d in wl["obs"].values()):
        size = min(size, 16384)          # pixel rows are 12 KB: keep the synthetic fill quick (ring is still > L2 and full)
    rb = AsyncReplayBuffer(size, n_envs, device=dev, sequential=True)
    torch.manual_seed(1234 + rank)
    chunk = 8192 if size > 16384 else 2048
    filled = 0
    while filled < size + size // 2:
        L = min(chunk, size + size // 2 - filled)
        d = synthetic_steps(L, n_envs, wl["obs"], wl["actions"], seed=1234 + rank * 1000 + filled)
        rb.add(d)
        filled += L
    agg = Agg()

    def step_device():
        batch = rb.sample(B, sequence_length=T, n_samples=1)
        train(fab, wm, actor, critic, target, wo, ao, co, batch[0], agg, cfg, False, wl["actions"], mom)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    W, K = max(args.warmup, 3), args.steps
    for _ in range(W):
        step_device()
    barrier()
    l0 = eng.ops.launches
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step_device()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = (eng.ops.launches - l0) + K   # + the gather kernel of every sample
    graphed = getattr(eng, "_graph", None) is not None
    if graphed:   # replayed kernels are not counted by the eager counter: graph nodes + the eager batch-load kernels + gather
        launches = K * (eng._graph_launches + 1) + (eng.ops.launches - l0)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t)
    value = world * K / (ms / 1000.0)

    # ---- e2e: pinned host batch -> H2D -> train() -> D2H metrics, every step
    host_batches = []
    for _ in range(4):
        b = rb.sample(B, sequence_length=T, n_samples=1)[0]
        host_batches.append({k: v.cpu().pin_memory() for k, v in b.items()})
    h2d = sum(v.numel() * v.element_size() for v in host_batches[0].values())
    dbuf = {k: torch.empty_like(v, device=dev) for k, v in host_batches[0].items()}
    hmet = torch.empty(16, dtype=torch.float32).pin_memory()

    def step_e2e(i):
        hb = host_batches[i % len(host_batches)]
        for k, v in hb.items():
            dbuf[k].copy_(v, non_blocking=True)
        train(fab, wm, actor, critic, target, wo, ao, co, dbuf, agg, cfg, False, wl["actions"], mom)
        hmet.copy_(eng.metrics, non_blocking=False)          # D2H read of the step's result (synchronises)
        return float(hmet[0])

    for i in range(3):
        step_e2e(i)
    barrier()
    e0.record()
    for i in range(K):
        step_e2e(i)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * K / (float(t) / 1000.0)

    # ---- roofline of the dominant kernel (the dense contraction), timed live with CUDA events on its stream
    roof = None
    if True:   # every rank runs the diagnostic steps (they contain the gradient all-reduces); only rank 0 reports
        pk = peaks()
        saved = os.environ.get("DV3_CUDA_GRAPH")
        os.environ["DV3_CUDA_GRAPH"] = "0"
        eng.ops.gemm_log = []
        for _ in range(2):
            # keep the GPU busy (~30 ms spin) while the host enqueues the eager step, so that the event pairs around each contraction
            # measure the kernel on a full queue and not the host's launch latency on an idle GPU
            torch.cuda._sleep(60_000_000)
            step_device()
        torch.cuda.synchronize()
        log, eng.ops.gemm_log = eng.ops.gemm_log, None
        if saved is None:
            os.environ.pop("DV3_CUDA_GRAPH", None)
        else:
            os.environ["DV3_CUDA_GRAPH"] = saved
        table = {}
        for (s_, e_, f, M, N, Kk, mode, kern) in log:
            t = table.setdefault(f"{mode} M{M} N{N} K{Kk} [{kern}]", [0, 0.0, f])
            t[0] += 1
            t[1] += s_.elapsed_time(e_)
        rows = sorted(((k, c, ms_ / c, f / (ms_ / c * 1e-3) / 1e12) for k, (c, ms_, f) in table.items()), key=lambda r: -r[1] * r[2])
        if rank == 0 and os.path.isdir(os.path.join(ROOT, "gpurun_out")):
            with open(os.path.join(ROOT, "gpurun_out", f"gemm_table_{args.size}_{args.precision}.txt"), "w") as fh:
                fh.write("shape | launches (2 steps) | avg ms | TFLOP/s\n")
                for k, c, ms_, tf in rows:
                    fh.write(f"{k} | {c} | {ms_:.4f} | {tf:.1f}\n")
        phases = eng.timed_step()
        if rank == 0 and os.path.isdir(os.path.join(ROOT, "gpurun_out")):
            with open(os.path.join(ROOT, "gpurun_out", f"phases_{args.size}_{args.precision}.json"), "w") as fh:
                json.dump(phases, fh, indent=1)
        # Roofline of the dense contractions: every tensor-core GEMM launch of one step is re-issued, in order and on the same
        # buffers, inside a CUDA graph that contains nothing else; CUDA events around the replays give the kernels' duration under
        # the launch conditions of the real step (graph replay), without the per-call event overhead of the eager table above.
        eng.ops.gemm_replay = []
        os.environ["DV3_CUDA_GRAPH"] = "0"
        step_device()
        torch.cuda.synchronize()
        if saved is None:
            os.environ.pop("DV3_CUDA_GRAPH", None)
        else:
            os.environ["DV3_CUDA_GRAPH"] = saved
        calls, eng.ops.gemm_replay = eng.ops.gemm_replay, None
        tot_fl = sum(f for _, f in calls)
        tot_ms = 0.0
        if calls:
            gg = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gg):
                for c, _ in calls:
                    c()
            gg.replay()
            torch.cuda.synchronize()
            r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            r0.record()
            for _ in range(3):
                gg.replay()
            r1.record()
            torch.cuda.synchronize()
            tot_ms = r0.elapsed_time(r1) / 3.0
        else:       # fp32 precision: CUDA-core contractions, eager event table
            sel = [(s_.elapsed_time(e_), f) for (s_, e_, f, M, N, Kk, mode, kern) in log if M > 32]
            tot_ms, tot_fl, calls = sum(x[0] for x in sel) / 2.0, sum(x[1] for x in sel) / 2.0, sel[: len(sel) // 2]
        big = calls
        all_ms = sum(s.elapsed_time(e) for (s, e, *_r) in log) / 2.0
        ach = tot_fl / (tot_ms * 1e-3) / 1e12 if tot_ms > 0 else 0.0
        traffic, traffic_of = None, None
        try:   # dram__bytes_read + dram__bytes_write of one `ncu --set full` launch of the dominant kernel (largest shape of the step)
            with open(os.path.join(ROOT, "profiles", "r01_gemm_tma_ncu.json")) as fh:
                nc = json.load(fh)
            if args.precision == "bf16" and args.size == "S":
                traffic, traffic_of = nc["dram_bytes_read"] + nc["dram_bytes_write"], nc["shape"] + "; algorithmic bytes %d" % nc["algorithmic_bytes"]
        except (OSError, KeyError, ValueError):
            pass
        roof = {"bound": "tensor", "kernel": eng.ops.gemm_kernel_name, "achieved": ach, "peak": pk["tc_sustained"], "unit": "TFLOP/s",
                "frac": ach / pk["tc_sustained"], "traffic": traffic, "traffic_of": traffic_of, "launches_per_step": len(big),
                "avg_launch_ms": tot_ms / max(1, len(big)), "kernel_ms_per_step": tot_ms, "all_gemm_ms_per_step_eager_events": all_ms,
                "method": "all tensor-core contraction launches of one step replayed alone in a CUDA graph, CUDA events around 3 replays",
                "peak_source": pk["src"] + " bf16_tflops_sustained (of measured)",
                "whole_step_tc_frac": GFLOP_PER_STEP[args.size] * (value / world) / 1e3 / pk["tc_sustained"]}

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sample = cpu_reference(args.size, 2, 1)
            cpu = {"value": v, "unit": "steps/s", "cores": cores, "kind": "port", "sample": sample}
        ws_gb = torch.cuda.memory_allocated(dev) / 1e9
        line = {"metric": "DreamerV3 grad steps/sec (B16xT64)", "value": value, "unit": "steps/s", "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": eng.ops.dtype_name, "data": "synthetic",
                "config": {"workload": wl["desc"] + f", B{B}xT{T}, H{H}", "global_batch": B * world, "seq_len": T,
                           "parallelism": f"dp{world}", "sequences_per_s": value * B, "cuda_graph": graphed,
                           "replay": f"{n_envs} envs x {size} steps per rank, device resident",
                           "l2": f"no explicit flush: the step streams ~{ws_gb:.1f} GB of resident activations/weights/optimizer state, >> 126 MB L2"},
                "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": "steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 64},
                "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu,
                "metrics_last_step": {k: float(v) for k, v in agg.v.items()}}
        print(json.dumps(line), flush=True)
    if world > 1:
        # the captured CUDA graph holds NCCL work: tearing the communicator down underneath it can block, so leave through a
        # barrier and a hard exit once the result line is flushed
        sys.stdout.flush()
        dist.barrier()
        torch.cuda.synchronize()
        os._exit(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", default="S", choices=sorted(WORKLOADS))
    ap.add_argument("--buffer", type=int, default=100000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"],
                    help="bf16: tcgen05 tensor-core contractions (north-star 2e-2 mode); fp32: CUDA-core strict-parity mode")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    os.environ["DV3_PRECISION"] = args.precision
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
