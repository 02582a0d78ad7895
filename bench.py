#!/usr/bin/env python
"""bench.py -- DreamerV3 gradient steps/sec (per-rank batch 16 x T64, H15) on N B200s, one JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--size S|XS|M|L] [--precision bf16|fp32]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --microbench          (BASELINE config 5: scan / imagination / lambda-return sweep -> gpurun_out/microbench.json)

A "step" = one replay-sample (device gather) + one full train() (world model, actor, critic updates).
`value`  : steps/s with the replay shard resident in HBM, CUDA-event timed, max over ranks.
`e2e`    : the same metric through the reference-facing train() API with the batch in pinned HOST memory:
           H2D copy of the batch + train() + D2H read of the 13 metrics inside the timed region, every step (the copy of the next
           batch overlaps the current step; the host reads a step's metrics one step later).
`roofline`: the kernel class with the largest share of the step (top level) + `kernels`: one entry per class SURVEY 8(d) names
           (dense contractions, forward / reverse RSSM scan in both 8(d) forms, fused Adam, replay gather), each timed alone with
           CUDA events on the stream it is launched on.
`also`   : config 3 (M, 64x64x3 pixels) in bf16 and config 2 in fp32 mode, a few steps each in the same run (N = 1 only).
`--impl reference`: the reference's algorithm on the host cores (oracle/dv3_oracle.py: the CPU restatement pinned
           against the reference's own train(); the reference itself is Python and is not present on the GPU box), at the
           same GLOBAL batch as the GPU arm (16 x N sequences per step).
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

T_SEQ, HORIZON = 64, 15
WORKLOADS = {
    # BASELINE.json configs[1]: DreamerV3 S, prey_env vector obs (Box(14,)), Discrete(100)
    "S": dict(size="S", B=16, obs={"state": 14}, actions=[100], desc="DreamerV3-S prey_env vector obs 14 / Discrete(100)"),
    # BASELINE.json configs[0]: DreamerV3 XS, CartPole-v1 (Box(4,), Discrete(2)) -- the reference's CPU-runnable case
    "XS": dict(size="XS", B=16, obs={"state": 4}, actions=[2], desc="DreamerV3-XS CartPole-v1 vector obs 4 / Discrete(2)"),
    # BASELINE.json configs[2]
    "M": dict(size="M", B=16, obs={"rgb": (3, 64, 64)}, actions=[18], desc="DreamerV3-M synthetic 64x64x3 pixel obs (Atari-shaped) / Discrete(18)"),
    # BASELINE.json configs[3]: multi-encoder, per-rank batch 32
    "L": dict(size="L", B=32, obs={"rgb": (3, 64, 64), "state": 64}, actions=[18],
              desc="DreamerV3-L multi-encoder 64x64x3 pixels + 64-dim vector / Discrete(18)"),
}
# the REFERENCE's op count per gradient step (discrete actions; counts the second actor pass, the third critic
# pass and the doubled transition the engine does not execute)
GFLOP_PER_STEP = {"S": 542.0, "XS": 174.9, "M": 1913.4, "L": 7675.3}
METRIC = "DreamerV3 grad steps/sec (B16xT64)"


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
def cpu_reference(workload: str, steps: int, warmup: int, n_ranks: int = 1):
    """Times the oracle (CPU restatement of the reference train()) on the host cores at the GPU arm's GLOBAL batch (per-rank batch x
    n_ranks).  Returns (steps/s, cores, sample description)."""
    import torch

    from oracle import dv3_oracle as O
    from oracle import harness
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from util import random_params_like

    wl = WORKLOADS[workload]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    B = wl["B"] * n_ranks
    cfg = cfg_for(wl, B, T_SEQ, HORIZON)
    params = random_params_like(cfg, wl["obs"], wl["actions"], seed=0)
    tr = O.OracleTrainer(cfg, wl["obs"], wl["actions"], params["wm"], params["actor"], params["critic"], params["target"])
    data = harness.synthetic_batch(T_SEQ, B, wl["obs"], wl["actions"], seed=1234)
    noise = harness.synthetic_noise(T_SEQ, B, HORIZON, 32, len(wl["actions"]), seed=99)
    for _ in range(warmup):
        tr.step({k: v.clone() for k, v in data.items()}, noise, keep=False)
    t0 = time.perf_counter()
    for _ in range(steps):
        tr.step({k: v.clone() for k, v in data.items()}, noise, keep=False)
    dt = time.perf_counter() - t0
    return steps / dt, cores, (f"{steps} full train() steps of the {wl['size']} workload at global batch {B} x T{T_SEQ} after {warmup} warm-up, "
                               f"torch {torch.__version__} CPU, {torch.get_num_threads()} threads")


def run_reference(args, rank):
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 3)), max(1, min(args.warmup, 1))
    wl = WORKLOADS[args.size]
    v, cores, sample = cpu_reference(args.size, steps, warmup, n_ranks=args.gpus)
    gb = wl["B"] * args.gpus
    # the GPU arm's `value` counts per-rank B16 x T64 steps summed over ranks; one CPU step at the global batch is worth n_gpus of them
    v_job = v * args.gpus
    line = {"impl": "reference", "metric": METRIC, "value": v_job, "unit": "steps/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": 1000.0 / v, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"] + f", B{wl['B']}xT{T_SEQ}, H{HORIZON}", "global_batch": gb, "seq_len": T_SEQ,
                       "sequences_per_s": v * gb, "global_steps_per_s": v,
                       "note": "one CPU process at the GPU arm's global batch (per-rank batch x n_gpus): equal work per step; value = global steps/s x n_gpus"},
            "cpu_baseline": {"value": v_job, "unit": "steps/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v_job, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
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


class Run:
    """Models + optimizers + replay shard of one workload on this rank's GPU, and the two step functions (device / end-to-end)."""

    def __init__(self, size, precision, dev, world, rank, buffer):
        import torch

        os.environ["DV3_PRECISION"] = precision
        from sheeprl_prey_b200.algos.dreamer_v3 import build_models, train
        from sheeprl_prey_b200.algos.dreamer_v3.dreamer_v3 import Moments, get_engine
        from sheeprl_prey_b200.configs import ObsSpace
        from sheeprl_prey_b200.data.buffers import AsyncReplayBuffer

        self.torch, self.train = torch, train
        self.wl = wl = WORKLOADS[size]
        self.size, self.precision, self.dev, self.world, self.rank = size, precision, dev, world, rank
        self.B, self.T, self.H = wl["B"], T_SEQ, HORIZON
        self.cfg = cfg = cfg_for(wl, self.B, self.T, self.H)
        self.fab = Fab(dev, world, rank)
        torch.manual_seed(0)                                    # identical initial weights on every rank
        self.models = build_models(self.fab, wl["actions"], False, cfg, ObsSpace({k: ((d,) if isinstance(d, int) else d) for k, d in wl["obs"].items()}))
        wm, actor, critic, target = self.models
        a = cfg.algo
        self.opts = (torch.optim.Adam(wm.parameters(), lr=a.world_model.optimizer.lr, eps=a.world_model.optimizer.eps),
                     torch.optim.Adam(actor.parameters(), lr=a.actor.optimizer.lr, eps=a.actor.optimizer.eps),
                     torch.optim.Adam(critic.parameters(), lr=a.critic.optimizer.lr, eps=a.critic.optimizer.eps))
        self.mom = Moments(self.fab, a.actor.moments.decay, a.actor.moments.max, a.actor.moments.percentile.low, a.actor.moments.percentile.high).to(dev)
        self.eng = get_engine(wm)
        # per-rank replay shard, device resident, ring full with the write head mid-ring
        self.n_envs, size_ = 4, buffer
        if any(not isinstance(d, int) for d in wl["obs"].values()):
            size_ = min(size_, 16384)          # pixel rows are 12 KB: keep the synthetic fill quick (ring is still > L2 and full)
        self.rb_size = size_
        self.rb = AsyncReplayBuffer(size_, self.n_envs, device=dev, sequential=True)
        torch.manual_seed(1234 + rank)
        chunk = 8192 if size_ > 16384 else 2048
        filled = 0
        while filled < size_ + size_ // 2:
            L = min(chunk, size_ + size_ // 2 - filled)
            self.rb.add(synthetic_steps(L, self.n_envs, wl["obs"], wl["actions"], seed=1234 + rank * 1000 + filled))
            filled += L
        self.agg = Agg()

    def step_device(self):
        batch = self.rb.sample(self.B, sequence_length=self.T, n_samples=1)
        self.train(self.fab, *self.models, *self.opts, batch[0], self.agg, self.cfg, False, self.wl["actions"], self.mom)

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist

            dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, fn, K):
        """K calls of fn bracketed by barrier + synchronize, CUDA events, MAX over ranks; returns milliseconds for the K calls."""
        torch = self.torch
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(K):
            fn(i)
        e1.record()
        self.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=self.dev)
        if self.world > 1:
            import torch.distributed as dist

            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def e2e_setup(self):
        """End-to-end loop as a training script would pipeline it: batches wait in pinned host memory; the H2D copy of batch i + 1 runs
        on a copy stream while step i computes (two device batches), train() is the public API call, and the 13 metrics of every
        step are copied to pinned host memory and read on the host one step later, after step i + 1 has been enqueued (the
        reference's loop hands metric tensors to its aggregator and only synchronises when it logs)."""
        torch = self.torch
        self.host_batches = []
        for _ in range(4):
            b = self.rb.sample(self.B, sequence_length=self.T, n_samples=1)[0]
            self.host_batches.append({k: v.cpu().pin_memory() for k, v in b.items()})
        self.h2d = sum(v.numel() * v.element_size() for v in self.host_batches[0].values())
        self.dbufs = [{k: torch.empty_like(v, device=self.dev) for k, v in self.host_batches[0].items()} for _ in range(2)]
        self.hmets = [torch.zeros(16, dtype=torch.float32).pin_memory() for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(self.dev)
        self.h2d_ev = [torch.cuda.Event() for _ in range(2)]
        self.free_ev = [torch.cuda.Event() for _ in range(2)]
        self.met_ev = [torch.cuda.Event() for _ in range(2)]
        self._e2e_n = 0
        self._e2e_last = None
        self._enqueue_h2d(0)

    def _enqueue_h2d(self, n):
        torch = self.torch
        slot = n % 2
        hb = self.host_batches[n % len(self.host_batches)]
        with torch.cuda.stream(self.copy_stream):
            if n >= 2:
                self.copy_stream.wait_event(self.free_ev[slot])      # the step that read this device batch has been enqueued and finished
            for k, v in hb.items():
                self.dbufs[slot][k].copy_(v, non_blocking=True)
            self.h2d_ev[slot].record(self.copy_stream)

    def step_e2e(self, i):
        torch = self.torch
        n, slot = self._e2e_n, self._e2e_n % 2
        main = torch.cuda.current_stream(self.dev)
        main.wait_event(self.h2d_ev[slot])
        self.train(self.fab, *self.models, *self.opts, self.dbufs[slot], self.agg, self.cfg, False, self.wl["actions"], self.mom)
        self.free_ev[slot].record(main)
        self._enqueue_h2d(n + 1)                                        # next batch travels while this step computes
        self.hmets[slot].copy_(self.eng.metrics, non_blocking=True)    # D2H read of this step's result ...
        self.met_ev[slot].record(main)
        if n > 0:                                                       # ... consumed on the host one step later
            self.met_ev[1 - slot].synchronize()
            self._e2e_last = float(self.hmets[1 - slot][0])
        self._e2e_n = n + 1
        return self._e2e_last


def event_time(torch, fn, reps, warm=2):
    """Average milliseconds of fn() over `reps` back-to-back launches on the current stream (CUDA events, synchronize on both sides)."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def scan_algorithmic(eng):
    """F_step per row of the observe scan (transition counted once) and the minimum traffic of one scan."""
    S, A, U, D, Hh, E = eng.S, eng.A, eng.U, eng.D, eng.Hh, eng.E
    f_step = 2.0 * ((S + A) * U + (U + D) * 3 * D + (D * Hh + Hh * S) + (D + E) * Hh + Hh * S)
    w_in_loop = S * U + (D + U) * 3 * D + D * 2 * Hh + 2 * Hh * S          # elements (App. G), bf16 in the benchmarked mode
    rows = eng.T * eng.B
    min_bytes = w_in_loop * 2 + rows * (E + A + 1 + D + 3 * S) * 4
    executed = 2.0 * ((U + D) * 3 * D + D * Hh + Hh * S)                     # contractions the kernel's tensor-core MMAs execute, per row
    return f_step, min_bytes, executed


def ncu_traffic(name):
    """dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` launch, if a capture summary is committed under profiles/."""
    for fn in (f"r02_{name}_ncu.json", f"r01_{name}_ncu.json"):
        try:
            with open(os.path.join(ROOT, "profiles", fn)) as fh:
                nc = json.load(fh)
            return nc["dram_bytes_read"] + nc["dram_bytes_write"], nc.get("shape", name) + f" ({fn})", nc.get("algorithmic_bytes")
        except (OSError, KeyError, ValueError):
            continue
    return None, None, None


def kernel_rooflines(run, value):
    """One roofline entry per kernel class of SURVEY 8(d), each measured live and alone (the step's other work is not running)."""
    torch = run.torch
    eng, ops, pk = run.eng, run.eng.ops, peaks()
    saved = os.environ.get("DV3_CUDA_GRAPH")

    def eager_step():
        os.environ["DV3_CUDA_GRAPH"] = "0"
        run.step_device()
        torch.cuda.synchronize()
        if saved is None:
            os.environ.pop("DV3_CUDA_GRAPH", None)
        else:
            os.environ["DV3_CUDA_GRAPH"] = saved

    out, notes = [], {}
    # ---- dense contractions: every tensor-core GEMM launch of one step re-issued, in order and on the same buffers, inside a CUDA
    # graph that contains nothing else; CUDA events around the replays
    ops.gemm_log, ops.gemm_replay = [], []
    eager_step()
    log, calls = ops.gemm_log, ops.gemm_replay
    ops.gemm_log = ops.gemm_replay = None
    table = {}
    for (s_, e_, f, M, N, Kk, mode, kern) in log:
        t = table.setdefault(f"{mode} M{M} N{N} K{Kk} [{kern}]", [0, 0.0, f])
        t[0] += 1
        t[1] += s_.elapsed_time(e_)
    rows = sorted(((k, c, ms_ / c, f / (ms_ / c * 1e-3) / 1e12) for k, (c, ms_, f) in table.items()), key=lambda r: -r[1] * r[2])
    if run.rank == 0 and os.path.isdir(os.path.join(ROOT, "gpurun_out")):
        with open(os.path.join(ROOT, "gpurun_out", f"gemm_table_{run.size}_{run.precision}.txt"), "w") as fh:
            fh.write("shape | launches (1 step, eager, event pairs) | avg ms | TFLOP/s\n")
            for k, c, ms_, tf in rows:
                fh.write(f"{k} | {c} | {ms_:.4f} | {tf:.1f}\n")
    gemm_flops_all = sum(f for (_s, _e, f, *_r) in log)
    tot_fl = sum(f for _, f in calls)
    gemm_ms = 0.0
    if calls:
        gg = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gg):
            for c, _ in calls:
                c()
        gemm_ms = event_time(torch, gg.replay, 3, warm=1)
        traffic, traffic_of, alg = ncu_traffic("gemm_tma")
        ach = tot_fl / (gemm_ms * 1e-3) / 1e12
        out.append({"class": "dense contractions", "kernel": ops.gemm_kernel_name, "bound": "tensor", "achieved": ach, "peak": pk["tc_sustained"],
                    "unit": "TFLOP/s", "frac": ach / pk["tc_sustained"], "traffic": traffic, "traffic_of": traffic_of,
                    "launches_per_step": len(calls), "avg_launch_ms": gemm_ms / len(calls), "kernel_ms_per_step": gemm_ms,
                    "algorithmic": f"sum of 2*M*N*K over the {len(calls)} tensor-core contraction launches of one step = {tot_fl / 1e9:.1f} GFLOP",
                    "method": "all tensor-core contraction launches of one step replayed alone in a CUDA graph, CUDA events around 3 replays"})
    # ---- RSSM scans (one launch each), both 8(d) forms
    f_step, min_bytes, executed = scan_algorithmic(eng)
    rows_tb = eng.T * eng.B
    scan_ms = {}
    if eng.use_fused_scan and ops.scan_kernel is not None:
        ops.reuse_packed_weights = True
        for name, fn in (("forward", lambda: ops.rssm_observe_fwd(eng, None)), ("reverse", lambda: ops.rssm_observe_bwd(eng))):
            kern = ops.scan_kernel if name == "forward" else ops.bwd_scan_kernel
            if name == "reverse" and kern is None:
                continue
            ms = event_time(torch, fn, 5)
            if name == "reverse":      # rssm_observe_bwd also launches dv3_unimix_bwd_prep (softmax quantities of all rows): timed with it
                pass
            scan_ms[name] = ms
            mult = 1.0 if name == "forward" else 2.0
            tc = mult * rows_tb * f_step / (ms * 1e-3) / 1e12
            hb = mult * min_bytes / (ms * 1e-3) / 1e9
            traffic, traffic_of, _ = ncu_traffic(f"scan_{name}")
            out.append({"class": f"RSSM observe scan, {name}", "kernel": f"rssm_observe_{'fwd' if name == 'forward' else 'bwd'}_{'tc' if kern == 'tcgen05' else kern}_kernel",
                        "bound": "latency (dependent chain of T steps on one 16-CTA cluster per 16 sequences); both 8(d) forms reported",
                        "achieved": tc, "peak": pk["tc_sustained"], "unit": "TFLOP/s", "frac": tc / pk["tc_sustained"],
                        "hbm_form": {"achieved": hb, "peak": pk["hbm"], "unit": "GB/s", "frac": hb / pk["hbm"],
                                     "algorithmic_bytes": mult * min_bytes},
                        "traffic": traffic, "traffic_of": traffic_of, "launches_per_step": 1, "avg_launch_ms": ms, "kernel_ms_per_step": ms,
                        "algorithmic": f"T*B*F_step = {rows_tb} rows x {mult * f_step / 1e6:.2f} MFLOP (8(d); x2 for the reverse scan); "
                                       f"the kernel's own MMAs execute {mult * executed * rows_tb / 1e9:.2f} GFLOP of it (prior head and hoisted projections run batched outside)",
                        "method": "5 launches back to back with the packed weights reused, CUDA events"})
        ops.reuse_packed_weights = False
    # ---- imagination rollout (one persistent launch: H steps of prior RSSM + actor on N = T*B rows)
    if getattr(ops, "use_rollout_tc", False) and ops.rollout_kernel is not None:
        S, U, D, Hh, A, lat, H, N = eng.S, eng.U, eng.D, eng.Hh, eng.A, eng.lat, eng.H, eng.N
        st = eng.act_stack
        Ua, La = st.layers[0]["w"].shape[0], len(st.layers)
        f_actor = 2.0 * (lat * Ua + (La - 1) * Ua * Ua + Ua * A)
        f_img = 2.0 * (S * U + (U + D) * 3 * D + D * Hh + Hh * S)       # the action part of the first layer is a row gather (one-hot actions)
        fl = N * (H * (f_img + f_actor) + f_actor)
        fn = lambda: eng.imagine()  # noqa: E731
        ms = event_time(torch, fn, 5)
        tc = fl / (ms * 1e-3) / 1e12
        traffic, traffic_of, _ = ncu_traffic("rollout")
        saved_bytes = N * (H + 1) * (lat * 6 + La * Ua * 10 + A * 4) + N * H * (3 * D + Hh + S) * 4.0
        out.append({"class": "imagination rollout", "kernel": "rollout_tc_kernel", "bound": "tensor", "achieved": tc, "peak": pk["tc_sustained"],
                    "unit": "TFLOP/s", "frac": tc / pk["tc_sustained"],
                    "hbm_form": {"achieved": saved_bytes / (ms * 1e-3) / 1e9, "peak": pk["hbm"], "unit": "GB/s",
                                 "frac": saved_bytes / (ms * 1e-3) / 1e9 / pk["hbm"], "algorithmic_bytes": saved_bytes},
                    "traffic": traffic, "traffic_of": traffic_of, "launches_per_step": 1, "avg_launch_ms": ms, "kernel_ms_per_step": ms,
                    "algorithmic": f"N x (H x (prior-RSSM step + actor) + actor) = {N} rows x {(H * (f_img + f_actor) + f_actor) / 1e6:.1f} MFLOP = {fl / 1e9:.1f} GFLOP; "
                                   "bytes form = trajectory (fp32 + bf16 mirror) and the activations kept for the behaviour backward",
                    "method": "5 x Engine.imagine() (start-state copy + weight-mirror refresh + the persistent kernel), CUDA events"})
    # ---- fused clip + Adam on the world-model arena (32 B / parameter)
    a = eng.wm
    hp = eng.opt_hparams["world"]
    keep = (a.flat.clone(), a.m.clone(), a.v.clone(), a.step.clone())
    ms = event_time(torch, lambda: ops.adam_clip(a.flat, a.grad, a.m, a.v, a.step, 0.0, hp["beta1"], hp["beta2"], hp["eps"], 1000.0, eng.metrics, 8), 5)
    a.flat.copy_(keep[0]); a.m.copy_(keep[1]); a.v.copy_(keep[2]); a.step.copy_(keep[3])
    gb = 32.0 * a.size / (ms * 1e-3) / 1e9
    out.append({"class": "fused clip + Adam", "kernel": "sumsq_partial + adam_prepare + adam_apply (csrc/optim.cu)", "bound": "hbm", "achieved": gb,
                "peak": pk["hbm"], "unit": "GB/s", "frac": gb / pk["hbm"], "traffic": None, "launches_per_step": 9, "avg_launch_ms": ms / 3,
                "kernel_ms_per_step": ms * (eng.wm.size + eng.actor.size + eng.critic.size) / a.size,
                "algorithmic": f"32 B x {a.size} parameters of the world-model arena = {32.0 * a.size / 1e6:.1f} MB per call",
                "method": "5 calls on the world-model arena (lr 0), CUDA events; per-step time scaled to the three arenas"})
    # ---- replay gather
    rb = run.rb
    torch.manual_seed(7)
    plan = rb.plan(run.B, run.T, 1)
    ms = event_time(torch, lambda: rb.gather(*plan, run.B, run.T, 1), 10)
    row_bytes = sum(v[0, 0].numel() * v.element_size() for v in rb._store.values())
    gbytes = 2.0 * run.B * run.T * row_bytes
    gb = gbytes / (ms * 1e-3) / 1e9
    out.append({"class": "replay sequence gather", "kernel": "seq_gather_kernel (csrc/gather.cu)", "bound": "hbm", "achieved": gb, "peak": pk["hbm"],
                "unit": "GB/s", "frac": gb / pk["hbm"], "traffic": None, "launches_per_step": 1, "avg_launch_ms": ms, "kernel_ms_per_step": ms,
                "algorithmic": f"2 x B*T x sum(row bytes) = 2 x {run.B * run.T} x {row_bytes} B = {gbytes / 1e6:.2f} MB (read + write)",
                "method": "10 gathers of a fixed plan (incl. the 5 x n_seq int32 index upload), CUDA events"})
    phases = eng.timed_step()
    if run.rank == 0 and os.path.isdir(os.path.join(ROOT, "gpurun_out")):
        with open(os.path.join(ROOT, "gpurun_out", f"phases_{run.size}_{run.precision}.json"), "w") as fh:
            json.dump(phases, fh, indent=1)
    # executed tensor-core flops of one step: every contraction launch + the two scans' MMAs
    executed_step = gemm_flops_all + (3.0 * executed * rows_tb if scan_ms else 0.0)
    step_ms = 1000.0 / (value / run.world)
    notes["whole_step_tc_frac_reference_op_count"] = GFLOP_PER_STEP[run.size] * (value / run.world) / 1e3 / pk["tc_sustained"]
    notes["whole_step_tc_frac_executed_flops"] = executed_step / 1e12 / (step_ms * 1e-3) / pk["tc_sustained"]
    notes["executed_gflop_per_step"] = executed_step / 1e9
    notes["peak_source"] = pk["src"] + " (bf16_tflops_sustained for kernels timed inside / next to a long step, hbm_gbs)"
    return out, notes


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    dev = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(dev)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"          # NCCL prints its version banner on stdout: keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    run = Run(args.size, args.precision, dev, world, rank, args.buffer)
    eng, wl = run.eng, run.wl
    W, K = max(args.warmup, 3), args.steps
    for _ in range(W):
        run.step_device()
    run.barrier()
    l0 = eng.ops.launches
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms = run.timed(lambda i: run.step_device(), K)
    launches = (eng.ops.launches - l0) + K   # + the gather kernel of every sample
    graphed = getattr(eng, "_graph", None) is not None
    if graphed:   # replayed kernels are not counted by the eager counter: graph nodes + the eager batch-load kernels + gather
        launches = K * (eng._graph_launches + 1) + (eng.ops.launches - l0)
    clocks = sampler.stop() if rank == 0 else None
    value = world * K / (ms / 1000.0)

    # ---- e2e: pinned host batch -> H2D -> train() -> D2H metrics, every step
    run.e2e_setup()
    for i in range(3):
        run.step_e2e(i)
    e2e_value = world * K / (run.timed(run.step_e2e, K) / 1000.0)
    h2d_bytes = run.h2d

    # ---- per-kernel rooflines (every rank runs the diagnostic steps: they contain the gradient all-reduces; rank 0 reports)
    kernels, notes = kernel_rooflines(run, value)
    top = max(kernels, key=lambda r: r["kernel_ms_per_step"])
    roof = dict(top)
    roof["share_of_step"] = top["kernel_ms_per_step"] / (ms / K)
    roof["why_top_level"] = "kernel class with the largest share of the step's time"
    roof.update(notes)
    roof["kernels"] = kernels

    # ---- also: config 3 (M pixels, bf16) and the fp32 mode of this workload, a few steps each in the same run
    also = None
    if world == 1 and not args.no_also:
        also = {}
        metrics_last = {k: float(v) for k, v in run.agg.v.items()}
        del run
        torch.cuda.empty_cache()
        for tag, size, prec in (("config3_M_pixels_bf16", "M", "bf16"), (f"{args.size}_fp32_mode", args.size, "fp32")):
            if size == args.size and prec == args.precision:
                continue
            try:
                r2 = Run(size, prec, dev, 1, 0, args.buffer)
                for _ in range(3):
                    r2.step_device()
                k2 = 5
                ms2 = r2.timed(lambda i: r2.step_device(), k2)
                also[tag] = {"workload": WORKLOADS[size]["desc"] + f", B{WORKLOADS[size]['B']}xT{T_SEQ}, H{HORIZON}", "precision": prec,
                             "value": k2 / (ms2 / 1000.0), "unit": "steps/s", "ms_per_step": ms2 / k2, "steps": k2, "warmup": 3,
                             "scan_kernels": [r2.eng.ops.scan_kernel, r2.eng.ops.bwd_scan_kernel], "rollout_kernel": r2.eng.ops.rollout_kernel}
                del r2
                torch.cuda.empty_cache()
            except Exception as ex:      # the headline line must survive a failure of the extras
                also[tag] = {"error": repr(ex)[:200]}
        os.environ["DV3_PRECISION"] = args.precision
    else:
        metrics_last = {k: float(v) for k, v in run.agg.v.items()}

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sample = cpu_reference(args.size, 16 if args.size in ("S", "XS") else 2, 2)      # ~10 s of host work at config S
            cpu = {"value": v, "unit": "steps/s", "cores": cores, "kind": "port", "sample": sample}
        ws_gb = torch.cuda.max_memory_allocated(dev) / 1e9
        B, T, H = wl["B"], T_SEQ, HORIZON
        line = {"metric": METRIC, "value": value, "unit": "steps/s", "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": eng.ops.dtype_name, "data": "synthetic",
                "config": {"workload": wl["desc"] + f", B{B}xT{T}, H{H}", "global_batch": B * world, "seq_len": T,
                           "parallelism": f"dp{world}", "sequences_per_s": value * B, "cuda_graph": graphed,
                           "scan_kernels": [eng.ops.scan_kernel, eng.ops.bwd_scan_kernel], "rollout_kernel": eng.ops.rollout_kernel,
                           "replay": f"4 envs x {args.buffer if 'rgb' not in wl['obs'] else min(args.buffer, 16384)} steps per rank, device resident",
                           "l2": f"no explicit flush: the step streams ~{ws_gb:.1f} GB of resident activations/weights/optimizer state, >> 126 MB L2"},
                "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": "steps/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 64,
                        "pipeline": "H2D of batch i+1 on a copy stream during step i; metrics of step i copied D2H every step, read on the host after step i+1 is enqueued"},
                "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu, "also": also,
                "metrics_last_step": metrics_last}
        print(json.dumps(line), flush=True)
    if world > 1:
        # the captured CUDA graph holds NCCL work: tearing the communicator down underneath it can block, so leave through a
        # barrier and a hard exit once the result line is flushed
        sys.stdout.flush()
        dist.barrier()
        torch.cuda.synchronize()
        os._exit(0)


# --------------------------------------------------------------------------------------------- microbench (BASELINE config 5)
def run_microbench(args):
    """Isolated RSSM scan + imagination + lambda-return timings: deter {512..8192} x batch {16..1024}, T64, H15, A18, Hh = U = min(D, 1024),
    vector observations embedded to E = U.  Writes gpurun_out/microbench.json; rows that do not fit the 180 GB are skipped."""
    import torch

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from sheeprl_prey_b200._lib import CudaOps
    from sheeprl_prey_b200.configs import make_cfg
    from util import engine_from_params, random_params_like

    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    deters = [int(x) for x in args.micro_deter.split(",")]
    batches = [int(x) for x in args.micro_batch.split(",")]
    res = []
    for D in deters:
        HU = min(D, 1024)
        cfg0 = make_cfg("S", batch_size=16, sequence_length=T_SEQ, horizon=HORIZON, deter=D, hidden=HU, units=HU, layers=2)
        t0 = time.time()
        params = random_params_like(cfg0, {"state": 64}, [18], seed=3)
        print(f"[microbench] D={D}: parameters in {time.time() - t0:.1f} s", file=sys.stderr, flush=True)
        for Bm in batches:
            rows = T_SEQ * Bm
            lat = 1024 + D
            need = (HORIZON + 1) * rows * (lat * 6 + 8 * HU * 6) + rows * (3 * D * 12 + 1024 * 40)
            row = {"deter": D, "hidden_units": HU, "batch": Bm, "T": T_SEQ, "H": HORIZON}
            if need > 120e9:
                row["skipped"] = f"~{need / 1e9:.0f} GB of activations"
                res.append(row)
                continue
            try:
                cfg = make_cfg("S", batch_size=Bm, sequence_length=T_SEQ, horizon=HORIZON, deter=D, hidden=HU, units=HU, layers=2)
                ops = CudaOps(dev, precision="bf16", scratch_bytes=64 << 20)
                eng = engine_from_params(cfg, {"state": 64}, [18], {n: {k: v.to(dev) for k, v in d.items()} for n, d in params.items()}, ops, dev)
                g = torch.Generator().manual_seed(1)
                data = {"state": torch.randn(T_SEQ, Bm, 64, generator=g), "actions": torch.nn.functional.one_hot(torch.randint(0, 18, (T_SEQ, Bm), generator=g), 18).float(),
                        "rewards": torch.randn(T_SEQ, Bm, 1, generator=g), "dones": torch.zeros(T_SEQ, Bm, 1), "is_first": torch.zeros(T_SEQ, Bm, 1)}
                eng.load_batch({k: v.to(dev) for k, v in data.items()})
                eng.load_noise(None)
                eng.world_forward()
                torch.cuda.synchronize()
                row["scan_kernel"] = ops.scan_kernel or "per-step composition"
                if ops.scan_kernel is not None:
                    ops.reuse_packed_weights = True
                    row["scan_fwd_ms"] = event_time(torch, lambda: ops.rssm_observe_fwd(eng, None), 3, warm=1)
                    ops.reuse_packed_weights = False
                else:
                    row["world_forward_ms (scan not isolated)"] = event_time(torch, eng.world_forward, 2, warm=0)
                f_step, min_bytes, _ = scan_algorithmic(eng)
                if "scan_fwd_ms" in row:
                    row["scan_fwd_tflops_8d"] = rows * f_step / (row["scan_fwd_ms"] * 1e-3) / 1e12
                row["imagination_ms"] = event_time(torch, eng.imagine, 2, warm=1)
                S, U, A = eng.S, eng.U, eng.A
                f_img = 2.0 * ((S + A) * U + (U + D) * 3 * D + D * HU + HU * S + (S + D) * U + U * U + U * A)
                row["imagination_tflops"] = HORIZON * rows * f_img / (row["imagination_ms"] * 1e-3) / 1e12
                a = cfg.algo
                row["lambda_returns_ms"] = event_time(torch, lambda: ops.lambda_returns(eng.values, eng.rewards_img, eng.cont_img.out.view(eng.M1), eng.dones, eng.H, eng.N,
                                                                                         a.gamma, a.lmbda, eng.lam, eng.discount), 10)
                row["lambda_returns_gbs"] = (4 * eng.M1 + 2 * eng.M0 + eng.N) * 4 / (row["lambda_returns_ms"] * 1e-3) / 1e9
                del eng, ops
            except torch.OutOfMemoryError:
                row["skipped"] = "does not fit the 180 GB of one B200 (the engine keeps every activation of the step resident)"
            except Exception as ex:
                row["error"] = repr(ex)[:300]
            torch.cuda.empty_cache()
            res.append(row)
            print("[microbench]", json.dumps(row), file=sys.stderr, flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "microbench.json"), "w") as fh:
        json.dump({"config": "BASELINE config 5: isolated observe scan (forward) + imagination rollout + lambda-returns, bf16 mode, 1xB200",
                   "notes": "scan_fwd_ms: the one-launch scan kernel alone (CUDA events, packed weights reused); imagination_ms: eng.imagine() launched EAGERLY "
                            "(H x ~15 launches, so small shapes are launch-bound here; inside train() the step is a CUDA graph); tflops use the 8(d) algorithmic count",
                   "rows": res}, fh, indent=1)
    print(json.dumps({"microbench_rows": len(res), "file": "gpurun_out/microbench.json"}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", default="S", choices=sorted(WORKLOADS))
    ap.add_argument("--buffer", type=int, default=100000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the config-3 / fp32-mode extras of the default line")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"],
                    help="bf16: tcgen05 tensor-core contractions (north-star 2e-2 mode); fp32: CUDA-core strict-parity mode")
    ap.add_argument("--microbench", action="store_true")
    ap.add_argument("--micro-deter", default="512,1024,2048,4096,8192")
    ap.add_argument("--micro-batch", default="16,64,256,1024")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    os.environ["DV3_PRECISION"] = args.precision
    if args.microbench:
        run_microbench(args)
    elif args.impl == "reference":
        run_reference(args, rank)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
