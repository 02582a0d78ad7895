"""bench.py contract pieces that run without a GPU: the reference arm's JSON line (the CPU restatement of the reference on the host
cores, per-rank-step units at N ranks) and the command-line defaults the driver relies on."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(600)
def test_reference_arm_prints_one_contract_line():
    env = dict(os.environ, OMP_NUM_THREADS="4")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--size", "XS", "--steps", "1",
                          "--warmup", "1"], capture_output=True, text=True, cwd=ROOT, env=env, timeout=580)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1                                   # ONE JSON line on stdout
    b = json.loads(lines[0])
    assert b["impl"] == "reference" and b["metric"] == "DreamerV3 grad steps/sec (B16xT64)" and b["unit"] == "steps/s"
    assert b["n_gpus"] == 2 and b["higher_is_better"] is True and b["scaling"] == "weak" and b["vs_baseline"] is None
    assert b["config"]["global_batch"] == 32                 # the GPU arm's global batch: equal work per step
    # value counts per-rank B16 x T64 steps (as the GPU arm does): global steps/s x n_gpus
    assert abs(b["value"] - 2 * b["config"]["global_steps_per_s"]) < 1e-9 * max(1.0, b["value"])
    assert b["cpu_baseline"]["kind"] == "port" and b["cpu_baseline"]["cores"] >= 1 and b["cpu_baseline"]["value"] == b["value"]
    assert b["e2e"] == {"value": b["value"], "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert b["gpu_launches"] == 0


def test_bench_defaults():
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert 'add_argument("--gpus", type=int, default=1)' in src
    assert 'add_argument("--warmup", type=int, default=3)' in src
    assert 'add_argument("--impl", default="ours", choices=["ours", "reference"])' in src
