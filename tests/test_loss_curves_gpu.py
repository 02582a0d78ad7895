"""-m gpu: loss curves overlaid (north_star) -- 30 free-running gradient steps of DreamerV3-XS (B16 x T64, H15) by the CPU oracle, the
CUDA library in fp32 mode and in bf16 tensor-core mode from the same parameters, batches and uniforms (tools/loss_curves.py; the
200-step run is profiles/r02_loss_curves.json).  Free-running draws separate wherever a uniform sits on a CDF step, so the statement
is about the smoothed curves: losses / entropies within 10 %, gradient norms within 30 % of the oracle's at every step."""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))


def test_smoothed_loss_curves_follow_the_oracle():
    import loss_curves as lc

    curves = lc.run_curves(30)
    g = lc.gaps(curves, window=10, skip=10)
    for impl, d in g.items():
        for k, v in d.items():
            assert v < (0.3 if k.startswith("Grads/") else 0.1), (impl, k, v)
    # and the world-model loss actually went down on every arm
    for impl, rows in curves.items():
        xs = rows["Loss/world_model_loss"]
        assert sum(xs[-5:]) / 5 < 0.8 * xs[0], (impl, xs[0], xs[-5:])
