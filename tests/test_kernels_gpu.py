"""-m gpu: every C-ABI kernel against its plain-PyTorch fp32 statement (oracle/torch_ops.py) on the same inputs.
Tolerances: 1e-5 relative (to the output's max magnitude) for fp32 arithmetic; exact for indices / byte copies."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 1e-5


def rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-20))


def rnd(*shape, scale=1.0, seed=None):
    g = torch.Generator(device="cpu")
    g.manual_seed(hash(shape) % 100000 if seed is None else seed)
    return (torch.randn(*shape, generator=g) * scale).to(DEV)


@pytest.mark.parametrize("mode", ["NT", "NN", "TN"])
@pytest.mark.parametrize("M,N,K", [(16, 512, 1124), (1, 1024, 256), (8, 100, 40), (32, 1536, 1024), (33, 77, 19),
                                   (1024, 512, 1536), (1024, 255, 512), (1024, 1, 512), (257, 130, 14), (128, 128, 16),
                                   (14, 512, 1024), (512, 14, 1024), (1, 512, 4096), (300, 3, 77), (16384, 1, 512), (1000, 14, 510),
                                   (16, 16, 15360), (1024, 18, 640), (1000, 31, 130)])
def test_gemm(cuda_ops, torch_ops, mode, M, N, K):
    A = rnd(M, K) if mode != "TN" else rnd(K, M)
    B = rnd(N, K) if mode == "NT" else rnd(K, N)
    bias = rnd(N)
    for acc in (False, True):
        C0 = rnd(M, N, seed=7)
        C1 = C0.clone()
        C0_d = C0.double()
        torch_ops.gemm(mode, A.double(), B.double(), C0_d, bias.double(), acc)
        cuda_ops.gemm(mode, A, B, C1, bias, acc)
        assert rel(C1.double(), C0_d) < TOL, (mode, M, N, K, acc)


def test_gemm_strided(cuda_ops, torch_ops):
    big_a, big_b, big_c = rnd(64, 300), rnd(200, 300), rnd(64, 400)
    A, B = big_a[:, 20:120], big_b[10:90, 50:150]
    C0 = big_c.clone()
    C1 = big_c.clone()
    torch_ops.gemm("NT", A, B, C0[:, 100:180])
    cuda_ops.gemm("NT", A, B, C1[:, 100:180])
    assert rel(C1, C0) < TOL
    assert torch.equal(C1[:, :100], big_c[:, :100]) and torch.equal(C1[:, 180:], big_c[:, 180:])


def test_transpose_copy(cuda_ops):
    s = rnd(100, 70)
    d = torch.zeros(50, 100, device=DEV)
    cuda_ops.transpose(s[:, 10:60], d)
    assert torch.equal(d, s[:, 10:60].t())
    d2 = torch.zeros(100, 90, device=DEV)
    cuda_ops.copy2d(s[:, 5:65], d2[:, 20:80])
    assert torch.equal(d2[:, 20:80], s[:, 5:65]) and float(d2[:, :20].abs().sum()) == 0


@pytest.mark.parametrize("M,U,act,eps", [(16, 512, 1, 1e-3), (1024, 640, 1, 1e-5), (3, 40, 0, 1e-5), (16, 1536, 1, 1e-3), (300, 512, 1, 1e-3),
                                         (257, 1024, 0, 1e-5), (256, 24, 1, 1e-3), (4099, 32, 1, 1e-3), (1031, 64, 1, 1e-3), (517, 100, 0, 1e-3),
                                         (300, 128, 1, 1e-3), (259, 8, 1, 1e-3)])
def test_ln_act(cuda_ops, torch_ops, M, U, act, eps):
    x, g, b = rnd(M, U, scale=2.0) + 0.3, rnd(U) * 0.2 + 1, rnd(U) * 0.1
    outs = []
    for ops in (torch_ops, cuda_ops):
        y, mean, rstd = torch.zeros(M, U + 8, device=DEV), torch.zeros(M, device=DEV), torch.zeros(M, device=DEV)
        ops.ln_act_fwd(x, g, b, eps, act, y[:, 4:4 + U], mean, rstd)
        outs.append((y, mean, rstd))
    for a, r in zip(outs[1], outs[0]):
        assert rel(a, r) < TOL
    y, mean, rstd = outs[0]
    dy = rnd(M, U, seed=3)
    res = []
    for ops in (torch_ops, cuda_ops):
        dx, dln = torch.zeros(M, U, device=DEV), torch.zeros(M, U, device=DEV)
        ops.ln_act_bwd(dy, x, g, b, mean, rstd, act, dx, dln)
        dg, db, db2 = torch.zeros(U, device=DEV), torch.zeros(U, device=DEV), torch.zeros(U, device=DEV)
        ops.ln_param_grad(dln, x, mean, rstd, dg, db)
        ops.colsum(dln, db2)
        res.append((dx, dln, dg, db, db2))
    for a, r in zip(res[1], res[0]):
        assert rel(a, r) < 2e-5
    # against autograd
    xx = x.clone().requires_grad_(True)
    gg, bb = g.clone().requires_grad_(True), b.clone().requires_grad_(True)
    ln = torch.nn.functional.layer_norm(xx, (U,), gg, bb, eps)
    out = torch.nn.functional.silu(ln) if act else ln
    out.backward(dy)
    assert rel(res[1][0], xx.grad) < 5e-5 and rel(res[1][2], gg.grad) < 5e-5 and rel(res[1][3], bb.grad) < 5e-5


@pytest.mark.parametrize("R,D", [(16, 512), (1024, 256), (5, 24)])
def test_gru_gate(cuda_ops, torch_ops, R, D):
    gp, lnw, lnb, h = rnd(R, 3 * D, scale=1.5), rnd(3 * D) * 0.2 + 1, rnd(3 * D) * 0.1, rnd(R, D)
    fw = []
    for ops in (torch_ops, cuda_ops):
        ho, mean, rstd = torch.zeros(R, D, device=DEV), torch.zeros(R, device=DEV), torch.zeros(R, device=DEV)
        ops.gru_gate_fwd(gp, lnw, lnb, h, ho, mean, rstd)
        fw.append((ho, mean, rstd))
    for a, r in zip(fw[1], fw[0]):
        assert rel(a, r) < TOL
    dho = rnd(R, D, seed=5)
    bw = []
    for ops in (torch_ops, cuda_ops):
        dgp, dln, dhi = torch.zeros(R, 3 * D, device=DEV), torch.zeros(R, 3 * D, device=DEV), torch.zeros(R, D, device=DEV)
        ops.gru_gate_bwd(dho, gp, lnw, lnb, fw[0][1], fw[0][2], h, dgp, dln, dhi)
        bw.append((dgp, dln, dhi))
    for a, r in zip(bw[1], bw[0]):
        assert rel(a, r) < 2e-5


@pytest.mark.parametrize("R,groups,K", [(16, 32, 32), (1024, 1, 100), (7, 1, 2), (64, 4, 18), (9, 1, 200)])
def test_unimix_sample(cuda_ops, torch_ops, R, groups, K):
    raw = rnd(R, groups * K, scale=3.0)
    lo = []
    for ops in (torch_ops, cuda_ops):
        l = torch.zeros(R, groups * K, device=DEV)
        ops.unimix_fwd(raw, groups, K, 0.01, l)
        lo.append(l)
    assert rel(lo[1], lo[0]) < TOL
    u = torch.rand(R, groups, device=DEV)
    out = []
    for ops in (torch_ops, cuda_ops):
        idx = torch.zeros(R, groups, dtype=torch.int32, device=DEV)
        oh = torch.zeros(R, groups * K, device=DEV)
        ops.sample_onehot(lo[0], groups, K, u, None, 0, idx, oh)
        idx_m = torch.zeros(R, groups, dtype=torch.int32, device=DEV)
        oh_m = torch.zeros(R, groups * K, device=DEV)
        ops.sample_onehot(lo[0], groups, K, None, None, 1, idx_m, oh_m)
        out.append((idx, oh, idx_m, oh_m))
    flips = float((out[0][0] != out[1][0]).float().mean())
    assert flips <= 2e-3, flips          # boundary draws may flip on 1-ulp prob differences
    assert torch.equal(out[0][2], out[1][2])
    assert torch.equal(out[1][1].reshape(R, groups, K).argmax(-1).int(), out[1][0])
    assert float(out[1][1].sum()) == R * groups
    # forced indices
    idx_f = torch.zeros(R, groups, dtype=torch.int32, device=DEV)
    oh_f = torch.zeros(R, groups * K, device=DEV)
    cuda_ops.sample_onehot(lo[0], groups, K, None, out[0][0], 0, idx_f, oh_f)
    assert torch.equal(idx_f, out[0][0]) and torch.equal(oh_f, out[0][1])
    # backward
    dl, dz = rnd(R, groups * K, seed=11), rnd(R, groups * K, seed=12)
    for a_dl, a_dz in ((dl, dz), (dl, None), (None, dz)):
        r = []
        for ops in (torch_ops, cuda_ops):
            d = torch.zeros(R, groups * K, device=DEV)
            ops.unimix_bwd(raw, groups, K, 0.01, a_dl, a_dz, d)
            r.append(d)
        assert rel(r[1], r[0]) < 2e-5


def test_kl_balance(cuda_ops, torch_ops):
    R, G, K = 64, 32, 32
    post, prior = rnd(R, G * K, scale=0.1), rnd(R, G * K, scale=0.1, seed=9)
    post[:32] *= 40   # rows well above the free-nats threshold; the rest below it
    res = []
    for ops in (torch_ops, cuda_ops):
        rows, dp, dq = torch.zeros(R, 4, device=DEV), torch.zeros(R, G * K, device=DEV), torch.zeros(R, G * K, device=DEV)
        ops.kl_balance(post, prior, G, K, 1.0, 0.5, 0.1, 1.0 / R, rows, dp, dq)
        res.append((rows, dp, dq))
    assert float(res[0][0][:, 0].max()) > 1.0 > float(res[0][0][:, 0].min())
    for a, r in zip(res[1], res[0]):
        assert rel(a, r) < 2e-5


@pytest.mark.parametrize("M,K", [(1024, 255), (37, 255), (16, 41)])
def test_twohot(cuda_ops, torch_ops, M, K):
    logits = rnd(M, K, scale=2.0)
    x = rnd(M, scale=30.0)
    x[0], x[1], x[2], x[3] = 0.0, 1e9, -1e9, math.expm1(20.0 / (K - 1) * 2)  # exact-bin / clipped edge cases
    w = torch.rand(M, device=DEV)
    res = []
    for ops in (torch_ops, cuda_ops):
        lr, dl, mean = torch.zeros(M, device=DEV), torch.zeros(M, K, device=DEV), torch.zeros(M, device=DEV)
        ops.twohot_loss(logits, x, None, 0.5, lr, dl, False)
        ops.twohot_loss(logits, -x, w, 0.25, lr, dl, True)
        ops.twohot_mean(logits, mean)
        res.append((lr, dl, mean))
    for a, r in zip(res[1], res[0]):
        assert rel(a, r) < 2e-5


def test_small_losses(cuda_ops, torch_ops):
    M, Dd = 1024, 14
    mode, tgt = rnd(M, Dd), rnd(M, Dd, seed=2)
    mode[0, :3] = tgt[0, :3] + 5e-5     # below the 1e-8 squared-distance tolerance
    logit, dones = rnd(M, scale=3.0), (torch.rand(M, device=DEV) < 0.2).float()
    x = torch.zeros(M, device=DEV)
    res = []
    for ops in (torch_ops, cuda_ops):
        lr, dm = torch.zeros(M, device=DEV), torch.zeros(M, Dd, device=DEV)
        ops.symlog_mse(mode, tgt, 1.0 / M, lr, dm)
        lb, dlb = torch.zeros(M, device=DEV), torch.zeros(M, 1, device=DEV)
        ops.bernoulli_loss(logit.view(M, 1), dones, 1.0 / M, lb, dlb)
        sy = torch.zeros(M, Dd, device=DEV)
        ops.symlog(mode * 10, sy)
        met = torch.zeros(4, device=DEV)
        ops.reduce_mean(dm, 3, met, 1, 0.5, False)
        ops.reduce_mean(dm, 4, met, 1, 0.5, True)
        res.append((lr, dm, lb, dlb, sy, met))
    for a, r in zip(res[1], res[0]):
        assert rel(a, r) < 2e-5


def test_lambda_moments(cuda_ops, torch_ops):
    H, N = 15, 1024
    v, r, c = rnd((H + 1) * N, scale=3.0), rnd((H + 1) * N, seed=4), rnd((H + 1) * N, scale=2.0, seed=5)
    dones = (torch.rand(N, device=DEV) < 0.1).float()
    res = []
    for ops in (torch_ops, cuda_ops):
        lam, disc = torch.zeros(H * N, device=DEV), torch.zeros((H + 1) * N, device=DEV)
        ops.lambda_returns(v, r, c, dones, H, N, 0.997, 0.95, lam, disc)
        st, out = torch.tensor([0.1, 0.7], device=DEV), torch.zeros(2, device=DEV)
        ops.moments(lam, st, out, 0.99, 1.0, 0.05, 0.95)
        ops.moments(lam * 3 + 1, st, out, 0.99, 1.0, 0.05, 0.95)
        res.append((lam, disc, st, out))
    assert rel(res[1][0], res[0][0]) < TOL and rel(res[1][1], res[0][1]) < TOL
    assert rel(res[1][2], res[0][2]) < 1e-6 and rel(res[1][3], res[0][3]) < 1e-6
    # exact order statistics
    x = rnd(15360, scale=5.0)
    for q in (0.05, 0.95, 0.5):
        st, out = torch.zeros(2, device=DEV), torch.zeros(2, device=DEV)
        cuda_ops.moments(x, st, out, 0.0, 1e8, q, q)
        assert float(st[0]) == float(torch.quantile(x, q)), q
    # ties, sizes that are not multiples of the warp, the all-gathered size of 8 ranks, narrow value ranges (shared exponent bytes)
    g = torch.Generator().manual_seed(9)
    cases = [torch.randint(0, 7, (1001,), generator=g).float(), torch.full((333,), 2.5), torch.randn(122880, generator=g) * 0.01 + 3.0,
             torch.cat((torch.zeros(50), -torch.zeros(50), torch.randn(31, generator=g))), torch.randn(2, generator=g), torch.randn(1, generator=g)]
    for xc in cases:
        xc = xc.to(DEV)
        for ql, qh in ((0.05, 0.95), (0.0, 1.0), (0.5, 0.123)):
            st, out = torch.zeros(2, device=DEV), torch.zeros(2, device=DEV)
            cuda_ops.moments(xc, st, out, 0.0, 1e8, ql, qh)
            assert float(st[0]) == float(torch.quantile(xc, ql)) and float(st[1]) == float(torch.quantile(xc, qh)), (xc.numel(), ql, qh)


@pytest.mark.parametrize("heads", [[100], [3, 5], [18], [200, 2]])
def test_actor_loss(cuda_ops, torch_ops, heads):
    M, A = 2048, sum(heads)
    raw = rnd(M, A, scale=2.0)
    idx = torch.stack([torch.randint(0, k, (M,)) for k in heads], -1).int().to(DEV)
    lam, val, disc = rnd(M, seed=1), rnd(M, seed=2), torch.rand(M, device=DEV)
    mom = torch.tensor([0.05, 1.3], device=DEV)
    res = []
    for ops in (torch_ops, cuda_ops):
        lr, dr = torch.zeros(M, device=DEV), torch.zeros(M, A, device=DEV)
        ops.actor_loss(raw, heads, idx, lam, val, disc, mom, 0.01, 3e-4, 1.0 / M, lr, dr)
        res.append((lr, dr))
    for a, r in zip(res[1], res[0]):
        assert rel(a, r) < 2e-5


def test_adam_clip_ema(cuda_ops, torch_ops):
    n = 1_000_003 + 61
    n = (n // 64) * 64
    p0, g0 = rnd(n), rnd(n, scale=0.01, seed=3)
    res = []
    for ops in (torch_ops, cuda_ops):
        p, g, m, v = p0.clone(), g0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
        step, norm = torch.zeros(1, device=DEV), torch.zeros(4, device=DEV)
        for it in range(3):
            ops.adam_clip(p, g * (it + 1), m, v, step, 1e-4, 0.9, 0.999, 1e-8, 5.0 if it < 2 else 1e9, norm, 2)
        tgt = p0.clone()
        ops.ema(tgt, p, 0.02)
        res.append((p, m, v, step, norm, tgt))
    # reference: torch.optim.Adam + clip_grad_norm_
    pr = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([pr], lr=1e-4, eps=1e-8)
    for it in range(3):
        pr.grad = g0 * (it + 1)
        torch.nn.utils.clip_grad_norm_([pr], 5.0 if it < 2 else 1e9)
        opt.step()
    assert rel(res[1][0], pr.detach()) < 1e-6
    st = opt.state[pr]
    assert rel(res[1][1], st["exp_avg"]) < 1e-6 and rel(res[1][2], st["exp_avg_sq"]) < 1e-6   # kernel follows torch's lerp form
    for a, r in zip(res[1], res[0]):
        assert rel(a, r) < 1e-5


def test_scan_plumbing(cuda_ops, torch_ops):
    B, D, S, A, T = 16, 64, 96, 7, 5
    first = (torch.rand(T, B, device=DEV) < 0.3).float()
    hp, zp, zr = rnd(B, D + 8), rnd(B, S + 4), rnd(1, S)
    for t_first in (True, False):
        res = []
        for ops in (torch_ops, cuda_ops):
            hi, zi = torch.zeros(B, D + 3, device=DEV), torch.zeros(B, S + 5, device=DEV)
            ops.scan_prep(first[1], None if t_first else hp[:, :D], None if t_first else zp[:, :S], zr, hi[:, :D], zi[:, :S])
            res.append((hi, zi))
        assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])
    res = []
    for ops in (torch_ops, cuda_ops):
        dzd, dhd = rnd(B, S + 2, seed=1), rnd(B, D + 2, seed=2)
        ops.scan_carry(first[2], zp[:, :S], hp[:, :D], hp[:, 8:], dzd[:, :S], dhd[:, :D])
        res.append((dzd, dhd))
    assert rel(res[1][0], res[0][0]) < 1e-6 and rel(res[1][1], res[0][1]) < 1e-6
    acts = rnd(T, B, A)
    res = []
    for ops in (torch_ops, cuda_ops):
        out = torch.zeros(T, B, S + A, device=DEV)
        ops.mask_shift_actions(acts, first, out[:, :, S:])
        res.append(out)
    assert torch.equal(res[0], res[1])


def test_uniform_fill(cuda_ops):
    out = torch.zeros(1 << 20, device=DEV)
    ctr = torch.zeros(1, dtype=torch.int64, device=DEV)
    cuda_ops.uniform_fill(out, 7, ctr)
    a = out.clone()
    cuda_ops.uniform_fill(out, 7, ctr)
    assert int(ctr) == 2 and not torch.equal(a, out)
    assert 0.0 <= float(out.min()) and float(out.max()) < 1.0
    assert abs(float(out.mean()) - 0.5) < 2e-3 and abs(float(out.var()) - 1 / 12) < 2e-3


@pytest.mark.parametrize("mode", ["NT", "NN", "TN"])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 128, 128), (1024, 512, 1536), (16384, 512, 512), (1024, 255, 512),
                                   (300, 100, 70), (512, 1536, 1024), (15360, 512, 1536), (512, 1536, 15360), (100, 512, 15360),
                                   (255, 512, 4100), (136, 72, 1000)])
def test_gemm_bf16_tensor_core(mode, M, N, K):
    """tcgen05 path: inputs are rounded to bf16 exactly once, products accumulate in fp32 -> must match a float64 product of
    the bf16-rounded inputs to fp32 accumulation accuracy, and the fp32 product to the north-star's 2e-2."""
    from sheeprl_prey_b200._lib import CudaOps

    ops = CudaOps(DEV, precision="bf16")
    A = rnd(M, K) if mode != "TN" else rnd(K, M)
    B = rnd(N, K) if mode == "NT" else rnd(K, N)
    bias = rnd(N)
    Ab, Bb = A.bfloat16().double(), B.bfloat16().double()
    ref = {"NT": lambda a, b: a @ b.t(), "NN": lambda a, b: a @ b, "TN": lambda a, b: a.t() @ b}[mode]
    for acc in (False, True):
        C0 = rnd(M, N, seed=7)
        C1 = C0.clone()
        exact = ref(Ab, Bb) + bias.double() + (C0.double() if acc else 0)
        ops.gemm(mode, A, B, C1, bias, acc)
        torch.cuda.synchronize()
        full = ref(A.double(), B.double()) + bias.double() + (C0.double() if acc else 0)
        on_tc = N >= 32 and K >= 32 and (M >= 128 or (mode == "TN" and M >= 16 and K >= 2048))   # CudaOps.gemm dispatch
        if on_tc:
            assert rel(C1.double(), exact) < 2e-5, (mode, M, N, K, acc, rel(C1.double(), exact))
            assert rel(C1.double(), full) < 2e-2
        else:       # small problems stay on the fp32 CUDA-core kernel
            assert rel(C1.double(), full) < 1e-5


def test_gemm_bf16_strided_views():
    from sheeprl_prey_b200._lib import CudaOps

    ops = CudaOps(DEV, precision="bf16")
    big_a, big_b, big_c = rnd(512, 300), rnd(400, 300), rnd(512, 600)
    A, B = big_a[:, 20:276], big_b[10:266, 40:296]            # unaligned column offsets -> scalar load path
    C1 = big_c.clone()
    ops.gemm("NT", A, B, C1[:, 101:357])
    exact = A.bfloat16().double() @ B.bfloat16().double().t()
    assert rel(C1[:, 101:357].double(), exact) < 2e-5
    assert torch.equal(C1[:, :101], big_c[:, :101]) and torch.equal(C1[:, 357:], big_c[:, 357:])


def test_colsum_narrow(cuda_ops, torch_ops):
    """Column sums / LayerNorm parameter sums of a tall, 3-column matrix (the image-channel bias gradient): the row-per-thread kernel."""
    M, U = 70001, 3
    d, x = rnd(M, U, seed=1), rnd(M, U, seed=2)
    mean, rstd = x.mean(-1), (x.var(-1, unbiased=False) + 1e-3).rsqrt()
    res = []
    for ops in (torch_ops, cuda_ops):
        db, dg, db2 = torch.zeros(U, device=DEV), torch.zeros(U, device=DEV), torch.zeros(U, device=DEV)
        ops.colsum(d, db)
        ops.ln_param_grad(d, x, mean, rstd, dg, db2)
        res.append((db, dg, db2))
    for a, r in zip(res[1], res[0]):
        assert rel(a.double(), r.double()) < 2e-5
    assert torch.allclose(res[1][0].double(), d.double().sum(0), rtol=1e-4, atol=1e-3)


def test_im2col_bf16_only(cuda_ops_bf16):
    """bf16 precision: im2col writes only the bf16 mirror of `col` (vector and scalar kernels) == bf16(unfold)."""
    from oracle.torch_ops import TorchOps

    for N, H, C in [(3, 16, 32), (2, 64, 3), (5, 8, 24)]:
        x = rnd(N * H * H, C)
        col = torch.full((N * (H // 2) ** 2, 16 * C), 7.0, device=DEV)
        cuda_ops_bf16.mirror(col)
        cuda_ops_bf16.im2col(x, N, H, H, C, col)
        ref = torch.zeros_like(col)
        TorchOps().im2col(x, N, H, H, C, ref)
        assert torch.equal(cuda_ops_bf16._m16_view(col).view_as(col), ref.bfloat16())
        assert torch.all(col == 7.0)       # the fp32 copy is left alone


@pytest.mark.parametrize("N,H,C", [(3, 64, 3), (5, 16, 24), (2, 8, 96), (7, 4, 40)])
def test_conv_data_movement(cuda_ops, torch_ops, N, H, C):
    x = rnd(N * H * H, C)
    res = []
    for ops in (torch_ops, cuda_ops):
        col = torch.zeros(N * (H // 2) ** 2, 16 * C, device=DEV)
        ops.im2col(x, N, H, H, C, col)
        y = torch.zeros(N * 4 * H * H, C, device=DEV)
        bias = rnd(C, seed=3)
        colT = rnd(N * H * H, 16 * C, seed=4)
        ops.col2im(colT, N, H, H, C, bias, 0.5, y)
        y2 = torch.zeros(N * 4 * H * H, C, device=DEV)
        ops.col2im(colT, N, H, H, C, None, 0.0, y2)
        res.append((col, y, y2))
    assert torch.equal(res[0][0], res[1][0])
    assert rel(res[1][1], res[0][1]) < 1e-6 and rel(res[1][2], res[0][2]) < 1e-6


@pytest.mark.parametrize("N,H,C,O", [(3, 64, 3, 48), (2, 16, 3, 32), (5, 8, 4, 24), (1, 64, 1, 96), (130, 4, 3, 48), (2, 32, 3, 64)])
def test_conv_narrow_direct_kernels(cuda_ops, torch_ops, N, H, C, O):
    """csrc/conv_direct.cu against im2col / col2im + matmul: forward (+ fused channel LayerNorm + SiLU), weight gradient, and the
    transposed convolution to a narrow-channel image."""
    rows = N * (H // 2) ** 2
    x, w = rnd(N * H * H, C), rnd(O, 16 * C, seed=2) * 0.2
    g, b = rnd(O, seed=3) * 0.2 + 1, rnd(O, seed=4) * 0.1
    col = torch.zeros(rows, 16 * C, device=DEV)
    torch_ops.im2col(x, N, H, H, C, col)
    y_ref = col.double() @ w.double().t()
    # plain convolution (the decoder's data gradient uses it like this)
    y = torch.zeros(rows, O, device=DEV)
    assert cuda_ops.conv_narrow_fwd(x, N, H, H, C, w, y)
    assert rel(y.double(), y_ref) < 1e-6
    # with LayerNorm + SiLU
    y2, act, mean, rstd = torch.zeros(rows, O, device=DEV), torch.zeros(rows, O, device=DEV), torch.zeros(rows, device=DEV), torch.zeros(rows, device=DEV)
    assert cuda_ops.conv_narrow_fwd(x, N, H, H, C, w, y2, g, b, 1e-3, act, mean, rstd)
    ln = torch.nn.functional.layer_norm(y_ref, (O,), g.double(), b.double(), 1e-3)
    assert torch.equal(y2, y)
    assert rel(act.double(), torch.nn.functional.silu(ln)) < 5e-6
    assert rel(mean.double(), y_ref.mean(-1)) < 2e-6 and rel(rstd.double(), (y_ref.var(-1, unbiased=False) + 1e-3).rsqrt()) < 5e-6
    # weight gradient
    if O <= 64:
        gy = rnd(rows, O, seed=5)
        dw = torch.zeros(O, 16 * C, device=DEV)
        assert cuda_ops.conv_narrow_wgrad(x, N, H, H, C, gy, dw)
        assert rel(dw.double(), gy.double().t() @ col.double()) < 1e-5
        dw2 = torch.zeros(O, 16 * C, device=DEV)
        cuda_ops.conv_narrow_wgrad(x, N, H, H, C, gy, dw2)
        assert torch.equal(dw, dw2)                       # two-stage reduction in a fixed order
    # transposed convolution Ci -> C image channels
    Ci, Hi = (O + 3) // 4 * 4, H // 2
    xin, wr, bias = rnd(N * Hi * Hi, Ci, seed=6), rnd(Ci, 16 * C, seed=7) * 0.2, rnd(C, seed=8)
    colT = (xin.double() @ wr.double()).float()
    ref = torch.zeros(N * H * H, C, device=DEV)
    torch_ops.col2im(colT, N, Hi, Hi, C, bias, 0.5, ref)
    out = torch.zeros(N * H * H, C, device=DEV)
    assert cuda_ops.deconv_narrow_fwd(xin, N, Hi, Hi, wr, C, bias, 0.5, out)
    assert rel(out, ref) < 5e-6


def test_conv_equals_torch_conv(cuda_ops):
    """im2col + GEMM == F.conv2d(k4,s2,p1);  GEMM + col2im == F.conv_transpose2d(k4,s2,p1)  (fp32 mode)."""
    N, H, Cin, Cout = 4, 16, 12, 20
    x_nchw = rnd(N, Cin, H, H)
    w = rnd(Cout, Cin, 4, 4, seed=2) * 0.1
    x = x_nchw.permute(0, 2, 3, 1).contiguous().view(N * H * H, Cin)
    wt = torch.zeros(Cout, 16 * Cin, device=DEV)
    cuda_ops.permute_last2(w, Cout, Cin, 16, wt)
    col = torch.zeros(N * (H // 2) ** 2, 16 * Cin, device=DEV)
    cuda_ops.im2col(x, N, H, H, Cin, col)
    y = torch.zeros(N * (H // 2) ** 2, Cout, device=DEV)
    cuda_ops.gemm("NT", col, wt, y)
    # float64 reference: torch's fp32 GPU convolutions run on TF32 tensor cores by default
    ref = torch.nn.functional.conv2d(x_nchw.double(), w.double(), None, stride=2, padding=1).permute(0, 2, 3, 1).reshape(-1, Cout)
    assert rel(y.double(), ref) < 1e-5
    wT = rnd(Cin, Cout, 4, 4, seed=5) * 0.1
    wr = torch.zeros(Cin, 16 * Cout, device=DEV)
    cuda_ops.permute_last2(wT, Cin, Cout, 16, wr)
    colT = torch.zeros(N * H * H, 16 * Cout, device=DEV)
    cuda_ops.gemm("NN", x, wr, colT)
    b = rnd(Cout, seed=6)
    out = torch.zeros(N * 4 * H * H, Cout, device=DEV)
    cuda_ops.col2im(colT, N, H, H, Cout, b, 0.5, out)
    refT = torch.nn.functional.conv_transpose2d(x_nchw.double(), wT.double(), b.double(), stride=2, padding=1) + 0.5
    assert rel(out.double(), refT.permute(0, 2, 3, 1).reshape(-1, Cout)) < 1e-5


def test_pixel_small_kernels(cuda_ops, torch_ops):
    N, C, HW = 6, 3, 64 * 64
    img = torch.randint(0, 256, (N, C, HW), dtype=torch.uint8).to(DEV)
    res = []
    for ops in (torch_ops, cuda_ops):
        dst = torch.zeros(N * HW, 5, device=DEV)
        ops.u8_to_nhwc(img, N, C, HW, 1 / 255.0, dst, 5, 2)
        a, b = rnd(N, 777), rnd(N, 777, seed=8)
        lr, dm = torch.zeros(N, device=DEV), torch.zeros(N, 777, device=DEV)
        ops.mse_rows(a, b, 0.25, lr, dm)
        src = rnd(4, 6 * 10 + 3)
        d2 = torch.zeros(4, 6 * 10 + 5, device=DEV)
        ops.permute_last2(src, 4, 6, 10, d2, lds=63, ldd=65)
        res.append((dst, lr, dm, d2))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][3], res[1][3])
    assert rel(res[1][1], res[0][1]) < 1e-5 and rel(res[1][2], res[0][2]) < 1e-6



@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["NT", "NN", "TN"])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (1024, 512, 512), (1024, 512, 1536), (16384, 512, 1536), (1000, 255, 520),
                                   (130, 40, 72), (512, 1536, 16384), (4, 512, 15360), (255, 512, 4096),
                                   # more tiles than 2 x 148 CTAs: the persistent kernel, with ragged M / N / K, one k-block, 64-wide N
                                   (40001, 200, 100), (65536, 48, 48), (20000, 130, 72), (9000, 640, 200)])
def test_gemm_bf16_tma(mode, M, N, K):
    """TMA-fed tcgen05 GEMM on bf16-resident operands: K-major and MN-major operands, ragged sizes (zero-filled by the TMA
    unit), split-K (long K with few output tiles) and accumulate."""
    from sheeprl_prey_b200._lib import load_library

    lib = load_library()
    pad = lambda n: (n + 7) // 8 * 8  # noqa: E731
    if mode == "NT":
        a_shape, b_shape = (M, K), (N, K)
    elif mode == "NN":
        a_shape, b_shape = (M, K), (K, N)
    else:
        a_shape, b_shape = (K, M), (K, N)
    A, B, bias = rnd(*a_shape), rnd(*b_shape, seed=3), rnd(N, seed=5)
    # bf16 operands with the row stride padded to a multiple of 8 elements (the TMA constraint)
    Ab = torch.zeros(a_shape[0], pad(a_shape[1]), dtype=torch.bfloat16, device=DEV)[:, :a_shape[1]]
    Bb = torch.zeros(b_shape[0], pad(b_shape[1]), dtype=torch.bfloat16, device=DEV)[:, :b_shape[1]]
    Ab.copy_(A)
    Bb.copy_(B)
    st = torch.cuda.current_stream().cuda_stream
    Ad, Bd = Ab.double(), Bb.double()
    prod = Ad @ Bd.t() if mode == "NT" else (Ad @ Bd if mode == "NN" else Ad.t() @ Bd)
    for acc in (0, 1):
        C0 = rnd(M, N, seed=7)
        C1 = C0.clone()
        rc = lib.dv3_gemm_bf16_tma({"NT": 0, "NN": 1, "TN": 2}[mode], M, N, K, Ab.data_ptr(), Ab.stride(0), Bb.data_ptr(), Bb.stride(0),
                                   C1.data_ptr(), N, bias.data_ptr(), acc, st)
        assert rc == 0, rc
        torch.cuda.synchronize()
        exact = prod + bias.double() + (C0.double() if acc else 0)
        assert rel(C1.double(), exact) < 2e-5, (mode, M, N, K, acc)


@pytest.mark.gpu
@pytest.mark.parametrize("R,C,lds,ldd", [(1, 4096 * 33, 4096 * 33, 4096 * 33), (1, 1001, 1001, 1001), (300, 512, 512, 512),
                                         (300, 512, 1536, 1536), (77, 255, 256, 256), (16, 31, 40, 33)])
def test_f32_to_bf16(R, C, lds, ldd):
    """Mirror refresh kernel == round-to-nearest-even on every (vectorised / strided / ragged) path."""
    from sheeprl_prey_b200._lib import load_library

    lib = load_library()
    src = rnd(R, lds)
    dst = torch.zeros(R, ldd, dtype=torch.bfloat16, device=DEV)
    assert lib.dv3_f32_to_bf16(R, C, src.data_ptr(), lds, dst.data_ptr(), ldd, torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    assert torch.equal(dst[:, :C], src[:, :C].bfloat16())
    assert not dst[:, C:].any()


@pytest.mark.gpu
def test_bf16_side_outputs(cuda_ops_bf16):
    """Producer kernels with a `*_bf16` pointer write bf16(output) into the mirror in the same pass."""
    ops = cuda_ops_bf16
    M, U, D, G, K = 300, 512, 256, 8, 32
    x, g, b = rnd(M, U), rnd(U, seed=2), rnd(U, seed=3)
    y, dx, dln = torch.zeros(M, U, device=DEV), torch.zeros(M, U, device=DEV), torch.zeros(M, U, device=DEV)
    h_out, onehot = torch.zeros(M, D, device=DEV), torch.zeros(M, G * K, device=DEV)
    for t in (y, dx, h_out, onehot):
        ops.mirror(t)
    mean, rstd = torch.zeros(M, device=DEV), torch.zeros(M, device=DEV)
    for rows in (M, 16):       # warp-per-row and block-per-row variants
        ops.ln_act_fwd(x[:rows], g, b, 1e-3, 1, y[:rows], mean[:rows], rstd[:rows])
        ops.ln_act_bwd(rnd(rows, U, seed=9), x[:rows], g, b, mean[:rows], rstd[:rows], 1, dx[:rows], dln[:rows])
        torch.cuda.synchronize()
        assert torch.equal(ops._m16_view(y[:rows]), y[:rows].bfloat16())
        assert torch.equal(ops._m16_view(dx[:rows]), dx[:rows].bfloat16())
    ops.gru_gate_fwd(rnd(M, 3 * D, seed=4), rnd(3 * D, seed=5), rnd(3 * D, seed=6), rnd(M, D, seed=7), h_out, mean, rstd)
    idx = torch.zeros(M, G, dtype=torch.int32, device=DEV)
    ops.sample_onehot(rnd(M, G * K, seed=8), G, K, torch.rand(M, G, device=DEV), None, 0, idx, onehot)
    src = rnd(M, D, seed=11)
    ops.copy2d(src, h_out[:, :D])
    torch.cuda.synchronize()
    assert torch.equal(ops._m16_view(onehot), onehot.bfloat16()) and onehot.sum() == M * G
    assert torch.equal(ops._m16_view(h_out), src.bfloat16())
    # a contraction on two mirrored operands takes the TMA kernel and equals the exact product of the bf16 values
    W = rnd(U, U, seed=12)
    ops.mirror(W)
    ops.refresh(W)
    out = torch.zeros(M, U, device=DEV)
    ops.gemm_log = []
    ops.gemm("NT", y, W, out)
    torch.cuda.synchronize()
    assert ops.gemm_log[0][-1] == "tma"
    ops.gemm_log = None
    assert rel(out.double(), y.bfloat16().double() @ W.bfloat16().double().t()) < 2e-5


@pytest.mark.parametrize("R,groups,K", [(1024, 32, 32), (1024, 1, 100), (7, 1, 2), (64, 4, 18)])
def test_unimix_sample_fused(cuda_ops, cuda_ops_bf16, R, groups, K):
    """dv3_unimix_sample == dv3_unimix_fwd followed by dv3_sample_onehot (same indices, one-hots and logits); the bf16-mode
    variant (warp prefix sum instead of the sequential cumsum) may only differ where the uniform sits on a CDF step."""
    raw = rnd(R, groups * K, scale=3.0)
    u = torch.rand(R, groups, device=DEV)
    lg = torch.zeros(R, groups * K, device=DEV)
    cuda_ops.unimix_fwd(raw, groups, K, 0.01, lg)
    idx0 = torch.zeros(R, groups, dtype=torch.int32, device=DEV)
    oh0 = torch.zeros(R, groups * K, device=DEV)
    cuda_ops.sample_onehot(lg, groups, K, u, None, 0, idx0, oh0)
    idx1, oh1, lg1 = torch.zeros_like(idx0), torch.zeros_like(oh0), torch.zeros_like(lg)
    cuda_ops.unimix_sample(raw, groups, K, 0.01, u, None, idx1, oh1, logits=lg1)
    assert torch.equal(idx1, idx0) and torch.equal(oh1, oh0) and torch.equal(lg1, lg)
    idx2, oh2 = torch.zeros_like(idx0), torch.zeros_like(oh0)
    cuda_ops.unimix_sample(raw, groups, K, 0.01, None, idx0, idx2, oh2)          # forced
    assert torch.equal(idx2, idx0) and torch.equal(oh2, oh0)
    idx3, oh3 = torch.zeros_like(idx0), torch.zeros_like(oh0)
    cuda_ops_bf16.unimix_sample(raw, groups, K, 0.01, u, None, idx3, oh3)
    assert float((idx3 != idx0).float().mean()) < 1e-3
    assert torch.equal(oh3.reshape(R, groups, K).argmax(-1).int(), idx3) and float(oh3.sum()) == R * groups


@pytest.mark.parametrize("M,U,act,bias", [(1024, 512, 1, True), (16384, 512, 1, False), (300, 640, 1, True), (257, 1024, 0, True), (256, 24, 1, True),
                                          (16, 512, 1, True), (1024, 510, 1, True), (40003, 32, 1, True), (9001, 64, 1, False), (517, 100, 0, True),
                                          (300, 128, 1, True), (2000, 256, 1, True), (259, 8, 1, True)])
def test_ln_act_bwd_fused(cuda_ops, torch_ops, M, U, act, bias):
    """LayerNorm(+SiLU) backward with fused column sums == the three separate ops (fallback shapes included)."""
    x, g, b = rnd(M, U, scale=2.0) + 0.3, rnd(U) * 0.2 + 1, rnd(U) * 0.1
    mean, rstd = x.mean(-1), (x.var(-1, unbiased=False) + 1e-3).rsqrt()
    dy = rnd(M, U, seed=3)
    res = []
    for ops in (torch_ops, cuda_ops):
        dx, dln = torch.zeros(M, U + 8, device=DEV)[:, :U], torch.zeros(M, U, device=DEV)
        dg, db, dbias = torch.zeros(U, device=DEV), torch.zeros(U, device=DEV), torch.zeros(U, device=DEV)
        ops.ln_act_bwd_fused(dy, x, g, b, mean, rstd, act, dx, dln, dg, db, dbias if bias else None)
        res.append((dx, dg, db, dbias))
    torch.cuda.synchronize()
    for i, (a, r) in enumerate(zip(res[1], res[0])):
        assert rel(a.double(), r.double()) < 3e-5, (i, rel(a.double(), r.double()))


@pytest.mark.parametrize("M,U,heads", [(1024, 512, [100]), (300, 640, [7, 3]), (256, 1024, [2, 2, 2, 2]), (257, 24, [18])])
def test_ln_act_fwd_gather(cuda_ops, torch_ops, M, U, heads):
    """LayerNorm(+SiLU) forward with the one-hot input block added as a row gather == the plain statement."""
    A = sum(heads)
    x, g, b = rnd(M, U, scale=2.0) + 0.3, rnd(U) * 0.2 + 1, rnd(U) * 0.1
    table = torch.zeros(A, (U + 3) // 4 * 4, device=DEV)[:, :U]
    table.copy_(rnd(A, U, seed=21))
    idx = torch.stack([torch.randint(0, k, (M,), device=DEV) for k in heads], 1).int()
    offs = [sum(heads[:j]) for j in range(len(heads))]
    outs = []
    for ops in (torch_ops, cuda_ops):
        y, mean, rstd = torch.zeros(M, U, device=DEV), torch.zeros(M, device=DEV), torch.zeros(M, device=DEV)
        assert ops.ln_act_fwd_gather(x, g, b, 1e-3, 1, y, mean, rstd, table, idx, offs)
        outs.append((y, mean, rstd))
    for a, r in zip(outs[1], outs[0]):
        assert rel(a, r) < TOL


@pytest.mark.parametrize("M,N,K,bias", [(1024, 512, 512, True), (16384, 512, 1536, True), (1000, 640, 520, False), (130, 1024, 64, True),
                                        (1024, 128, 256, True)])
def test_gemm_ln_fused(cuda_ops_bf16, M, N, K, bias):
    """Linear + LayerNorm + SiLU in one clustered tcgen05 kernel == TMA GEMM followed by the LayerNorm kernel (same bf16 operands)."""
    ops = cuda_ops_bf16
    ops.use_gemm_ln = True          # opt-in path (see CudaOps.__init__)
    pad = (K + 7) // 8 * 8
    A = torch.zeros(M, pad, device=DEV)[:, :K]
    W = torch.zeros(N, pad, device=DEV)[:, :K]
    A.copy_(rnd(M, K))
    W.copy_(rnd(N, K, seed=3) * 0.1)
    b = rnd(N, seed=4) if bias else None
    g, be = rnd(N, seed=5) * 0.2 + 1, rnd(N, seed=6) * 0.1
    for t in (A._base, W._base):
        ops.mirror(t)
        ops.refresh(t)
    pre0, y0, mean0, rstd0 = torch.zeros(M, N, device=DEV), torch.zeros(M, N, device=DEV), torch.zeros(M, device=DEV), torch.zeros(M, device=DEV)
    ops.gemm("NT", A, W, pre0, bias=b)
    ops.ln_act_fwd(pre0, g, be, 1e-3, 1, y0, mean0, rstd0)
    pre1, y1, mean1, rstd1 = torch.zeros_like(pre0), torch.zeros_like(y0), torch.zeros_like(mean0), torch.zeros_like(rstd0)
    ops.mirror(y1)
    assert ops.gemm_ln(A, W, b, g, be, 1e-3, 1, pre1, y1, mean1, rstd1)
    torch.cuda.synchronize()
    assert rel(pre1, pre0) < 2e-5 and rel(y1, y0) < 5e-5 and rel(mean1, mean0) < 2e-5 and rel(rstd1, rstd0) < 2e-5
    assert torch.equal(ops._m16_view(y1), y1.bfloat16())
