"""-m gpu, collected last on purpose.

`dv3_col_reduce_noalloc` is the one entry point for hosts that cannot allocate device memory for us (a plugin that only
receives its inputs and outputs, a latency-critical loop that cannot take cudaMalloc's device-wide synchronisation, a
runtime that owns the whole device through its own pool).  It must give exactly what `dv3_col_reduce` gives with a
caller-provided workspace.  It is called here through the raw C-ABI, twice in a row with different data, because every call
reuses the same built-in array and the second call sees the first call's leftovers."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _call_pair(lib, M, U, seed):
    g = torch.Generator().manual_seed(seed)
    d = torch.randn(M, U + 3, generator=g).to(DEV)[:, :U]            # strided view: ld = U + 3
    x = (torch.randn(M, U, generator=g) * 2 + 0.5).to(DEV)
    mean = x.mean(-1).contiguous()
    rstd = torch.rsqrt(x.var(-1, unbiased=False) + 1e-5).contiguous()
    stream = torch.cuda.current_stream().cuda_stream
    out = {}
    nbytes = lib.dv3_col_reduce_scratch_bytes(M, U)
    scratch = torch.empty(nbytes // 4 + 16, dtype=torch.float32, device=DEV)
    a0, b0 = torch.zeros(U, device=DEV), torch.zeros(U, device=DEV)
    rc = lib.dv3_col_reduce(M, U, d.data_ptr(), d.stride(0), x.data_ptr(), x.stride(0), mean.data_ptr(), rstd.data_ptr(),
                            a0.data_ptr(), b0.data_ptr(), scratch.data_ptr(), scratch.numel() * 4, stream)
    assert rc == 0
    a1, b1 = torch.full((U,), 7.0, device=DEV), torch.full((U,), -7.0, device=DEV)
    rc = lib.dv3_col_reduce_noalloc(M, U, d.data_ptr(), d.stride(0), x.data_ptr(), x.stride(0), mean.data_ptr(), rstd.data_ptr(),
                                    a1.data_ptr(), b1.data_ptr(), stream)
    assert rc == 0
    torch.cuda.synchronize()
    xhat = (x - mean[:, None]) * rstd[:, None]
    return (a0, b0), (a1, b1), ((d * xhat).sum(0), d.sum(0))


def test_noalloc_matches_workspace_entry_point_twice_in_a_row():
    from sheeprl_prey_b200._lib import load_library

    lib = load_library()
    for call, (M, U, seed) in enumerate([(1024, 512, 1), (16384, 1536, 2)]):     # second call: bigger, different data
        (a0, b0), (a1, b1), (ta, tb) = _call_pair(lib, M, U, seed)
        tol = 1e-6 * float(ta.abs().max())
        assert float((a1 - a0).abs().max()) <= tol and float((b1 - b0).abs().max()) <= 1e-6 * float(tb.abs().max()), call
        assert float((a1 - ta).abs().max()) <= 2e-5 * float(ta.abs().max())       # and both are the right answer
        assert float((b1 - tb).abs().max()) <= 2e-5 * float(tb.abs().max())
    # bias-only form (x == NULL) and the size limit
    d = torch.randn(300, 4096, device=DEV)
    b = torch.zeros(4096, device=DEV)
    rc = lib.dv3_col_reduce_noalloc(300, 4096, d.data_ptr(), 4096, None, 0, None, None, None, b.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert rc == 0 and float((b - d.sum(0)).abs().max()) <= 2e-5 * float(d.sum(0).abs().max())
    big = torch.zeros(4, 4097, device=DEV)
    rc = lib.dv3_col_reduce_noalloc(4, 4097, big.data_ptr(), 4097, None, 0, None, None, None, b.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream)
    assert rc == -4      # DV3_ERR_TOO_LARGE: refused, nothing launched
