import sys, torch
sys.path.insert(0, ".")
from sheeprl_prey_b200._lib import CudaOps
ops = CudaOps("cuda:0")
for n in (15360, 122880):
    x = torch.randn(n, device="cuda") * 0.3 + 2.0
    st, out = torch.zeros(2, device="cuda"), torch.zeros(2, device="cuda")
    for _ in range(3): ops.moments(x, st, out, 0.99, 1.0, 0.05, 0.95)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): ops.moments(x, st, out, 0.99, 1.0, 0.05, 0.95)
    e1.record(); torch.cuda.synchronize()
    print("moments n", n, "us", e0.elapsed_time(e1) / 20 * 1e3)
