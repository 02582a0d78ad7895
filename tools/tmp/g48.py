import sys, torch
sys.path.insert(0, ".")
from sheeprl_prey_b200._lib import CudaOps
ops = CudaOps("cuda:0", precision="bf16")
M, N, K = 1048576, 48, 48
A = torch.randn(M, K, device="cuda"); B = torch.randn(N, K, device="cuda"); C = torch.zeros(M, N, device="cuda")
ops.mirror(A); ops.mirror(B); ops.refresh(A); ops.refresh(B)
for _ in range(3):
    ops.gemm("NT", A, B, C)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    ops.gemm("NT", A, B, C)
e1.record(); torch.cuda.synchronize()
print("ms", e0.elapsed_time(e1) / 10, ops.gemm_log)
