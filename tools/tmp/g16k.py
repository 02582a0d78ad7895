import sys, torch
sys.path.insert(0, ".")
from sheeprl_prey_b200._lib import CudaOps
ops = CudaOps("cuda:0", precision="bf16")
M, N, K = 16384, 512, 1536
A = torch.randn(M, K, device="cuda"); B = torch.randn(N, K, device="cuda"); C = torch.zeros(M, N, device="cuda")
ops.mirror(A); ops.mirror(B); ops.refresh(A); ops.refresh(B)
for _ in range(5):
    ops.gemm("NT", A, B, C)
torch.cuda.synchronize()
print("ok")
