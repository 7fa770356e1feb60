import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_jordan_crazy_acceleration_b200.ops import tc05_gemm_update
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
K = int(sys.argv[2]) if len(sys.argv) > 2 else 128
L = torch.randn(n, K, dtype=torch.float64, device="cuda"); R = torch.randn(K, n, dtype=torch.float64, device="cuda")
Cm = torch.zeros(n, n, dtype=torch.float64, device="cuda")
for _ in range(3):
    tc05_gemm_update(Cm, L, R, 8)
torch.cuda.synchronize()
print("done")
