import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_jordan_crazy_acceleration_b200.ops import tc05_gemm_update
torch.manual_seed(0)
for rows, cols, K, S in [(128, 64, 128, 8), (256, 192, 128, 8), (1024, 1000, 128, 8), (512, 512, 256, 8), (512, 512, 128, 6), (4096, 4096, 128, 8)]:
    L = torch.randn(rows, K, dtype=torch.float64, device="cuda") * torch.exp(3 * torch.randn(rows, 1, dtype=torch.float64, device="cuda"))
    R = torch.randn(K, cols, dtype=torch.float64, device="cuda")
    C0 = torch.randn(rows, cols + (cols % 2), dtype=torch.float64, device="cuda")[:, :cols]
    ref = C0 + L @ R
    C1 = C0.clone() if cols % 2 == 0 else None
    if C1 is None:
        buf = torch.zeros(rows, cols + 1, dtype=torch.float64, device="cuda"); buf[:, :cols] = C0; C1 = buf[:, :cols]
        ref = C1 + L @ R
    tc05_gemm_update(C1, L, R, S)
    torch.cuda.synchronize()
    scale = (L.abs() @ R.abs()).clamp_min(1e-300)
    err = ((C1 - ref).abs() / scale).max().item()
    print(f"rows={rows} cols={cols} K={K} S={S} max rel-to-|L||R| err = {err:.3e}", flush=True)
n = 16384
L = torch.randn(n, 128, dtype=torch.float64, device="cuda"); R = torch.randn(128, n, dtype=torch.float64, device="cuda")
Cm = torch.zeros(n, n, dtype=torch.float64, device="cuda")
# time the MMA kernel inclusive of slicing (the tool entry allocates; it is a functional check)
for _ in range(2):
    tc05_gemm_update(Cm, L, R, 8)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); tc05_gemm_update(Cm, L, R, 8); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print(f"n={n} K=128 S=8: {ms:.3f} ms (incl. slicing + alloc) -> {2*n*n*128/ms/1e9:.1f} TFLOP/s fp64-equivalent")
