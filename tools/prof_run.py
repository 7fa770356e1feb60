"""Short single-GPU run that touches every shipped kernel, for `ncu` captures (profiles/ncu/):
    python tools/prof_run.py fp64 8192 128      fill, K1, fused step kernels, un-permute, residual
    python tools/prof_run.py tc05 4096 128      + slicers and the tcgen05 update kernel"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mpi_jordan_crazy_acceleration_b200.models.jordan import JordanSolver, _SingleComm  # noqa: E402

prec, n, m = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
s = JordanSolver(n, m, comm=_SingleComm(), precision=prec)
for _ in range(2):
    s.fill_formula(1)
    sec = s.invert()
res = s.residual()
print(f"{prec} n={n} m={m} seconds={sec:.5f} residual={res:.3e}")
s.close()
torch.cuda.synchronize()
